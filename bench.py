#!/usr/bin/env python
"""bench.py -- the Minimal_RAG retrieval step (score every stored embedding against one query, top-k)
on N B200s of one node.  One JSON line on stdout (rank 0).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one pass of the hot path over one batch of synthetic input: ONE query scored against the
whole device-resident corpus shard(s) + fused top-k + exact rescoring (+ all-gather/merge for N > 1).
Workloads (BASELINE.json configs):
  target10m 10M x 768 fp32, top-10      <- default at N = 1: north_star's stated target (>= 85 % of HBM on 1 GPU)
  config4   bf16 x 768, top-100, 12.5M rows per GPU (100M over 8)  <- default at N > 1: north_star's sharded corpus
  config1   10k x 768 fp32, top-5       (the reference's own CPU-runnable case; L2-resident, latency-bound)   } measured as
  config2   1M  x 768 fp32, top-10      (configs[1])                                                          } `extra.*`
  config3   10M x 1024 fp32, 256 queries per step, top-10: the tcgen05/TMEM GEMM path (single GPU)            } at N = 1
  config5   streaming: 64k-row host batches appended while 1k-query batches are served (any N; block-cyclic)
N > 1 runs the chosen workload row-sharded with the per-GPU shard size fixed (weak scaling).
Every line carries `parity`: the GPU hits of query 0 (config3: 8 of the 256 queries) against the CPU oracle run over the
WHOLE corpus of that line (rows regenerated from the counter hash chunk by chunk on the host, every rank its own shard).

value   = corpus vectors scanned per second, whole job, inputs resident in HBM when the clock starts
e2e     = the same through the reference-facing C-ABI call with HOST buffers (query H2D, hits D2H inside)
roofline= algorithmic bytes of the dominant kernel (rows * dim * sizeof(elem)) / its CUDA-event duration
cpu_baseline = the oracle (CPU restatement of RagIndex.QueryAsync) timed on this box, rank 0, N = 1 only
--impl reference times that CPU restatement with all host threads (the reference itself is C#; no dotnet here).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (rows per GPU, dim, dtype, top_k, description)
    "config1": (10_000, 768, "f32", 5, "10k x 768 fp32 synthetic normalised, single query, top-5 cosine"),
    "config2": (1_000_000, 768, "f32", 10, "1M x 768 fp32 corpus, single query, top-10"),
    "target10m": (10_000_000, 768, "f32", 10, "10M x 768 fp32 corpus, single query, top-10 (north_star target)"),
    "config4": (12_500_000, 768, "bf16", 100, "100M x 768 bf16 row-sharded over 8 (12.5M rows per GPU), single query, top-100"),
    "config3": (10_000_000, 1024, "f32", 10, "10M x 1024 fp32 corpus, batched 256 queries, top-10 (tcgen05 tf32 GEMM + exact rescoring)"),
    "config3_shadow": (10_000_000, 1024, "f32", 10, "10M x 1024 fp32 corpus + bf16 shadow image (RAGSTORE_CREATE_BF16_SHADOW), batched 256 queries, top-10 "
                                                    "(tcgen05 bf16 GEMM nominates 128 candidates per query from the shadow, exact rescoring from the fp32 rows)"),
    "config4_batch": (12_500_000, 768, "bf16", 100, "one config4 shard (12.5M x 768 bf16) answering 256 queries per sweep, top-100 (tensor-core path with 128-256 candidates per query)"),
}
STORE_FLAGS = {"config3_shadow": 4}    # RAGSTORE_CREATE_BF16_SHADOW
STREAM = {"name": "config5", "total_rows": 10_000_000, "dim": 768, "top_k": 10, "batch_rows": 65536, "queries_per_batch": 1024,
          "desc": "streaming ingest + query: 10M x 768 fp32 appended in 64k-row host batches through the upload/normalise kernel "
                  "while 1k-query batches (top-10) are served; rows dealt block-cyclically over the GPUs"}
BATCH_QUERIES = {"config3": 256, "config3_shadow": 256, "config4_batch": 256}   # queries per step for the batched (tensor-core) workloads
CORPUS_SEED, QUERY_SEED = 1234, 5678


_REAL_STDOUT = None


def quiet_stdout():
    """Libraries (NCCL prints its version banner from C) must not add lines to stdout: the contract is ONE JSON line.
    File descriptor 1 is pointed at stderr for the run; emit() writes the line to the original stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        os.write(1, line)
    else:
        os.write(_REAL_STDOUT, line)


def cpu_model():
    """(model name, logical cores) of the host the CPU baseline runs on."""
    name = "unknown"
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.lower().startswith("model name"):
                    name = line.split(":", 1)[1].strip()
                    break
    except OSError:
        pass
    return name, os.cpu_count()


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            with open(p) as f:
                return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_tensor_peaks():
    """(burst, sustained) dense bf16 TFLOP/s; tf32 runs at half the bf16 rate on the tensor pipe."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            j = json.load(f)
        return float(j["bf16_tflops"]), float(j.get("bf16_tflops_sustained", j["bf16_tflops"])), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 1590.0, 1400.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index, enabled=True):
        self.samples, self.reasons, self.max_mhz, self._stop = [], set(), None, threading.Event()
        self.power, self.power_limit = [], None
        self.nv = None
        if not enabled:          # only rank 0 polls NVML: N processes hammering its global lock perturb each other
            self.t = None
            return
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            try:
                self.power_limit = pynvml.nvmlDeviceGetEnforcedPowerLimit(self.h) / 1000.0
            except Exception:
                self.power_limit = None
        except Exception:
            self.nv = None
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake": 0x80, "sync_boost": 0x10}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            try:
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            except Exception:
                pass
            time.sleep(0.005)     # short timed regions (20 steps of a few ms) still get a handful of samples

    def __enter__(self):
        if self.nv:
            self.t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self.nv:
            self.t.join(timeout=1)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        s = sorted(self.samples)
        out = {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}
        if self.power:
            p = sorted(self.power)
            out.update({"power_w_median": p[len(p) // 2], "power_w_max": p[-1], "power_limit_w": self.power_limit})
        return out


def default_workload(world):
    """north_star's two stated targets: 10M x 768 fp32 on one GPU; the 100M x 768 bf16 top-100 corpus sharded over N."""
    return "target10m" if world == 1 else "config4"


def config_dict(name, world, rows_override=0):
    """The `config` object of the JSON line.  Built in ONE place so that our arm and the reference arm describe the
    workload with the same keys and values."""
    rows, dim, dtype, k, desc = WORKLOADS[name]
    if rows_override:
        rows = rows_override
    esz = 4 if dtype == "f32" else 2
    return {"workload": name, "desc": desc, "rows_per_gpu": rows, "total_rows": rows * world, "dim": dim, "top_k": k, "storage": dtype,
            "queries_per_step": BATCH_QUERIES.get(name, 1) if world == 1 else 1,
            "parallelism": f"row-sharded x{world}" if world > 1 else "single GPU",
            "l2": "corpus shard is larger than L2 (no flush needed)" if rows * dim * esz > 512e6 else
                  "L2-resident workload: latency config, HBM fraction not meaningful"}


def oracle_topk_streamed(row0, n_rows, dim, bf16, queries, k, n_threads, id_base=None, chunk_rows=250_000):
    """CPU oracle over rows [row0, row0 + n_rows) of the synthetic corpus WITHOUT holding them all: the rows are
    regenerated from the counter hash chunk by chunk (oracle/ragindex_oracle.c:rag_synth_rows, bit-identical to the
    device generator), every chunk is scored with the reference's arithmetic (rag_oracle_query) and the per-chunk hit
    lists are merged under the reference's order (score descending, lower row first).  Returns [(ids, scores)] per
    query; ids are global (id_base + position in the range; id_base defaults to row0)."""
    import numpy as np
    from oracle import oracle_c as oc
    if id_base is None:
        id_base = row0
    best = [(np.zeros(0, dtype=np.int64), np.zeros(0, dtype=np.float64)) for _ in queries]
    for c0 in range(0, n_rows, chunk_rows):
        m = min(chunk_rows, n_rows - c0)
        rows = oc.synth_rows(row0 + c0, m, dim, CORPUS_SEED, round_bf16=bf16, n_threads=n_threads)
        for qi, q in enumerate(queries):
            idx, sc = oc.query(rows, q, k, n_threads=n_threads)
            ids = np.concatenate([best[qi][0], idx + (id_base + c0)])
            scs = np.concatenate([best[qi][1], sc])
            order = np.lexsort((ids, -scs))[:max(1, k)]     # synthetic scores are finite: plain descending order
            best[qi] = (ids[order], scs[order])
        del rows
    return best


def merge_hit_lists(lists, k):
    """Merge per-shard (ids, scores) lists into the global top-k, (score desc, id asc)."""
    import numpy as np
    ids = np.concatenate([l[0] for l in lists])
    scs = np.concatenate([l[1] for l in lists])
    order = np.lexsort((ids, -scs))[:max(1, k)]
    return ids[order], scs[order]


def compare_hits(gpu_idx, gpu_score, ora_idx, ora_score):
    import numpy as np
    gi, gs = np.asarray(gpu_idx, dtype=np.int64), np.asarray(gpu_score, dtype=np.float64)
    oi, os_ = np.asarray(ora_idx, dtype=np.int64), np.asarray(ora_score, dtype=np.float64)
    same_len = gi.shape == oi.shape
    ok_idx = bool(same_len and np.array_equal(gi, oi))
    ok_sc = bool(same_len and np.array_equal(gs.view(np.uint64), os_.view(np.uint64)))
    rel = float(np.max(np.abs(gs - os_) / np.maximum(np.abs(os_), 1e-300))) if same_len and gi.size else None
    return ok_idx, ok_sc, rel


def run_reference(args, rank, world):
    """--impl reference: the CPU restatement of RagIndex.QueryAsync (oracle/) with all host threads."""
    if rank != 0:
        return
    import numpy as np
    from oracle import oracle_c as oc
    rows, dim, dtype, k, desc = WORKLOADS[args.workload]
    if args.rows:
        rows = args.rows
    cores = os.cpu_count() or 1
    sample_rows = min(rows, 1_000_000)     # bounded sample of the workload: the first 1M rows
    corpus = oc.synth_rows(0, sample_rows, dim, CORPUS_SEED, round_bf16=(dtype == "bf16"))
    qs = oc.synth_rows(0, max(args.steps + args.warmup, 1), dim, QUERY_SEED)
    for i in range(args.warmup):
        oc.query(corpus, qs[i], k, n_threads=cores)
    t0 = time.perf_counter()
    for i in range(args.steps):
        oc.query(corpus, qs[args.warmup + i], k, n_threads=cores)
    dt = time.perf_counter() - t0
    v = sample_rows * args.steps / dt
    out = {"impl": "reference", "metric": "corpus vectors scanned/sec (single query, top-k)", "value": v, "unit": "vectors/s",
           "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 products, f64 accumulate",
           "data": "synthetic", "queries_per_s": args.steps / dt * (rows / sample_rows) ** -1 if rows else None,
           "config": config_dict(args.workload, max(world, args.gpus), args.rows),
           "cpu_baseline": {"value": v, "unit": "vectors/s", "cores": cores, "kind": "port", "cpu": cpu_model()[0], "host_cores": cpu_model()[1],
                            "sample": f"one query per step over the first {sample_rows} rows of the workload, all {cores} host threads (row-parallel, per-row arithmetic unchanged)"},
           "e2e": {"value": v, "unit": "vectors/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0,
           "note": "the reference is C# (net9.0); no dotnet toolchain here, so this arm times oracle/ragindex_oracle.c, the CPU restatement of RagIndex.cs:59-67,124-135"}
    emit(out)


def fill_store(st, native, torch, rows, dim, dtype, row0, chunk=1 << 18):
    """Synthetic unit-norm rows generated on the device, ingested through the store's add call (K1)."""
    bf = dtype == "bf16"
    buf = torch.empty((min(chunk, rows), dim), dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    for r0 in range(0, rows, chunk):
        m = min(chunk, rows - r0)
        native.synth_fill_device(buf, row0 + r0, m, dim, dim, CORPUS_SEED, round_bf16=bf, device=torch.cuda.current_device(), stream=stream)
        torch.cuda.current_stream().synchronize()
        st.add_device(buf, m, dim, prenormalised=bf)
    del buf


def time_device_steps(torch, dist, world, step_fn, steps, warmup):
    for _ in range(warmup):
        step_fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step_fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        ms = float(t.item())
    torch.cuda.synchronize()
    return ms


def measure_workload(name, args, torch, dist, rb, native, rank, world, dev, do_e2e=True, do_cpu=True, do_parity=True):
    import numpy as np
    rows, dim, dtype, k, desc = WORKLOADS[name]
    if args.rows:
        rows = args.rows
    esz = 4 if dtype == "f32" else 2
    st = rb.RagStore(dim, rows, rb.RAGSTORE_F32 if dtype == "f32" else rb.RAGSTORE_BF16, dev, STORE_FLAGS.get(name, 0))
    shadow = bool(STORE_FLAGS.get(name, 0) & 4)
    st.set_row_base(rank * rows)
    ts = torch.cuda.Stream()
    with torch.cuda.stream(ts):
        t0 = time.perf_counter()
        fill_store(st, native, torch, rows, dim, dtype, rank * rows)
        torch.cuda.synchronize()
        ingest_s = time.perf_counter() - t0
        n_q = 64
        d_q = torch.empty((n_q, dim), dtype=torch.float32, device="cuda")
        native.synth_fill_device(d_q, 0, n_q, dim, dim, QUERY_SEED, device=dev, stream=ts.cuda_stream)
        d_idx = torch.zeros(k, dtype=torch.int64, device="cuda")
        d_score = torch.zeros(k, dtype=torch.float64, device="cuda")
        d_found = torch.zeros(1, dtype=torch.int32, device="cuda")
        d_flags = torch.zeros(1, dtype=torch.int32, device="cuda")
        qi = [0]
        nqb = BATCH_QUERIES.get(name, 1)
        if world == 1 and nqb > 1:
            n_q = nqb
            d_q = torch.empty((n_q, dim), dtype=torch.float32, device="cuda")
            native.synth_fill_device(d_q, 0, n_q, dim, dim, QUERY_SEED, device=dev, stream=ts.cuda_stream)
            d_idx = torch.zeros((nqb, k), dtype=torch.int64, device="cuda")
            d_score = torch.zeros((nqb, k), dtype=torch.float64, device="cuda")
            d_found = torch.zeros(nqb, dtype=torch.int32, device="cuda")
            d_flags = torch.zeros(nqb, dtype=torch.int32, device="cuda")

            def step():
                qi[0] += 1
                st.search_device(d_q, nqb, k, d_idx, d_score, d_found, d_flags, stream=ts.cuda_stream)
            launches_per_step = 5      # normalise queries, threshold pass (tcgen05) + thresholds, full pass (tcgen05), per-query finalize
        elif world == 1:
            def step():
                q = d_q[qi[0] % n_q]
                qi[0] += 1
                st.search_device(q, 1, k, d_idx, d_score, d_found, d_flags, stream=ts.cuda_stream)
            launches_per_step = 1
        elif nqb > 1:
            # batched + sharded: every shard sweeps its rows once for the whole batch (tensor-core path), the per-query
            # candidates are all-gathered (nq * k * 16 bytes per rank) and merged, one CTA per query
            n_q = nqb
            d_q = torch.empty((n_q, dim), dtype=torch.float32, device="cuda")
            native.synth_fill_device(d_q, 0, n_q, dim, dim, QUERY_SEED, device=dev, stream=ts.cuda_stream)
            sh = rb.ShardedSearcher(st, rank, world, rank * rows, max_k=max(k, 1), device=dev, exchange=args.exchange, max_queries=nqb)
            sh.search_batch_device(d_q, nqb, k)          # allocates the batch buffers
            d_cand, d_all, d_flags = sh._b_cand, sh._b_all, sh.d_flags
            d_idx, d_score, d_found = sh._b_idx, sh._b_score, sh._b_found
            d_flags = torch.ones(nqb, dtype=torch.int32, device="cuda")

            def step():
                qi[0] += 1
                sh.search_batch_device(d_q, nqb, k)
            launches_per_step = 6      # the five batched-search launches of the shard + the merge kernel (NCCL all-gather not counted)
        else:
            sh = rb.ShardedSearcher(st, rank, world, rank * rows, max_k=max(k, 1), device=dev, exchange=args.exchange, max_queries=1)
            d_cand, d_all, d_idx, d_score, d_found, d_flags = sh.d_cand, sh.d_all, sh.d_idx, sh.d_score, sh.d_found, sh.d_flags

            def step():
                q = d_q[qi[0] % n_q]
                qi[0] += 1
                sh.search_device(q, k)      # shard scan -> exchange (peer-memory kernel, or NCCL all-gather + merge kernel)
            launches_per_step = 1 if (sh.peer is not None and sh.fused) else 2

        with ClockSampler(dev, enabled=(rank == 0)) as clk:
            ms = time_device_steps(torch, dist, world, step, args.steps, args.warmup)
        clocks = clk.summary()
        total_rows = rows * world
        exchange_used = None
        if world > 1:
            exchange_used = ("p2p, fused: the shard scan's last CTA publishes, waits and merges over IPC-mapped peer memory (one launch per rank)"
                             if sh.fused and nqb == 1 else "p2p: one exchange kernel per rank over IPC-mapped peer memory (csrc/exchange.cu)") if sh.peer is not None else \
                ("nccl all-gather + merge kernel" + (" (p2p requested but peers could not be mapped)" if args.exchange == "p2p" else ""))
        res = {"ms_per_step": ms / args.steps, "value": total_rows * args.steps / (ms * 1e-3), "clocks": clocks, "exchange": exchange_used,
               "ingest_vectors_per_s": rows / ingest_s, "launches_per_step": launches_per_step,
               "rows_per_gpu": rows, "dim": dim, "top_k": k, "storage": dtype, "desc": desc}
        res["queries_per_s"] = nqb * args.steps / (ms * 1e-3)
        res["queries_per_step"] = nqb
        if nqb > 1:
            res["tensor_tflops"] = 2.0 * nqb * rows * dim * args.steps / (ms * 1e-3) / 1e12
        # dominant kernel: the fused scan; at world == 1 a step IS one launch of it
        if world == 1 or nqb > 1:
            kern_ms = ms / args.steps
        else:
            qi[0] = 0
            kern_ms = time_device_steps(torch, dist, 1, lambda: st.search_shard_device(d_q[0], k, d_cand, d_flags, stream=ts.cuda_stream),
                                        max(10, args.steps // 4), 3) / max(10, args.steps // 4)
        res["kernel_ms"] = kern_ms
        if world > 1:
            # every rank's own shard scan, timed alone on its GPU: the step is paced by the slowest of them plus the exchange
            km = torch.tensor([kern_ms], dtype=torch.float64, device="cuda")
            allk = torch.empty(world, dtype=torch.float64, device="cuda")
            dist.all_gather_into_tensor(allk, km)
            res["kernel_ms_per_rank"] = [float(x) for x in allk.cpu().tolist()]
        if world == 1:
            # per-step latency distribution (median and p99), a separate pass with an event between steps
            n_lat = min(max(args.steps, 20), 400)
            evs = [torch.cuda.Event(enable_timing=True) for _ in range(n_lat + 1)]
            evs[0].record()
            for i in range(n_lat):
                step()
                evs[i + 1].record()
            torch.cuda.synchronize()
            lat = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(n_lat))
            res["latency_ms"] = {"n": n_lat, "min": lat[0], "median": lat[n_lat // 2],
                                 "p99": lat[min(n_lat - 1, int(n_lat * 0.99))], "max": lat[-1]}
        # a shadowed store's batches read the bf16 image: 2 bytes per element are the algorithmic bytes of that sweep
        res["achieved_gbs"] = rows * dim * (2 if shadow else esz) / (kern_ms * 1e-3) / 1e9
        if shadow:
            res["operands"] = "bf16"
        last_q_index = (qi[0] - 1) % n_q
        torch.cuda.synchronize()
        res["flags"] = int(d_flags.flatten()[0].item())
        res["certified_fraction"] = float(((d_flags & 1) != 0).float().mean().item())
        res["top_idx"] = [int(x) for x in d_idx.reshape(-1, k)[0].cpu().tolist()]
        res["top_score"] = [float(x) for x in d_score.reshape(-1, k)[0].cpu().tolist()]
        res["last_query_index"] = last_q_index

    # e2e: the call the C# RagIndex makes -- host query in, host hits out, synchronous
    if do_e2e and world == 1:
        hq = np.ascontiguousarray(d_q.cpu().numpy())
        hq_rows = [np.ascontiguousarray(hq[i]) for i in range(n_q)] if nqb == 1 else None
        hq_step = (lambda i: hq) if nqb > 1 else (lambda i: hq_rows[i % n_q])
        # ragstore_search through preallocated host buffers: what a compiled host (the C# shim) does per request
        call, idx, sc, found, flags = st.search_call(nqb, k)
        for i in range(max(3, args.warmup // 4)):
            call(hq_step(i))
        n_e2e = max(100, args.steps // 4)     # >= 100 calls whatever --steps says: a p99 of five samples is not a p99
        per_call = []
        t0 = time.perf_counter()
        for i in range(n_e2e):
            tc = time.perf_counter()
            call(hq_step(i))
            per_call.append(time.perf_counter() - tc)
        dt = time.perf_counter() - t0
        per_call.sort()
        take = min(rows, max(1, k))
        res["e2e"] = {"value": rows * n_e2e / dt, "unit": "vectors/s", "h2d_bytes_per_step": nqb * dim * 4,
                      "d2h_bytes_per_step": nqb * (take * 16 + 8), "ms_per_step": dt / n_e2e * 1e3, "steps": n_e2e,
                      "queries_per_s": nqb * n_e2e / dt,
                      "median_ms": per_call[n_e2e // 2] * 1e3, "p99_ms": per_call[min(n_e2e - 1, int(n_e2e * 0.99))] * 1e3,
                      "api": "ragstore_search (C ABI, host buffers in, host buffers out, synchronous; results written by the kernel into page-locked host memory)"}
        res["e2e_top_idx"] = [int(x) for x in idx[0, :found[0]]]
    elif do_e2e and nqb == 1:
        # sharded: host query -> every rank's device, hits back to the host on every rank
        hq_rows = [np.ascontiguousarray(r) for r in d_q.cpu().numpy()]
        host_path = sh.peer is not None and dim <= 1024
        hq_pin = torch.empty(dim, dtype=torch.float32).pin_memory()
        h_idx = torch.empty(k, dtype=torch.int64).pin_memory()
        h_sc = torch.empty(k, dtype=torch.float64).pin_memory()
        d_q1 = torch.empty(dim, dtype=torch.float32, device="cuda")
        with torch.cuda.stream(ts):
            def e2e_step(i):
                if host_path:
                    sh.search_host(hq_rows[i % n_q], k)     # query in the kernel parameters, hits written to page-locked memory
                    return
                hq_pin.copy_(torch.from_numpy(hq_rows[i % n_q]))
                d_q1.copy_(hq_pin, non_blocking=True)
                sh.search_device(d_q1, k)
                h_idx.copy_(d_idx[:k], non_blocking=True)
                h_sc.copy_(d_score[:k], non_blocking=True)
                ts.synchronize()
            for i in range(3):
                e2e_step(i)
            dist.barrier()
            n_e2e = max(100, args.steps // 4)
            t0 = time.perf_counter()
            for i in range(n_e2e):
                e2e_step(i)
            dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            dt = float(dt.item())
        res["e2e"] = {"value": rows * world * n_e2e / dt, "unit": "vectors/s", "h2d_bytes_per_step": dim * 4,
                      "d2h_bytes_per_step": k * 16, "ms_per_step": dt / n_e2e * 1e3, "steps": n_e2e,
                      "api": (("ShardedSearcher.search_host: ragstore_search_shard_exchange -- query in the kernel parameters, the scan's last CTA exchanges "
                               "and merges over peer memory and writes the hits into page-locked host memory; ONE launch and one stream wait per rank") if sh.fused else
                              ("ShardedSearcher.search_host: ragstore_search_shard_hostq (query in the kernel parameters) + ragstore_exchange_merge writing "
                               "into page-locked host memory; two launches and one stream wait per rank")) if host_path else
                             "host query -> ShardedSearcher.search_device (shard scan + exchange + merge) -> host hits, per rank"}

    # CPU baseline (rank 0, N = 1 only): the oracle, one thread, on a bounded sample of the workload
    if do_cpu and world == 1 and rank == 0:
        from oracle import oracle_c as oc
        sample_rows = min(rows, 1_000_000)
        corpus = oc.synth_rows(0, sample_rows, dim, CORPUS_SEED, round_bf16=(dtype == "bf16"))
        q = oc.synth_rows(0, 1, dim, QUERY_SEED)[0]
        reps, t_used = 0, 0.0
        while t_used < 8.0 and reps < 50:
            t0 = time.perf_counter()
            oc.query(corpus, q, k, n_threads=1)
            t_used += time.perf_counter() - t0
            reps += 1
        del corpus
        res["cpu_baseline"] = {"value": sample_rows * reps / t_used, "unit": "vectors/s", "cores": 1, "kind": "port", "cpu": cpu_model()[0], "host_cores": cpu_model()[1],
                               "sample": f"{reps} single-thread queries over the first {sample_rows} rows of the workload "
                                         "(oracle/ragindex_oracle.c; the reference's LINQ pipeline is single-threaded)"}

    # Parity at the REAL size, inside the line: the GPU hits against the CPU oracle over the whole corpus of this run.
    # The oracle regenerates the rows from the counter hash chunk by chunk (never all resident); at N > 1 every rank
    # scores its own shard with the host cores it is entitled to and rank 0 merges the per-shard lists.
    if do_parity:
        import numpy as np
        t_par = time.perf_counter()
        threads = max(1, (os.cpu_count() or 1) // world)
        n_check = 1 if (nqb == 1 or world > 1) else min(8, nqb)
        hq_par = np.ascontiguousarray(d_q[:n_check].cpu().numpy())   # the very floats the GPU searched with
        with torch.cuda.stream(ts):
            if world == 1:
                p_idx = torch.zeros((max(nqb, 1), k), dtype=torch.int64, device="cuda")
                p_sc = torch.zeros((max(nqb, 1), k), dtype=torch.float64, device="cuda")
                p_found = torch.zeros(max(nqb, 1), dtype=torch.int32, device="cuda")
                p_flags = torch.zeros(max(nqb, 1), dtype=torch.int32, device="cuda")
                st.search_device(d_q[:nqb] if nqb > 1 else d_q[0], nqb, k, p_idx, p_sc, p_found, p_flags, stream=ts.cuda_stream)
                ts.synchronize()
                g_idx, g_sc = p_idx[:n_check].cpu().numpy(), p_sc[:n_check].cpu().numpy()
                g_flags = p_flags[:n_check].cpu().numpy()
            else:
                sh.search_device(d_q[0], k)
                ts.synchronize()
                nf = int(sh.d_found.item())
                g_idx, g_sc = sh.d_idx[:nf].cpu().numpy()[None, :], sh.d_score[:nf].cpu().numpy()[None, :]
                g_flags = sh.global_flags(sh.d_flags, sh.d_gflags)     # certified / exact only if every shard's part is
        mine = oracle_topk_streamed(rank * rows, rows, dim, dtype == "bf16", list(hq_par), k, threads)
        if world > 1:
            # fixed-size exchange of the per-shard oracle lists (k ids + k scores per rank) through the process group
            o_ids = torch.full((k,), -1, dtype=torch.int64, device="cuda")
            o_scs = torch.full((k,), float("-inf"), dtype=torch.float64, device="cuda")
            n_m = len(mine[0][0])
            o_ids[:n_m] = torch.from_numpy(mine[0][0]).cuda()
            o_scs[:n_m] = torch.from_numpy(mine[0][1]).cuda()
            all_ids = torch.empty(world * k, dtype=torch.int64, device="cuda")
            all_scs = torch.empty(world * k, dtype=torch.float64, device="cuda")
            dist.all_gather_into_tensor(all_ids, o_ids)
            dist.all_gather_into_tensor(all_scs, o_scs)
            ai, asc = all_ids.cpu().numpy(), all_scs.cpu().numpy()
            keep = ai >= 0
            oracle_hits = [merge_hit_lists([(ai[keep], asc[keep])], k)]
        else:
            oracle_hits = mine
        oks = [compare_hits(g_idx[i], g_sc[i], oracle_hits[i][0], oracle_hits[i][1]) for i in range(n_check)]
        res["parity"] = {"checked": f"{n_check} quer{'y' if n_check == 1 else 'ies'} x the full corpus of this run ({rows * world} rows) vs the CPU oracle "
                                    "(oracle/ragindex_oracle.c), rows regenerated chunk by chunk" + (", every rank its own shard" if world > 1 else ""),
                         "indices_bit_exact": all(o[0] for o in oks), "scores_bit_exact": all(o[1] for o in oks),
                         "max_rel_score_err": max((o[2] or 0.0) for o in oks), "queries_checked": n_check, "rows_checked": rows * world,
                         "gpu_flags": [int(x) for x in g_flags.reshape(-1)[:n_check]], "oracle_threads_per_rank": threads,
                         "seconds": time.perf_counter() - t_par}
    st.close()
    return res


def measure_single_process(name, args, torch, rb, native, devices, do_parity=True):
    """The same sharded workload through ONE handle in ONE process (ragstore_create_sharded) -- the shape the reference's
    singleton RagIndex has (Program.cs:35,43): host query in, host hits out, all GPUs of the box behind one call."""
    import numpy as np
    rows, dim, dtype, k, desc = WORKLOADS[name]
    if args.rows:
        rows = args.rows
    G, bf = len(devices), dtype == "bf16"
    total = rows * G
    torch.cuda.set_device(devices[0])
    st = rb.RagStore.create_sharded(dim, total, devices, dtype=rb.RAGSTORE_BF16 if bf else rb.RAGSTORE_F32, block_rows=65536)
    try:
        t0 = time.perf_counter()
        chunk = 1 << 19
        buf = torch.empty((chunk, dim), dtype=torch.float32, device="cuda")
        for r0 in range(0, total, chunk):          # rows are generated on the first GPU; every shard's upload kernel reads its blocks over NVLink
            m = min(chunk, total - r0)
            native.synth_fill_device(buf, r0, m, dim, dim, CORPUS_SEED, round_bf16=bf, device=devices[0])
            torch.cuda.synchronize()
            st.add_device(buf, m, dim, prenormalised=bf)
        del buf
        ingest_s = time.perf_counter() - t0
        n_q = 64
        d_q = torch.empty((n_q, dim), dtype=torch.float32, device="cuda")
        native.synth_fill_device(d_q, 0, n_q, dim, dim, QUERY_SEED, device=devices[0])
        torch.cuda.synchronize()
        hq = [np.ascontiguousarray(r) for r in d_q.cpu().numpy()]
        call, idx, sc, found, flags = st.search_call(1, k)
        for i in range(max(5, args.warmup)):
            call(hq[i % n_q])
        n_calls = max(100, args.steps)
        per = []
        t0 = time.perf_counter()
        for i in range(n_calls):
            tc = time.perf_counter()
            call(hq[i % n_q])
            per.append(time.perf_counter() - tc)
        dt = time.perf_counter() - t0
        per.sort()
        out = {"desc": desc, "api": "ragstore_create_sharded + ragstore_search: one process, one handle, host query in / host hits out; the shards' last CTAs "
                                    "store their hits into the first GPU's memory over NVLink peer access, one merge kernel writes to page-locked memory",
               "devices": devices, "rows_total": total, "rows_per_gpu": rows, "dim": dim, "top_k": k, "storage": dtype,
               "ms_per_query": dt / n_calls * 1e3, "median_ms": per[n_calls // 2] * 1e3, "p99_ms": per[min(n_calls - 1, int(n_calls * 0.99))] * 1e3,
               "vectors_per_s": total * n_calls / dt, "queries_per_s": n_calls / dt, "calls": n_calls, "ingest_vectors_per_s": total / ingest_s,
               "h2d_bytes_per_step": dim * 4, "d2h_bytes_per_step": k * 16 + 8}
        if do_parity:
            t_par = time.perf_counter()
            call(hq[0])
            g_idx, g_sc, g_flags = idx[0, :found[0]].copy(), sc[0, :found[0]].copy(), int(flags[0])
            ora = oracle_topk_streamed(0, total, dim, bf, [hq[0]], k, os.cpu_count() or 1)[0]
            ok = compare_hits(g_idx, g_sc, ora[0], ora[1])
            out["parity"] = {"checked": f"1 query x all {total} rows vs the CPU oracle", "indices_bit_exact": ok[0], "scores_bit_exact": ok[1],
                             "gpu_flags": g_flags, "seconds": time.perf_counter() - t_par}
        return out
    finally:
        st.close()


def run_streaming_single_process(args, torch, rb, native, devices):
    """config5 through ONE handle in ONE process: a writer thread appends host batches (one 64k-row block per GPU per call, so
    every GPU's upload pipeline runs at once) while the main thread serves 1k-query batches through the same handle."""
    import threading
    import numpy as np
    c = STREAM
    dim, k, B, QB = c["dim"], c["top_k"], c["batch_rows"], c["queries_per_batch"]
    G = len(devices)
    total = args.rows * G if args.rows else c["total_rows"]
    group = B * G                                   # rows per append call: one block for every shard
    n_groups = max(2, total // group)
    torch.cuda.set_device(devices[0])
    st = rb.RagStore.create_sharded(dim, n_groups * group, devices, block_rows=B)
    try:
        if not args.no_reserve:
            st.reserve_rows(n_groups * group)
        # The whole stream as DISTINCT rows in host memory (generated on the device, copied back once): repeating one batch
        # would fill the corpus with exact duplicates, i.e. ties at every top-k boundary, which no margin can certify.
        h_all = torch.empty((n_groups * group, dim), dtype=torch.float32)
        if args.pinned:
            h_all = h_all.pin_memory()
        d_batch = torch.empty((group, dim), dtype=torch.float32, device="cuda")
        for g in range(n_groups):
            native.synth_fill_device(d_batch, g * group, group, dim, dim, CORPUS_SEED, device=devices[0])
            torch.cuda.synchronize()
            h_all[g * group:(g + 1) * group].copy_(d_batch)
        del d_batch
        h_np = h_all.numpy()
        h_groups = [h_np[g * group:(g + 1) * group] for g in range(n_groups)]
        d_q = torch.empty((QB, dim), dtype=torch.float32, device="cuda")
        native.synth_fill_device(d_q, 0, QB, dim, dim, QUERY_SEED, device=devices[0])
        torch.cuda.synchronize()
        hq = np.ascontiguousarray(d_q.cpu().numpy())
        del d_q
        call, idx, sc, found, flags = st.search_call(QB, k)
        st.add(h_groups[0])                          # first group in before serving starts (also warms up)
        for _ in range(3):
            call(hq)
        stop = threading.Event()
        t_ingest = [0.0]

        def writer():
            torch.cuda.set_device(devices[0])
            t0 = time.perf_counter()
            for g in range(1, n_groups):
                st.add(h_groups[g])
            t_ingest[0] = time.perf_counter() - t0
            stop.set()

        w = threading.Thread(target=writer)
        qbatches = 0
        with ClockSampler(devices[0]) as clk:
            t0 = time.perf_counter()
            w.start()
            while True:
                done = stop.is_set()
                call(hq)
                qbatches += 1
                if done:
                    break
            t_srv = time.perf_counter() - t0
        w.join()
        rows_streamed = (n_groups - 1) * group
        # all rows are resident now: query 0 of the last batch served must equal the oracle over the whole stream
        call(hq)
        ora = oracle_topk_streamed(0, n_groups * group, dim, False, [hq[0]], k, os.cpu_count() or 1)[0]
        ok_idx, ok_sc, _ = compare_hits(idx[0, :found[0]], sc[0, :found[0]], ora[0], ora[1])
        info = st.info()
        return {"metric": "streaming ingest vectors/sec while serving 1k-query batches (top-10), one process / one handle", "value": rows_streamed / t_ingest[0],
                "unit": "vectors/s", "n_gpus": G, "steps": n_groups - 1, "warmup": 1, "ms_per_step": t_ingest[0] / (n_groups - 1) * 1e3,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "queries_per_s": qbatches * QB / t_srv, "query_batches_served": qbatches,
                "config": {"workload": "config5", "desc": c["desc"], "rows_total": n_groups * group, "dim": dim, "top_k": k, "batch_rows": B,
                           "rows_per_append_call": group, "queries_per_batch": QB, "parallelism": f"one handle over {G} GPUs (ragstore_create_sharded), block-cyclic",
                           "host_batches": "page-locked (direct PCIe copy)" if args.pinned else "pageable (staged through the library's pinned ring)"},
                "roofline": {"bound": "pcie (host->device copy of the appended rows)", "achieved": rows_streamed * dim * 4 / t_ingest[0] / 1e9 / G, "peak": 64.0,
                             "unit": "GB/s per GPU", "frac": rows_streamed * dim * 4 / t_ingest[0] / 1e9 / G / 64.0, "traffic": None, "peak_source": "nominal PCIe Gen5 x16"},
                "cpu_baseline": None,
                "e2e": {"value": rows_streamed / t_ingest[0], "unit": "vectors/s", "h2d_bytes_per_step": group * dim * 4, "d2h_bytes_per_step": 0,
                        "api": "ragstore_add + ragstore_search on one multi-device handle, from two threads of one process"},
                "gpu_launches": int(info["batch_launches"]) * 5 + int(info["scan_launches"]) + (n_groups - 1) * G * 2, "clocks": clk.summary(),
                "parity": {"checked": f"query 0 after the stream (all {n_groups * group} rows resident) vs the CPU oracle over the whole stream",
                           "indices_bit_exact": bool(ok_idx), "scores_bit_exact": bool(ok_sc)},
                "final_count": info["count"]}
    finally:
        st.close()


def measure_shadow_single(args, torch, rb, native, dev, rows=10_000_000, dim=768, k=10, do_parity=True):
    """Single queries on an fp32 store WITH the opt-in bf16 shadow (RAGSTORE_CREATE_BF16_SHADOW): ragstore_search scans the bf16
    image (d * 2 bytes per row) and rescoring reads the fp32 rows, so hits and scores are those of the plain store.  Host query in,
    host hits out (the device-only entry point keeps scanning the fp32 rows: it cannot retry what the image cannot certify)."""
    import numpy as np
    st = rb.RagStore(dim, rows, rb.RAGSTORE_F32, dev, 4)
    try:
        fill_store(st, native, torch, rows, dim, "f32", 0)
        d_q = torch.empty((64, dim), dtype=torch.float32, device="cuda")
        native.synth_fill_device(d_q, 0, 64, dim, dim, QUERY_SEED, device=dev)
        torch.cuda.synchronize()
        hq = [np.ascontiguousarray(r) for r in d_q.cpu().numpy()]
        call, idx, sc, found, flags = st.search_call(1, k)
        for i in range(10):
            call(hq[i])
        n_calls, per, cert = max(100, args.steps // 4), [], 0
        i0 = st.info()
        t0 = time.perf_counter()
        for i in range(n_calls):
            tc = time.perf_counter()
            call(hq[i % 64])
            per.append(time.perf_counter() - tc)
            cert += int(flags[0] & 1)
        dt = time.perf_counter() - t0
        per.sort()
        i1 = st.info()
        out = {"desc": f"{rows} x {dim} fp32 store + bf16 shadow, single query, top-{k}: scan of the bf16 image, exact rescoring from the fp32 rows",
               "api": "ragstore_search (host buffers)", "ms_per_query": dt / n_calls * 1e3, "median_ms": per[n_calls // 2] * 1e3,
               "p99_ms": per[min(n_calls - 1, int(n_calls * 0.99))] * 1e3, "calls": n_calls, "vectors_per_s": rows * n_calls / dt,
               "image_gbs": rows * dim * 2 / (per[n_calls // 2]) / 1e9, "algorithmic_bytes_per_query": rows * dim * 2,
               "scan_launches_per_query": (i1["scan_launches"] - i0["scan_launches"]) / n_calls, "certified_fraction": cert / n_calls}
        if do_parity:
            call(hq[0])
            ora = oracle_topk_streamed(0, rows, dim, False, [hq[0]], k, os.cpu_count() or 1)[0]
            ok = compare_hits(idx[0, :found[0]], sc[0, :found[0]], ora[0], ora[1])
            out["parity"] = {"checked": f"1 query x all {rows} rows vs the CPU oracle", "indices_bit_exact": ok[0], "scores_bit_exact": ok[1], "gpu_flags": int(flags[0])}
        return out
    finally:
        st.close()


def measure_k1(torch, rb, native, dev, rows=1 << 20, dim=768, reps=8):
    """K1 alone: device-resident fp32 rows -> normalised store rows (+ reciprocal norms, flags), read + write = 2*d*4 B/row."""
    buf = torch.empty((rows, dim), dtype=torch.float32, device="cuda")
    native.synth_fill_device(buf, 0, rows, dim, dim, CORPUS_SEED, device=dev)
    torch.cuda.synchronize()
    st = rb.RagStore(dim, rows * (reps + 1), rb.RAGSTORE_F32, dev)
    st.add_device(buf, rows, dim)          # warm-up (also maps the first chunk of memory)
    t0 = time.perf_counter()
    kern_ns = 0
    for _ in range(reps):
        st.add_device(buf, rows, dim)      # synchronous: returns when the rows are published
        kern_ns += st.info()["last_add_kernel_ns"]
    dt = time.perf_counter() - t0
    st.close()
    peak, _ = measured_peaks()
    gbs_call = 2.0 * rows * dim * 4 * reps / dt / 1e9
    gbs_kern = 2.0 * rows * dim * 4 * reps / max(kern_ns, 1)
    return {"kernel": "normalise_rows_vec_kernel (K1), device-resident fp32 source, 8 x 1M rows x 768",
            "vectors_per_s_call": rows * reps / dt, "achieved_gbs_call": gbs_call,
            "vectors_per_s_kernel": rows * reps / (kern_ns * 1e-9), "achieved_gbs": gbs_kern, "frac_of_measured_hbm": gbs_kern / peak,
            "bytes_per_vector": 2 * dim * 4,
            "note": "achieved_gbs = CUDA-event time of the K1 launches; *_call also includes mapping HBM behind the new rows (VMM) and the publish sync"}


def run_streaming(args, torch, dist, rb, native, rank, world, dev):
    """config5: host batches appended (K1) by a writer thread while the main thread serves 1k-query batches."""
    import threading
    c = STREAM
    dim, k, B, QB = c["dim"], c["top_k"], c["batch_rows"], c["queries_per_batch"]
    total = args.rows * world if args.rows else c["total_rows"]
    n_batches = max(1, total // B // world)                 # batches appended by THIS rank
    st = rb.RagStore(dim, n_batches * B, rb.RAGSTORE_F32, dev)
    st.set_row_blocking(B, world, rank)
    if not args.no_reserve:
        st.reserve_rows(n_batches * B)      # HBM for the announced capacity mapped at start-up, not inside the stream
    ts = torch.cuda.Stream()
    with torch.cuda.stream(ts):
        # This rank's part of the stream as DISTINCT rows in host memory: batch b of rank r is block b * world + r of the global
        # insertion order (rows [g*B, (g+1)*B) of the synthetic corpus), generated on the device and copied back once.  Repeating
        # one batch would fill the corpus with exact duplicates -- ties at every top-k boundary, which no margin can certify --
        # and would leave nothing to check the final answer against.
        h_all = torch.empty((n_batches * B, dim), dtype=torch.float32)
        if args.pinned:
            h_all = h_all.pin_memory()
        d_batch = torch.empty((B, dim), dtype=torch.float32, device="cuda")
        for b in range(n_batches):
            native.synth_fill_device(d_batch, (b * world + rank) * B, B, dim, dim, CORPUS_SEED, device=dev, stream=ts.cuda_stream)
            ts.synchronize()
            h_all[b * B:(b + 1) * B].copy_(d_batch)
        del d_batch
        h_np = h_all.numpy()
        h_batches = [h_np[b * B:(b + 1) * B] for b in range(n_batches)]   # the host-side batches the "application" appends
        d_q = torch.empty((QB, dim), dtype=torch.float32, device="cuda")
        native.synth_fill_device(d_q, 0, QB, dim, dim, QUERY_SEED, device=dev, stream=ts.cuda_stream)
        local_only = bool(os.environ.get("BENCH_STREAM_LOCAL"))    # experiment: no exchange between the ranks
        sh = rb.ShardedSearcher(st, rank, 1 if local_only else world, 0, max_k=k, device=dev)
        st.add(h_batches[0])                                 # first batch in before serving starts (also warms up)
        for _ in range(3):
            # the serving loop below, once, outside the timed region: NCCL sets its connections up lazily per collective
            # (device-wide allocations and mappings that stall the ingest stream as well)
            wflag = torch.zeros(1, dtype=torch.int32, device="cuda")
            if world > 1 and not local_only:
                dist.all_reduce(wflag, op=dist.ReduceOp.MIN)
            sh.search_batch_device(d_q, QB, k)
            ts.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        stop = threading.Event()
        t_ingest = [0.0]

        def writer():
            torch.cuda.set_device(dev)      # CUDA's current device is per thread; a new thread starts on device 0
            t0 = time.perf_counter()
            # one native call for the whole stream of batches (ragstore_add_many): every batch is still uploaded,
            # normalised and published on its own, but the producer loop does not compete for the Python GIL with the
            # serving loop of this process -- that contention, not the device, bounded earlier multi-GPU runs
            st.add_many(h_batches[1:])
            t_ingest[0] = time.perf_counter() - t0
            stop.set()

        w = threading.Thread(target=writer)
        qbatches = 0
        t0 = time.perf_counter()
        w.start()
        with ClockSampler(dev, enabled=(rank == 0)) as clk:
            trace = [0.0, 0.0, 0.0, 0.0] if os.environ.get("BENCH_STREAM_TRACE") else None
            while True:
                # every rank must agree on when to stop: the flag is reduced with the batch
                ta = time.perf_counter()
                flag = torch.tensor([1 if stop.is_set() else 0], dtype=torch.int32, device="cuda")
                if world > 1 and not local_only:
                    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                tb = time.perf_counter()
                sh.search_batch_device(d_q, QB, k)
                tc = time.perf_counter()
                ts.synchronize()
                td = time.perf_counter()
                qbatches += 1
                done = int(flag.item()) == 1
                if trace is not None:
                    te = time.perf_counter()
                    for i, v in enumerate((tb - ta, tc - tb, td - tc, te - td)):
                        trace[i] += v
                if done:
                    break
            if trace is not None and rank == 0:
                print("[stream trace] per iteration ms: flag+all_reduce %.2f, enqueue search %.2f, stream sync %.2f, flag read %.2f" %
                      tuple(1e3 * v / qbatches for v in trace), file=sys.stderr, flush=True)
        w.join()
        t_serve = time.perf_counter() - t0
        tt = torch.tensor([t_ingest[0], t_serve], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_ing, t_srv = float(tt[0].item()), float(tt[1].item())
        # all rows are resident now: query 0 against the oracle over the whole stream, every rank scoring its own blocks
        parity = None
        if not args.no_parity and not local_only:
            import numpy as np
            from oracle import oracle_c as oc
            bi, bs, bf = sh.search_batch_device(d_q, QB, k)
            ts.synchronize()
            g_idx, g_sc = bi[0].cpu().numpy(), bs[0].cpu().numpy()
            q0 = np.ascontiguousarray(d_q[0].cpu().numpy())
            threads = max(1, (os.cpu_count() or 1) // world)
            lists = []
            for b in range(n_batches):
                g = b * world + rank
                rows_b = oc.synth_rows(g * B, B, dim, CORPUS_SEED, n_threads=threads)
                li, ls = oc.query(rows_b, q0, k, n_threads=threads)
                lists.append((li + g * B, ls))
            mi, ms_ = merge_hit_lists(lists, k)
            o_ids = torch.full((k,), -1, dtype=torch.int64, device="cuda")
            o_scs = torch.full((k,), float("-inf"), dtype=torch.float64, device="cuda")
            o_ids[:len(mi)] = torch.from_numpy(mi).cuda()
            o_scs[:len(mi)] = torch.from_numpy(ms_).cuda()
            if world > 1:
                all_ids = torch.empty(world * k, dtype=torch.int64, device="cuda")
                all_scs = torch.empty(world * k, dtype=torch.float64, device="cuda")
                dist.all_gather_into_tensor(all_ids, o_ids)
                dist.all_gather_into_tensor(all_scs, o_scs)
            else:
                all_ids, all_scs = o_ids, o_scs
            ai, asc = all_ids.cpu().numpy(), all_scs.cpu().numpy()
            oi, osc = merge_hit_lists([(ai[ai >= 0], asc[ai >= 0])], k)
            ok_idx, ok_sc, _ = compare_hits(g_idx, g_sc, oi, osc)
            parity = {"checked": f"query 0 after the stream (all {n_batches * B * world} rows resident) vs the CPU oracle over the whole stream, every rank its own blocks",
                      "indices_bit_exact": ok_idx, "scores_bit_exact": ok_sc}
    rows_streamed = (n_batches - 1) * B * world
    info = st.info()
    out = None
    if rank == 0:
        out = {"metric": "streaming ingest vectors/sec while serving 1k-query batches (top-10)", "value": rows_streamed / t_ing,
               "unit": "vectors/s", "n_gpus": world, "steps": n_batches - 1, "warmup": 1, "ms_per_step": t_ing / max(1, n_batches - 1) * 1e3,
               "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "queries_per_s": qbatches * QB / t_srv, "query_batches_served": qbatches,
               "config": {"workload": "config5", "desc": c["desc"], "rows_total": n_batches * B * world, "rows_per_gpu": n_batches * B,
                          "dim": dim, "top_k": k, "batch_rows": B, "queries_per_batch": QB, "parallelism": f"block-cyclic x{world}",
                          "start": "one batch per GPU preloaded, the rest appended during the timed region",
                          "device_memory": "mapped on demand during the appends" if args.no_reserve else "mapped for the whole capacity before the timed region (ragstore_reserve_rows)",
                          "host_batches": "page-locked (direct PCIe copy)" if args.pinned else "pageable (staged through the library's pinned ring)"},
               "roofline": {"bound": "pcie (host->device copy of the appended rows; not an HBM or tensor bound)",
                            "achieved": rows_streamed * dim * 4 / t_ing / 1e9 / world, "peak": 64.0, "unit": "GB/s per GPU",
                            "frac": rows_streamed * dim * 4 / t_ing / 1e9 / world / 64.0, "traffic": None,
                            "peak_source": "nominal PCIe Gen5 x16"},
               "cpu_baseline": None,
               "e2e": {"value": rows_streamed / t_ing, "unit": "vectors/s", "h2d_bytes_per_step": B * dim * 4 * world, "d2h_bytes_per_step": 0,
                       "api": "ragstore_add (host rows) concurrent with ragstore_search_shard_batch_device + all-gather + merge"},
               "gpu_launches": int(info["batch_launches"]) * 3 + (n_batches - 1) * 2, "clocks": clk.summary(),
               "parity": parity, "final_count_rank0": info["count"], "sidecar_rows": info["sidecar_rows"]}
    st.close()
    return out


def tensor_block(r):
    burst, sust, src = measured_tensor_peaks()
    bf = r["storage"] != "f32" or r.get("operands") == "bf16"
    div = 1.0 if bf else 2.0    # tf32 runs at half the bf16 rate on the tensor pipe
    return {"achieved_tflops": r["tensor_tflops"],
            "kind": "bf16 (tcgen05.mma kind::f16, fp32 accumulate in TMEM)" if bf else "tf32 (tcgen05.mma kind::tf32, fp32 accumulate in TMEM)",
            "peak_tflops_bf16_burst": burst, "peak_tflops_bf16_sustained": sust, "peak_source": src,
            "frac_of_burst": r["tensor_tflops"] / (burst / div), "frac_of_sustained": r["tensor_tflops"] / (sust / div),
            "note": "tf32 peak taken as half the measured bf16 cuBLAS rate; the kernel runs inside a long step (power-capped clocks), so the "
                    "sustained figure is the fair denominator; see DESIGN.md K3",
            "certified_fraction": r.get("certified_fraction")}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS) + ["config5"],
                    help="default: target10m at N = 1, config4 at N > 1 (north_star's two stated targets)")
    ap.add_argument("--rows", type=int, default=0, help="override rows per GPU")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "p2p-2launch", "nccl"],
                    help="N > 1: how the per-shard hits meet -- inside the shard scan's last CTA over peer memory (default: one launch per "
                         "rank and query), the same exchange as a second launch, or NCCL all-gather + merge kernel")
    ap.add_argument("--no-reserve", action="store_true", help="config5: let appends map device memory on demand")
    ap.add_argument("--pinned", action="store_true", help="config5: the application's batches live in page-locked memory")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra config1 / config2 / config3 measurements")
    ap.add_argument("--no-parity", action="store_true", help="skip the full-corpus oracle check (profiling runs)")
    ap.add_argument("--single-process", action="store_true",
                    help="run the sharded workload through ONE handle over --gpus devices in this process (ragstore_create_sharded); "
                         "start with plain `python bench.py`, not torchrun")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    quiet_stdout()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload is None:
        # the reference arm may be started without torchrun: --gpus then says which line it is the partner of
        args.workload = default_workload(max(world, args.gpus) if args.impl == "reference" else world)

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import ragstore_b200 as rb
    from ragstore_b200 import native

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU oracle)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    if args.single_process and args.workload == "config5":
        emit(run_streaming_single_process(args, torch, rb, native, list(range(args.gpus))))
        return
    if args.single_process:
        name = args.workload if args.workload in WORKLOADS else "config4"
        r = measure_single_process(name, args, torch, rb, native, list(range(args.gpus)), do_parity=not args.no_parity)
        emit({"metric": "corpus vectors scanned/sec (single query, top-k), one process over all GPUs", "value": r["vectors_per_s"], "unit": "vectors/s",
              "n_gpus": args.gpus, "steps": r["calls"], "warmup": args.warmup, "ms_per_step": r["ms_per_query"], "higher_is_better": True,
              "scaling": "weak", "vs_baseline": None, "dtype": "f32" if r["storage"] == "f32" else "bf16 storage, f32 accumulate", "data": "synthetic",
              "config": config_dict(name, args.gpus, args.rows), "single_process": r,
              "e2e": {"value": r["vectors_per_s"], "unit": "vectors/s", "h2d_bytes_per_step": r["h2d_bytes_per_step"],
                      "d2h_bytes_per_step": r["d2h_bytes_per_step"], "ms_per_step": r["ms_per_query"], "steps": r["calls"], "api": r["api"]},
              "gpu_launches": r["calls"] * (args.gpus + 1), "parity": r.get("parity")})
        return
    if args.workload == "config5":
        out = run_streaming(args, torch, dist, rb, native, rank, world, local)
        if rank == 0:
            emit(out)
        if world > 1:
            dist.destroy_process_group()
        return
    main_res = measure_workload(args.workload, args, torch, dist, rb, native, rank, world, local, do_parity=not args.no_parity)
    k1 = None
    if world == 1 and not args.no_extra:
        try:
            k1 = measure_k1(torch, rb, native, local)
        except Exception as ex:
            k1 = {"error": repr(ex)[:200]}
    extra = {}
    if world == 1 and not args.no_extra and not args.rows:
        peak, _ = measured_peaks()
        for name in ("config2", "config1", "config3", "config3_shadow", "target10m"):
            if name == args.workload:
                continue
            try:
                sub = argparse.Namespace(**vars(args))
                sub.steps = max(100, args.steps) if name in ("config1", "config2") else max(50, min(args.steps, 200))
                sub.warmup = max(args.warmup, 5)
                r = measure_workload(name, sub, torch, dist, rb, native, rank, world, local, do_e2e=(name != "target10m"), do_cpu=False,
                                     do_parity=not args.no_parity)
                extra[name] = {"desc": r["desc"], "vectors_per_s": r["value"], "queries_per_s": r["queries_per_s"], "ms_per_step": r["ms_per_step"],
                               "achieved_gbs": r["achieved_gbs"], "frac_of_measured_hbm": r["achieved_gbs"] / peak,
                               "frac_of_8TBs_spec": r["achieved_gbs"] / 8000.0, "steps": sub.steps, "flags": r["flags"], "latency_ms": r.get("latency_ms"),
                               "clocks": r["clocks"], "parity": r.get("parity"), "certified_fraction": r.get("certified_fraction"),
                               "algorithmic_bytes_per_sweep": r["rows_per_gpu"] * r["dim"] * (4 if (r["storage"] == "f32" and r.get("operands") != "bf16") else 2)}
                if "e2e" in r:
                    extra[name]["e2e"] = r["e2e"]
                if "tensor_tflops" in r:
                    extra[name]["tensor"] = tensor_block(r)
            except Exception as ex:  # an extra must never take the headline down
                extra[name] = {"error": repr(ex)[:300]}
        try:
            extra["target10m_shadow_single_query"] = measure_shadow_single(args, torch, rb, native, local, do_parity=not args.no_parity)
        except Exception as ex:
            extra["target10m_shadow_single_query"] = {"error": repr(ex)[:300]}

    if rank == 0:
        peak, peak_src = measured_peaks()
        r = main_res
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                with open(tp) as f:
                    traffic = json.load(f).get(f"{args.workload}:{r['rows_per_gpu']}")
            except Exception:
                traffic = None
        batched = r.get("queries_per_step", 1) > 1
        out = {
            "metric": "corpus vectors scanned/sec (single query, top-k)" if not batched else "corpus vectors scanned/sec (per sweep of a 256-query batch, top-k)",
            "value": r["value"], "unit": "vectors/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if r["storage"] == "f32" else "bf16 storage, f32 accumulate", "data": "synthetic",
            "queries_per_s": r["queries_per_s"],
            "config": config_dict(args.workload, world, args.rows),
            "exchange": r.get("exchange"),
            "shard_scan_ms_per_rank": r.get("kernel_ms_per_rank"),
            # what ONE GPU does on the same shard without any exchange (rank 0's scan, timed alone in this run): the N = 1 point
            # of this workload's weak-scaling curve (the N = 1 default line is a different workload, 10M x 768 fp32)
            "single_gpu_same_shard": ({"ms_per_step": r["kernel_ms"], "value": r["rows_per_gpu"] / (r["kernel_ms"] * 1e-3), "unit": "vectors/s"}
                                      if world > 1 else None),
            "rescoring": "exact fp64 (reference arithmetic) for k' >= k candidates inside the same launch",
            "roofline": {"bound": "hbm", "achieved": r["achieved_gbs"], "peak": peak, "unit": "GB/s", "frac": r["achieved_gbs"] / peak,
                         "traffic": traffic, "peak_source": peak_src, "frac_of_8TBs_spec": r["achieved_gbs"] / 8000.0,
                         "kernel": "batch_score_kernel_2sm (threshold pass + full pass; prep, thresholds and finalize included: whole step timed)" if batched else "scan_topk_kernel",
                         "kernel_ms": r["kernel_ms"],
                         "algorithmic_bytes_per_launch": r["rows_per_gpu"] * r["dim"] * (4 if (r["storage"] == "f32" and r.get("operands") != "bf16") else 2),
                         "note": "peak = measured read+write copy bandwidth; a read-only stream can exceed it, so both denominators are given"},
            "tensor": tensor_block(r) if "tensor_tflops" in r else None,
            "cpu_baseline": r.get("cpu_baseline"),
            "e2e": {k: v for k, v in r.get("e2e", {}).items()} if "e2e" in r else None,
            "gpu_launches": args.steps * r["launches_per_step"],
            "clocks": r["clocks"],
            "parity": r.get("parity"),
            "latency_ms": r.get("latency_ms"),
            "result_flags": r["flags"],
            "ingest_vectors_per_s": r["ingest_vectors_per_s"],
            "k1_upload_normalise": k1,
            "extra": extra,
        }
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        if world > 1 and not args.no_extra:
            # The same corpus behind ONE handle in ONE process -- what a singleton RagIndex uses.  Measured by rank 0 after
            # the process group is gone: the other ranks have left their GPUs (no collective kernel spins there).
            time.sleep(2.0)
            try:
                out["extra"]["single_process_sharded"] = measure_single_process(args.workload, args, torch, rb, native, list(range(world)),
                                                                                do_parity=not args.no_parity)
            except Exception as ex:
                out["extra"]["single_process_sharded"] = {"error": repr(ex)[:300]}
        emit(out)


if __name__ == "__main__":
    main()
