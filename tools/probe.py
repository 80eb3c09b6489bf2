#!/usr/bin/env python
"""Scratch timings on one GPU (not the bench contract -- see bench.py).  One file, sub-commands:

  python tools/probe.py scan [N]              single-query scan on N x 768 fp32 (default 10M), 30 launches, phase times
  python tools/probe.py scan-sizes            the scan on the shapes used while tuning (10k .. 4M rows, k = 5/10/100, fp32 / bf16)
  python tools/probe.py widths                scan bandwidth across embedding widths 100 .. 4096 (profiles/r2_scan_widths.txt)
  python tools/probe.py batch [small|sizes|bf16]   the batched tensor-core search (256 / 64 / 100 queries)
  python tools/probe.py few                   2 .. 15 queries at once: scans against one tensor-core sweep
  python tools/probe.py pair [N] [f32|bf16]   the CTA-pair kernel (RAGSTORE_BATCH_2SM=0 selects the single-CTA form)
  python tools/probe.py shadow [N]            256 queries over N x 1024 fp32 with and without the bf16 shadow
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch  # noqa: E402
import ragstore_b200 as rb  # noqa: E402
from ragstore_b200 import native  # noqa: E402
from quick_bench import fill, run as run_scan  # noqa: E402


def run_batch(n, d, nq, k, iters=5, bf16=False, flags=0):
    st = rb.RagStore(d, n, rb.RAGSTORE_BF16 if bf16 else rb.RAGSTORE_F32, 0, flags)
    fill(st, n, d, 1234, bf16)
    q = torch.empty((nq, d), dtype=torch.float32, device="cuda")
    native.synth_fill_device(q, 0, nq, d, d, 5678)
    idx = torch.zeros((nq, k), dtype=torch.int64, device="cuda")
    sc = torch.zeros((nq, k), dtype=torch.float64, device="cuda")
    fnd = torch.zeros(nq, dtype=torch.int32, device="cuda")
    flg = torch.zeros(nq, dtype=torch.int32, device="cuda")
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)
    s = ts.cuda_stream
    for _ in range(2):
        st.search_device(q, nq, k, idx, sc, fnd, flg, stream=s)
    torch.cuda.synchronize()
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    for i in range(iters):
        e0[i].record()
        st.search_device(q, nq, k, idx, sc, fnd, flg, stream=s)
        e1[i].record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in zip(e0, e1))
    med = ms[len(ms) // 2]
    shadow = bool(flags & rb.RAGSTORE_CREATE_BF16_SHADOW)
    esz = 2 if (bf16 or shadow) else 4
    gbs, tf = n * d * esz / med / 1e6, 2.0 * nq * n * d / med / 1e9
    cert = ((flg & 1) != 0).float().mean().item()
    kind = "bf16" if (bf16 or shadow) else "tf32"
    print(f"n={n} d={d} nq={nq} k={k}{' shadow' if shadow else ''}: median {med:.3f} ms -> operand bytes {gbs:.0f} GB/s, {tf:.0f} TFLOP/s {kind}, "
          f"{nq / med * 1e3:.0f} queries/s, certified {cert:.3f}", flush=True)
    st.close()


def main():
    cmd = sys.argv[1] if len(sys.argv) > 1 else "scan"
    arg = sys.argv[2:] + [None, None]
    F, B = rb.RAGSTORE_F32, rb.RAGSTORE_BF16
    if cmd == "scan":
        run_scan(int(arg[0] or 10_000_000), 768, 10, F, iters=30)
    elif cmd == "scan-sizes":
        for (n, d, k, t) in [(10000, 768, 5, F), (10000, 768, 100, F), (10000, 768, 100, B), (1000000, 768, 10, F), (4000000, 768, 10, F),
                             (2000000, 768, 100, F), (2000000, 768, 10, B), (2000000, 768, 100, B), (1000000, 1024, 10, F), (300000, 4096, 10, F)]:
            run_scan(n, d, k, t)
    elif cmd == "widths":
        for (d, t) in [(384, F), (768, F), (1024, F), (1536, F), (2048, F), (3072, F), (4096, F), (1536, B), (3072, B), (4096, B), (1000, F), (100, F)]:
            run_scan(int(6e9 // (d * (4 if t == F else 2))), d, 10, t)       # ~6 GB per corpus: far beyond L2
    elif cmd == "batch":
        if arg[0] == "small":
            run_batch(40000, 768, 256, 10, iters=3)
            run_batch(40000, 768, 64, 10, iters=3)
        elif arg[0] == "sizes":
            for n in (40000, 150000, 500000, 1000000, 2000000, 4000000):
                run_batch(n, 1024, 256, 10)
                run_batch(n, 768, 64, 10)
                run_batch(n, 768, 256, 10, bf16=True)
        elif arg[0] == "bf16":
            for nq in (256, 128, 32):
                run_batch(10000000, 768, nq, 10, bf16=True)
        else:
            run_batch(2000000, 1024, 256, 10)
            run_batch(10000000, 1024, 256, 10, iters=3)
            run_batch(6000000, 1024, 100, 10, iters=3)
    elif cmd == "few":      # RAGSTORE_BATCH_MINQ=99 forces per-query scans for comparison
        for n in (10000, 100000, 1000000, 4000000):
            for nq in (2, 3, 4, 8, 15):
                run_batch(n, 768, nq, 10, iters=7)
    elif cmd == "pair":
        n, which = int(arg[0] or 4000000), arg[1] or "all"
        if which in ("all", "f32"):
            run_batch(n, 1024, 256, 10)
        if which in ("all", "bf16"):
            run_batch(n, 768, 256, 10, bf16=True)
    elif cmd == "shadow":
        n = int(arg[0] or 4000000)
        for nq in (256, 128, 16):
            run_batch(n, 1024, nq, 10)
            run_batch(n, 1024, nq, 10, flags=rb.RAGSTORE_CREATE_BF16_SHADOW)
        run_batch(n, 1024, 256, 100, flags=rb.RAGSTORE_CREATE_BF16_SHADOW)
    else:
        sys.exit(__doc__)


if __name__ == "__main__":
    main()
