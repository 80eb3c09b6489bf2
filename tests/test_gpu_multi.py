"""Several GPUs: (i) one process per GPU -- the peer-memory exchange (csrc/exchange.cu) against the NCCL all-gather +
merge form and against the CPU oracle over the whole corpus; (ii) one process, one handle over all GPUs
(ragstore_create_sharded).  Skipped on boxes with one GPU (run with `gpurun --gpus 2` / `--gpus 8`)."""
import os
import sys
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CASES = {
    # name: n_total, dim, k, nq, bf16, block (0 = contiguous shards), consecutive single-query calls, delayed rank
    "f32_top10": dict(n=30001, d=256, k=10, nq=24, bf16=False, block=0, calls=6, delay_rank=None),
    # top-100 needs candidate lists of 128 (16x the exchange payload of top-10); bf16 shards; many calls: both inbox
    # slots and many sequence-number wraps of the slot index; one rank launches 50 ms late: everybody else WAITS for its
    # records inside the kernel (flag polling over NVLink), and no status word may report a time-out
    "bf16_top100_200calls_one_rank_late": dict(n=40003, d=128, k=100, nq=8, bf16=True, block=0, calls=200, delay_rank=1),
    # streaming layout (config 5): blocks of rows dealt round-robin, global id = place in the insertion order
    "f32_block_cyclic": dict(n=26000, d=192, k=10, nq=40, bf16=False, block=1000, calls=6, delay_rank=None),
    # rows crowding around query 0 closer together than the tensor-core pass can tell apart (eps ~2e-3) but far enough
    # for the fp32 scan: the per-shard certificate must travel with the hits, and uncertified queries must be redone
    "f32_near_duplicates": dict(n=24000, d=256, k=10, nq=16, bf16=False, block=0, calls=3, delay_rank=None, near_dup=True),
}


def make_corpus(case):
    from oracle import oracle_c as oc
    rows = oc.synth_rows(0, case["n"], case["d"], 1234, round_bf16=case["bf16"])
    if case.get("near_dup"):
        q = oc.synth_rows(0, 1, case["d"], 5678)[0].astype(np.float64)
        for j in range(220):        # 20 rows at cosine ~1 - 1e-4, 200 rows at ~1 - 1e-3, spread over all shards
            eps = 0.014 if j < 20 else 0.045
            r = 97 * j + 11
            v = q + eps * rows[r].astype(np.float64)
            rows[r] = (v / np.linalg.norm(v)).astype(np.float32)
    return rows


def _shard_rows_of(case, world, rank):
    """(global row ids of this rank's rows, in local order)"""
    n, block = case["n"], case["block"]
    if block == 0:
        per = (n + world - 1) // world
        lo = min(n, rank * per)
        return np.arange(lo, min(n, lo + per), dtype=np.int64)
    ids = [np.arange(b * block, min(n, (b + 1) * block), dtype=np.int64) for b in range((n + block - 1) // block) if b % world == rank]
    return np.concatenate(ids) if ids else np.zeros(0, dtype=np.int64)


def _worker(rank, world, port, name, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import ragstore_b200 as rb
    from oracle import oracle_c as oc
    case = CASES[name]
    d, k, nq, bf = case["d"], case["k"], case["nq"], case["bf16"]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    ids = _shard_rows_of(case, world, rank)
    corpus = make_corpus(case)
    rows = corpus[ids]
    qs = torch.from_numpy(oc.synth_rows(0, nq, d, 5678)).cuda()
    res = {}
    with rb.RagStore(d, max(1, len(ids)), rb.RAGSTORE_BF16 if bf else rb.RAGSTORE_F32, rank) as st:
        if len(ids):
            st.add(rows, prenormalised=bf)
        for mode in ("p2p", "p2p-2launch", "nccl"):     # fused into the scan's last CTA / separate exchange kernel / NCCL
            sh = rb.ShardedSearcher(st, rank, world, int(ids[0]) if (len(ids) and case["block"] == 0) else 0, max_k=k, device=rank,
                                    exchange=mode, max_queries=nq)
            if case["block"]:
                st.set_row_blocking(case["block"], world, rank)
            if mode != "nccl":
                assert sh.peer is not None, "peer inboxes could not be mapped"
            single = []
            n_calls = case["calls"] if mode == "p2p" else min(case["calls"], 12)
            for i in range(n_calls):                # consecutive calls: both inbox slots, many times
                if mode == "p2p" and case["delay_rank"] == rank and i == n_calls // 2:
                    time.sleep(0.05)                # this rank is late: the others wait for it inside their kernels
                idx, sc = sh.search(qs[i % nq], k)
                single.append((idx.copy(), sc.copy()))
            if mode != "nccl" and d <= 1024:        # host query in (kernel parameters), host hits out (page-locked memory)
                hq = np.ascontiguousarray(qs.cpu().numpy())
                for i in range(min(6, n_calls)):
                    hidx, hsc = sh.search_host(hq[i % nq], k)
                    assert np.array_equal(hidx, single[i][0]) and np.array_equal(hsc.view(np.uint64), single[i][1].view(np.uint64))
            bi, bs, bfnd = sh.search_batch_device(qs, nq, k)
            torch.cuda.synchronize()
            assert int(sh.d_status.abs().sum().item()) == 0          # no peer ever timed out
            res[mode] = (single[:6], bi.cpu().numpy().copy(), bs.cpu().numpy().copy(), bfnd.cpu().numpy().copy(), single)
            if case.get("near_dup"):
                ci, cs, cf, cflags = sh.search_batch(qs, nq, k)     # certificate carried across ranks, uncertified queries redone
                res[mode] = (single[:6], ci, cs, cf, single)
                np.save(os.path.join(out_dir, f"r{rank}_{mode}_flags.npy"), cflags)
                np.save(os.path.join(out_dir, f"r{rank}_{mode}_localflags.npy"), sh._b_flags[:nq].cpu().numpy())
            dist.barrier()
    np.save(os.path.join(out_dir, f"r{rank}_p2p_idx.npy"), res["p2p"][1])
    np.save(os.path.join(out_dir, f"r{rank}_p2p_sc.npy"), res["p2p"][2])
    np.save(os.path.join(out_dir, f"r{rank}_single_idx.npy"), np.stack([s[0] for s in res["p2p"][4]]))
    np.save(os.path.join(out_dir, f"r{rank}_single_sc.npy"), np.stack([s[1] for s in res["p2p"][4]]))
    for m in ("p2p", "p2p-2launch"):
        for a, b in zip(res[m][0], res["nccl"][0]):
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint64), b[1].view(np.uint64)), m
        assert np.array_equal(res[m][1], res["nccl"][1]), m
        assert np.array_equal(res[m][2].view(np.uint64), res["nccl"][2].view(np.uint64)), m
        assert np.array_equal(res[m][3], res["nccl"][3]), m
    dist.destroy_process_group()


def _worlds():
    import torch
    g = torch.cuda.device_count()
    return sorted({w for w in (2, g) if 2 <= w <= g})


@pytest.mark.parametrize("name", sorted(CASES))
def test_peer_memory_exchange_between_real_gpus(tmp_path, name):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    from oracle import oracle_c as oc
    case = CASES[name]
    for wi, world in enumerate(_worlds()):
        out = tmp_path / f"w{world}"
        out.mkdir()
        mp.spawn(_worker, args=(world, 29533 + wi, name, str(out)), nprocs=world, join=True)
        rows = make_corpus(case)
        qs = oc.synth_rows(0, case["nq"], case["d"], 5678)
        if case.get("near_dup"):
            flags = np.load(out / "r0_p2p_flags.npy")
            assert (flags & 1).all()                                  # after the redo every query is certified
            assert not (np.load(out / "r0_p2p_localflags.npy")[0] & 1) or world > 2   # the tensor-core pass alone was not enough for query 0
        idx0, sc0 = np.load(out / "r0_p2p_idx.npy"), np.load(out / "r0_p2p_sc.npy")
        for r in range(1, world):                   # every rank: the same answer
            assert np.array_equal(idx0, np.load(out / f"r{r}_p2p_idx.npy"))
            assert np.array_equal(sc0.view(np.uint64), np.load(out / f"r{r}_p2p_sc.npy").view(np.uint64))
        expect = [oc.query(rows, qs[i], case["k"], n_threads=4) for i in range(case["nq"])]
        for i in range(case["nq"]):
            assert list(idx0[i]) == list(expect[i][0]), (world, i)
            assert np.array_equal(sc0[i].view(np.uint64), expect[i][1].view(np.uint64)), (world, i)
        sidx, ssc = np.load(out / "r0_single_idx.npy"), np.load(out / "r0_single_sc.npy")
        for c in range(sidx.shape[0]):              # every one of the consecutive calls, not just the last
            e = expect[c % case["nq"]]
            assert list(sidx[c]) == list(e[0]) and np.array_equal(ssc[c].view(np.uint64), e[1].view(np.uint64)), (world, c)


@pytest.mark.parametrize("dtype,k,block", [(0, 10, 512), (1, 100, 4096), (0, 300, 1000)])
def test_one_handle_over_all_gpus(dtype, k, block):
    """ragstore_create_sharded over every GPU of the box, from ONE process: the shards' last CTAs store their hits into
    GPU 0's memory over NVLink and GPU 0 merges.  Against the oracle over the whole corpus."""
    import torch
    import ragstore_b200 as rb
    from oracle import oracle_c as oc
    g = torch.cuda.device_count()
    if g < 2:
        pytest.skip("needs two GPUs")
    bf = dtype == 1
    n, d, nq = 60000, 256, 70
    rows = oc.synth_rows(0, n, d, 77, round_bf16=bf)
    qs = oc.synth_rows(0, nq, d, 78)
    st = rb.RagStore.create_sharded(d, n, list(range(g)), dtype=rb.RAGSTORE_BF16 if bf else rb.RAGSTORE_F32, block_rows=block)
    try:
        st.add(rows[: n // 3], prenormalised=bf)
        st.add(rows[n // 3:], prenormalised=bf)
        for rep in range(50):                       # many consecutive calls over the peers
            i = rep % nq
            idx, sc, found, flags = st.search(qs[i], k)
            if rep < 5 or rep == 49:
                eidx, esc = oc.query(rows, qs[i], k, n_threads=4)
                assert list(idx[0, :found[0]]) == list(eidx)
                assert np.array_equal(sc[0, :found[0]].view(np.uint64), esc.view(np.uint64))
        idx, sc, found, flags = st.search(qs, k)    # batch: tensor-core path on every device (k <= 24), scans otherwise
        for i in range(0, nq, 3):
            eidx, esc = oc.query(rows, qs[i], k, n_threads=4)
            assert list(idx[i, :found[i]]) == list(eidx), i
            assert np.array_equal(sc[i, :found[i]].view(np.uint64), esc.view(np.uint64)), i
        assert st.info()["scratch_overlaps"] == 0
    finally:
        st.close()
