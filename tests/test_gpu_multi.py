"""Two GPUs, two processes: the peer-memory exchange (csrc/exchange.cu) against the NCCL all-gather + merge form and
against the CPU oracle over the whole corpus.  Skipped on boxes with one GPU (run with `gpurun --gpus 2`)."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n_total, d, k, nq, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import ragstore_b200 as rb
    from oracle import oracle_c as oc
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    lo, hi = rb.shard_range(n_total, world, rank)
    rows = oc.synth_rows(lo, hi - lo, d, 1234)
    qs = torch.from_numpy(oc.synth_rows(0, nq, d, 5678)).cuda()
    res = {}
    with rb.RagStore(d, hi - lo, rb.RAGSTORE_F32, rank) as st:
        st.add(rows)
        for mode in ("p2p", "nccl"):
            sh = rb.ShardedSearcher(st, rank, world, lo, max_k=k, device=rank, exchange=mode, max_queries=nq)
            if mode == "p2p":
                assert sh.peer is not None, "peer inboxes could not be mapped"
            single = []
            for i in range(6):                      # consecutive calls: both inbox slots, several times
                idx, sc = sh.search(qs[i], k)
                single.append((idx.copy(), sc.copy()))
            if mode == "p2p":                       # host query in (kernel parameters), host hits out (page-locked memory)
                hq = np.ascontiguousarray(qs.cpu().numpy())
                for i in range(6):
                    hidx, hsc = sh.search_host(hq[i], k)
                    assert np.array_equal(hidx, single[i][0]) and np.array_equal(hsc.view(np.uint64), single[i][1].view(np.uint64))
            bi, bs, bf = sh.search_batch_device(qs, nq, k)
            torch.cuda.synchronize()
            assert int(sh.d_status.abs().sum().item()) == 0
            res[mode] = (single, bi.cpu().numpy().copy(), bs.cpu().numpy().copy(), bf.cpu().numpy().copy())
            dist.barrier()
    np.save(os.path.join(out_dir, f"r{rank}_p2p_idx.npy"), res["p2p"][1])
    np.save(os.path.join(out_dir, f"r{rank}_p2p_sc.npy"), res["p2p"][2])
    for a, b in zip(res["p2p"][0], res["nccl"][0]):
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint64), b[1].view(np.uint64))
    assert np.array_equal(res["p2p"][1], res["nccl"][1])
    assert np.array_equal(res["p2p"][2].view(np.uint64), res["nccl"][2].view(np.uint64))
    assert np.array_equal(res["p2p"][3], res["nccl"][3])
    dist.destroy_process_group()


def test_peer_memory_exchange_two_gpus(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    from oracle import oracle_c as oc
    n_total, d, k, nq = 30001, 256, 10, 24
    mp.spawn(_worker, args=(2, 29533, n_total, d, k, nq, str(tmp_path)), nprocs=2, join=True)
    rows = oc.synth_rows(0, n_total, d, 1234)
    qs = oc.synth_rows(0, nq, d, 5678)
    idx0, sc0 = np.load(tmp_path / "r0_p2p_idx.npy"), np.load(tmp_path / "r0_p2p_sc.npy")
    idx1, sc1 = np.load(tmp_path / "r1_p2p_idx.npy"), np.load(tmp_path / "r1_p2p_sc.npy")
    assert np.array_equal(idx0, idx1) and np.array_equal(sc0.view(np.uint64), sc1.view(np.uint64))   # every rank: same answer
    for i in range(nq):
        eidx, esc = oc.query(rows, qs[i], k, n_threads=4)
        assert list(idx0[i]) == list(eidx), i
        assert np.array_equal(sc0[i].view(np.uint64), esc.view(np.uint64)), i
