"""World-size-2 gloo test (CPU) of the host side of the row-sharded search: block partition, global ids,
all-gather payload layout and the (score desc, id asc) merge order -- with the oracle standing in for the
per-shard GPU search, and checked against the oracle over the whole corpus."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import ragstore_b200 as rb
from oracle import oracle_c as oc

REC = np.dtype([("score", "<f8"), ("id", "<i8")])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _merge_np(recs, k):
    """numpy restatement of merge_shards_kernel: recs = [(score, id)], id -1 = empty."""
    live = [(s, i) for s, i in recs if i >= 0]
    live.sort(key=lambda t: (-t[0] if t[0] == t[0] else float("inf"), t[1]))
    return live[:max(1, k)]


def _shard_rows(lo, hi, dim):
    rows = oc.synth_rows(lo, hi - lo, dim, 1234)
    rows[: min(3, hi - lo)] = oc.synth_rows(0, 1, dim, 1234)[0]        # duplicates across shards -> id tie-break
    return rows


def _worker(rank, world, port, n, dim, k, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = rb.shard_range(n, world, rank)
    rows = _shard_rows(lo, hi, dim)
    q = oc.synth_rows(0, 1, dim, 5678)[0]
    idx, sc = oc.query(rows, q, k)                                       # stands in for search_shard_device
    per, total = rb.gather_layout(world, k)
    local = np.zeros(per, dtype=np.uint8)
    rec = local.view(REC)
    rec["score"], rec["id"] = -np.inf, -1
    rec["score"][: len(idx)], rec["id"][: len(idx)] = sc, idx + lo       # global id = row_base + local row
    gathered = torch.zeros(total, dtype=torch.uint8)
    dist.all_gather_into_tensor(gathered, torch.from_numpy(local))
    allrec = gathered.numpy().view(REC)
    out[rank] = _merge_np([(float(s), int(i)) for s, i in zip(allrec["score"], allrec["id"])], k)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n,k", [(1001, 10), (7, 100)])
def test_two_rank_sharded_search_matches_single_index(n, k):
    dim, world = 64, 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), n, dim, k, out), nprocs=world, join=True)
    full = np.concatenate([_shard_rows(*rb.shard_range(n, world, r), dim) for r in range(world)])
    q = oc.synth_rows(0, 1, dim, 5678)[0]
    eidx, esc = oc.query(full, q, k)
    for r in range(world):
        got = out[r]
        assert [i for _, i in got] == [int(i) for i in eidx]
        assert [s for s, _ in got] == [float(s) for s in esc]


def test_shard_range_covers_everything_once():
    for n in (0, 1, 7, 100, 1001):
        for world in (1, 2, 3, 8):
            seen = []
            for r in range(world):
                lo, hi = rb.shard_range(n, world, r)
                assert 0 <= lo <= hi <= n
                seen.extend(range(lo, hi))
            assert seen == list(range(n))
    assert rb.gather_layout(8, 100) == (1600, 12800)
