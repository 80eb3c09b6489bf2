"""The parity machinery inside bench.py is itself checked on the CPU: the chunked oracle (rows regenerated chunk by
chunk, per-chunk hit lists merged under the reference's order) must equal the oracle run over the whole corpus at once,
and the per-shard merge must equal the oracle over the concatenation -- otherwise a green `parity` in a bench line would
prove nothing.  Also: both arms of the bench describe their workload with the same `config` object."""
import importlib.util
import os
import subprocess
import sys
import json

import numpy as np

from oracle import oracle_c as oc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(bench)


def test_streamed_oracle_equals_the_oracle_over_the_whole_corpus():
    n, d, k = 5003, 96, 100
    qs = oc.synth_rows(0, 3, d, bench.QUERY_SEED)
    for bf in (False, True):
        rows = oc.synth_rows(0, n, d, bench.CORPUS_SEED, round_bf16=bf)
        got = bench.oracle_topk_streamed(0, n, d, bf, list(qs), k, 2, chunk_rows=777)
        for i in range(3):
            eidx, esc = oc.query(rows, qs[i], k, n_threads=2)
            assert list(got[i][0]) == list(eidx)
            assert np.array_equal(got[i][1].view(np.uint64), esc.view(np.uint64))


def test_streamed_oracle_breaks_ties_by_lower_global_id_across_chunks_and_shards():
    n, d, k, world = 4000, 64, 10, 3
    q = oc.synth_rows(0, 1, d, bench.QUERY_SEED)[0]
    rows = oc.synth_rows(0, n, d, bench.CORPUS_SEED)
    eidx, esc = oc.query(rows, q, k, n_threads=2)
    per = (n + world - 1) // world
    lists = []
    for r in range(world):
        lo, hi = r * per, min(n, (r + 1) * per)
        lists.append(bench.oracle_topk_streamed(lo, hi - lo, d, False, [q], k, 1, chunk_rows=500)[0])
    ids, scs = bench.merge_hit_lists(lists, k)
    assert list(ids) == list(eidx) and np.array_equal(scs.view(np.uint64), esc.view(np.uint64))
    ok_idx, ok_sc, rel = bench.compare_hits(ids, scs, eidx, esc)
    assert ok_idx and ok_sc and rel == 0.0
    ok_idx, ok_sc, _ = bench.compare_hits(ids[::-1], scs, eidx, esc)
    assert not ok_idx


def test_both_arms_describe_the_workload_identically():
    for world, name in ((1, "target10m"), (8, "config4")):
        assert bench.default_workload(world) == name
    ours = bench.config_dict("config4", 8)
    env = dict(os.environ, WORLD_SIZE="1", RANK="0")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "8", "--steps", "1", "--warmup", "1", "--rows", "20000"],
                       capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["config"] == bench.config_dict("config4", 8, 20000)
    assert set(line["config"]) == set(ours)
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["cpu_baseline"]["kind"] == "port"
