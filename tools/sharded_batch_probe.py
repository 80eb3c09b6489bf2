"""Where does a sharded 1024-query batch spend its time?  torchrun --nproc-per-node N tools/sharded_batch_probe.py [rows_per_gpu]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
import torch.distributed as dist
import ragstore_b200 as rb
from ragstore_b200 import native

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
dev = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", dev))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2500000
d, QB, k = 768, 1024, 10
st = rb.RagStore(d, n, rb.RAGSTORE_F32, dev)
buf = torch.empty((1 << 18, d), dtype=torch.float32, device="cuda")
for r0 in range(0, n, 1 << 18):
    m = min(1 << 18, n - r0)
    native.synth_fill_device(buf, rank * n + r0, m, d, d, 1234, device=dev)
    torch.cuda.synchronize()
    st.add_device(buf, m, d)
del buf
q = torch.empty((QB, d), dtype=torch.float32, device="cuda"); native.synth_fill_device(q, 0, QB, d, d, 5678, device=dev)
ts = torch.cuda.Stream()
with torch.cuda.stream(ts):
    sh = rb.ShardedSearcher(st, rank, world, rank * n, max_k=k, device=dev)
    for _ in range(3):
        sh.search_batch_device(q, QB, k)
    ts.synchronize()
    if world > 1:
        dist.barrier()
    need = QB * k * 16
    def timed(fn, reps=10):
        ts.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record(); ts.synchronize()
        return e0.elapsed_time(e1) / reps
    t_shard = timed(lambda: st.search_shard_batch_device(q, QB, k, sh._b_cand, None, stream=ts.cuda_stream))
    t_full = timed(lambda: sh.search_batch_device(q, QB, k))
    t_gather = timed(lambda: dist.all_gather_into_tensor(sh._b_all[: world * need], sh._b_cand[:need])) if world > 1 else 0.0
    t_merge = timed(lambda: native.merge_shards_batch_device(sh._b_all if world > 1 else sh._b_cand, world, QB, k, k, sh._b_idx, sh._b_score, sh._b_found, device=dev, stream=ts.cuda_stream))
    if rank == 0:
        print(f"world {world}, {n} rows/GPU, {QB} queries: shard search {t_shard:.2f} ms, all-gather {t_gather:.3f} ms, merge {t_merge:.3f} ms, whole step {t_full:.2f} ms", flush=True)
st.close()
if world > 1:
    dist.destroy_process_group()
