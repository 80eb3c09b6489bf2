"""Per-call time of ragstore_add for a stream of 64k-row page-locked batches (where do the spikes come from?)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ragstore_b200 as rb
from ragstore_b200 import native

n_batches = int(sys.argv[1]) if len(sys.argv) > 1 else 60
world = int(os.environ.get("WORLD_SIZE", 1))
if world > 1:   # under torchrun: same measurement with an initialised NCCL process group (peer access between the GPUs)
    import torch.distributed as dist
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    dist.barrier()
    if len(sys.argv) > 2 and sys.argv[2] == "allgather":
        x = torch.zeros(1024, device="cuda"); y = torch.zeros(1024 * world, device="cuda")
        dist.all_gather_into_tensor(y, x); torch.cuda.synchronize()
B, dim = 65536, 768
d = torch.empty((B, dim), dtype=torch.float32, device="cuda")
native.synth_fill_device(d, 0, B, dim, dim, 1234, device=torch.cuda.current_device())
h = d.cpu().pin_memory().numpy()
st = rb.RagStore(dim, n_batches * B, rb.RAGSTORE_F32, torch.cuda.current_device())
ts = []
for i in range(n_batches):
    t0 = time.perf_counter(); st.add(h); ts.append((time.perf_counter() - t0) * 1e3)
print(f"[rank {os.environ.get('RANK', 0)}] per-call ms:", " ".join(f"{t:.1f}" for t in ts))
s = sorted(ts)
print(f"median {s[len(s)//2]:.2f} ms ({B*dim*4/s[len(s)//2]/1e6:.1f} GB/s), total {sum(ts):.1f} ms, rows/s overall {n_batches*B/sum(ts)*1e3/1e6:.2f} M")
st.close()
