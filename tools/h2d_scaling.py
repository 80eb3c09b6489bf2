"""How does page-locked host -> device copy bandwidth scale with the number of GPUs copying at once on this box?
(explains config5's ingest rate on multi-GPU boxes)   torchrun --nproc-per-node N tools/h2d_scaling.py"""
import os, time
import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
nbytes = 256 << 20
h = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
d = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
for _ in range(3):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
t0 = time.perf_counter()
reps = 20
for _ in range(reps):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
mine = nbytes * reps / dt / 1e9
t = torch.tensor([mine], dtype=torch.float64, device="cuda")
if world > 1:
    allv = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allv, t)
    vals = [float(x.item()) for x in allv]
else:
    vals = [mine]
if rank == 0:
    print(f"H2D pinned, {world} GPUs copying at once: per GPU {['%.1f' % v for v in vals]} GB/s, aggregate {sum(vals):.1f} GB/s", flush=True)
if world > 1:
    dist.destroy_process_group()
