"""Scratch timing of the batched (tensor-core) search."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ragstore_b200 as rb
from ragstore_b200 import native
from quick_bench import fill

def run(n, d, nq, k, iters=5, bf16=False):
    st = rb.RagStore(d, n, rb.RAGSTORE_BF16 if bf16 else rb.RAGSTORE_F32)
    fill(st, n, d, 1234, bf16)
    q = torch.empty((nq, d), dtype=torch.float32, device="cuda")
    native.synth_fill_device(q, 0, nq, d, d, 5678)
    idx = torch.zeros((nq, k), dtype=torch.int64, device="cuda"); sc = torch.zeros((nq, k), dtype=torch.float64, device="cuda")
    fnd = torch.zeros(nq, dtype=torch.int32, device="cuda"); flg = torch.zeros(nq, dtype=torch.int32, device="cuda")
    ts = torch.cuda.Stream(); torch.cuda.set_stream(ts); s = ts.cuda_stream
    for _ in range(2):
        st.search_device(q, nq, k, idx, sc, fnd, flg, stream=s)
    torch.cuda.synchronize()
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]; e1 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    for i in range(iters):
        e0[i].record(); st.search_device(q, nq, k, idx, sc, fnd, flg, stream=s); e1[i].record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in zip(e0, e1)); med = ms[len(ms) // 2]
    gbs = n * d * (2 if bf16 else 4) / med / 1e6; tf = 2.0 * nq * n * d / med / 1e9
    cert = ((flg & 1) != 0).float().mean().item()
    print(f"n={n} d={d} nq={nq} k={k}: median {med:.3f} ms -> corpus {gbs:.0f} GB/s ({gbs/6542.1:.3f} of measured HBM), {tf:.0f} TFLOP/s {'bf16' if bf16 else 'tf32'}, "
          f"{nq/med*1e3:.0f} queries/s, {n*nq/med/1e6:.2f} G scores/s, certified {cert:.3f}", flush=True)
    st.close()

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "small":
        run(40000, 768, 256, 10, iters=3)
        run(40000, 768, 64, 10, iters=3)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "sizes":
        for n in (40000, 150000, 500000, 1000000, 2000000, 4000000):
            run(n, 1024, 256, 10, iters=5)
            run(n, 768, 64, 10, iters=5)
            run(n, 768, 256, 10, iters=5, bf16=True)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "bf16":
        run(10000000, 768, 256, 10, iters=5, bf16=True)
        run(10000000, 768, 128, 10, iters=5, bf16=True)
        run(10000000, 768, 32, 10, iters=5, bf16=True)
        run(1000000, 768, 256, 10, iters=5, bf16=True)
        sys.exit(0)
    run(2000000, 1024, 256, 10)
    run(6000000, 1024, 256, 10, iters=3)
    run(10000000, 1024, 256, 10, iters=3)
    run(6000000, 1024, 100, 10, iters=3)
