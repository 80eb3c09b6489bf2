"""Scratch timing of the device-resident search (not the bench contract; see bench.py)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ragstore_b200 as rb
from ragstore_b200 import native

def fill(st, n, d, seed, bf16=False, chunk=1 << 18):
    buf = torch.empty((chunk, d), dtype=torch.float32, device="cuda")
    for r0 in range(0, n, chunk):
        m = min(chunk, n - r0)
        native.synth_fill_device(buf, r0, m, d, d, seed, round_bf16=bf16)
        torch.cuda.synchronize()
        st.add_device(buf, m, d, prenormalised=bf16)

def run(n, d, k, dtype, iters=20):
    st = rb.RagStore(d, n, dtype)
    t = time.time(); fill(st, n, d, 1234, dtype == rb.RAGSTORE_BF16); t_fill = time.time() - t
    q = torch.empty((1, d), dtype=torch.float32, device="cuda")
    native.synth_fill_device(q, 0, 1, d, d, 5678)
    idx = torch.zeros(max(1, k), dtype=torch.int64, device="cuda"); sc = torch.zeros(max(1, k), dtype=torch.float64, device="cuda")
    fnd = torch.zeros(1, dtype=torch.int32, device="cuda"); flg = torch.zeros(1, dtype=torch.int32, device="cuda")
    ts = torch.cuda.Stream(); torch.cuda.set_stream(ts)
    s = ts.cuda_stream
    for _ in range(3):
        st.search_device(q, 1, k, idx, sc, fnd, flg, stream=s)
    torch.cuda.synchronize()
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]; e1 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    for i in range(iters):
        e0[i].record(); st.search_device(q, 1, k, idx, sc, fnd, flg, stream=s); e1[i].record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in zip(e0, e1))
    med = ms[len(ms) // 2]
    esz = 4 if dtype == rb.RAGSTORE_F32 else 2
    gbs = n * d * esz / med / 1e6
    print(f"n={n} d={d} k={k} dtype={dtype}: fill {t_fill:.1f}s  median {med:.4f} ms  min {ms[0]:.4f}  -> {gbs:.0f} GB/s ({gbs/6542.1:.3f} of measured copy peak), "
          f"{n/med/1e6:.3f} G vec/s; idx={idx[:3].tolist()} flags={flg.item()}", flush=True)
    t = st.debug_stamps()
    print("   phases us: scan %.1f | grid-merge %.1f | stage rows %.1f | chains %.1f | rank+out %.1f" % tuple((t[i+1]-t[i])/1e3 for i in range(5)), flush=True)
    st.close()

if __name__ == "__main__":
    F, B = rb.RAGSTORE_F32, rb.RAGSTORE_BF16
    for (n, d, k, t) in [(10000, 768, 5, F), (10000, 768, 100, F), (10000, 768, 100, B),
                         (1000000, 768, 10, F), (4000000, 768, 10, F), (2000000, 768, 100, F), (2000000, 768, 10, B), (2000000, 768, 100, B),
                         (1000000, 1024, 10, F), (1000000, 384, 10, F), (1000000, 1000, 10, F), (300000, 4096, 10, F)]:
        run(n, d, k, t)
