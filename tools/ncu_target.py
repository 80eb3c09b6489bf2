"""Tiny driver for ncu captures: builds a synthetic store on the device and launches the search kernel a few times.
usage: python tools/ncu_target.py N DIM K DTYPE(f32|bf16) [ITERS]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ragstore_b200 as rb
from ragstore_b200 import native

n, d, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
dtype = rb.RAGSTORE_BF16 if sys.argv[4] == "bf16" else rb.RAGSTORE_F32
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 4
st = rb.RagStore(d, n, dtype)
chunk = 1 << 18
buf = torch.empty((chunk, d), dtype=torch.float32, device="cuda")
for r0 in range(0, n, chunk):
    m = min(chunk, n - r0)
    native.synth_fill_device(buf, r0, m, d, d, 1234, round_bf16=(dtype == rb.RAGSTORE_BF16))
    torch.cuda.synchronize()
    st.add_device(buf, m, d, prenormalised=(dtype == rb.RAGSTORE_BF16))
q = torch.empty((1, d), dtype=torch.float32, device="cuda")
native.synth_fill_device(q, 0, 1, d, d, 5678)
idx = torch.zeros(max(1, k), dtype=torch.int64, device="cuda"); sc = torch.zeros(max(1, k), dtype=torch.float64, device="cuda")
fnd = torch.zeros(1, dtype=torch.int32, device="cuda"); flg = torch.zeros(1, dtype=torch.int32, device="cuda")
ts = torch.cuda.Stream()
for _ in range(iters):
    st.search_device(q, 1, k, idx, sc, fnd, flg, stream=ts.cuda_stream)
torch.cuda.synchronize()
print("ok", idx[:3].tolist(), flg.item())
