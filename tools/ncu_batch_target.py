"""ncu driver for the batched path: python tools/ncu_batch_target.py N DIM NQ K [ITERS] [STORE_FLAGS (4 = bf16 shadow)]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
import ragstore_b200 as rb
from ragstore_b200 import native
from quick_bench import fill
n, d, nq, k = (int(x) for x in sys.argv[1:5]); iters = int(sys.argv[5]) if len(sys.argv) > 5 else 3
st = rb.RagStore(d, n, rb.RAGSTORE_F32, 0, int(sys.argv[6]) if len(sys.argv) > 6 else 0); fill(st, n, d, 1234)
q = torch.empty((nq, d), dtype=torch.float32, device="cuda"); native.synth_fill_device(q, 0, nq, d, d, 5678)
idx = torch.zeros((nq, k), dtype=torch.int64, device="cuda"); sc = torch.zeros((nq, k), dtype=torch.float64, device="cuda")
fnd = torch.zeros(nq, dtype=torch.int32, device="cuda"); flg = torch.zeros(nq, dtype=torch.int32, device="cuda")
ts = torch.cuda.Stream()
for _ in range(iters):
    st.search_device(q, nq, k, idx, sc, fnd, flg, stream=ts.cuda_stream)
torch.cuda.synchronize(); print("ok", idx[0, :3].tolist(), int(flg[0]))
