"""ragstore_search_noalloc -- the search entry point for hosts that may not allocate device memory
(an arena-managed server that owns all of HBM, or a latency-bound request thread that cannot take the
device-wide stall of cudaMalloc).  It must return exactly what ragstore_search_device returns for the
same matrix and query, call after call, although every call reuses one fixed scratch array that still
holds the previous call's bytes.  Collected last on purpose (file name), so a failure here cannot hide
the other GPU tests from a run that stops at the first failure."""
import numpy as np
import pytest

import ragstore_b200 as rb
from ragstore_b200 import native

pytestmark = pytest.mark.gpu


def _both(st, mat_ptr, rn_ptr, n, dim, dtype, q, k, stream):
    import torch
    out = []
    for which in ("existing", "noalloc"):
        idx = torch.full((k,), -7, dtype=torch.int64, device="cuda")
        sc = torch.full((k,), float("nan"), dtype=torch.float64, device="cuda")
        fnd = torch.zeros(1, dtype=torch.int32, device="cuda")
        flg = torch.zeros(1, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        if which == "existing":
            st.search_device(q, 1, k, idx, sc, fnd, flg, stream=stream.cuda_stream)
        else:
            native.search_noalloc(mat_ptr, rn_ptr, n, dim, dtype, q, k, idx, sc, fnd, flg, device=0, stream=stream.cuda_stream)
        stream.synchronize()
        out.append((idx.cpu().numpy(), sc.cpu().numpy(), int(fnd.item()), int(flg.item())))
    return out


@pytest.mark.parametrize("dtype,n,dim,k", [(rb.RAGSTORE_F32, 30000, 768, 10), (rb.RAGSTORE_BF16, 20000, 512, 100), (rb.RAGSTORE_F32, 9, 64, 20)])
def test_noalloc_matches_existing_entry_point_twice(dtype, n, dim, k):
    import torch
    bf = dtype == rb.RAGSTORE_BF16
    stream = torch.cuda.Stream()
    with rb.RagStore(dim, n, dtype) as st:
        buf = torch.empty((n, dim), dtype=torch.float32, device="cuda")
        native.synth_fill_device(buf, 0, n, dim, dim, 4321, round_bf16=bf)
        torch.cuda.synchronize()
        st.add_device(buf, n, dim, prenormalised=bf)
        assert st.info()["sidecar_rows"] == 0          # unit rows: stored exactly, nothing outside the matrix
        mat_ptr, rn_ptr, _ = st.device_views()
        qs = torch.empty((2, dim), dtype=torch.float32, device="cuda")
        native.synth_fill_device(qs, 0, 2, dim, dim, 8765)
        torch.cuda.synchronize()
        # two calls in a row: the second starts from a scratch array that still holds the first call's lists
        for call in range(2):
            (i0, s0, f0, g0), (i1, s1, f1, g1) = _both(st, mat_ptr, rn_ptr, n, dim, dtype, qs[call], k, stream)
            take = min(n, k)
            assert f0 == f1 == take
            assert np.array_equal(i0[:take], i1[:take])
            assert np.array_equal(s0[:take].view(np.uint64), s1[:take].view(np.uint64))   # same tolerance as the parity tests: bit-equal
            assert (g0 & rb.RAGSTORE_HIT_CERTIFIED) == (g1 & rb.RAGSTORE_HIT_CERTIFIED)


def test_noalloc_rejects_sizes_beyond_its_fixed_scratch():
    import torch
    q = torch.zeros(64, dtype=torch.float32, device="cuda")
    m = torch.zeros((8, 64), dtype=torch.float32, device="cuda")
    rn = torch.ones(8, dtype=torch.float32, device="cuda")
    idx = torch.zeros(300, dtype=torch.int64, device="cuda")
    sc = torch.zeros(300, dtype=torch.float64, device="cuda")
    fnd = torch.zeros(1, dtype=torch.int32, device="cuda")
    with pytest.raises(rb.RagStoreError) as ei:
        native.search_noalloc(m, rn, 8, 64, rb.RAGSTORE_F32, q, 300, idx, sc, fnd)
    assert ei.value.code == 4 and "256" in str(ei.value)          # RAGSTORE_ERR_UNSUPPORTED, names the limit
    with pytest.raises(rb.RagStoreError):
        native.search_noalloc(m, rn, 8, 62, rb.RAGSTORE_F32, q, 5, idx, sc, fnd)   # rows not a multiple of 16 bytes
    assert 300_000 < rb.lib().ragstore_noalloc_scratch_bytes() < 400_000


def test_noalloc_tiny_query_negative_scores_against_the_oracle():
    """A query of norm 1e-6 whose scores are all negative is where the reference's `+1e-9` (RagIndex.cs:134) moves scores
    the most.  The hits must be the oracle's, bit for bit, and the no-allocation entry point -- which cannot know the
    corpus' smallest row norm and assumes 1/2 -- must never claim a certificate that the store-backed entry point,
    which does know it, withholds.  Run on a second stream as well: the fixed scratch belongs to no stream."""
    import torch
    from oracle import oracle_c as oc
    n, dim, k = 30000, 768, 10
    rows = np.abs(oc.synth_rows(0, n, dim, 4321))          # unit norm, all components >= 0
    q = -np.abs(oc.synth_rows(0, 1, dim, 8765)[0]) * np.float32(1e-6)
    eidx, esc = oc.query(rows, q, k, n_threads=4)
    assert (esc < 0).all()
    with rb.RagStore(dim, n, rb.RAGSTORE_F32) as st:
        st.add(rows)
        assert st.info()["sidecar_rows"] == 0
        mat_ptr, rn_ptr, _ = st.device_views()
        dq = torch.from_numpy(q).cuda()
        for stream in (torch.cuda.Stream(), torch.cuda.Stream()):
            (i0, s0, f0, g0), (i1, s1, f1, g1) = _both(st, mat_ptr, rn_ptr, n, dim, rb.RAGSTORE_F32, dq, k, stream)
            for i, s, f in ((i0, s0, f0), (i1, s1, f1)):
                assert f == k and list(i[:k]) == list(eidx)
                assert np.array_equal(s[:k].view(np.uint64), esc.view(np.uint64))
            assert (g1 & rb.RAGSTORE_HIT_CERTIFIED) <= (g0 & rb.RAGSTORE_HIT_CERTIFIED)
