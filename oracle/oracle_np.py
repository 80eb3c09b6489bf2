"""numpy / pure-Python restatement of the reference search -- TEST INFRASTRUCTURE ONLY.

Independent of oracle/ragindex_oracle.c on purpose: the two are cross-checked in
tests/test_oracle.py.  PARITY UNPINNED (no reference tests / golden vectors exist;
the C# cannot run here).

Follows /root/reference/Minimal_Agent_RAG/Minimal_RAG/Helpers/RagIndex.cs:
  :124-135  Cosine   -- fp32 products, widened, summed sequentially in fp64
  :62-66    QueryAsync -- Select(Cosine).OrderByDescending(Score).Take(max(1,k))
"""
from __future__ import annotations

import math

import numpy as np


def cosine_py(a, b) -> float:
    """Pure-Python loop, one row (RagIndex.cs:124-135).  Small cases only."""
    a = np.asarray(a, dtype=np.float32)
    b = np.asarray(b, dtype=np.float32)
    if a.shape[0] != b.shape[0]:
        return -1.0                                   # :126
    dot = 0.0
    na = 0.0
    nb = 0.0
    for i in range(a.shape[0]):                       # :128
        dot += float(np.float32(a[i] * b[i]))         # fp32 product then widened, :130
        na += float(np.float32(a[i] * a[i]))          # :131
        nb += float(np.float32(b[i] * b[i]))          # :132
    return dot / (math.sqrt(na) * math.sqrt(nb) + 1e-9)   # :134


def _seq_sum64(prod32: np.ndarray) -> np.ndarray:
    """Strictly left-to-right fp64 sum along axis 1 (np.cumsum is sequential;
    np.sum is pairwise and would NOT match the reference)."""
    if prod32.shape[1] == 0:
        return np.zeros(prod32.shape[0], dtype=np.float64)
    # `double dot = 0; dot += p0; ...` (:127-132): the accumulator STARTS at +0.0, so a run of -0.0 products sums to
    # +0.0 (np.cumsum alone would start at p0 and keep -0.0); adding +0.0 afterwards restores exactly that:
    # x + (+0.0) == x for every x except -0.0, which becomes +0.0 -- and a sequential sum can only end at -0.0 when
    # every term was -0.0.
    return np.cumsum(prod32.astype(np.float64), axis=1)[:, -1] + 0.0


def score_all(rows: np.ndarray, q: np.ndarray, chunk: int = 4096) -> np.ndarray:
    """fp64 score of every row (vectorised over rows; per-row arithmetic as :124-135)."""
    rows = np.ascontiguousarray(rows, dtype=np.float32)
    q = np.ascontiguousarray(q, dtype=np.float32)
    n = rows.shape[0]
    if rows.ndim != 2:
        raise ValueError("rows must be 2-D")
    if rows.shape[1] != q.shape[0]:
        return np.full(n, -1.0, dtype=np.float64)     # :126 for every row
    na = _seq_sum64((q * q)[None, :])[0]
    out = np.empty(n, dtype=np.float64)
    for r0 in range(0, n, chunk):
        blk = rows[r0:r0 + chunk]
        dot = _seq_sum64(blk * q[None, :])
        nb = _seq_sum64(blk * blk)
        out[r0:r0 + chunk] = dot / (math.sqrt(na) * np.sqrt(nb) + 1e-9)
    return out


def _dotnet_desc_key(scores: np.ndarray) -> np.ndarray:
    """Sort key such that ascending key == OrderByDescending with double.CompareTo:
    NaN is below every number (so last when descending); -0.0 == +0.0."""
    s = np.where(np.isnan(scores), -np.inf, scores)
    # NaN must also rank below a genuine -inf: push NaN one step further with a second key
    return -(s + 0.0)


def select(scores: np.ndarray, top_k: int):
    """Stable descending order, first max(1,k) entries (RagIndex.cs:64-66)."""
    n = scores.shape[0]
    take = min(n, max(1, int(top_k)))                 # :65
    isn = np.isnan(scores)
    key = _dotnet_desc_key(scores)
    # lexsort: last key is primary.  (nan flag asc) puts numbers before NaN; then key; stable on index.
    order = np.lexsort((np.arange(n), key, isn.astype(np.int8)))
    idx = order[:take].astype(np.int64)
    return idx, scores[idx]


def query(rows: np.ndarray, q: np.ndarray, top_k: int = 5):
    """RagIndex.QueryAsync after the embedding call; returns (idx int64[], score fp64[])."""
    rows = np.asarray(rows, dtype=np.float32)
    if rows.shape[0] == 0:
        return np.zeros(0, dtype=np.int64), np.zeros(0, dtype=np.float64)
    return select(score_all(rows, np.asarray(q, dtype=np.float32)), top_k)


# ---------------------------------------------------------------------------
# synthetic generator (bit-identical to ragindex_oracle.c:rag_synth_rows and csrc/synth.cu)
# ---------------------------------------------------------------------------
_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(z: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M
        return z ^ (z >> np.uint64(31))


def bf16_round(x: np.ndarray) -> np.ndarray:
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).copy()
    finite = (u & np.uint32(0x7F800000)) != np.uint32(0x7F800000)
    with np.errstate(over="ignore"):
        u[finite] = u[finite] + np.uint32(0x7FFF) + ((u[finite] >> np.uint32(16)) & np.uint32(1))
    u &= np.uint32(0xFFFF0000)
    return u.view(np.float32)


def synth_rows(row0: int, n_rows: int, dim: int, seed: int, round_bf16: bool = False) -> np.ndarray:
    with np.errstate(over="ignore"):
        r = (np.arange(row0, row0 + n_rows, dtype=np.uint64) + np.uint64(1))
        row_key = _splitmix64((np.uint64(seed) + np.uint64(0x632BE59BD9B4E019) * r) & _M)
        c = np.arange(dim, dtype=np.uint64)
        ck = (c * np.uint64(0xD1B54A32D192ED03) + np.uint64(0x8CB92BA72F3D8DD7)) & _M
        u = _splitmix64(row_key[:, None] ^ ck[None, :])
    s = ((u & np.uint64(0xFFFF)) + ((u >> np.uint64(16)) & np.uint64(0xFFFF)) +
         ((u >> np.uint64(32)) & np.uint64(0xFFFF)) + ((u >> np.uint64(48)) & np.uint64(0xFFFF))).astype(np.int64) - 131070
    n2 = (s * s).sum(axis=1)
    norm = np.sqrt(n2.astype(np.float64))
    with np.errstate(divide="ignore", invalid="ignore"):
        v = np.where(n2[:, None] != 0, s.astype(np.float64) / norm[:, None], 0.0)
    f = v.astype(np.float32)
    return bf16_round(f) if round_bf16 else f
