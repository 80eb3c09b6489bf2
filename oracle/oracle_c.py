"""ctypes loader for oracle/liboracle.so -- TEST INFRASTRUCTURE ONLY (see ragindex_oracle.c)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "ragindex_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        fp, ip, dp = C.POINTER(C.c_float), C.POINTER(C.c_int64), C.POINTER(C.c_double)
        L.rag_oracle_cosine.restype = C.c_double
        L.rag_oracle_cosine.argtypes = [fp, C.c_int64, fp, C.c_int64]
        L.rag_oracle_score_all_mt.restype = None
        L.rag_oracle_score_all_mt.argtypes = [fp, C.c_int64, C.c_int64, fp, C.c_int64, dp, C.c_int]
        L.rag_oracle_select.restype = C.c_int64
        L.rag_oracle_select.argtypes = [dp, C.c_int64, C.c_int64, ip, dp]
        L.rag_oracle_query.restype = C.c_int64
        L.rag_oracle_query.argtypes = [fp, C.c_int64, C.c_int64, fp, C.c_int64, C.c_int64, ip, dp, C.c_int]
        L.rag_synth_rows_mt.restype = None
        L.rag_synth_rows_mt.argtypes = [fp, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_uint64, C.c_int, C.c_int]
        _LIB = L
    return _LIB


def _f(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def cosine(a, b) -> float:
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    return float(lib().rag_oracle_cosine(_f(a), a.shape[0], _f(b), b.shape[0]))


def score_all(rows, q, n_threads: int = 1) -> np.ndarray:
    rows = np.ascontiguousarray(rows, dtype=np.float32)
    q = np.ascontiguousarray(q, dtype=np.float32)
    n, d = rows.shape
    out = np.empty(n, dtype=np.float64)
    lib().rag_oracle_score_all_mt(_f(rows), n, d, _f(q), q.shape[0], out.ctypes.data_as(C.POINTER(C.c_double)), n_threads)
    return out


def query(rows, q, top_k: int = 5, n_threads: int = 1):
    rows = np.ascontiguousarray(rows, dtype=np.float32)
    q = np.ascontiguousarray(q, dtype=np.float32)
    n, d = rows.shape if rows.ndim == 2 else (0, 0)
    cap = max(1, min(max(n, 1), max(1, int(top_k))))
    idx = np.zeros(cap, dtype=np.int64)
    sc = np.zeros(cap, dtype=np.float64)
    got = lib().rag_oracle_query(_f(rows), n, d, _f(q), q.shape[0], int(top_k),
                                 idx.ctypes.data_as(C.POINTER(C.c_int64)), sc.ctypes.data_as(C.POINTER(C.c_double)), n_threads)
    return idx[:got].copy(), sc[:got].copy()


def synth_rows(row0: int, n_rows: int, dim: int, seed: int, round_bf16: bool = False, n_threads: int = 0) -> np.ndarray:
    out = np.empty((n_rows, dim), dtype=np.float32)
    nt = n_threads or (os.cpu_count() or 1)
    lib().rag_synth_rows_mt(_f(out), row0, n_rows, dim, dim, seed, int(round_bf16), nt)
    return out
