"""ctypes binding of include/ragstore.h (the same entry points the C# P/Invoke layer binds;
see INTEGRATION.md).  Loads the in-tree libragstore.so; never falls back to anything else."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

RAGSTORE_F32, RAGSTORE_BF16 = 0, 1
RAGSTORE_ADD_HOST, RAGSTORE_ADD_DEVICE, RAGSTORE_ADD_PRENORMALISED = 0, 1, 2
RAGSTORE_HIT_CERTIFIED, RAGSTORE_HIT_EXACT_ROWS = 1, 2
RAGSTORE_CREATE_SIDECAR_ON, RAGSTORE_CREATE_SIDECAR_OFF, RAGSTORE_CREATE_BF16_SHADOW = 1, 2, 4

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class RagStoreError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"ragstore error {code}: {msg}")
        self.code = code


class Cand(C.Structure):
    _fields_ = [("score", C.c_double), ("id", C.c_int64)]


def lib_path() -> str:
    return os.path.join(_HERE, "libragstore.so")


def lib():
    """The loaded library.  Raises if it has not been built (python __graft_entry__.py build)."""
    global _LIB
    if _LIB is None:
        p = lib_path()
        if not os.path.exists(p):
            raise RagStoreError(-1, f"{p} not built: run `make -C {os.path.join(_HERE, 'csrc')}`; there is no CPU fallback")
        L = C.CDLL(p)
        vp, i64p, f64p, i32p, f32p = C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_float)
        sig = {
            "ragstore_version": (C.c_int, []),
            "ragstore_last_error": (C.c_int, [C.c_char_p, C.c_int]),
            "ragstore_create": (C.c_int, [C.c_int, C.c_int, C.c_int64, C.c_int, C.c_uint32, C.POINTER(vp)]),
            "ragstore_create_sharded": (C.c_int, [C.c_int, C.c_int, C.c_int64, C.POINTER(C.c_int), C.c_int, C.c_int64, C.c_uint32, C.POINTER(vp)]),
            "ragstore_destroy": (C.c_int, [vp]),
            "ragstore_add": (C.c_int, [vp, vp, C.c_int64, C.c_int, C.c_uint32, i64p]),
            "ragstore_count": (C.c_int, [vp, i64p]),
            "ragstore_reserve_rows": (C.c_int, [vp, C.c_int64]),
            "ragstore_add_many": (C.c_int, [vp, C.POINTER(vp), i64p, C.c_int, C.c_int, C.c_uint32, i64p]),
            "ragstore_set_row_base": (C.c_int, [vp, C.c_int64]),
            "ragstore_search": (C.c_int, [vp, f32p, C.c_int, C.c_int, C.c_int, i64p, f64p, i32p, i32p]),
            "ragstore_hit_filters": (C.c_int, [f64p, C.c_int, C.c_double, C.c_double, C.c_double, C.POINTER(C.c_uint8), f64p, i32p]),
            "ragstore_save": (C.c_int, [vp, C.c_char_p]),
            "ragstore_load": (C.c_int, [C.c_char_p, C.c_int, C.c_int64, C.POINTER(vp)]),
            "ragstore_load_sharded": (C.c_int, [C.c_char_p, C.POINTER(C.c_int), C.c_int, C.c_int64, C.POINTER(vp)]),
            "ragstore_search_coalesced": (C.c_int, [vp, f32p, C.c_int, C.c_int, i64p, f64p, i32p, i32p]),
            "ragstore_set_coalescing": (C.c_int, [vp, C.c_int, C.c_int]),
            "ragstore_search_device": (C.c_int, [vp, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp]),
            "ragstore_search_shard_device": (C.c_int, [vp, vp, C.c_int, vp, vp, vp]),
            "ragstore_search_shard_hostq": (C.c_int, [vp, vp, C.c_int, vp, vp, vp]),
            "ragstore_set_row_blocking": (C.c_int, [vp, C.c_int64, C.c_int, C.c_int]),
            "ragstore_search_shard_batch_device": (C.c_int, [vp, vp, C.c_int, C.c_int, vp, vp, vp]),
            "ragstore_merge_shards_batch_device": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int, vp]),
            "ragstore_merge_shards_device": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int, vp]),
            "ragstore_exchange_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(vp), vp]),
            "ragstore_exchange_connect": (C.c_int, [vp, vp]),
            "ragstore_exchange_merge": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp]),
            "ragstore_exchange_merge_flags": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp]),
            "ragstore_search_shard_exchange": (C.c_int, [vp, vp, vp, vp, C.c_int, vp, vp, vp, vp, vp, vp, vp]),
            "ragstore_exchange_destroy": (C.c_int, [vp]),
            "ragstore_info": (C.c_int, [vp, i64p, C.c_int]),
            "ragstore_debug_reload_env": (C.c_int, []),
            "ragstore_search_noalloc": (C.c_int, [vp, vp, C.c_int64, C.c_int, C.c_int, vp, C.c_int, vp, vp, vp, vp, C.c_int, vp]),
            "ragstore_noalloc_scratch_bytes": (C.c_int64, []),
            "ragstore_noalloc_last_error": (C.c_char_p, []),
            "ragstore_debug_stamps": (C.c_int, [vp, C.POINTER(C.c_uint64), C.c_int]),
            "ragstore_device_views": (C.c_int, [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]),
            "ragstore_synth_fill_device": (C.c_int, [vp, C.c_int64, C.c_int64, C.c_int, C.c_int64, C.c_uint64, C.c_int, C.c_int, vp]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _LIB = L
    return _LIB


def last_error() -> str:
    buf = C.create_string_buffer(1024)
    lib().ragstore_last_error(buf, 1024)
    return buf.value.decode("utf-8", "replace")


def check(rc: int):
    if rc != 0:
        raise RagStoreError(rc, last_error())


def _ptr(x):
    """Device pointer of a torch tensor / int / None."""
    if x is None:
        return None
    if isinstance(x, int):
        return C.c_void_p(x)
    return C.c_void_p(x.data_ptr())


class RagStore:
    """Owning wrapper of one ragstore_t (one device-resident shard)."""

    def __init__(self, dim: int, capacity_rows: int, dtype: int = RAGSTORE_F32, device: int = 0, flags: int = 0, _handle=None):
        if _handle is not None:
            self._h = _handle
        else:
            self._h = C.c_void_p()
            check(lib().ragstore_create(dim, dtype, capacity_rows, device, flags, C.byref(self._h)))
        self.dim, self.dtype, self.device, self.capacity = dim, dtype, device, capacity_rows

    @classmethod
    def create_sharded(cls, dim: int, capacity_rows: int, devices, dtype: int = RAGSTORE_F32, block_rows: int = 0, flags: int = 0) -> "RagStore":
        """One handle over several GPUs of this process (ragstore_create_sharded): same add/search/count calls."""
        devs = (C.c_int * len(devices))(*[int(d) for d in devices])
        h = C.c_void_p()
        check(lib().ragstore_create_sharded(dim, dtype, capacity_rows, devs, len(devices), block_rows, flags, C.byref(h)))
        st = cls(dim, capacity_rows, dtype, int(devices[0]), _handle=h)
        st.devices = [int(d) for d in devices]
        return st

    def add_many(self, blocks, prenormalised: bool = False) -> int:
        """Append a list of C-contiguous float32 host arrays (rows x dim each) with ONE native call; each block is
        published as soon as it is in.  Returns the row id of the first appended row."""
        arrs = [np.ascontiguousarray(b, dtype=np.float32) for b in blocks]
        for a in arrs:
            if a.ndim != 2 or a.shape[1] != self.dim:
                raise ValueError("every block must be rows x dim")
        n = len(arrs)
        ptrs = (C.c_void_p * max(1, n))(*[a.ctypes.data for a in arrs])
        rows = (C.c_int64 * max(1, n))(*[a.shape[0] for a in arrs])
        first = C.c_int64(0)
        flags = RAGSTORE_ADD_HOST | (RAGSTORE_ADD_PRENORMALISED if prenormalised else 0)
        check(lib().ragstore_add_many(self._h, ptrs, rows, n, self.dim, flags, C.byref(first)))
        return first.value

    def reserve_rows(self, n_rows: int):
        """Map device memory for the first n_rows rows now (appends otherwise map on demand)."""
        check(lib().ragstore_reserve_rows(self._h, int(n_rows)))

    def save(self, path: str):
        check(lib().ragstore_save(self._h, path.encode()))

    @classmethod
    def load(cls, path: str, device: int = 0, capacity_rows: int = 0) -> "RagStore":
        h = C.c_void_p()
        check(lib().ragstore_load(path.encode(), device, capacity_rows, C.byref(h)))
        st = cls(0, 0, 0, device, _handle=h)
        i = st.info()
        st.dim, st.dtype, st.capacity = i["dim"], i["dtype"], i["capacity"]
        return st

    @classmethod
    def load_sharded(cls, path: str, devices, capacity_rows: int = 0) -> "RagStore":
        devs = (C.c_int * len(devices))(*[int(d) for d in devices])
        h = C.c_void_p()
        check(lib().ragstore_load_sharded(path.encode(), devs, len(devices), capacity_rows, C.byref(h)))
        st = cls(0, 0, 0, int(devices[0]), _handle=h)
        i = st.info()
        st.dim, st.dtype, st.capacity = i["dim"], i["dtype"], i["capacity"]
        st.devices = [int(d) for d in devices]
        return st

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().ragstore_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- store side
    def add(self, rows: np.ndarray, prenormalised: bool = False) -> int:
        rows = np.ascontiguousarray(rows, dtype=np.float32)
        if rows.ndim != 2:
            raise ValueError("rows must be [n, dim]")
        first = C.c_int64(-1)
        flags = RAGSTORE_ADD_HOST | (RAGSTORE_ADD_PRENORMALISED if prenormalised else 0)
        check(lib().ragstore_add(self._h, C.c_void_p(rows.ctypes.data), rows.shape[0], rows.shape[1], flags, C.byref(first)))
        return first.value

    def add_device(self, d_rows, n_rows: int, row_dim: int, prenormalised: bool = False) -> int:
        """d_rows: torch CUDA fp32 tensor (or raw device pointer) [n_rows, row_dim], contiguous."""
        first = C.c_int64(-1)
        flags = RAGSTORE_ADD_DEVICE | (RAGSTORE_ADD_PRENORMALISED if prenormalised else 0)
        check(lib().ragstore_add(self._h, _ptr(d_rows), n_rows, row_dim, flags, C.byref(first)))
        return first.value

    def count(self) -> int:
        n = C.c_int64(0)
        check(lib().ragstore_count(self._h, C.byref(n)))
        return n.value

    def set_row_base(self, base: int):
        check(lib().ragstore_set_row_base(self._h, base))

    def set_row_blocking(self, block_rows: int, n_shards: int, shard_index: int):
        check(lib().ragstore_set_row_blocking(self._h, block_rows, n_shards, shard_index))

    def search_shard_batch_device(self, d_queries, n_queries: int, top_k: int, d_cand, d_flags=None, stream: int = 0):
        check(lib().ragstore_search_shard_batch_device(self._h, _ptr(d_queries), n_queries, int(top_k), _ptr(d_cand), _ptr(d_flags),
                                                       C.c_void_p(stream)))

    def info(self) -> dict:
        v = (C.c_int64 * 16)()
        check(lib().ragstore_info(self._h, v, 16))
        keys = ["dim", "ld", "dtype", "count", "capacity", "n_inexact_rows", "device", "sm_count",
                "has_sidecar", "sidecar_rows", "matrix_bytes_mapped", "uses_vmm", "scan_launches", "batch_launches", "last_add_kernel_ns",
                "scratch_overlaps"]
        return dict(zip(keys, list(v)))

    def debug_stamps(self):
        v = (C.c_uint64 * 8)()
        check(lib().ragstore_debug_stamps(self._h, v, 8))
        return list(v)

    def device_views(self):
        m, r, e = C.c_void_p(), C.c_void_p(), C.c_void_p()
        check(lib().ragstore_device_views(self._h, C.byref(m), C.byref(r), C.byref(e)))
        return m.value, r.value, e.value

    # ---- search side (host buffers: the call the C# RagIndex makes)
    def search(self, queries: np.ndarray, top_k: int):
        q = np.ascontiguousarray(queries, dtype=np.float32)
        if q.ndim == 1:
            q = q[None, :]
        nq, qd = q.shape
        req = max(1, int(top_k))
        idx = np.full((nq, req), -1, dtype=np.int64)
        score = np.full((nq, req), np.nan, dtype=np.float64)
        found = np.zeros(nq, dtype=np.int32)
        flags = np.zeros(nq, dtype=np.int32)
        check(lib().ragstore_search(self._h, q.ctypes.data_as(C.POINTER(C.c_float)), nq, qd, int(top_k),
                                    idx.ctypes.data_as(C.POINTER(C.c_int64)), score.ctypes.data_as(C.POINTER(C.c_double)),
                                    found.ctypes.data_as(C.POINTER(C.c_int32)), flags.ctypes.data_as(C.POINTER(C.c_int32))))
        return idx, score, found, flags

    def search_call(self, n_queries: int, top_k: int):
        """Preallocated form of `search`, the way a compiled host uses the ABI (buffers pinned once, reused per request):
        returns (call, idx, score, found, flags); `call(q)` takes a C-contiguous float32 array of n_queries x dim, runs
        ragstore_search into the four arrays and returns nothing.  Per-call Python work is one ctypes call."""
        req = max(1, int(top_k))
        idx = np.full((n_queries, req), -1, dtype=np.int64)
        score = np.full((n_queries, req), np.nan, dtype=np.float64)
        found = np.zeros(n_queries, dtype=np.int32)
        flags = np.zeros(n_queries, dtype=np.int32)
        fn, h, dim, k = lib().ragstore_search, self._h, self.dim, int(top_k)
        p_idx, p_sc = idx.ctypes.data_as(C.POINTER(C.c_int64)), score.ctypes.data_as(C.POINTER(C.c_double))
        p_fnd, p_flg = found.ctypes.data_as(C.POINTER(C.c_int32)), flags.ctypes.data_as(C.POINTER(C.c_int32))
        f32p = C.POINTER(C.c_float)

        def call(q: np.ndarray):
            rc = fn(h, C.cast(q.ctypes.data, f32p), n_queries, dim, k, p_idx, p_sc, p_fnd, p_flg)
            if rc:
                check(rc)
        return call, idx, score, found, flags

    def search_coalesced(self, query: np.ndarray, top_k: int):
        """One query per call; concurrent callers share a corpus sweep (ragstore_search_coalesced)."""
        q = np.ascontiguousarray(query, dtype=np.float32).reshape(-1)
        req = max(1, int(top_k))
        idx = np.full(req, -1, dtype=np.int64)
        score = np.full(req, np.nan, dtype=np.float64)
        found, flags = C.c_int32(0), C.c_int32(0)
        check(lib().ragstore_search_coalesced(self._h, q.ctypes.data_as(C.POINTER(C.c_float)), q.shape[0], int(top_k),
                                              idx.ctypes.data_as(C.POINTER(C.c_int64)), score.ctypes.data_as(C.POINTER(C.c_double)),
                                              C.byref(found), C.byref(flags)))
        return idx[:found.value], score[:found.value], flags.value

    def set_coalescing(self, window_us: int, max_batch: int = 256):
        check(lib().ragstore_set_coalescing(self._h, window_us, max_batch))

    # ---- device-resident variants (torch tensors or raw pointers), asynchronous on `stream`
    def search_device(self, d_queries, n_queries: int, top_k: int, d_idx, d_score, d_found=None, d_flags=None, stream: int = 0):
        check(lib().ragstore_search_device(self._h, _ptr(d_queries), n_queries, int(top_k), _ptr(d_idx), _ptr(d_score),
                                           _ptr(d_found), _ptr(d_flags), C.c_void_p(stream)))

    def search_shard_device(self, d_query, top_k: int, d_cand, d_flags=None, stream: int = 0):
        check(lib().ragstore_search_shard_device(self._h, _ptr(d_query), int(top_k), _ptr(d_cand), _ptr(d_flags), C.c_void_p(stream)))

    def search_shard_exchange(self, peer: "PeerExchange", top_k: int, d_mine, d_idx, d_score, d_found, d_flags=None, d_status=None,
                              d_query=None, h_query: np.ndarray = None, stream: int = 0):
        """One launch: shard scan + peer-memory exchange + merge (ragstore_search_shard_exchange).  Give d_query (device
        tensor) or h_query (C-contiguous float32 host array, rides in the kernel parameters)."""
        hq = C.c_void_p(h_query.ctypes.data) if h_query is not None else None
        check(lib().ragstore_search_shard_exchange(self._h, peer._h, _ptr(d_query), hq, int(top_k), _ptr(d_mine), _ptr(d_idx), _ptr(d_score),
                                                   _ptr(d_found), _ptr(d_flags), _ptr(d_status), C.c_void_p(stream)))

    def search_shard_hostq(self, h_query: np.ndarray, top_k: int, d_cand, d_flags=None, stream: int = 0):
        """Shard search with the query in host memory (C-contiguous float32 of `dim` values); it rides in the kernel parameters."""
        check(lib().ragstore_search_shard_hostq(self._h, C.c_void_p(h_query.ctypes.data), int(top_k), _ptr(d_cand), _ptr(d_flags),
                                                C.c_void_p(stream)))


def search_noalloc(d_matrix, d_rnorm, n_rows: int, dim: int, dtype: int, d_query, top_k: int, d_idx, d_score, d_found,
                   d_flags=None, device: int = 0, stream: int = 0):
    """ragstore_search_noalloc: caller-owned device buffers in and out, fixed in-library scratch, no allocation."""
    rc = lib().ragstore_search_noalloc(_ptr(d_matrix), _ptr(d_rnorm), n_rows, dim, dtype, _ptr(d_query), int(top_k),
                                       _ptr(d_idx), _ptr(d_score), _ptr(d_found), _ptr(d_flags), device, C.c_void_p(stream))
    if rc != 0:
        raise RagStoreError(rc, lib().ragstore_noalloc_last_error().decode("utf-8", "replace"))


def hit_filters(scores, evidence_min_score: float = 0.35, best_factor: float = 0.6, low_score_floor: float = 0.20):
    """(cited mask, best score, low-score flag) of a hit list: Program.cs:77-89,125-128 in one call."""
    sc = np.ascontiguousarray(scores, dtype=np.float64)
    cited = np.zeros(sc.shape[0], dtype=np.uint8)
    best, low = C.c_double(0.0), C.c_int32(0)
    check(lib().ragstore_hit_filters(sc.ctypes.data_as(C.POINTER(C.c_double)), sc.shape[0], evidence_min_score, best_factor, low_score_floor,
                                     cited.ctypes.data_as(C.POINTER(C.c_uint8)), C.byref(best), C.byref(low)))
    return cited.astype(bool), best.value, bool(low.value)


def merge_shards_device(d_gathered, n_shards: int, per_shard: int, top_k: int, d_idx, d_score, d_found, device: int = 0, stream: int = 0):
    check(lib().ragstore_merge_shards_device(_ptr(d_gathered), n_shards, per_shard, int(top_k), _ptr(d_idx), _ptr(d_score),
                                             _ptr(d_found), device, C.c_void_p(stream)))


def merge_shards_batch_device(d_gathered, n_shards: int, n_queries: int, per_shard: int, top_k: int, d_idx, d_score, d_found,
                              device: int = 0, stream: int = 0):
    check(lib().ragstore_merge_shards_batch_device(_ptr(d_gathered), n_shards, n_queries, per_shard, int(top_k), _ptr(d_idx),
                                                   _ptr(d_score), _ptr(d_found), device, C.c_void_p(stream)))


EXCHANGE_HANDLE_BYTES = 64


class PeerExchange:
    """This rank's end of the peer-memory exchange (ragstore_exchange_*): an inbox every peer writes into directly.
    `handle` (64 bytes) must reach every rank; `connect(all_handles)` takes them back in rank order."""

    def __init__(self, device: int, world: int, rank: int, max_records: int, max_queries: int = 1):
        self._h = C.c_void_p()
        buf = C.create_string_buffer(EXCHANGE_HANDLE_BYTES)
        check(lib().ragstore_exchange_create(device, world, rank, max_records, max_queries, C.byref(self._h), buf))
        self.handle = buf.raw
        self.world, self.rank, self.device = world, rank, device
        self.max_records, self.max_queries = max_records, max_queries

    def connect(self, all_handles: bytes):
        if len(all_handles) != self.world * EXCHANGE_HANDLE_BYTES:
            raise ValueError("all_handles must hold world x 64 bytes, in rank order")
        check(lib().ragstore_exchange_connect(self._h, C.c_char_p(all_handles)))

    def merge(self, d_mine, n_queries: int, per_shard: int, top_k: int, d_idx, d_score, d_found, d_status=None, stream: int = 0,
              d_my_flags=None, d_flags=None):
        """d_my_flags: this rank's per-query RAGSTORE_HIT_* bits; d_flags receives their AND over all ranks."""
        check(lib().ragstore_exchange_merge_flags(self._h, _ptr(d_mine), _ptr(d_my_flags), n_queries, per_shard, int(top_k), _ptr(d_idx),
                                                  _ptr(d_score), _ptr(d_found), _ptr(d_flags), _ptr(d_status), C.c_void_p(stream)))

    def close(self):
        if self._h:
            lib().ragstore_exchange_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def synth_fill_device(d_out, row0: int, n_rows: int, dim: int, ld: int, seed: int, round_bf16: bool = False, device: int = 0, stream: int = 0):
    check(lib().ragstore_synth_fill_device(_ptr(d_out), row0, n_rows, dim, ld, seed, int(round_bf16), device, C.c_void_p(stream)))
