"""Row-sharded search across the GPUs of one node: one process per GPU, each owning a
contiguous block of rows; a query is scored by every shard, each shard's exactly-rescored top-k
(k * 16 bytes per rank) reaches every rank, and every rank merges the G * k records on its GPU.
The order (score desc, global id asc) is what a single RagIndex over the whole corpus returns
(reference Helpers/RagIndex.cs:64-65).

Two forms of the exchange:
  * "p2p"  -- ONE launch per rank and query: the shard scan's last CTA stores its records into every peer's inbox over
              NVLink, waits for the peers' arrival words and merges (ragstore_search_shard_exchange); batches use the
              stand-alone exchange kernel after the shard's batched search (csrc/exchange.cu).  No NCCL on the query path;
  * "p2p-2launch" -- the same exchange as a second launch behind every shard scan (what "p2p" was before the fusion);
  * "nccl" -- torch.distributed.all_gather_into_tensor + the merge kernel (also what tests inject a fake into)."""
from __future__ import annotations

from typing import Callable, Optional, Tuple

CAND_BYTES = 16  # struct ragstore_cand { double score; int64 id; }


def shard_range(n_total: int, world: int, rank: int) -> Tuple[int, int]:
    """Rows [lo, hi) owned by `rank`: contiguous blocks of ceil(N/G) rows; global id == insertion order."""
    per = (n_total + world - 1) // world
    lo = min(n_total, rank * per)
    return lo, min(n_total, lo + per)


def gather_layout(world: int, k: int) -> Tuple[int, int]:
    """(bytes per rank, total bytes) of the all-gather payload."""
    per = max(1, k) * CAND_BYTES
    return per, per * world


def _current_stream_handle(torch) -> int:
    """cudaStream_t of torch's current stream.  The C ABI reads NULL as "the store's own stream", so torch's default
    stream (handle 0) is passed as cudaStreamLegacy (0x1): scan, exchange and merge must share ONE stream."""
    h = torch.cuda.current_stream().cuda_stream
    return h if h else 1


class ShardedSearcher:
    """One shard (a RagStore on this rank's GPU) + the exchange.  `all_gather(out, inp)` defaults to
    torch.distributed.all_gather_into_tensor; tests inject their own."""

    WIDE_K = 200    # a request for 200 hits makes every shard keep its widest candidate lists (256 keys)

    def __init__(self, store, rank: int, world: int, row_base: int, max_k: int = 256,
                 all_gather: Optional[Callable] = None, device: Optional[int] = None, exchange: str = "nccl",
                 max_queries: int = 256):
        import torch
        from . import native
        self.torch, self.native = torch, native
        self.store, self.rank, self.world = store, rank, world
        self.device = torch.cuda.current_device() if device is None else device
        store.set_row_base(row_base)
        self.max_k = max_k
        max_k = max(max_k, self.WIDE_K)      # room for the wide redo of an uncertified query (search_batch)
        self.d_cand = torch.zeros(max_k * CAND_BYTES, dtype=torch.uint8, device="cuda")
        self.d_all = torch.zeros(world * max_k * CAND_BYTES, dtype=torch.uint8, device="cuda")
        self.d_idx = torch.zeros(max_k, dtype=torch.int64, device="cuda")
        self.d_score = torch.zeros(max_k, dtype=torch.float64, device="cuda")
        self.d_found = torch.zeros(1, dtype=torch.int32, device="cuda")
        self.d_flags = torch.zeros(1, dtype=torch.int32, device="cuda")
        self.d_gflags = torch.zeros(1, dtype=torch.int32, device="cuda")    # d_flags ANDed over all shards (peer exchange)
        self.last_flags = None      # RAGSTORE_HIT_* of the last synchronous search, ANDed over all shards
        if all_gather is None:
            import torch.distributed as dist
            all_gather = dist.all_gather_into_tensor
        self._all_gather = all_gather
        self.peer = None
        self.d_status = torch.zeros(max_queries, dtype=torch.int32, device="cuda")
        self.fused = exchange == "p2p"
        if exchange in ("p2p", "p2p-2launch"):
            self.peer = self._connect_peers(max_k, max_queries)
        elif exchange != "nccl":
            raise ValueError("exchange must be 'p2p', 'p2p-2launch' or 'nccl'")

    def _connect_peers(self, max_k: int, max_queries: int):
        """Create this rank's inbox, swap the 64-byte IPC handles (one all-gather at set-up time), map the peers.
        Returns None -- on EVERY rank -- when any rank cannot map its peers, so that all ranks fall back together."""
        torch, native = self.torch, self.native
        px, ok = None, 1
        try:
            px = native.PeerExchange(self.device, self.world, self.rank, max_records=max(max_k * max_queries, self.WIDE_K), max_queries=max_queries)
            mine = torch.frombuffer(bytearray(px.handle), dtype=torch.uint8).cuda()
        except native.RagStoreError:
            ok, mine = 0, torch.zeros(native.EXCHANGE_HANDLE_BYTES, dtype=torch.uint8, device="cuda")
        if self.world > 1:
            import torch.distributed as dist
            everyone = torch.zeros(self.world * native.EXCHANGE_HANDLE_BYTES, dtype=torch.uint8, device="cuda")
            dist.all_gather_into_tensor(everyone, mine)
            handles = bytes(everyone.cpu().numpy().tobytes())
        else:
            handles = bytes(mine.cpu().numpy().tobytes())
        if ok:
            try:
                px.connect(handles)
            except native.RagStoreError:
                ok = 0
        if self.world > 1:
            import torch.distributed as dist
            flag = torch.tensor([ok], dtype=torch.int32, device="cuda")
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            ok = int(flag.item())
        if not ok and px is not None:
            px.close()
        return px if ok else None

    def search_device(self, d_query, k: int):
        """Enqueue shard search -> all-gather -> merge on the current torch stream; results stay on the device
        (self.d_idx / d_score / d_found).  Every rank ends with the same global top-k."""
        req = max(1, int(k))
        if req > max(self.max_k, self.WIDE_K):
            raise ValueError("k above max_k")
        stream = _current_stream_handle(self.torch)
        per, total = gather_layout(self.world, req)
        if self.peer is not None and self.fused:
            self.store.search_shard_exchange(self.peer, k, self.d_cand, self.d_idx, self.d_score, self.d_found, self.d_gflags, self.d_status,
                                             d_query=d_query, stream=stream)
            return
        self.store.search_shard_device(d_query, k, self.d_cand, self.d_flags, stream=stream)
        if self.peer is not None:
            self.peer.merge(self.d_cand, 1, req, k, self.d_idx, self.d_score, self.d_found, self.d_status, stream=stream,
                            d_my_flags=self.d_flags, d_flags=self.d_gflags)
            return
        if self.world > 1:
            self._all_gather(self.d_all[:total], self.d_cand[:per])
            src = self.d_all
        else:
            src = self.d_cand
        self.native.merge_shards_device(src, self.world, req, k, self.d_idx, self.d_score, self.d_found,
                                        device=self.device, stream=stream)

    def search_host(self, h_query, k: int):
        """Host query in, host hits out (what a host process serving requests does): the query rides in the scan
        kernel's parameter block, the exchange kernel writes the merged hits into page-locked host memory; one stream
        wait.  Needs the peer-memory exchange (or world == 1) and dim <= 1024.  Returns (idx, score) numpy views that
        the next call overwrites."""
        torch = self.torch
        req = max(1, int(k))
        if req > self.max_k:
            raise ValueError("k above max_k")
        if self.peer is None:
            raise RuntimeError("search_host needs exchange='p2p'")
        if getattr(self, "_h_idx", None) is None:
            self._h_idx = torch.zeros(self.max_k, dtype=torch.int64).pin_memory()
            self._h_score = torch.zeros(self.max_k, dtype=torch.float64).pin_memory()
            self._h_meta = torch.zeros(4, dtype=torch.int32).pin_memory()          # [found, status, flags ANDed over the shards]
            self._h_np = (self._h_idx.numpy(), self._h_score.numpy(), self._h_meta.numpy())   # views of the same memory
            self._h_ptr = (self._h_idx.data_ptr(), self._h_score.data_ptr(), self._h_meta.data_ptr(), self._h_meta.data_ptr() + 4)
        cur = torch.cuda.current_stream()
        stream = cur.cuda_stream or 1
        p_idx, p_sc, p_found, p_status = self._h_ptr
        if self.fused and self.store.dim <= 1024:
            self.store.search_shard_exchange(self.peer, k, self.d_cand, p_idx, p_sc, p_found, self._h_meta.data_ptr() + 8, p_status,
                                             h_query=h_query, stream=stream)
        else:
            self.store.search_shard_hostq(h_query, k, self.d_cand, self.d_flags, stream=stream)
            self.peer.merge(self.d_cand, 1, req, k, p_idx, p_sc, p_found, p_status, stream=stream,
                            d_my_flags=self.d_flags, d_flags=self._h_meta.data_ptr() + 8)
        cur.synchronize()
        np_idx, np_sc, np_meta = self._h_np
        if np_meta[1] != 0:
            raise RuntimeError("peer exchange timed out: a rank did not publish its hits within 5 s")
        n = int(np_meta[0])
        self.last_flags = int(np_meta[2])
        return np_idx[:n], np_sc[:n]

    def global_flags(self, d_flags, d_gflags=None):
        """RAGSTORE_HIT_* bits ANDed over all shards: a result is certified / exact only if every shard's part of it is.
        With the peer exchange the AND was formed inside the exchange kernel (d_gflags); with NCCL it is one more small
        collective here (a bitwise AND is a MIN per bit).  Returns a host numpy array."""
        if self.peer is not None and d_gflags is not None:
            return d_gflags.cpu().numpy()
        f = d_flags.clone()
        if self.world > 1:
            import torch.distributed as dist
            bits = self.torch.stack([f & 1, (f >> 1) & 1])
            dist.all_reduce(bits, op=dist.ReduceOp.MIN)
            f = bits[0] | (bits[1] << 1)
        return f.cpu().numpy()

    def search_batch(self, d_queries, n_queries: int, k: int):
        """Synchronous batched search WITH the certificate handled: queries that some shard's tensor-core pass could not
        certify (tf32 / bf16 scores, eps ~2e-3) are run again through the fp32 scan on every shard, as ragstore_search does
        on a single store; what is still uncertified comes back with the flag clear.  Returns host arrays
        (idx [nq, k], score [nq, k], found [nq], flags [nq])."""
        torch = self.torch
        req = max(1, int(k))
        bi, bs, bf = self.search_batch_device(d_queries, n_queries, k)
        torch.cuda.current_stream().synchronize()
        if self.peer is not None and int(self.d_status[:n_queries].abs().sum().item()) != 0:
            raise RuntimeError("peer exchange timed out: a rank did not publish its hits within 5 s")
        flags = self.global_flags(self._b_flags[:n_queries], self._b_gflags[:n_queries] if self._b_via_peer else None).copy()
        idx, sc, found = bi.cpu().numpy().copy(), bs.cpu().numpy().copy(), bf.cpu().numpy().copy()
        for q in range(n_queries):
            if flags[q] & 1:
                continue
            # every rank sees the same ANDed flags, so every rank takes this branch for the same queries.  Asking for
            # WIDE_K hits runs the fp32 scan with 256-key candidate lists on every shard; the first k of that answer
            # are the top k (the order is total), certified by a far wider margin.
            i1, s1 = self.search(d_queries[q], max(req, self.WIDE_K))
            m = min(len(i1), req)
            idx[q, :m], sc[q, :m], found[q] = i1[:m], s1[:m], m
            flags[q] = self.last_flags
        return idx, sc, found, flags

    def search_batch_device(self, d_queries, n_queries: int, k: int):
        """Batched form (tensor-core path on each shard): returns device tensors (idx [nq, k], score, found).  The
        per-shard RAGSTORE_HIT_* flags of this rank land in self._b_flags; global_flags() combines them across ranks
        (search_batch does, and redoes what is uncertified)."""
        torch = self.torch
        req = max(1, int(k))
        need = n_queries * req * CAND_BYTES
        if getattr(self, "_b_cap", 0) < need:
            self._b_cand = torch.zeros(need, dtype=torch.uint8, device="cuda")
            self._b_all = torch.zeros(self.world * need, dtype=torch.uint8, device="cuda")
            self._b_idx = torch.zeros(n_queries * req, dtype=torch.int64, device="cuda")
            self._b_score = torch.zeros(n_queries * req, dtype=torch.float64, device="cuda")
            self._b_found = torch.zeros(n_queries, dtype=torch.int32, device="cuda")
            self._b_flags = torch.zeros(n_queries, dtype=torch.int32, device="cuda")
            self._b_gflags = torch.zeros(n_queries, dtype=torch.int32, device="cuda")
            self._b_cap = need
        stream = _current_stream_handle(torch)
        self.store.search_shard_batch_device(d_queries, n_queries, k, self._b_cand, self._b_flags, stream=stream)
        self._b_via_peer = self.peer is not None and n_queries <= self.peer.max_queries and n_queries * req <= self.peer.max_records
        if self._b_via_peer:
            self.peer.merge(self._b_cand, n_queries, req, k, self._b_idx, self._b_score, self._b_found, self.d_status, stream=stream,
                            d_my_flags=self._b_flags, d_flags=self._b_gflags)
            return self._b_idx[: n_queries * req].view(n_queries, req), self._b_score[: n_queries * req].view(n_queries, req), self._b_found[:n_queries]
        if self.world > 1:
            self._all_gather(self._b_all[: self.world * need], self._b_cand[:need])
            src = self._b_all
        else:
            src = self._b_cand
        self.native.merge_shards_batch_device(src, self.world, n_queries, req, k, self._b_idx, self._b_score, self._b_found,
                                              device=self.device, stream=stream)
        return self._b_idx[: n_queries * req].view(n_queries, req), self._b_score[: n_queries * req].view(n_queries, req), self._b_found[:n_queries]

    def search(self, d_query, k: int):
        self.search_device(d_query, k)
        self.torch.cuda.current_stream().synchronize()
        if self.peer is not None and int(self.d_status[0].item()) != 0:
            raise RuntimeError("peer exchange timed out: a rank did not publish its hits within 5 s")
        n = int(self.d_found.item())
        self.last_flags = int(self.global_flags(self.d_flags, self.d_gflags)[0])
        return self.d_idx[:n].cpu().numpy(), self.d_score[:n].cpu().numpy()
