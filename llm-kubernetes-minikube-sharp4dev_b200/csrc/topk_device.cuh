// topk_device.cuh -- device code shared by the single-query scan (scan_topk.cu) and the batched
// tensor-core path (batch_gemm.cu): sorted candidate lists, the exact fp64 rescoring chains
// (reference Helpers/RagIndex.cs:124-135) and the merge + rescore + rank + output stage (:64-66).
#pragma once
#include <math_constants.h>

#include "common.cuh"
#include "exchange_device.cuh"
#include "ragstore_internal.h"

namespace rag {

struct ScanMisc {  // lives in shared memory at off_misc
    double na;
    double ek;
    double red[kConsumerWarps];
    int is_last;
    int inexact;
    double ex[kMaxKp];
    double dot[kMaxKp];
    double nb[kMaxKp];
    uint8_t raw_fmt[kMaxKp];  // 1 = the staged row is the fp32 original from the sidecar
    unsigned long long stage_bar;  // mbarrier for the bulk copies that stage candidate rows
    int n_cand;                    // survivors actually present in the merged list
};

// ---------------------------------------------------------------- warp-level sorted lists
// Lists live in shared memory (kp = 32*S keys, sorted descending, position j = slot*32 + lane) but are
// merged in registers: A <- top kp of A u B.  Element-wise max of A with reversed B leaves a bitonic
// sequence holding the kp largest keys; a bitonic merge (slot strides in registers, lane strides by
// shuffle) sorts it.  No shared-memory round trips inside the network.
template <int S>
__device__ __forceinline__ void warp_merge_desc_t(uint64_t* A, const uint64_t* B, int lane) {
    uint64_t a[S];
#pragma unroll
    for (int s = 0; s < S; ++s) {
        const uint64_t x = A[s * 32 + lane];
        const uint64_t y = B[(S - 1 - s) * 32 + 31 - lane];
        a[s] = x > y ? x : y;
    }
#pragma unroll
    for (int ss = S >> 1; ss > 0; ss >>= 1) {
#pragma unroll
        for (int s = 0; s < S; ++s) {
            if ((s & ss) == 0) {
                const uint64_t x = a[s], y = a[s + ss];
                a[s] = x > y ? x : y;
                a[s + ss] = x > y ? y : x;
            }
        }
    }
#pragma unroll
    for (int st = 16; st > 0; st >>= 1) {
        const bool lower = (lane & st) == 0;
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const uint64_t o = __shfl_xor_sync(0xffffffffu, a[s], st);
            a[s] = lower ? (a[s] > o ? a[s] : o) : (a[s] < o ? a[s] : o);
        }
    }
#pragma unroll
    for (int s = 0; s < S; ++s) A[s * 32 + lane] = a[s];
    __syncwarp();
}

__device__ __forceinline__ void warp_merge_desc(uint64_t* A, const uint64_t* B, int kp, int lane) {
    switch (kp) {
        case 32: warp_merge_desc_t<1>(A, B, lane); break;
        case 64: warp_merge_desc_t<2>(A, B, lane); break;
        case 128: warp_merge_desc_t<4>(A, B, lane); break;
        default: warp_merge_desc_t<8>(A, B, lane); break;
    }
}

// one key per lane -> sorted descending across lanes (bitonic network on shuffles)
__device__ __forceinline__ uint64_t warp_sort_desc(uint64_t v, int lane) {
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            const uint64_t o = __shfl_xor_sync(0xffffffffu, v, j);
            const bool desc_block = (lane & k) == 0;
            const bool lower = (lane & j) == 0;
            const bool keep_max = (lower == desc_block);
            v = keep_max ? (v > o ? v : o) : (v < o ? v : o);
        }
    }
    return v;
}

// Merge the cnt (<= 32) buffered keys of this warp into its sorted list L; B is warp-private scratch.
__device__ __forceinline__ uint64_t warp_flush(uint64_t* L, uint64_t* B, const uint64_t* buf, int cnt, int kp,
                                               int lane) {
    __syncwarp();
    uint64_t v = lane < cnt ? buf[lane] : 0ull;
    v = warp_sort_desc(v, lane);
    B[lane] = v;
    for (int j = 32 + lane; j < kp; j += 32) B[j] = 0ull;
    __syncwarp();
    warp_merge_desc(L, B, kp, lane);
    return L[kp - 1];
}

template <typename T>
__device__ __forceinline__ float elem_at(const uint8_t* row, int i) {
    if constexpr (sizeof(T) == 4)
        return reinterpret_cast<const float*>(row)[i];
    else
        return bf16_bits_to_f32(reinterpret_cast<const uint16_t*>(row)[i]);
}

// Sequential fp64 accumulation of fp32 products over one staged row (RagIndex.cs:127-133).
// kind 0: sum q[i]*x[i]   kind 1: sum x[i]*x[i]   (row == nullptr: x := q, i.e. the query norm)
// The additions form one dependent chain (that IS the reference's arithmetic): 8.1 cycles per DADD on
// B200 (tools/ubench/fp64_lat.cu).  Loads, products and the fp32->fp64 conversions of a 4-element
// block are independent of the chain and overlap with it; measured ~9 cycles per element.
template <typename T>
__device__ __forceinline__ double exact_chain(const uint8_t* row, bool f32row, const float* q, int dim, int kind) {
    double acc = 0.0;  // :127
    int i = 0;
    for (; i + 4 <= dim; i += 4) {
        float x[4];
        if (row == nullptr) {
#pragma unroll
            for (int j = 0; j < 4; ++j) x[j] = q[i + j];
        } else if (f32row) {
            const float4 v = *reinterpret_cast<const float4*>(row + static_cast<size_t>(i) * 4);
            x[0] = v.x, x[1] = v.y, x[2] = v.z, x[3] = v.w;
        } else {
            const uint2 v = *reinterpret_cast<const uint2*>(row + static_cast<size_t>(i) * 2);
            x[0] = __uint_as_float(v.x << 16), x[1] = __uint_as_float(v.x & 0xFFFF0000u);
            x[2] = __uint_as_float(v.y << 16), x[3] = __uint_as_float(v.y & 0xFFFF0000u);
        }
        float pr[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) pr[j] = kind ? __fmul_rn(x[j], x[j]) : __fmul_rn(q[i + j], x[j]);  // :130-132
#pragma unroll
        for (int j = 0; j < 4; ++j) acc = __dadd_rn(acc, static_cast<double>(pr[j]));
    }
    for (; i < dim; ++i) {
        const float x = row == nullptr ? q[i] : (f32row ? reinterpret_cast<const float*>(row)[i] : elem_at<T>(row, i));
        const float pr = kind ? __fmul_rn(x, x) : __fmul_rn(q[i], x);
        acc = __dadd_rn(acc, static_cast<double>(pr));
    }
    return acc;
}

// ---------------------------------------------------------------- merge + exact rescoring + output
// Shared by the single-query scan (its last CTA) and the batched path (one CTA per query): merges
// `nlists` sorted candidate lists of kp keys from global memory, rescores the survivors with the
// reference's fp64 arithmetic (RagIndex.cs:124-135), orders them (:64-66) and writes the hits.
// Global id of a local row.  Contiguous shards: row_base + local.  Streaming ingest spreads batches of
// `row_block` rows round-robin over `row_nshards` shards; the global id is then the row's position in the
// global insertion order, so (score desc, id asc) is the order of one index fed with the same stream.
__device__ __forceinline__ int64_t global_row_id(const ScanArgs& a, uint32_t local) {
    if (a.row_block <= 0) return a.row_base + static_cast<int64_t>(local);
    const int64_t blk = static_cast<int64_t>(local) / a.row_block, off = static_cast<int64_t>(local) % a.row_block;
    return a.row_base + (blk * a.row_nshards + a.row_shard) * a.row_block + off;
}

// Merges `nlists` sorted lists of list_len keys (global memory) into lists[0..kp) (shared, 2*kConsumerWarps*kp keys of
// scratch), kp >= list_len (a shorter list is staged padded with empty keys); `stage` is idle shared memory of
// stage_bytes.  Called by all kScanThreads threads of the CTA.
__device__ __forceinline__ void merge_global_lists(const uint64_t* glists, int nlists, int list_len, int kp, uint64_t* lists,
                                                   uint64_t* stage, uint32_t stage_bytes) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int cap_lists = max(1, static_cast<int>(stage_bytes / (static_cast<uint32_t>(kp) * 8u)));
    if (warp < kConsumerWarps) {
        for (int j = lane; j < kp; j += 32) lists[warp * kp + j] = 0ull;
    }
    for (int l0 = 0; l0 < nlists; l0 += cap_lists) {
        const int nl = min(cap_lists, nlists - l0);
        const uint64_t* src = glists + static_cast<size_t>(l0) * list_len;
        const int total = nl * kp;
        if (list_len == kp) {
#pragma unroll 8
            for (int i = tid; i < total; i += kScanThreads) stage[i] = __ldcg(src + i);  // all loads in flight
        } else {
#pragma unroll 4
            for (int i = tid; i < total; i += kScanThreads) {
                const int l = i / kp, j = i - l * kp;
                stage[i] = j < list_len ? __ldcg(src + static_cast<size_t>(l) * list_len + j) : 0ull;
            }
        }
        __syncthreads();
        if (warp < kConsumerWarps)
            for (int c = warp; c < nl; c += kConsumerWarps) warp_merge_desc(lists + warp * kp, stage + c * kp, kp, lane);
        __syncthreads();
    }
    for (int w = 1; w < kConsumerWarps; w <<= 1) {
        if (warp < kConsumerWarps && (warp % (2 * w)) == 0 && warp + w < kConsumerWarps)
            warp_merge_desc(lists + warp * kp, lists + (warp + w) * kp, kp, lane);
        __syncthreads();
    }
}

struct FinalizeCtx {
    uint8_t* ring;          // idle shared memory used as staging space
    uint32_t ring_bytes;
    uint32_t row_bytes;     // bytes of one stored row
    uint64_t* lists;        // shared: 2 * kConsumerWarps * kp keys
    ScanMisc* misc;         // shared
    const float* qraw;      // shared: the raw query, dim floats (16-byte aligned)
    const uint64_t* glists; // global: nlists * list_len keys
    int nlists;
    int list_len;           // keys per global list (<= a.kp; 0 = a.kp).  Shorter than a.kp: the batched path keeps 32 keys
                            // per CTA and query but merges them into a wider candidate set (a.kp)
    float bound_floor;      // every row that was never offered to a list scored <= this (-inf if none)
};

__device__ __forceinline__ void stamp(const ScanArgs& a, int i) {
    if (a.scratch.stamps) a.scratch.stamps[i] = globaltimer_ns();
}

template <typename T>
__device__ __forceinline__ void finalize_candidates(const ScanArgs& a, const FinalizeCtx& f) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int kp = a.kp;
    uint64_t* lists = f.lists;
    ScanMisc* misc = f.misc;
    const float* qraw = f.qraw;
    const uint8_t* mat = static_cast<const uint8_t*>(a.mat);
    const uint32_t ring_bytes = f.ring_bytes;
    const int list_len = f.list_len > 0 ? f.list_len : kp;
    // Lists shorter than the merged set: a row can have been pushed out of ITS list by 32 better rows of the same CTA
    // while still beating the merged set's last entry.  Every such row scored no higher than the last key of the (full)
    // list that dropped it, so the largest of those keys joins the bound below.
    if (tid == 0) misc->n_cand = 0;   // reused as the max of the full lists' last score bits until the count below
    __syncthreads();
    if (list_len < kp) {
        unsigned int mx = 0u;
        for (int l = tid; l < f.nlists; l += kScanThreads) {
            const uint64_t last = __ldcg(f.glists + static_cast<size_t>(l) * list_len + (list_len - 1));
            mx = max(mx, static_cast<unsigned int>(last >> 32));
        }
        if (mx) atomicMax(reinterpret_cast<unsigned int*>(&misc->n_cand), mx);
    }
    __syncthreads();
    const unsigned int dropped_bits = static_cast<unsigned int>(misc->n_cand);
    __syncthreads();
    merge_global_lists(f.glists, f.nlists, list_len, kp, lists, reinterpret_cast<uint64_t*>(f.ring), ring_bytes);
    // lists[0..kp) : the kp best approximate keys of the whole shard, descending
    if (tid == 0) stamp(a, 2);

    // survivors = leading non-empty keys (the list is sorted, empty slots are 0 and sit at the end)
    int n_cand = static_cast<int>(min(static_cast<int64_t>(kp), a.n_rows));
    {
        int cnt = 0;
        for (int j = tid; j < n_cand; j += kScanThreads) cnt += lists[j] != 0ull;
        if (tid == 0) misc->n_cand = 0;
        __syncthreads();
        if (cnt) atomicAdd(&misc->n_cand, cnt);
        __syncthreads();
        n_cand = misc->n_cand;
    }
    const int k_take = min(a.k_take, n_cand);
    // Rows are staged in the (now idle) tile ring.  A row that was stored inexactly (normalised with
    // scale != 1, or rounded to bf16) is rescored from its ORIGINAL fp32 values kept in the sidecar,
    // so the fp64 score below is the reference's value for the row as it was added.
    // A scan over the bf16 shadow of an fp32 store (a.rescore_rows) rescoring from the fp32 rows themselves: same mechanism.
    const uint32_t raw_bytes = a.raw_slot ? static_cast<uint32_t>(a.raw_ld) * 4u : 0u;
    const uint32_t res_bytes = a.rescore_rows ? static_cast<uint32_t>(a.rescore_ld) * 4u : 0u;
    const uint32_t rs = max(max(f.row_bytes, raw_bytes), res_bytes) + 16;  // padded smem row stride: spreads rows over banks
    const int cap = max(1, static_cast<int>(ring_bytes / rs));
    uint8_t* rowbuf = f.ring;

    for (int b0 = 0; b0 < n_cand; b0 += cap) {
        const int nb = min(cap, n_cand - b0);
        // stage: one bulk async copy per candidate row, all in flight at once, one mbarrier for the batch
        if (warp == 0) {
            fence_proxy_async();  // the ring was last touched through the generic proxy (list staging / previous batch)
            uint64_t* sbar = reinterpret_cast<uint64_t*>(&misc->stage_bar);
            uint32_t my_bytes = 0;
            for (int c = lane; c < nb; c += 32) {
                const uint32_t row = key_row(lists[b0 + c]);
                const int32_t slot = a.raw_slot ? a.raw_slot[row] : -1;
                my_bytes += slot >= 0 ? raw_bytes : (a.rescore_rows ? res_bytes : f.row_bytes);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) my_bytes += __shfl_xor_sync(0xffffffffu, my_bytes, o);
            if (lane == 0) mbar_arrive_expect_tx(sbar, my_bytes);
            __syncwarp();
            for (int c = lane; c < nb; c += 32) {
                const uint32_t row = key_row(lists[b0 + c]);
                const int32_t slot = a.raw_slot ? a.raw_slot[row] : -1;
                const void* src = slot >= 0 ? static_cast<const void*>(a.raw_rows + static_cast<size_t>(slot) * a.raw_ld)
                                  : a.rescore_rows ? static_cast<const void*>(a.rescore_rows + static_cast<size_t>(row) * a.rescore_ld)
                                                   : static_cast<const void*>(mat + static_cast<size_t>(row) * f.row_bytes);
                bulk_g2s(rowbuf + static_cast<size_t>(c) * rs, src, slot >= 0 ? raw_bytes : (a.rescore_rows ? res_bytes : f.row_bytes), sbar);
                misc->raw_fmt[c] = (slot >= 0 || a.rescore_rows) ? 1 : 0;
            }
        }
        mbar_wait(reinterpret_cast<uint64_t*>(&misc->stage_bar), static_cast<uint32_t>(b0 / cap) & 1u);
        __syncthreads();
        if (tid == 0 && b0 == 0) stamp(a, 3);
        // chains: threads [0, nbp) run the dot products, [nbp, 2*nbp) the row norms, thread 2*nbp the query
        // norm -- every warp runs one kind only, so the two kinds never serialise inside a warp
        const int nbp = (nb + 31) & ~31;
        const int nslots = 2 * nbp + (b0 == 0 ? 1 : 0);
        for (int id = tid; id < nslots; id += kScanThreads) {
            if (id == 2 * nbp) {
                misc->na = exact_chain<T>(nullptr, true, qraw, a.dim, 1);  // :131
            } else {
                const int kind = id >= nbp ? 1 : 0;
                const int c = id - kind * nbp;
                if (c < nb) {
                    const uint8_t* row = rowbuf + static_cast<size_t>(c) * rs;
                    const bool f32row = (sizeof(T) == 4) || misc->raw_fmt[c];
                    const double v = exact_chain<T>(row, f32row, qraw, a.dim, kind);
                    if (kind)
                        misc->nb[c] = v;  // :132
                    else
                        misc->dot[c] = v;  // :130
                }
            }
        }
        __syncthreads();
        for (int c = tid; c < nb; c += kScanThreads) {
            const double den = __dadd_rn(__dmul_rn(__dsqrt_rn(misc->na), __dsqrt_rn(misc->nb[c])), 1e-9);
            misc->ex[b0 + c] = __ddiv_rn(misc->dot[c], den);  // :134
        }
        __syncthreads();
    }

    if (tid == 0) stamp(a, 4);
    // final order: (fp64 score desc with double.CompareTo semantics, row asc); rank by counting
    for (int c = tid; c < n_cand; c += kScanThreads) {
        const double sc = misc->ex[c];
        const uint64_t kc = ord_f64(sc);
        const uint32_t rc = key_row(lists[c]);
        int rank = 0;
        for (int u = 0; u < n_cand; ++u) {
            const uint64_t ku = ord_f64(misc->ex[u]);
            const uint32_t ru = key_row(lists[u]);
            rank += (ku > kc) || (ku == kc && ru < rc);
        }
        const int64_t gid = global_row_id(a, rc);
        if (rank < k_take) {
            if (a.out_idx) a.out_idx[rank] = gid;
            if (a.out_score) a.out_score[rank] = sc;
            if (a.exact && a.exact[rc] == 0 && !(a.raw_slot && a.raw_slot[rc] >= 0)) atomicAdd(&misc->inexact, 1);
            if (rank == k_take - 1) misc->ek = sc;
        }
        if (a.out_cand && rank < a.cand_slots && rank < k_take) {
            a.out_cand[rank].score = sc;
            a.out_cand[rank].id = gid;
        }
    }
    if (a.out_cand) {
        for (int j = k_take + tid; j < a.cand_slots; j += kScanThreads) {
            a.out_cand[j].score = -CUDART_INF;
            a.out_cand[j].id = -1;
        }
    }
    __syncthreads();
    if (tid == 0) {
        bool cert;
        double bound_total = -CUDART_INF;
        if (a.n_rows <= static_cast<int64_t>(kp) && a.key_below == 0ull) {
            cert = true;  // every row was rescored
        } else {
            // Rows outside the survivors have approximate score <= bound.  Their reference score is at
            // most bound + eps, where eps covers the fp32 rounding of the scan AND the reference's
            // "+1e-9" in the denominator (RagIndex.cs:134), which moves a score by at most
            // 1e-9 / (|q| * smallest row norm) towards zero.
            double bound = lists[kp - 1] ? static_cast<double>(key_score(lists[kp - 1])) : -CUDART_INF;
            if (static_cast<double>(f.bound_floor) > bound) bound = static_cast<double>(f.bound_floor);
            if (dropped_bits) bound = fmax(bound, static_cast<double>(unord_f32(dropped_bits)));
            double shrink = 0.0;
            if (a.min_norm_bits || a.min_norm_assumed > 0.f) {
                const double mn = a.min_norm_bits ? static_cast<double>(__uint_as_float(*a.min_norm_bits)) : static_cast<double>(a.min_norm_assumed);
                const double d = sqrt(misc->na) * mn;
                shrink = 1e-9 / d;  // +inf when the query is zero: then only a full rescore certifies
            }
            bound_total = bound + static_cast<double>(a.eps) + shrink;
            // the list did not fill up and nothing was skipped by an initial threshold: no row was left out below it
            const bool complete = n_cand < kp && !(f.bound_floor > -CUDART_INF_F) && dropped_bits == 0u;
            if (complete) bound_total = -CUDART_INF;
            cert = misc->ek > bound_total || complete;  // false if either side is NaN
            if (misc->na == 0.0) {
                // All-zero query: every reference score is exactly +0.0 (or NaN, ranked last), any k rows in index order
                // are right.  A NON-zero query whose fp32 squares all flush to zero (|q_i| < ~3e-23) also has na == 0,
                // but the reference then ranks by dot / 1e-9, which the cosine scan knows nothing about: not certified.
                bool all_zero = true;
                for (int i = 0; i < a.dim; ++i) all_zero = all_zero && (qraw[i] == 0.0f);
                cert = all_zero;
            }
            if (k_take < a.k_take) cert = false;  // fewer survivors than requested hits: nothing to certify against
        }
        int32_t flags = 0;
        if (cert) flags |= RAGSTORE_HIT_CERTIFIED;
        if (misc->inexact == 0) flags |= RAGSTORE_HIT_EXACT_ROWS;
        if (a.out_flags) *a.out_flags = flags;
        if (a.out_found) *a.out_found = k_take;
        if (a.out_bound) *a.out_bound = bound_total;
        if (a.out_last_key) *a.out_last_key = n_cand > 0 ? lists[n_cand - 1] : 0ull;
        stamp(a, 5);
        if (a.out_done) {
            // Hits, found and flags were stored by threads of this CTA before the barrier above (or by this thread just now):
            // the system-scope fence orders all of them before the word the host polls.
            __threadfence_system();
            *reinterpret_cast<volatile unsigned int*>(a.out_done) = a.done_seq;
        }
    }
    if (a.xch == nullptr) return;

    // ---- the exchange step, fused: this CTA is the only one of the launch still running.  Publish this shard's records to
    // every rank (own inbox included), raise the arrival words, wait for everybody else's, rank the G x k records.
    // Liveness and the two-slot argument are those of exchange.cu: every rank publishes BEFORE it waits for anything,
    // and a rank can only reach call s+1 once every rank has published call s.
    __syncthreads();   // out_cand / out_flags of this launch were written by threads of this CTA
    const ExchangeView& v = *a.xch;
    const int per = a.cand_slots;
    const int slot = static_cast<int>(a.xseq & 1ull);
    const size_t rec_base = (static_cast<size_t>(slot) * v.world + v.rank) * v.max_records;
    const size_t flag_base = (static_cast<size_t>(slot) * v.world + v.rank) * v.max_queries;
    {
        const uint4* src = reinterpret_cast<const uint4*>(a.out_cand);
        for (int i = tid; i < v.world * per; i += kScanThreads) {
            const int r = i / per, j = i - r * per;
            reinterpret_cast<uint4*>(v.rec[r] + rec_base)[j] = src[j];   // plain stores over NVLink into the peer's inbox
        }
    }
    __threadfence_system();
    if (tid == 0) {
        misc->is_last = 0;   // reused: "a peer timed out"
        misc->inexact = 3;   // reused: flags ANDed over the ranks
    }
    __syncthreads();
    const int my_flags = a.out_flags ? *a.out_flags : 3;
    if (tid < v.world)
        st_release_sys(v.flag[tid] + flag_base, (a.xseq << 2) | static_cast<unsigned long long>(my_flags & 3));
    if (tid < v.world) {
        const unsigned long long* fw = v.flag[v.rank] + (static_cast<size_t>(slot) * v.world + tid) * v.max_queries;
        const unsigned long long t0 = globaltimer_ns();
        unsigned long long w;
        while (((w = ld_acquire_sys(fw)) >> 2) != a.xseq) {
            if (globaltimer_ns() - t0 > 5000000000ull) {   // 5 s: reported, never hung on
                misc->is_last = 1;
                w = 0;
                break;
            }
            __nanosleep(64);
        }
        atomicAnd(&misc->inexact, static_cast<int>(w & 3ull));
    }
    __syncthreads();
    if (tid == 0) {
        if (a.x_status) *a.x_status = misc->is_last;
        if (a.x_flags) *a.x_flags = misc->inexact;
    }
    const ragstore_cand* inbox = v.rec[v.rank] + static_cast<size_t>(slot) * v.world * v.max_records;
    merge_query_block<true>(inbox, static_cast<size_t>(v.max_records), v.world, 0, per, per, a.x_idx, a.x_score, a.x_found, f.ring,
                            &misc->n_cand);
}

}  // namespace rag
