// scan_topk.cu -- K2+K4+K5: single-query scoring of the whole corpus, fused with top-k selection
// and the exact fp64 rescoring of the surviving candidates.  One launch per query.
//
// Replaces (reference, Minimal_Agent_RAG/Minimal_RAG/Helpers/RagIndex.cs):
//   :62-63   _chunks.Select(c => new RagHit(.., Cosine(q, c.Embedding)))     N dot products
//   :64-66   .OrderByDescending(h => h.Score).Take(Math.Max(1, topK))        stable top-k
//   :124-135 Cosine: fp32 products summed sequentially in fp64                (rescoring stage)
//
// Shape of the kernel (B200 / sm_100a):
//   * persistent grid, one CTA per SM; CTA = 8 consumer warps + 1 producer warp;
//   * the corpus is row-major and contiguous, so a tile of R rows is ONE contiguous span: the
//     producer lane streams tiles (plus the R reciprocal norms) into a multi-stage shared-memory
//     ring with 1-D bulk async copies (cp.async.bulk -> TMA engine, SASS UBLKCP) that complete on
//     mbarriers; nothing is staged through registers, ~170 KB per SM are in flight;
//   * a consumer warp owns whole rows: 128-bit conflict-free LDS, the normalised query in
//     registers, a fixed 4-accumulator + xor-butterfly reduction (identical tree for every row, so
//     duplicate rows score identically and ties fall to the lower row);
//   * the score never leaves the SM: it is packed with the row id into an order-preserving 64-bit
//     key and offered to the warp's sorted candidate list (threshold test first; inserts are
//     O(k log N) per warp over the whole scan);
//   * lists are merged warp -> CTA -> (last CTA to finish, atomic ticket) grid with bitonic merges;
//   * the last CTA then recomputes the reference's exact fp64 arithmetic for the k' >= k survivors,
//     orders them by (fp64 score desc, row asc), emits the top k and a certificate that no row
//     outside the survivors could have ranked (approximate score of the worst survivor + eps is
//     below the k-th exact score).
// HBM traffic: every stored element is read exactly once (d * sizeof(elem) + 4 bytes per row).
#include <math_constants.h>

#include "common.cuh"
#include "ragstore_internal.h"

namespace rag {

struct ScanLayout {
    uint32_t rows_per_tile, n_stages, row_bytes, tile_bytes, stage_bytes;
    uint32_t nchunks;    // 16-byte chunks per row
    uint32_t qpad;       // floats in each query buffer (zero padded)
    uint32_t off_bars, off_qraw, off_qhat, off_lists, off_misc, total;
};

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
};

// ---------------------------------------------------------------- warp-level sorted lists in smem
// L sorted descending, length kp (multiple of 32); precondition x > L[kp-1]; x is warp-uniform.
__device__ __forceinline__ void warp_insert_desc(uint64_t* L, int kp, uint64_t x, int lane) {
    for (int base = kp - 32; base >= 0; base -= 32) {
        const int j = base + lane;
        const uint64_t cur = L[j];
        const uint64_t prev = j > 0 ? L[j - 1] : ~0ull;
        __syncwarp();
        L[j] = cur > x ? cur : (prev > x ? x : prev);
        __syncwarp();
        const uint64_t head_prev = __shfl_sync(0xffffffffu, prev, 0);
        if (head_prev > x) break;  // everything before this block already ranks above x
    }
}

// A, B sorted descending (kp a power of two >= 32).  A <- top kp of A u B, sorted descending.
__device__ __forceinline__ void warp_merge_desc(uint64_t* A, const uint64_t* B, int kp, int lane) {
    for (int j = lane; j < kp; j += 32) {
        const uint64_t x = A[j], y = B[kp - 1 - j];
        A[j] = x > y ? x : y;  // bitonic sequence holding the kp largest
    }
    __syncwarp();
    for (int stride = kp >> 1; stride > 0; stride >>= 1) {
        for (int t = lane; t < (kp >> 1); t += 32) {
            const int i = ((t & ~(stride - 1)) << 1) | (t & (stride - 1));
            const int j = i + stride;
            const uint64_t x = A[i], y = A[j];
            if (x < y) {
                A[i] = y;
                A[j] = x;
            }
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------- per-row dot product from smem
template <typename T>
struct Elem;
template <>
struct Elem<float> {
    static constexpr int kPerChunk = 4;  // elements per 16-byte chunk
    static constexpr int kQV = 1;        // float4 of query per chunk
};
template <>
struct Elem<uint16_t> {  // bf16 bits
    static constexpr int kPerChunk = 8;
    static constexpr int kQV = 2;
};

template <typename T>
__device__ __forceinline__ void chunk_fma(const float4& v, const float4* q, float& a0, float& a1, float& a2,
                                          float& a3) {
    if constexpr (sizeof(T) == 4) {
        a0 = fmaf(v.x, q[0].x, a0);
        a1 = fmaf(v.y, q[0].y, a1);
        a2 = fmaf(v.z, q[0].z, a2);
        a3 = fmaf(v.w, q[0].w, a3);
    } else {
        const uint32_t w0 = __float_as_uint(v.x), w1 = __float_as_uint(v.y);
        const uint32_t w2 = __float_as_uint(v.z), w3 = __float_as_uint(v.w);
        a0 = fmaf(__uint_as_float(w0 << 16), q[0].x, a0);
        a1 = fmaf(__uint_as_float(w0 & 0xFFFF0000u), q[0].y, a1);
        a2 = fmaf(__uint_as_float(w1 << 16), q[0].z, a2);
        a3 = fmaf(__uint_as_float(w1 & 0xFFFF0000u), q[0].w, a3);
        a0 = fmaf(__uint_as_float(w2 << 16), q[1].x, a0);
        a1 = fmaf(__uint_as_float(w2 & 0xFFFF0000u), q[1].y, a1);
        a2 = fmaf(__uint_as_float(w3 << 16), q[1].z, a2);
        a3 = fmaf(__uint_as_float(w3 & 0xFFFF0000u), q[1].w, a3);
    }
}

// DV > 0: query chunks for this lane live in registers (qreg); DV == 0: read them from smem (qhat).
template <typename T, int DV>
__device__ __forceinline__ float row_dot(const uint8_t* row, const float4* qreg, const float* qhat, int lane,
                                         int nchunks) {
    constexpr int QV = Elem<T>::kQV;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    if constexpr (DV > 0) {
#pragma unroll
        for (int c = 0; c < DV; ++c) {
            const int ch = c * 32 + lane;
            if (ch < nchunks) {
                const float4 v = lds128(row + ch * 16);
                chunk_fma<T>(v, qreg + c * QV, a0, a1, a2, a3);
            }
        }
    } else {
        for (int ch = lane; ch < nchunks; ch += 32) {
            const float4 v = lds128(row + ch * 16);
            float4 q[QV];
#pragma unroll
            for (int u = 0; u < QV; ++u) q[u] = lds128(qhat + (ch * QV + u) * 4);
            chunk_fma<T>(v, q, a0, a1, a2, a3);
        }
    }
    return warp_sum((a0 + a1) + (a2 + a3));
}

template <typename T>
__device__ __forceinline__ float elem_at(const uint8_t* row, int i) {
    if constexpr (sizeof(T) == 4)
        return reinterpret_cast<const float*>(row)[i];
    else
        return bf16_bits_to_f32(reinterpret_cast<const uint16_t*>(row)[i]);
}

// ---------------------------------------------------------------- the kernel
template <typename T, int DV>
__global__ void __launch_bounds__(kScanThreads, 1) scan_topk_kernel(const ScanArgs a, const ScanLayout L) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int QV = Elem<T>::kQV;

    uint64_t* full = reinterpret_cast<uint64_t*>(smem + L.off_bars);
    uint64_t* empty = full + L.n_stages;
    float* qraw = reinterpret_cast<float*>(smem + L.off_qraw);
    float* qhat = reinterpret_cast<float*>(smem + L.off_qhat);
    uint64_t* lists = reinterpret_cast<uint64_t*>(smem + L.off_lists);
    ScanMisc* misc = reinterpret_cast<ScanMisc*>(smem + L.off_misc);

    const int kp = a.kp;
    const int R = static_cast<int>(L.rows_per_tile);
    const int64_t n_tiles = (a.n_rows + R - 1) / R;
    const uint8_t* mat = static_cast<const uint8_t*>(a.mat);

    if (tid == 0) {
        for (uint32_t s = 0; s < L.n_stages; ++s) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, kConsumerWarps);
        }
        fence_barrier_init();
    }
    __syncthreads();

    if (warp == kConsumerWarps) {
        // ---------------- producer: stream tiles of R rows (+ their reciprocal norms)
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            uint32_t it = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
                const uint32_t s = it % L.n_stages, ph = (it / L.n_stages) & 1u;
                mbar_wait(empty + s, ph ^ 1u);
                const int64_t row0 = tile * R;
                const int nrows = static_cast<int>(min(static_cast<int64_t>(R), a.n_rows - row0));
                const uint32_t bytes = static_cast<uint32_t>(nrows) * L.row_bytes;
                const uint32_t rnb = static_cast<uint32_t>((nrows + 3) & ~3) * 4u;
                uint8_t* dst = smem + static_cast<size_t>(s) * L.stage_bytes;
                mbar_arrive_expect_tx(full + s, bytes + rnb);
                bulk_g2s_hint(dst, mat + static_cast<size_t>(row0) * L.row_bytes, bytes, full + s, pol);
                bulk_g2s(dst + L.tile_bytes, a.rnorm + row0, rnb, full + s);
            }
        }
        __syncwarp();
    } else {
        // ---------------- consumers: query prologue (overlaps the first tiles already in flight)
        const int ctid = tid;  // 0 .. kConsumerWarps*32-1
        double part = 0.0;
        for (int i = ctid; i < static_cast<int>(L.qpad); i += kConsumerWarps * 32) {
            const float v = i < a.dim ? a.d_query[i] : 0.f;
            qraw[i] = v;
            part += static_cast<double>(v) * static_cast<double>(v);
        }
        part = warp_sum(part);
        if (lane == 0) misc->red[warp] = part;
        named_bar_sync(1, kConsumerWarps * 32);
        double nq = 0.0;
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w) nq += misc->red[w];
        const float rq = nq > 0.0 ? static_cast<float>(1.0 / sqrt(nq)) : 0.f;
        for (int i = ctid; i < static_cast<int>(L.qpad); i += kConsumerWarps * 32) qhat[i] = qraw[i] * rq;
        named_bar_sync(1, kConsumerWarps * 32);

        float4 qreg[(DV > 0 ? DV : 1) * QV];
        if constexpr (DV > 0) {
#pragma unroll
            for (int c = 0; c < DV; ++c)
#pragma unroll
                for (int u = 0; u < QV; ++u) qreg[c * QV + u] = lds128(qhat + ((c * 32 + lane) * QV + u) * 4);
        }

        uint64_t* myL = lists + warp * kp;
        for (int j = lane; j < kp; j += 32) myL[j] = 0ull;
        __syncwarp();
        uint64_t thr = 0ull;
        const int nchunks = static_cast<int>(L.nchunks);

        uint32_t it = 0;
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
            const uint32_t s = it % L.n_stages, ph = (it / L.n_stages) & 1u;
            mbar_wait(full + s, ph);
            const uint8_t* st = smem + static_cast<size_t>(s) * L.stage_bytes;
            const float* rn = reinterpret_cast<const float*>(st + L.tile_bytes);
            const int64_t row0 = tile * R;
            const int nrows = static_cast<int>(min(static_cast<int64_t>(R), a.n_rows - row0));
            for (int r = warp; r < nrows; r += kConsumerWarps) {
                const float d = row_dot<T, DV>(st + static_cast<size_t>(r) * L.row_bytes, qreg, qhat, lane, nchunks);
                const float sc = d * rn[r] + 0.0f;  // +0.0f: -0.0 -> +0.0 (the reference never yields -0.0)
                const uint64_t key = make_key(sc, static_cast<uint32_t>(row0 + r));
                if (key > thr) {
                    warp_insert_desc(myL, kp, key, lane);
                    thr = myL[kp - 1];
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty + s);
        }

        // ---------------- CTA-level merge of the warp lists (tree), then publish to global scratch
        named_bar_sync(1, kConsumerWarps * 32);
        for (int s = 1; s < kConsumerWarps; s <<= 1) {
            if ((warp % (2 * s)) == 0 && warp + s < kConsumerWarps)
                warp_merge_desc(lists + warp * kp, lists + (warp + s) * kp, kp, lane);
            named_bar_sync(1, kConsumerWarps * 32);
        }
        uint64_t* gl = a.scratch.cta_lists + static_cast<size_t>(blockIdx.x) * kp;
        for (int j = ctid; j < kp; j += kConsumerWarps * 32) gl[j] = lists[j];
        __threadfence();
        named_bar_sync(1, kConsumerWarps * 32);
        if (tid == 0) {
            const unsigned int t = atomicAdd(a.scratch.ticket, 1u);
            const int last = (t == gridDim.x - 1);
            if (last) *a.scratch.ticket = 0u;  // self-reset: every other CTA has already arrived
            misc->is_last = last;
            misc->inexact = 0;
        }
    }
    __syncthreads();
    if (!misc->is_last) return;

    // =====================================================================================
    // Last CTA: grid-level merge, exact rescoring (RagIndex.cs:124-135), final order (:64-66)
    // =====================================================================================
    __threadfence();
    if (warp < kConsumerWarps) {
        uint64_t* A = lists + warp * kp;
        uint64_t* B = lists + (kConsumerWarps + warp) * kp;
        for (int j = lane; j < kp; j += 32) A[j] = 0ull;
        __syncwarp();
        for (int c = warp; c < static_cast<int>(gridDim.x); c += kConsumerWarps) {
            const uint64_t* gl = a.scratch.cta_lists + static_cast<size_t>(c) * kp;
            for (int j = lane; j < kp; j += 32) B[j] = __ldcg(gl + j);
            __syncwarp();
            warp_merge_desc(A, B, kp, lane);
        }
    }
    __syncthreads();
    for (int s = 1; s < kConsumerWarps; s <<= 1) {
        if (warp < kConsumerWarps && (warp % (2 * s)) == 0 && warp + s < kConsumerWarps)
            warp_merge_desc(lists + warp * kp, lists + (warp + s) * kp, kp, lane);
        __syncthreads();
    }
    // lists[0..kp) : the kp best approximate keys of the whole shard, descending

    const int n_cand = static_cast<int>(min(static_cast<int64_t>(kp), a.n_rows));
    // Rows are staged in the (now idle) tile ring.  A row that was stored inexactly (normalised with
    // scale != 1, or rounded to bf16) is rescored from its ORIGINAL fp32 values kept in the sidecar,
    // so the fp64 score below is the reference's value for the row as it was added.
    const uint32_t raw_bytes = a.raw_slot ? static_cast<uint32_t>(a.raw_ld) * 4u : 0u;
    const uint32_t rs = max(L.row_bytes, raw_bytes) + 16;  // padded smem row stride: spreads rows over banks
    const int cap = max(1, static_cast<int>((L.n_stages * L.stage_bytes) / rs));
    const int nchunks = static_cast<int>(L.nchunks);
    uint8_t* rowbuf = smem;

    for (int b0 = 0; b0 < n_cand; b0 += cap) {
        const int nb = min(cap, n_cand - b0);
        for (int c = warp; c < nb; c += kScanThreads / 32) {
            const uint32_t row = key_row(lists[b0 + c]);
            const int32_t slot = a.raw_slot ? a.raw_slot[row] : -1;
            const uint4* src;
            int nch;
            if (slot >= 0) {
                src = reinterpret_cast<const uint4*>(a.raw_rows + static_cast<size_t>(slot) * a.raw_ld);
                nch = static_cast<int>(raw_bytes / 16);
            } else {
                src = reinterpret_cast<const uint4*>(mat + static_cast<size_t>(row) * L.row_bytes);
                nch = nchunks;
            }
            uint4* dst = reinterpret_cast<uint4*>(rowbuf + static_cast<size_t>(c) * rs);
            for (int ch = lane; ch < nch; ch += 32) dst[ch] = __ldg(src + ch);
            if (lane == 0) misc->raw_fmt[c] = slot >= 0 ? 1 : 0;
        }
        __syncthreads();
        const int nchains = 2 * nb + (b0 == 0 ? 1 : 0);
        for (int id = tid; id < nchains; id += kScanThreads) {
            double acc = 0.0;  // :127
            if (id == 2 * nb) {
                for (int i = 0; i < a.dim; ++i) acc = __dadd_rn(acc, static_cast<double>(__fmul_rn(qraw[i], qraw[i])));  // :131
                misc->na = acc;
            } else {
                const int c = id >> 1;
                const uint8_t* row = rowbuf + static_cast<size_t>(c) * rs;
                const bool f32row = (sizeof(T) == 4) || misc->raw_fmt[c];
                if (id & 1) {
                    for (int i = 0; i < a.dim; ++i) {
                        const float x = f32row ? reinterpret_cast<const float*>(row)[i] : elem_at<T>(row, i);
                        acc = __dadd_rn(acc, static_cast<double>(__fmul_rn(x, x)));  // :132
                    }
                    misc->nb[c] = acc;
                } else {
                    for (int i = 0; i < a.dim; ++i) {
                        const float x = f32row ? reinterpret_cast<const float*>(row)[i] : elem_at<T>(row, i);
                        acc = __dadd_rn(acc, static_cast<double>(__fmul_rn(qraw[i], x)));  // :130
                    }
                    misc->dot[c] = acc;
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
        const int64_t gid = a.row_base + static_cast<int64_t>(rc);
        if (rank < a.k_take) {
            if (a.out_idx) a.out_idx[rank] = gid;
            if (a.out_score) a.out_score[rank] = sc;
            if (a.exact && a.exact[rc] == 0 && !(a.raw_slot && a.raw_slot[rc] >= 0)) atomicAdd(&misc->inexact, 1);
            if (rank == a.k_take - 1) misc->ek = sc;
        }
        if (a.out_cand && rank < a.cand_slots && rank < a.k_take) {
            a.out_cand[rank].score = sc;
            a.out_cand[rank].id = gid;
        }
    }
    if (a.out_cand) {
        for (int j = a.k_take + tid; j < a.cand_slots; j += kScanThreads) {
            a.out_cand[j].score = -CUDART_INF;
            a.out_cand[j].id = -1;
        }
    }
    __syncthreads();
    if (tid == 0) {
        bool cert;
        if (a.n_rows <= static_cast<int64_t>(kp)) {
            cert = true;  // every row was rescored
        } else {
            // Rows outside the survivors have approximate score <= bound.  Their reference score is at
            // most bound + eps, where eps covers the fp32 rounding of the scan AND the reference's
            // "+1e-9" in the denominator (RagIndex.cs:134), which moves a score by at most
            // 1e-9 / (|q| * smallest row norm) towards zero.
            const double bound = static_cast<double>(key_score(lists[kp - 1]));
            double shrink = 0.0;
            if (a.min_norm_bits) {
                const double d = sqrt(misc->na) * static_cast<double>(__uint_as_float(*a.min_norm_bits));
                shrink = 1e-9 / d;  // +inf when the query is zero: then only a full rescore certifies
            }
            cert = misc->ek > bound + static_cast<double>(a.eps) + shrink;  // false if either side is NaN
            if (misc->na == 0.0) cert = true;  // zero query: every reference score is exactly +0.0 (or NaN, ranked last)
        }
        int32_t flags = 0;
        if (cert) flags |= RAGSTORE_HIT_CERTIFIED;
        if (misc->inexact == 0) flags |= RAGSTORE_HIT_EXACT_ROWS;
        if (a.out_flags) *a.out_flags = flags;
        if (a.out_found) *a.out_found = a.k_take;
    }
}

// ---------------------------------------------------------------- host side
static constexpr uint32_t kSmemBudget = 220 * 1024;  // of the 227 KB a CTA may opt in to

static bool make_layout(const ScanArgs& a, int dv, ScanLayout& L) {
    const uint32_t esz = a.dtype == RAGSTORE_F32 ? 4u : 2u;
    L.row_bytes = static_cast<uint32_t>(a.ld) * esz;
    if (L.row_bytes % 16) return false;
    L.nchunks = L.row_bytes / 16;
    const uint32_t epc = 16 / esz;
    L.qpad = dv > 0 ? static_cast<uint32_t>(dv) * 32u * epc : static_cast<uint32_t>(a.ld);
    if (L.qpad < static_cast<uint32_t>(a.ld)) return false;
    const uint32_t fixed_lists = sizeof(uint64_t) * 2u * kConsumerWarps * static_cast<uint32_t>(a.kp);
    const uint32_t fixed = 16u * 8u * 2u /*barriers*/ + 2u * L.qpad * 4u + fixed_lists + sizeof(ScanMisc) + 512u;
    if (fixed + 2u * (4u * L.row_bytes + 128u) > kSmemBudget) return false;
    if (a.raw_slot && static_cast<uint32_t>(a.raw_ld) * 4u + 16u > 2u * (4u * L.row_bytes + 128u)) return false;
    const uint32_t ring = kSmemBudget - fixed;
    // rows per tile: a multiple of 4 (16-byte granularity of the reciprocal-norm slice), ~32 KB
    uint32_t R = (32u * 1024u) / L.row_bytes;
    R = R < 4 ? 4 : (R > 64 ? 64 : (R & ~3u));
    while (R > 4 && 2u * (R * L.row_bytes + R * 4u + 128u) > ring) R -= 4;
    L.rows_per_tile = R;
    L.tile_bytes = R * L.row_bytes;
    L.stage_bytes = (L.tile_bytes + R * 4u + 127u) & ~127u;
    uint32_t ns = ring / L.stage_bytes;
    if (ns < 2) return false;
    L.n_stages = ns > 8 ? 8 : ns;
    uint32_t off = L.n_stages * L.stage_bytes;
    L.off_bars = off;
    off += 16u * 8u * 2u;
    L.off_qraw = off;
    off += L.qpad * 4u;
    L.off_qhat = off;
    off += L.qpad * 4u;
    off = (off + 15u) & ~15u;
    L.off_lists = off;
    off += fixed_lists;
    off = (off + 15u) & ~15u;
    L.off_misc = off;
    off += sizeof(ScanMisc);
    L.total = off;
    return L.total <= 227u * 1024u && L.n_stages <= 16;
}

template <typename T, int DV>
static int launch_one(const ScanArgs& a, const ScanLayout& L, int grid, cudaStream_t stream) {
    auto kfn = scan_topk_kernel<T, DV>;
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return static_cast<int>(e);
    kfn<<<grid, kScanThreads, L.total, stream>>>(a, L);
    return static_cast<int>(cudaGetLastError());
}

template <typename T>
static int dispatch_dv(const ScanArgs& a, int dv, const ScanLayout& L, int grid, cudaStream_t st) {
    switch (dv) {
        case 1: return launch_one<T, 1>(a, L, grid, st);
        case 2: return launch_one<T, 2>(a, L, grid, st);
        case 3: return launch_one<T, 3>(a, L, grid, st);
        case 4: return launch_one<T, 4>(a, L, grid, st);
        case 5: return launch_one<T, 5>(a, L, grid, st);
        case 6: return launch_one<T, 6>(a, L, grid, st);
        case 7: return launch_one<T, 7>(a, L, grid, st);
        case 8: return launch_one<T, 8>(a, L, grid, st);
        default: return launch_one<T, 0>(a, L, grid, st);
    }
}

int launch_scan(const ScanArgs& a, cudaStream_t stream) {
    if (a.n_rows <= 0 || a.n_rows >= 0xFFFFFFFFll || a.kp < 32 || a.kp > kMaxKp || (a.kp & (a.kp - 1)) ||
        a.k_take < 1 || a.k_take > a.kp)
        return static_cast<int>(cudaErrorInvalidValue);
    const uint32_t esz = a.dtype == RAGSTORE_F32 ? 4u : 2u;
    const uint32_t nchunks = static_cast<uint32_t>(a.ld) * esz / 16u;
    int dv = static_cast<int>((nchunks + 31u) / 32u);
    if (dv > 8) dv = 0;
    ScanLayout L;
    if (!make_layout(a, dv, L)) return static_cast<int>(cudaErrorInvalidConfiguration);
    const int64_t n_tiles = (a.n_rows + L.rows_per_tile - 1) / L.rows_per_tile;
    int grid = static_cast<int>(n_tiles < a.sm_count ? n_tiles : a.sm_count);
    if (grid > kMaxCtas) grid = kMaxCtas;
    if (a.dtype == RAGSTORE_F32) return dispatch_dv<float>(a, dv, L, grid, stream);
    return dispatch_dv<uint16_t>(a, dv, L, grid, stream);
}

int scan_configure_device() { return 0; }

}  // namespace rag
