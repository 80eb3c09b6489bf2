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
//     mbarriers; nothing is staged through registers, ~190 KB per SM are in flight;
//   * a consumer warp owns whole tiles (tile i of the CTA goes to warp i mod 8) and scores them four
//     rows at a time: 128-bit conflict-free LDS, the normalised query in registers, four FMA chains
//     per row and a transposed xor-butterfly that reduces four rows with six shuffles.  Every row
//     goes through the same reduction tree, so duplicate rows score identically and ties fall to
//     the lower row;
//   * the score never leaves the SM: it is packed with the row id into an order-preserving 64-bit
//     key; keys above the warp's threshold are buffered and merged (bitonic) into the warp's sorted
//     candidate list 32 at a time;
//   * lists are merged warp -> CTA -> (last CTA to finish, atomic ticket) grid with bitonic merges;
//   * the last CTA then recomputes the reference's exact fp64 arithmetic for the k' >= k survivors,
//     orders them by (fp64 score desc, row asc), emits the top k and a certificate that no row
//     outside the survivors could have ranked.
// HBM traffic: every stored element is read exactly once (d * sizeof(elem) + 4 bytes per row).
#include <math_constants.h>

#include <atomic>
#include <cstdlib>

#include <cstring>

#include "topk_device.cuh"

namespace rag {

constexpr int kGroup = 4;  // rows scored together by one warp

struct ScanLayout {
    uint32_t rows_per_tile, n_stages, row_bytes, tile_bytes, stage_bytes;
    uint32_t nchunks;    // 16-byte chunks per row
    uint32_t qpad;       // floats in each query buffer (zero padded)
    uint32_t off_bars, off_qraw, off_qhat, off_lists, off_buf, off_misc, total;
    uint32_t n_consumers;  // consumer warps that take tiles; divides n_stages, so a warp owns fixed stages
    uint32_t no_hint;      // experiments: 1 = plain bulk copies without the L2 evict_first policy
};

// ---------------------------------------------------------------- per-row dot products from smem
template <typename T>
struct Elem;
template <>
struct Elem<float> {
    static constexpr int kQV = 1;  // float4 of query per 16-byte chunk
};
template <>
struct Elem<uint16_t> {  // bf16 bits
    static constexpr int kQV = 2;
};

template <typename T>
__device__ __forceinline__ void chunk_fma(const float4& v, const float4* q, float (&acc)[4]) {
    if constexpr (sizeof(T) == 4) {
        acc[0] = fmaf(v.x, q[0].x, acc[0]);
        acc[1] = fmaf(v.y, q[0].y, acc[1]);
        acc[2] = fmaf(v.z, q[0].z, acc[2]);
        acc[3] = fmaf(v.w, q[0].w, acc[3]);
    } else {
        const uint32_t w0 = __float_as_uint(v.x), w1 = __float_as_uint(v.y);
        const uint32_t w2 = __float_as_uint(v.z), w3 = __float_as_uint(v.w);
        acc[0] = fmaf(__uint_as_float(w0 << 16), q[0].x, acc[0]);
        acc[1] = fmaf(__uint_as_float(w0 & 0xFFFF0000u), q[0].y, acc[1]);
        acc[2] = fmaf(__uint_as_float(w1 << 16), q[0].z, acc[2]);
        acc[3] = fmaf(__uint_as_float(w1 & 0xFFFF0000u), q[0].w, acc[3]);
        acc[0] = fmaf(__uint_as_float(w2 << 16), q[1].x, acc[0]);
        acc[1] = fmaf(__uint_as_float(w2 & 0xFFFF0000u), q[1].y, acc[1]);
        acc[2] = fmaf(__uint_as_float(w3 << 16), q[1].z, acc[2]);
        acc[3] = fmaf(__uint_as_float(w3 & 0xFFFF0000u), q[1].w, acc[3]);
    }
}

// Scores kGroup rows (p[0..3]) against the query.  On return every lane holds the full dot product
// of the row slot (lane>>4)*2 + ((lane>>3)&1).  The reduction tree is the standard xor butterfly
// (16, 8, 4, 2, 1) for every row, whatever its slot in the group.
template <typename T, int DV, bool FULL>
__device__ __forceinline__ float group_dot(const uint8_t* const (&p)[kGroup], const float4* qreg, const float* qhat,
                                           int lane, int nchunks) {
    constexpr int QV = Elem<T>::kQV;
    float acc[kGroup][4];
#pragma unroll
    for (int r = 0; r < kGroup; ++r) acc[r][0] = acc[r][1] = acc[r][2] = acc[r][3] = 0.f;
    if constexpr (DV > 0) {
#pragma unroll
        for (int c = 0; c < DV; ++c) {
            const int ch = c * 32 + lane;
            if (FULL || ch < nchunks) {
#pragma unroll
                for (int r = 0; r < kGroup; ++r) {
                    const float4 v = lds128(p[r] + ch * 16);
                    chunk_fma<T>(v, qreg + c * QV, acc[r]);
                }
            }
        }
    } else {
        for (int ch = lane; ch < nchunks; ch += 32) {
            float4 q[QV];
#pragma unroll
            for (int u = 0; u < QV; ++u) q[u] = lds128(qhat + (ch * QV + u) * 4);
#pragma unroll
            for (int r = 0; r < kGroup; ++r) {
                const float4 v = lds128(p[r] + ch * 16);
                chunk_fma<T>(v, q, acc[r]);
            }
        }
    }
    float s[kGroup];
#pragma unroll
    for (int r = 0; r < kGroup; ++r) s[r] = (acc[r][0] + acc[r][1]) + (acc[r][2] + acc[r][3]);
    // transposed butterfly: 4 rows -> 2 -> 1 while halving the lanes that carry each row
    const bool hi = (lane & 16) != 0;
    float k0 = hi ? s[2] : s[0];
    const float g0 = hi ? s[0] : s[2];
    float k1 = hi ? s[3] : s[1];
    const float g1 = hi ? s[1] : s[3];
    k0 += __shfl_xor_sync(0xffffffffu, g0, 16);
    k1 += __shfl_xor_sync(0xffffffffu, g1, 16);
    const bool hi2 = (lane & 8) != 0;
    float k = hi2 ? k1 : k0;
    const float g = hi2 ? k0 : k1;
    k += __shfl_xor_sync(0xffffffffu, g, 8);
    k += __shfl_xor_sync(0xffffffffu, k, 4);
    k += __shfl_xor_sync(0xffffffffu, k, 2);
    k += __shfl_xor_sync(0xffffffffu, k, 1);
    return k;
}

// ---------------------------------------------------------------- the kernel
// A raw query carried in the kernel's parameter block (see launch_scan): used when a.d_query is null.
struct QueryParam {
    float v[kParamQueryMax];
};

template <typename T, int DV, bool FULL>
__global__ void __launch_bounds__(kScanThreads, 1) scan_topk_kernel(const ScanArgs a, const ScanLayout L,
                                                                    const __grid_constant__ QueryParam qp) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int QV = Elem<T>::kQV;

    uint64_t* full = reinterpret_cast<uint64_t*>(smem + L.off_bars);
    uint64_t* empty = full + L.n_stages;
    float* qraw = reinterpret_cast<float*>(smem + L.off_qraw);
    float* qhat = reinterpret_cast<float*>(smem + L.off_qhat);
    uint64_t* lists = reinterpret_cast<uint64_t*>(smem + L.off_lists);
    uint64_t* cbuf = reinterpret_cast<uint64_t*>(smem + L.off_buf);
    ScanMisc* misc = reinterpret_cast<ScanMisc*>(smem + L.off_misc);

    const int kp = a.kp;
    const int R = static_cast<int>(L.rows_per_tile);
    const int n_stages = static_cast<int>(L.n_stages);
    const int64_t n_tiles = (a.n_rows + R - 1) / R;
    // tiles of this CTA: blockIdx.x, blockIdx.x + gridDim.x, ...   (local sequence number `it`)
    const int64_t my_tiles = n_tiles > blockIdx.x ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const uint8_t* mat = static_cast<const uint8_t*>(a.mat);

    if (tid == 0) {
        if (blockIdx.x == 0) {
            stamp(a, 0);
            // two launches overlapping on one scratch corrupt each other's lists silently: count it (ragstore_info)
            if (atomicExch(a.scratch.ticket + 1, 1u) != 0u) atomicAdd(a.scratch.ticket + 2, 1u);
        }
        for (int s = 0; s < n_stages; ++s) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, 1);  // one consumer warp drains a tile
        }
        mbar_init(reinterpret_cast<uint64_t*>(&misc->stage_bar), 1);
        fence_barrier_init();
    }
    __syncthreads();

    if (warp == kConsumerWarps) {
        // ---------------- producer: stream tiles of R rows (+ their reciprocal norms)
        if (lane == 0) {
            const uint64_t pol = l2_policy_evict_first();
            int s = 0;
            uint32_t ph = 0;
            for (int64_t it = 0; it < my_tiles; ++it) {
                mbar_wait(empty + s, ph ^ 1u);
                const int64_t row0 = (blockIdx.x + it * gridDim.x) * R;
                const int nrows = static_cast<int>(min(static_cast<int64_t>(R), a.n_rows - row0));
                const uint32_t bytes = static_cast<uint32_t>(nrows) * L.row_bytes;
                const uint32_t rnb = static_cast<uint32_t>((nrows + 3) & ~3) * 4u;
                uint8_t* dst = smem + static_cast<size_t>(s) * L.stage_bytes;
                mbar_arrive_expect_tx(full + s, bytes + rnb);
                if (L.no_hint)
                    bulk_g2s(dst, mat + static_cast<size_t>(row0) * L.row_bytes, bytes, full + s);
                else
                    bulk_g2s_hint(dst, mat + static_cast<size_t>(row0) * L.row_bytes, bytes, full + s, pol);
                bulk_g2s(dst + L.tile_bytes, a.rnorm + row0, rnb, full + s);
                if (++s == n_stages) {
                    s = 0;
                    ph ^= 1u;
                }
            }
        }
        __syncwarp();
    } else {
        // ---------------- consumers: query prologue (overlaps the first tiles already in flight)
        double part = 0.0;
        const float* qsrc = a.d_query ? a.d_query : qp.v;
        for (int i = tid; i < static_cast<int>(L.qpad); i += kConsumerWarps * 32) {
            const float v = i < a.dim ? qsrc[i] : 0.f;
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
        for (int i = tid; i < static_cast<int>(L.qpad); i += kConsumerWarps * 32) qhat[i] = qraw[i] * rq;
        named_bar_sync(1, kConsumerWarps * 32);

        float4 qreg[(DV > 0 ? DV : 1) * QV];
        if constexpr (DV > 0) {
#pragma unroll
            for (int c = 0; c < DV; ++c)
#pragma unroll
                for (int u = 0; u < QV; ++u) qreg[c * QV + u] = lds128(qhat + ((c * 32 + lane) * QV + u) * 4);
        }

        uint64_t* myL = lists + warp * kp;
        uint64_t* myB = lists + (kConsumerWarps + warp) * kp;
        uint64_t* mybuf = cbuf + warp * 32;
        for (int j = lane; j < kp; j += 32) myL[j] = 0ull;
        __syncwarp();
        uint64_t thr = 0ull;
        int cnt = 0;
        const int nchunks = static_cast<int>(L.nchunks);
        const int rsel = ((lane >> 4) << 1) | ((lane >> 3) & 1);  // row slot this lane ends up holding

        // Local tile `it` lives in stage it % n_stages and is scored by warp it % n_consumers.  n_consumers
        // divides n_stages, so a warp only ever waits on stages whose previous phase it drained itself
        // (an mbarrier parity wait can only tell adjacent phases apart).
        const int nc = static_cast<int>(L.n_consumers);
        int s = warp;
        uint32_t ph = 0;
        for (int64_t it = warp; warp < nc && it < my_tiles; it += nc) {
            mbar_wait(full + s, ph);
            const uint8_t* st = smem + static_cast<size_t>(s) * L.stage_bytes;
            const float* rn = reinterpret_cast<const float*>(st + L.tile_bytes);
            const int64_t row0 = (blockIdx.x + it * gridDim.x) * R;
            const int nrows = static_cast<int>(min(static_cast<int64_t>(R), a.n_rows - row0));
            const int last = nrows - 1;
            for (int g0 = 0; g0 < nrows; g0 += kGroup) {
                const uint8_t* p[kGroup];
#pragma unroll
                for (int r = 0; r < kGroup; ++r) p[r] = st + static_cast<size_t>(min(g0 + r, last)) * L.row_bytes;
                const float d = group_dot<T, DV, FULL>(p, qreg, qhat, lane, nchunks);
                const int r = g0 + rsel;
                const bool valid = r <= last;
                const float sc = d * rn[valid ? r : last] + 0.0f;  // +0.0f: -0.0 -> +0.0 (the reference never yields -0.0)
                const uint64_t key = make_key(sc, static_cast<uint32_t>(row0 + r));
                unsigned m = __ballot_sync(0xffffffffu, valid && key > thr && (a.key_below == 0ull || key < a.key_below));
                while (m) {  // rare: a row beat the warp's current k'-th best
                    const int src = __ffs(m) - 1;
                    const uint64_t kk = __shfl_sync(0xffffffffu, key, src);
                    m &= ~(0xFFu << (src & ~7));
                    if (lane == 0) mybuf[cnt] = kk;
                    if (++cnt == 32) {
                        thr = warp_flush(myL, myB, mybuf, cnt, kp, lane);
                        cnt = 0;
                        m &= __ballot_sync(0xffffffffu, valid && key > thr);  // (the key_below test already passed for these)
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty + s);
            s += nc;
            if (s >= n_stages) {
                s -= n_stages;
                ph ^= 1u;
            }
        }
        if (cnt > 0) warp_flush(myL, myB, mybuf, cnt, kp, lane);

        // ---------------- CTA-level merge of the warp lists (tree), then publish to global scratch
        named_bar_sync(1, kConsumerWarps * 32);
        for (int w = 1; w < kConsumerWarps; w <<= 1) {
            if ((warp % (2 * w)) == 0 && warp + w < kConsumerWarps)
                warp_merge_desc(lists + warp * kp, lists + (warp + w) * kp, kp, lane);
            named_bar_sync(1, kConsumerWarps * 32);
        }
        uint64_t* gl = a.scratch.cta_lists + static_cast<size_t>(blockIdx.x) * kp;
        for (int j = tid; j < kp; j += kConsumerWarps * 32) gl[j] = lists[j];
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
    if (tid == 0) stamp(a, 1);
    FinalizeCtx f;
    f.ring = smem;
    f.ring_bytes = L.n_stages * L.stage_bytes;  // the tile ring is idle now: reused as staging space
    f.row_bytes = L.row_bytes;
    f.lists = lists;
    f.misc = misc;
    f.qraw = qraw;
    f.glists = a.scratch.cta_lists;
    f.nlists = static_cast<int>(gridDim.x);
    f.list_len = 0;
    f.bound_floor = -CUDART_INF_F;
    finalize_candidates<T>(a, f);
    if (tid == 0) atomicExch(a.scratch.ticket + 1, 0u);  // scratch free again
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
    const uint32_t lists_bytes = sizeof(uint64_t) * 2u * kConsumerWarps * static_cast<uint32_t>(a.kp);
    const uint32_t buf_bytes = sizeof(uint64_t) * 32u * kConsumerWarps;
    const uint32_t bars_bytes = 16u * 8u * 2u;
    const uint32_t fixed = bars_bytes + 2u * L.qpad * 4u + lists_bytes + buf_bytes + sizeof(ScanMisc) + 512u;
    if (fixed + 2u * (4u * L.row_bytes + 128u) > kSmemBudget) return false;
    if (a.raw_slot && static_cast<uint32_t>(a.raw_ld) * 4u + 16u > 2u * (4u * L.row_bytes + 128u)) return false;
    if (a.rescore_rows && static_cast<uint32_t>(a.rescore_ld) * 4u + 16u > 2u * (4u * L.row_bytes + 128u)) return false;
    const uint32_t ring = kSmemBudget - fixed;
    // Rows per tile R: a multiple of 4 (row group of a warp; 16-byte granularity of the reciprocal-norm
    // slice).  Stages: 16, 8 or fewer, with min(8, stages) consumer warps so that each warp owns
    // fixed stages.  Pick the R (<= ~24 KB tiles) that keeps the most bytes in flight.
    static const uint32_t tile_kb = [] {  // tuning knob for experiments; default 24 KB tiles
        const char* e = std::getenv("RAGSTORE_SCAN_TILE_KB");
        const int v = e ? atoi(e) : 0;
        return static_cast<uint32_t>(v >= 4 && v <= 96 ? v : 24);
    }();
    uint32_t Rmax = (tile_kb * 1024u) / L.row_bytes;
    Rmax = Rmax < 4 ? 4 : (Rmax > 64 ? 64 : (Rmax & ~3u));
    uint32_t best_R = 0, best_ns = 0, best_bytes = 0;
    for (uint32_t R = Rmax; R >= 4; R -= 4) {
        const uint32_t sb = (R * L.row_bytes + R * 4u + 127u) & ~127u;
        uint32_t ns = ring / sb;
        ns = ns >= 16 ? 16 : (ns >= 8 ? 8 : ns);
        if (ns < 2) continue;
        if (ns * sb > best_bytes) best_bytes = ns * sb, best_R = R, best_ns = ns;
    }
    if (best_R == 0) return false;
    const uint32_t R = best_R;
    L.rows_per_tile = R;
    L.tile_bytes = R * L.row_bytes;
    L.stage_bytes = (L.tile_bytes + R * 4u + 127u) & ~127u;
    L.n_stages = best_ns;
    L.n_consumers = best_ns < kConsumerWarps ? best_ns : kConsumerWarps;
    static const uint32_t no_hint = std::getenv("RAGSTORE_SCAN_NOHINT") ? 1u : 0u;   // experiments; read once
    L.no_hint = no_hint;
    uint32_t off = L.n_stages * L.stage_bytes;
    L.off_bars = off;
    off += bars_bytes;
    L.off_qraw = off;
    off += L.qpad * 4u;
    L.off_qhat = off;
    off += L.qpad * 4u;
    off = (off + 15u) & ~15u;
    L.off_lists = off;
    off += lists_bytes;
    L.off_buf = off;
    off += buf_bytes;
    off = (off + 15u) & ~15u;
    L.off_misc = off;
    off += sizeof(ScanMisc);
    L.total = off;
    return L.total <= 227u * 1024u;
}

template <typename T, int DV, bool FULL>
static int launch_one(const ScanArgs& a, const ScanLayout& L, int grid, cudaStream_t stream, const QueryParam& qp) {
    auto kfn = scan_topk_kernel<T, DV, FULL>;
    // the shared-memory opt-in is per function AND per device: done once for each (bit = device ordinal), not per launch
    static std::atomic<uint64_t> opted_in{0};
    int dev = 0;
    cudaGetDevice(&dev);
    const uint64_t bit = 1ull << (dev & 63);
    if (!(opted_in.load(std::memory_order_acquire) & bit)) {
        cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return static_cast<int>(e);
        opted_in.fetch_or(bit, std::memory_order_release);
    }
    kfn<<<grid, kScanThreads, L.total, stream>>>(a, L, qp);
    return static_cast<int>(cudaGetLastError());
}

template <typename T, int DV>
static int launch_full(const ScanArgs& a, const ScanLayout& L, int grid, cudaStream_t st, bool full, const QueryParam& qp) {
    return full ? launch_one<T, DV, true>(a, L, grid, st, qp) : launch_one<T, DV, false>(a, L, grid, st, qp);
}

template <typename T>
static int dispatch_dv(const ScanArgs& a, int dv, bool full, const ScanLayout& L, int grid, cudaStream_t st, const QueryParam& qp) {
    switch (dv) {
        case 1: return launch_full<T, 1>(a, L, grid, st, full, qp);
        case 2: return launch_full<T, 2>(a, L, grid, st, full, qp);
        case 3: return launch_full<T, 3>(a, L, grid, st, full, qp);
        case 4: return launch_full<T, 4>(a, L, grid, st, full, qp);
        case 6: return launch_full<T, 6>(a, L, grid, st, full, qp);
        case 8: return launch_full<T, 8>(a, L, grid, st, full, qp);
        default: return launch_one<T, 0, false>(a, L, grid, st, qp);
    }
}

int launch_scan(const ScanArgs& a, cudaStream_t stream, const float* h_query) {
    if ((a.d_query == nullptr) != (h_query != nullptr) || (h_query && a.dim > kParamQueryMax)) return static_cast<int>(cudaErrorInvalidValue);
    QueryParam qp;  // left as it is unless the query travels in it: the launch copies the block either way
    if (h_query) memcpy(qp.v, h_query, sizeof(float) * static_cast<size_t>(a.dim));
    if (a.n_rows <= 0 || a.n_rows >= 0xFFFFFFFFll || a.kp < 32 || a.kp > kMaxKp || (a.kp & (a.kp - 1)) ||
        a.k_take < 1 || a.k_take > a.kp)
        return static_cast<int>(cudaErrorInvalidValue);
    const uint32_t esz = a.dtype == RAGSTORE_F32 ? 4u : 2u;
    const uint32_t nchunks = static_cast<uint32_t>(a.ld) * esz / 16u;
    int dv = static_cast<int>((nchunks + 31u) / 32u);
    if (dv == 5) dv = 6;  // instantiated register-query widths: 1,2,3,4,6,8 chunks per lane
    if (dv == 7) dv = 8;
    if (dv > 8) dv = 0;   // wider rows read the query from shared memory
    const bool full = dv > 0 && nchunks == static_cast<uint32_t>(dv) * 32u;
    ScanLayout L;
    if (!make_layout(a, dv, L)) return static_cast<int>(cudaErrorInvalidConfiguration);
    const int64_t n_tiles = (a.n_rows + L.rows_per_tile - 1) / L.rows_per_tile;
    int grid = static_cast<int>(n_tiles < a.sm_count ? n_tiles : a.sm_count);
    if (grid > kMaxCtas) grid = kMaxCtas;
    if (a.dtype == RAGSTORE_F32) return dispatch_dv<float>(a, dv, full, L, grid, stream, qp);
    return dispatch_dv<uint16_t>(a, dv, full, L, grid, stream, qp);
}

}  // namespace rag
