// batch_gemm.cu -- K3: batched-query scoring on the 5th-generation tensor cores (tcgen05 + TMEM),
// with the top-k filter fused into the accumulator read-back.  This is the one place on the path that
// is a dense contraction: scores[q][row] = sum_d Q[q][d] * C[row][d]   (SURVEY.md 2.1 "K3", config 3).
//
// Replaces the reference's per-request loop over `Cosine` (Helpers/RagIndex.cs:62-63,124-135) for a
// whole batch of queries at once; selection and exact rescoring follow the same rules as the
// single-query kernel (topk_device.cuh), so the returned hits are the reference's.
//
// Shape (B200 / sm_100a), one persistent CTA per SM, 384 threads, clusters of 2:
//   warp 0      TMA producer: 2-D tiled bulk tensor copies (cp.async.bulk.tensor, 128B swizzle) of a
//               256-row corpus tile slice and (multicast across the cluster) of the 128/256-query slice
//               into a 4..5-stage ring; one stage = 128 bytes of K per row;
//   warp 1      MMA issuer: one lane issues tcgen05.mma.cta_group::1 (M = 128 queries, N = 256 corpus
//               rows; kind::tf32 with K = 8 for fp32 stores, kind::f16 on bf16 operands with K = 16 for
//               bf16 stores), accumulators in TMEM, tcgen05.commit releases ring stages and publishes
//               finished accumulators;
//   warp 2      allocates / frees the 512 TMEM columns;
//   warps 4-11  epilogue: thread <-> query (TMEM lane), tcgen05.ld of its 256 scores per tile, a
//               register threshold and a thread-private candidate list (scores in registers, row ids
//               in shared memory) -- no atomics, no overflow, and the M x N score block is never
//               written anywhere.
// Queries are the M operand on purpose: a TMEM lane then carries ONE query's scores across corpus rows,
// which is what makes the per-query running top-k thread-private.
// Reduced-precision products differ from the reference's (tf32: two 10-bit mantissas; bf16 stores: the
// query is rounded to bf16, the stored rows are used as they are), so this kernel only nominates
// k' = 32 candidates per query; batch_finalize_kernel re-scores them with the reference's exact
// fp64 arithmetic and certifies the margin (eps ~ 2.1e-3); uncertified queries are redone by the
// scan kernel (host side, ragstore_api.cu).
#include <cuda.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <set>
#include <utility>

#include "topk_device.cuh"

// In-kernel cycle counters of the epilogue (RAGSTORE_BATCH_DEBUG=8 prints them) are compiled in only with
// -DRAGSTORE_BATCH_PROFILE (make EXTRA=-DRAGSTORE_BATCH_PROFILE): the production kernels carry no clock reads and no printf.
#ifdef RAGSTORE_BATCH_PROFILE
#define RAG_PROF(...) __VA_ARGS__
#else
#define RAG_PROF(...)
#endif

namespace rag {

constexpr int kBM = 128;            // queries per MMA (M)
constexpr int kBN = 256;            // corpus rows per tile (MMA N)
constexpr int kBKBytes = 128;       // bytes of K per ring stage row = one 128B swizzle atom row
constexpr int kBatchThreads = 384;
constexpr int kEpiWarp0 = 4;        // first epilogue warp
constexpr int kBatchKp = 32;        // candidates kept per query and CTA
constexpr int kCluster = 2;         // CTAs that share (multicast) the query operand
constexpr uint32_t kOperandBytes = 128 * kBKBytes;  // one 128-row operand slice: 16 KB

struct BatchLayout {
    uint32_t n_stages, stage_bytes, mblocks;
    uint32_t off_bars, off_tmem, off_lists, total;
};

// ---------------------------------------------------------------- PTX wrappers (tcgen05 / TMA)
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
// same, delivered to the same shared-memory offset (and mbarrier offset) of every CTA in cta_mask
__device__ __forceinline__ void tma_load_2d_mc(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1,
                                               uint16_t cta_mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
        " [%0], [%1, {%4, %5}], [%2], %3;" ::"r"(smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "h"(cta_mask), "r"(c0), "r"(c1)
        : "memory");
}
// L2 prefetch of a tensor box (no shared-memory destination): decouples DRAM latency from the ring
__device__ __forceinline__ void tma_prefetch_2d(const CUtensorMap* map, int c0, int c1) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global [%0, {%1, %2}];" ::"l"(reinterpret_cast<uint64_t>(map)), "r"(c0),
                 "r"(c1)
                 : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// arrives on the mbarrier once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// the same arrival on the barrier at this offset in every CTA of cta_mask
__device__ __forceinline__ void tc_commit_mc(uint64_t* bar, uint16_t cta_mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     smem_u32(bar)),
                 "h"(cta_mask)
                 : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, tf32 inputs, fp32 accumulate
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// same with bf16 inputs (kind::f16, K = 16 per instruction)
__device__ __forceinline__ void mma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// 32 lanes x 32 consecutive 32-bit columns: thread i of the warp gets lane (base + i)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor for a K-major operand slice written by TMA with 128B swizzle:
// rows are 128 bytes apart, 8-row groups 1024 bytes apart (SBO), LBO unused, descriptor version 1.
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFFu);        // start address, 16-byte units
    d |= static_cast<uint64_t>(1u) << 16;                          // leading byte offset (ignored for swizzled K-major)
    d |= static_cast<uint64_t>(1024u >> 4) << 32;                  // stride byte offset: 8 rows x 128 B
    d |= static_cast<uint64_t>(1u) << 46;                          // descriptor version (Blackwell)
    d |= static_cast<uint64_t>(2u) << 61;                          // layout type: SWIZZLE_128B
    return d;
}

// instruction descriptor: D fp32, A/B tf32, both K-major, M = 128, N = 128
constexpr uint32_t kIdescTf32 = (1u << 4) | (2u << 7) | (2u << 10) | ((kBN >> 3) << 17) | ((kBM >> 4) << 24);
// the same shape for kind::f16 with bf16 A and B (format code 1)
constexpr uint32_t kIdescBf16 = (1u << 4) | (1u << 7) | (1u << 10) | ((kBN >> 3) << 17) | ((kBM >> 4) << 24);

struct BatchArgs {
    int64_t n_rows;
    int nq;            // real queries in this launch (<= 128 * mblocks)
    int mblocks;       // 1 or 2 blocks of 128 queries
    int nkb;           // k-blocks of 128 bytes (32 fp32 / 64 bf16) per row
    const float* rnorm; // per-row reciprocal norms (bf16 stores: applied to the tensor-core dot products)
    uint64_t* glists;  // [nq][n_ctas][kBatchKp] candidate keys, written by this kernel
    int debug;         // profiling experiments: 1 = epilogue drains without scanning, 2 = no MMA issue
    int tile_stride;   // 1 = every tile; s > 1 = the sampling pass: tiles 0, s, 2s, ...
    const float* thr0; // [nq] initial thresholds from the sampling pass (rows scoring below can be skipped), or null
    int sample;        // 1 = sampling pass: keep only each thread's 4 best scores (no lists), written to `glists` as floats
};

// ---------------------------------------------------------------- thread-private candidate lists
// A TMEM lane (= epilogue thread) carries one query, so its running top-32 is private to the thread:
// the 32 approximate scores live in REGISTERS (unsorted, -inf = empty), the matching row ids in shared
// memory (rows[slot][query]).  An insert overwrites the current minimum and re-finds the minimum with a
// compare tree -- no atomics, no shared-memory traffic on the hot compare path, no overflow.  Which of
// several EQUAL minimum scores is evicted is arbitrary; that can only matter for rows tied with the
// worst survivor, a case the certificate (bound + eps < k-th exact score) rejects, so such a query is
// redone by the fp32 scan kernel.
struct ArgMin {
    float v;
    int i;
};
__device__ __forceinline__ ArgMin amin(const ArgMin& a, const ArgMin& b) { return b.v < a.v ? b : a; }

__device__ __forceinline__ void list_insert(float (&S)[kBatchKp], int& minpos, float& thr, uint32_t* rows, int NQ, int q,
                                            float sc, uint32_t row) {
#pragma unroll
    for (int p = 0; p < kBatchKp; ++p) S[p] = (p == minpos) ? sc : S[p];
    rows[minpos * NQ + q] = row;
    ArgMin t16[16], t8[8], t4[4], t2[2];
#pragma unroll
    for (int p = 0; p < 16; ++p) t16[p] = amin(ArgMin{S[2 * p], 2 * p}, ArgMin{S[2 * p + 1], 2 * p + 1});
#pragma unroll
    for (int p = 0; p < 8; ++p) t8[p] = amin(t16[2 * p], t16[2 * p + 1]);
#pragma unroll
    for (int p = 0; p < 4; ++p) t4[p] = amin(t8[2 * p], t8[2 * p + 1]);
#pragma unroll
    for (int p = 0; p < 2; ++p) t2[p] = amin(t4[2 * p], t4[2 * p + 1]);
    const ArgMin m = amin(t2[0], t2[1]);
    minpos = m.i;
    thr = m.v;
}

__device__ __forceinline__ void tmem_ld_32x16_nowait(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// bf16 stores: the smallest raw dot product that could still exceed `thr` once multiplied by a reciprocal norm
// <= rn_max.  Rounded down with slack, so the test `dot > raw_threshold` never misses a row; a negative threshold
// (lists still cold) admits everything.
__device__ __forceinline__ float raw_threshold(float thr, float rn_max) {
    if (thr < 0.f) return -CUDART_INF_F;
    return __fdiv_rd(thr, rn_max) * (1.0f - 2.4e-7f);   // thr > 0, rn_max == 0: +inf (nothing passes, all scores are 0)
}

// ---------------------------------------------------------------- epilogue pieces shared by both kernels
// smallest / largest reciprocal norm of a tile's valid rows (bf16 stores; fp32 stores keep unit rows: 1, 1)
template <bool BF>
__device__ __forceinline__ void tile_rnorm_range(const float* rnorm, int64_t row0, int nvalid, int lane, float& rn_lo, float& rn_hi) {
    rn_lo = 1.f, rn_hi = 1.f;
    if (BF) {
        float lo = CUDART_INF_F, hi = 0.f;
        for (int j = lane; j < nvalid; j += 32) {
            const float r = __ldg(rnorm + row0 + j);
            lo = fminf(lo, r), hi = fmaxf(hi, r);
        }
        rn_lo = __uint_as_float(__reduce_min_sync(0xffffffffu, __float_as_uint(lo)));
        rn_hi = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(hi)));
    }
}

// Sampling pass, one finished 256-row accumulator at TMEM address `tacc` (this warp's lanes): fold the thread's query's
// scores into its 4 best (t0 >= t1 >= t2 >= t3).  bf16 store: a sampled value only has to be a LOWER bound of a real
// row's score, so the tile's smallest / largest reciprocal norm stands in for the row's own.
template <bool BF>
__device__ __forceinline__ void sample_tile(uint32_t tacc, int nvalid, float rn_lo, float rn_hi, float& t0, float& t1, float& t2,
                                            float& t3) {
    for (int cb = 0; cb < kBN; cb += 32) {
        uint32_t v[32];
        tmem_ld_32x32(tacc + cb, v);
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            float sc = __uint_as_float(v[j]);
            if (BF) sc = sc > 0.f ? sc * rn_lo : sc * rn_hi;
            if (sc > t3 && cb + j < nvalid) {  // 4-slot sorted insert
                const float a0 = fmaxf(t0, sc), b0 = fminf(t0, sc);
                const float a1 = fmaxf(t1, b0), b1 = fminf(t1, b0);
                const float a2 = fmaxf(t2, b1), b2 = fminf(t2, b1);
                t0 = a0, t1 = a1, t2 = a2, t3 = fmaxf(t3, b2);
            }
        }
    }
}

// Main pass, one finished 256-row accumulator: feed the thread's candidate list.
// bf16 store: score = dot * rnorm[row].  The steady-state test compares the raw dot products with thr / (largest rnorm
// of the tile), a superset of the rows that can beat thr; only those get the exact product.
template <bool BF>
__device__ __forceinline__ void scan_tile(uint32_t tacc, int64_t row0, int nvalid, float rn_max, const float* rnorm, bool active,
                                          float (&S)[kBatchKp], int& minpos, float& thr, uint32_t* rows, int NQ, int q) {
    float thr_cmp = BF ? raw_threshold(thr, rn_max) : thr;
    for (int cb = 0; cb < kBN; cb += 64) {
        // four 16-column loads in flight per wait: the TMEM read latency is paid once per 64 columns
        uint32_t v[4][16];
#pragma unroll
        for (int u = 0; u < 4; ++u) tmem_ld_32x16_nowait(tacc + cb + u * 16, v[u]);
        tmem_ld_wait();
        // Hits are rare once the lists have warmed up.  Steady state: one max over the 64 scores and one vote; the
        // per-column mask is built only when some lane of the warp has a score above its threshold; the warp then
        // revisits only the marked columns (one-column TMEM load) and the lanes concerned insert.  One copy of the
        // insert code keeps the loop in the I-cache.
        float mx[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            mx[u] = __uint_as_float(v[u][0]);
#pragma unroll
            for (int j = 1; j < 16; ++j) mx[u] = fmaxf(mx[u], __uint_as_float(v[u][j]));
        }
        const bool hot = active && fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) > thr_cmp;
        unsigned long long mine = 0, any = 0;
        if (__any_sync(0xffffffffu, hot)) {
            if (hot) {
#pragma unroll
                for (int u = 0; u < 4; ++u)
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        mine |= static_cast<unsigned long long>(__uint_as_float(v[u][j]) > thr_cmp ? 1u : 0u) << (u * 16 + j);
            }
            const unsigned lo = __reduce_or_sync(0xffffffffu, static_cast<unsigned>(mine));
            const unsigned hi = __reduce_or_sync(0xffffffffu, static_cast<unsigned>(mine >> 32));
            any = (static_cast<unsigned long long>(hi) << 32) | lo;
        }
        while (any) {
            const int j = __ffsll(static_cast<long long>(any)) - 1;
            any &= any - 1;
            uint32_t one;
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(one) : "r"(tacc + cb + j) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            float sc = __uint_as_float(one);
            if (BF) sc = cb + j < nvalid ? sc * __ldg(rnorm + row0 + cb + j) + 0.0f : -CUDART_INF_F;
            if (((mine >> j) & 1ull) && sc > thr && cb + j < nvalid) {  // rows ascend, so an equal score never displaces
                list_insert(S, minpos, thr, rows, NQ, q, sc, static_cast<uint32_t>(row0 + cb + j));
                thr_cmp = BF ? raw_threshold(thr, rn_max) : thr;
            }
        }
    }
}

// End of the main pass.  The ring is idle (every stage was consumed before the last accumulator completed): reuse it
// to hand each query's 32 keys to a warp, which sorts them (one key per lane) and publishes them as
// glists[query][list_id][pos].  `ql` = the thread's query within this CTA (0 .. NQ-1), `qg0` = global index of the
// CTA's first query; called by all 4 * NQ / 128 epilogue warps.
__device__ __forceinline__ void publish_lists(uint8_t* smem, const float (&S)[kBatchKp], const uint32_t* rows, int NQ, int ql, int qg0,
                                              int nq, uint64_t* glists, int nlists, int list_id, int lane) {
    uint64_t* keys = reinterpret_cast<uint64_t*>(smem);  // [query][slot]
    named_bar_sync(2, NQ);
#pragma unroll
    for (int p = 0; p < kBatchKp; ++p) {
        const uint32_t r = rows[p * NQ + ql];
        keys[ql * kBatchKp + p] = (S[p] > -CUDART_INF_F && r != 0xFFFFFFFFu) ? make_key(S[p] + 0.0f, r) : 0ull;
    }
    __syncwarp();
    const int ql0 = ql & ~31;
    const int n_live = max(0, min(32, nq - (qg0 + ql0)));
    for (int i = 0; i < n_live; ++i) {
        uint64_t k = keys[(ql0 + i) * kBatchKp + lane];
        k = warp_sort_desc(k, lane);
        glists[(static_cast<size_t>(qg0 + ql0 + i) * nlists + list_id) * kBatchKp + lane] = k;
    }
}

// ---------------------------------------------------------------- the tensor-core kernel
// MB = blocks of 128 queries (1 or 2).  Tile = 256 corpus rows: one MMA is M=128 (queries) x N=256 (rows),
// so MB = 1 double-buffers its accumulator (2 x 256 TMEM columns) while MB = 2 uses all 512 columns for
// one tile.  Per ring stage a CTA ingests 32 KB of corpus + MB x 16 KB of queries: the L2->SM ingest per
// corpus byte (1.5x / 2x) is what bounds this kernel (measured ~46 B/clk/SM for 128-byte TMA rows), which
// is why the tile is as tall as TMEM allows.
// BF = bf16 store: operands are bf16 (64 elements per 128-byte slice), and because such a store keeps rows
// as given (their norm is 1 only to within bf16 rounding) the epilogue applies the row's reciprocal norm.
template <int MB, bool BF>
__global__ void __cluster_dims__(kCluster, 1, 1) __launch_bounds__(kBatchThreads, 1)
batch_score_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_c,
                   const BatchArgs a, const BatchLayout L) {
    extern __shared__ uint8_t smem_raw[];
    // 128B-swizzled operand tiles must start on a 1024-byte boundary
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_stages = static_cast<int>(L.n_stages);
    constexpr int AS = MB == 1 ? 2 : 1;       // accumulator stages
    constexpr int NQ = kBM * MB;
    constexpr int KE = BF ? 64 : 32;          // elements per 128-byte K slice
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + L.off_bars);
    uint64_t* empty = full + n_stages;
    uint64_t* tmem_full = empty + n_stages;   // [AS]
    uint64_t* tmem_empty = tmem_full + 2;     // [AS]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(smem + L.off_tmem);
    uint32_t* rows = reinterpret_cast<uint32_t*>(smem + L.off_lists);  // [slot][query]

    // Clusters of kCluster CTAs walk the corpus together: in round `it` cluster c takes tiles
    // (c + it * n_clusters) * kCluster + rank.  The CTAs of a cluster need the SAME query slices at the
    // same time, so each loads 1/kCluster of a slice and multicasts it to all of them: query reads
    // out of L2 drop by kCluster.  A CTA whose tile lies past the end scores zero rows (TMA zero fill).
    const uint32_t crank = cluster_ctarank();
    const int64_t all_tiles = (a.n_rows + kBN - 1) / kBN;
    const int64_t n_tiles = (all_tiles + a.tile_stride - 1) / a.tile_stride;  // tiles this launch visits
    const int64_t n_clusters = gridDim.x / kCluster, cluster_id = blockIdx.x / kCluster;
    const int64_t n_groups = (n_tiles + kCluster - 1) / kCluster;
    const int64_t my_tiles = n_groups > cluster_id ? (n_groups - cluster_id + n_clusters - 1) / n_clusters : 0;
    const uint16_t cmask = static_cast<uint16_t>((1u << kCluster) - 1u);

    if (tid == 0) {
        for (int s = 0; s < n_stages; ++s) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, kCluster);  // every CTA of the cluster must be done with the stage
        }
        for (int s = 0; s < AS; ++s) {
            mbar_init(tmem_full + s, 1);
            mbar_init(tmem_empty + s, 4 * MB);  // one arrival per epilogue warp
        }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, 512);
    for (int i = tid; i < kBatchKp * NQ; i += kBatchThreads) rows[i] = 0xFFFFFFFFu;
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();  // peers' barriers are initialised before anything is multicast at them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ------------------------------------------------ TMA producer
        if (lane == 0) {
            tma_prefetch_desc(&tmap_q);
            tma_prefetch_desc(&tmap_c);
            int s = 0;
            uint32_t ph = 0;
            constexpr uint32_t tx = MB * kOperandBytes + 2u * kOperandBytes;
            constexpr int q_rows = NQ / kCluster;  // this CTA's share of the query slice
            const uint32_t q_off = crank * static_cast<uint32_t>(q_rows) * kBKBytes;
            for (int64_t it = 0; it < my_tiles; ++it) {
                const int row0 = static_cast<int>(((cluster_id + it * n_clusters) * kCluster + crank) * a.tile_stride * kBN);
                for (int kb = 0; kb < a.nkb; ++kb) {
                    mbar_wait(empty + s, ph ^ 1u);
                    uint8_t* st = smem + static_cast<size_t>(s) * L.stage_bytes;
                    mbar_arrive_expect_tx(full + s, tx);
                    tma_load_2d_mc(st + q_off, &tmap_q, full + s, kb * KE, static_cast<int>(crank) * q_rows, cmask);
                    tma_load_2d(st + MB * kOperandBytes, &tmap_c, full + s, kb * KE, row0);
                    if (++s == n_stages) {
                        s = 0;
                        ph ^= 1u;
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ------------------------------------------------ MMA issuer (one lane)
        if (lane == 0) {
            int s = 0;
            uint32_t ph = 0;
            for (int64_t it = 0; it < my_tiles; ++it) {
                const uint32_t as = static_cast<uint32_t>(it % AS), aph = static_cast<uint32_t>((it / AS) & 1);
                mbar_wait(tmem_empty + as, aph ^ 1u);  // the epilogue has drained this accumulator
                tc_fence_after();
                for (int kb = 0; kb < a.nkb; ++kb) {
                    mbar_wait(full + s, ph);
                    tc_fence_after();
                    const uint32_t st = smem_u32(smem + static_cast<size_t>(s) * L.stage_bytes);
                    const uint32_t b_addr = st + MB * kOperandBytes;
#pragma unroll
                    for (int k4 = 0; k4 < 4; ++k4) {  // 4 x (K=8 tf32 | K=16 bf16) = 32 bytes each per 128-byte slice
                        const uint64_t bdesc = make_sw128_desc(b_addr + k4 * 32);
#pragma unroll
                        for (int m = 0; m < MB; ++m) {
                            const uint64_t adesc = make_sw128_desc(st + m * kOperandBytes + k4 * 32);
                            if (!(a.debug & 2)) {
                                if (BF)
                                    mma_bf16(tmem_base + (as * MB + m) * 256u, adesc, bdesc, kIdescBf16, (kb | k4) != 0);
                                else
                                    mma_tf32(tmem_base + (as * MB + m) * 256u, adesc, bdesc, kIdescTf32, (kb | k4) != 0);
                            }
                        }
                    }
                    tc_commit_mc(empty + s, cmask);  // stage reusable (here and in the peers) once these MMAs have read it
                    if (++s == n_stages) {
                        s = 0;
                        ph ^= 1u;
                    }
                }
                tc_commit(tmem_full + as);  // accumulator complete
            }
        }
        __syncwarp();
    } else if (warp >= kEpiWarp0 && warp < kEpiWarp0 + 4 * MB) {
        // ------------------------------------------------ epilogue: thread <-> query
        const int q = (warp - kEpiWarp0) * 32 + lane;  // 0 .. NQ-1
        const int mblock = q >> 7;
        const bool live = q < a.nq;
        const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;  // TMEM lanes this warp may read
        float S[kBatchKp];
#pragma unroll
        for (int p = 0; p < kBatchKp; ++p) S[p] = -CUDART_INF_F;
        float thr = -CUDART_INF_F;
        int minpos = 0;
        // Sampling-pass threshold: the k'-th best score among the sampled rows is a lower bound of the k'-th best
        // overall, so rows scoring BELOW it need never enter a list.  Ties with it are kept (nextafter), and the
        // list slots start at this value so that it stays the threshold until 32 better rows have been seen.
        const float floor0 = (a.thr0 && live) ? nextafterf(a.thr0[q], -CUDART_INF_F) : -CUDART_INF_F;
        if (floor0 > -CUDART_INF_F) {
#pragma unroll
            for (int p = 0; p < kBatchKp; ++p) S[p] = floor0;
            thr = floor0;
        }
        const uint32_t tacc0 = tmem_base + lane_base + mblock * 256u;
        if (a.sample) {
            // Sampling pass: only thresholds are wanted, so a thread keeps just its 4 best scores (registers, no row
            // ids, no lists): the warm-up that makes list maintenance expensive costs almost nothing here.
            float t0 = -CUDART_INF_F, t1 = -CUDART_INF_F, t2 = -CUDART_INF_F, t3 = -CUDART_INF_F;  // t0 >= t1 >= t2 >= t3
            for (int64_t it = 0; it < my_tiles; ++it) {
                const uint32_t as = static_cast<uint32_t>(it % AS), aph = static_cast<uint32_t>((it / AS) & 1);
                const int64_t row0 = ((cluster_id + it * n_clusters) * kCluster + crank) * a.tile_stride * kBN;
                const int nvalid = static_cast<int>(min(static_cast<int64_t>(kBN), a.n_rows - row0));
                float rn_lo, rn_hi;
                tile_rnorm_range<BF>(a.rnorm, row0, nvalid, lane, rn_lo, rn_hi);  // before the wait: overlaps the MMAs
                mbar_wait(tmem_full + as, aph);
                tc_fence_after();
                sample_tile<BF>(tacc0 + as * MB * 256u, nvalid, rn_lo, rn_hi, t0, t1, t2, t3);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(tmem_empty + as);
            }
            if (live) {
                float* smp = reinterpret_cast<float*>(a.glists) + (static_cast<size_t>(q) * gridDim.x + blockIdx.x) * 4;
                smp[0] = t0, smp[1] = t1, smp[2] = t2, smp[3] = t3;
            }
        } else {
            RAG_PROF(long long t_wait = 0, t_work = 0;)  // cycles spent waiting for / draining accumulators
            for (int64_t it = 0; it < my_tiles; ++it) {
                const uint32_t as = static_cast<uint32_t>(it % AS), aph = static_cast<uint32_t>((it / AS) & 1);
                RAG_PROF(const long long c0k = clock64();)
                const int64_t row0 = ((cluster_id + it * n_clusters) * kCluster + crank) * a.tile_stride * kBN;
                const int nvalid = static_cast<int>(min(static_cast<int64_t>(kBN), a.n_rows - row0));
                float rn_lo, rn_max;
                tile_rnorm_range<BF>(a.rnorm, row0, nvalid, lane, rn_lo, rn_max);
                mbar_wait(tmem_full + as, aph);
                tc_fence_after();
                RAG_PROF(const long long c1k = clock64(); t_wait += c1k - c0k;)
                scan_tile<BF>(tacc0 + as * MB * 256u, row0, nvalid, rn_max, a.rnorm, live && !(a.debug & 1), S, minpos, thr, rows, NQ, q);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(tmem_empty + as);
                RAG_PROF(t_work += clock64() - c1k;)
            }
            RAG_PROF(if ((a.debug & 8) && blockIdx.x == 0 && lane == 0 && my_tiles > 0)
                printf("[batch dbg] warp %d: tiles %lld, per tile: wait %lld cyc, drain+scan %lld cyc\n", warp, (long long)my_tiles,
                       t_wait / my_tiles, t_work / my_tiles);)
            publish_lists(smem, S, rows, NQ, q, 0, a.nq, a.glists, static_cast<int>(gridDim.x), static_cast<int>(blockIdx.x), lane);
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();  // no CTA leaves while a peer may still multicast into it
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------- the CTA-pair (cta_group::2) kernel
// For 129..256 queries.  Two CTAs (one TPC pair) form one MMA: M = 256 queries (each CTA keeps ITS 128
// queries as its half of A), N = 256 corpus rows (each CTA loads 128 of them as its half of B; the
// tensor cores read both halves).  Per ring stage a CTA ingests 16 KB of queries + 16 KB of corpus for
// 256 rows of the pair, i.e. half the L2->SM bytes per corpus row of the single-CTA kernel, and its
// TMEM holds 128 queries x 256 rows = 256 columns, so the accumulator is double buffered and the
// epilogue of tile t overlaps the MMAs of tile t+1.
// Barriers: both CTAs' TMA loads complete on the LEADER's full barrier (armed by the leader alone); the leader's MMA lane commits
// (multicast) to both CTAs' empty / tmem_full barriers; both epilogues arrive on the leader's tmem_empty.
constexpr int kPairThreads = 256;   // warp0 TMA, warp1 MMA (leader), warp2 TMEM, warp3 idle, warps 4-7 epilogue
constexpr uint32_t kIdescTf32Pair = (1u << 4) | (2u << 7) | (2u << 10) | ((256u >> 3) << 17) | ((256u >> 4) << 24);

__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    // executed by both CTAs of the pair; the peer bit of the barrier address is cleared so that the
    // transaction bytes are counted on the leader's (even rank) barrier
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::
            "r"(smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1)
        : "memory");
}
// Arrive on the barrier at this offset in CTA `target_rank` of the cluster.  Relaxed: a cluster-scope release costs the
// arriving thread ~1500 cycles (measured), and the hand-off it is used for carries no memory payload -- the epilogue
// giving a drained TMEM accumulator back, after tcgen05.wait::ld + fence::before_thread_sync.
__device__ __forceinline__ void mbar_arrive_cluster_relaxed(uint64_t* local_bar, uint32_t target_rank) {
    uint32_t raddr;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(raddr) : "r"(smem_u32(local_bar)), "r"(target_rank));
    asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(raddr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    }
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* smem_dst, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_commit_2sm_mc(uint64_t* bar, uint16_t cta_mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     smem_u32(bar)),
                 "h"(cta_mask)
                 : "memory");
}
__device__ __forceinline__ void mma_tf32_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ void mma_bf16_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
constexpr uint32_t kIdescBf16Pair = (1u << 4) | (1u << 7) | (1u << 10) | ((256u >> 3) << 17) | ((256u >> 4) << 24);

template <bool BF>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPairThreads, 1)
batch_score_kernel_2sm(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_c,
                       const BatchArgs a, const BatchLayout L) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_stages = static_cast<int>(L.n_stages);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + L.off_bars);  // used in the leader only
    uint64_t* empty = full + n_stages;
    uint64_t* tmem_full = empty + n_stages;   // [2]
    uint64_t* tmem_empty = tmem_full + 2;     // [2], used in the leader only
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(smem + L.off_tmem);
    uint32_t* rows = reinterpret_cast<uint32_t*>(smem + L.off_lists);  // [slot][local query]
    constexpr int NQL = kBM;  // queries handled by this CTA

    const uint32_t crank = cluster_ctarank();
    const bool leader = crank == 0;
    constexpr int KE = BF ? 64 : 32;
    const int64_t all_ptiles = (a.n_rows + kBN - 1) / kBN;  // pair tiles of 256 rows
    const int64_t n_ptiles = (all_ptiles + a.tile_stride - 1) / a.tile_stride;  // tiles this launch visits
    const int64_t n_pairs = gridDim.x / 2, pair_id = blockIdx.x / 2;
    const int64_t my_tiles = n_ptiles > pair_id ? (n_ptiles - pair_id + n_pairs - 1) / n_pairs : 0;

    if (tid == 0) {
        for (int s = 0; s < n_stages; ++s) {
            mbar_init(full + s, 1);   // the leader's expect_tx arrival (covers the bytes of both CTAs)
            mbar_init(empty + s, 1);  // the leader's multicast commit
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(tmem_full + s, 1);   // the leader's multicast commit
            mbar_init(tmem_empty + s, 8);  // four epilogue warps in each CTA
        }
        fence_barrier_init();
    }
    for (int i = tid; i < kBatchKp * NQL; i += kPairThreads) rows[i] = 0xFFFFFFFFu;
    __syncthreads();
    cluster_sync_all();
    if (warp == 2) tmem_alloc_2sm(tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ------------------------------------------------ TMA producer (both CTAs)
        if (lane == 0) {
            tma_prefetch_desc(&tmap_q);
            tma_prefetch_desc(&tmap_c);
            int s = 0;
            uint32_t ph = 0;
            for (int64_t it = 0; it < my_tiles; ++it) {
                const int row0 = static_cast<int>((pair_id + it * n_pairs) * a.tile_stride * kBN + crank * 128);
                for (int kb = 0; kb < a.nkb; ++kb) {
                    mbar_wait(empty + s, ph ^ 1u);
                    uint8_t* st = smem + static_cast<size_t>(s) * L.stage_bytes;
                    // Only the leader arms the barrier, with the bytes of both CTAs.  The peer's copies may land first:
                    // they can only be issued after the leader's MMA lane released this stage, i.e. after the barrier's
                    // previous phase completed, and the phase cannot complete before the leader's own arrival.  (A
                    // per-stage remote arrive with cluster-scope release semantics cost the peer ~1500 cycles per stage.)
                    if (leader) mbar_arrive_expect_tx(full + s, 4u * kOperandBytes);
                    tma_load_2d_2sm(st, &tmap_q, full + s, kb * KE, static_cast<int>(crank) * kBM);
                    tma_load_2d_2sm(st + kOperandBytes, &tmap_c, full + s, kb * KE, row0);
                    if (++s == n_stages) {
                        s = 0;
                        ph ^= 1u;
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ------------------------------------------------ MMA issuer: one lane of the leader CTA
        if (leader && lane == 0) {
            int s = 0;
            uint32_t ph = 0;
            for (int64_t it = 0; it < my_tiles; ++it) {
                const uint32_t as = static_cast<uint32_t>(it & 1), aph = static_cast<uint32_t>((it >> 1) & 1);
                mbar_wait_cluster(tmem_empty + as, aph ^ 1u);  // both epilogues have drained this accumulator
                tc_fence_after();
                for (int kb = 0; kb < a.nkb; ++kb) {
                    mbar_wait_cluster(full + s, ph);
                    tc_fence_after();
                    const uint32_t st = smem_u32(smem + static_cast<size_t>(s) * L.stage_bytes);
#pragma unroll
                    for (int k4 = 0; k4 < 4; ++k4) {
                        const uint64_t adesc = make_sw128_desc(st + k4 * 32);
                        const uint64_t bdesc = make_sw128_desc(st + kOperandBytes + k4 * 32);
                        if (!(a.debug & 2)) {
                            if (BF)
                                mma_bf16_2sm(tmem_base + as * 256u, adesc, bdesc, kIdescBf16Pair, (kb | k4) != 0);
                            else
                                mma_tf32_2sm(tmem_base + as * 256u, adesc, bdesc, kIdescTf32Pair, (kb | k4) != 0);
                        }
                    }
                    tc_commit_2sm_mc(empty + s, 0x3);  // both CTAs may refill the stage once these MMAs are done
                    if (++s == n_stages) {
                        s = 0;
                        ph ^= 1u;
                    }
                }
                tc_commit_2sm_mc(tmem_full + as, 0x3);  // accumulator complete, in both CTAs
            }
        }
        __syncwarp();
    } else if (warp >= kEpiWarp0) {
        // ------------------------------------------------ epilogue: thread <-> one of this CTA's 128 queries
        const int ql = (warp - kEpiWarp0) * 32 + lane;
        const int q = static_cast<int>(crank) * kBM + ql;
        const bool live = q < a.nq;
        const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
        float S[kBatchKp];
#pragma unroll
        for (int p = 0; p < kBatchKp; ++p) S[p] = -CUDART_INF_F;
        float thr = -CUDART_INF_F;
        int minpos = 0;
        const float floor0 = (a.thr0 && live) ? nextafterf(a.thr0[q], -CUDART_INF_F) : -CUDART_INF_F;  // see batch_score_kernel
        if (floor0 > -CUDART_INF_F) {
#pragma unroll
            for (int p = 0; p < kBatchKp; ++p) S[p] = floor0;
            thr = floor0;
        }
        float t0 = -CUDART_INF_F, t1 = -CUDART_INF_F, t2 = -CUDART_INF_F, t3 = -CUDART_INF_F;  // sampling pass: 4 best
        RAG_PROF(long long t_wait = 0, t_work = 0;)
        for (int64_t it = 0; it < my_tiles; ++it) {
            const uint32_t as = static_cast<uint32_t>(it & 1), aph = static_cast<uint32_t>((it >> 1) & 1);
            RAG_PROF(const long long c0k = clock64();)
            const int64_t row0 = (pair_id + it * n_pairs) * a.tile_stride * kBN;
            const int nvalid = static_cast<int>(min(static_cast<int64_t>(kBN), a.n_rows - row0));
            float rn_lo, rn_hi;
            tile_rnorm_range<BF>(a.rnorm, row0, nvalid, lane, rn_lo, rn_hi);
            mbar_wait(tmem_full + as, aph);
            tc_fence_after();
            RAG_PROF(const long long c1k = clock64(); t_wait += c1k - c0k;)
            const uint32_t tacc = tmem_base + lane_base + as * 256u;
            if (a.sample)
                sample_tile<BF>(tacc, nvalid, rn_lo, rn_hi, t0, t1, t2, t3);
            else
                scan_tile<BF>(tacc, row0, nvalid, rn_hi, a.rnorm, live && !(a.debug & 1), S, minpos, thr, rows, NQL, ql);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster_relaxed(tmem_empty + as, 0);
            RAG_PROF(t_work += clock64() - c1k;)
        }
        RAG_PROF(if ((a.debug & 8) && blockIdx.x == 0 && lane == 0 && my_tiles > 0)
            printf("[batch dbg pair] warp %d: tiles %lld, per tile: wait %lld cyc, drain+scan %lld cyc\n", warp, (long long)my_tiles,
                   t_wait / my_tiles, t_work / my_tiles);)
        if (a.sample) {
            if (live) {
                float* smp = reinterpret_cast<float*>(a.glists) + (static_cast<size_t>(q) * n_pairs + pair_id) * 4;
                smp[0] = t0, smp[1] = t1, smp[2] = t2, smp[3] = t3;
            }
        } else {
            publish_lists(smem, S, rows, NQL, ql, static_cast<int>(crank) * kBM, a.nq, a.glists, static_cast<int>(n_pairs),
                          static_cast<int>(pair_id), lane);
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc_2sm(tmem_base, 512);
    }
}

// ---------------------------------------------------------------- sampling-pass thresholds
// One CTA per query: the sampling pass left 4 best scores per CTA; the k'-th largest of these n values is the score of
// a real row that at least k' real rows reach, i.e. a lower bound of the k'-th best score over the whole corpus.
__global__ void __launch_bounds__(256) batch_thresholds_kernel(const float* smp, int n, int kth, float* thr0) {
    extern __shared__ float vals[];
    const int q = blockIdx.x;
    for (int i = threadIdx.x; i < n; i += blockDim.x) vals[i] = smp[static_cast<size_t>(q) * n + i];
    if (threadIdx.x == 0) thr0[q] = -CUDART_INF_F;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float x = vals[i];
        if (!(x > -CUDART_INF_F)) continue;
        int rank = 0;
        for (int u = 0; u < n; ++u) rank += (vals[u] > x) || (vals[u] == x && u < i);
        if (rank == kth - 1) thr0[q] = x;
    }
}

// ---------------------------------------------------------------- per-query merge + exact rescoring
struct BatchFinalizeArgs {
    ScanArgs scan;          // store views, k_take, kp, eps, row_base; d_query / out_* are per-launch bases
    const float* d_queries; // raw queries [nq][dim]
    const uint64_t* glists; // [nq][nlists][kp]
    const float* thr0;      // [nq] sampling-pass thresholds (null: single pass)
    int nlists;             // lists per query, kBatchKp keys each; merged into scan.kp candidates
    int out_stride;         // hits per query in out_idx / out_score
    uint32_t ring_bytes, row_bytes, qpad;
};

template <typename T>
__global__ void __launch_bounds__(kScanThreads, 1) batch_finalize_kernel(const BatchFinalizeArgs b) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int q = blockIdx.x, tid = threadIdx.x;
    uint8_t* ring = smem;
    float* qraw = reinterpret_cast<float*>(smem + b.ring_bytes);
    uint64_t* lists = reinterpret_cast<uint64_t*>(smem + b.ring_bytes + b.qpad * 4u);
    ScanMisc* misc = reinterpret_cast<ScanMisc*>(reinterpret_cast<uint8_t*>(lists) + sizeof(uint64_t) * 2u * kConsumerWarps * b.scan.kp);
    ScanArgs a = b.scan;
    a.d_query = b.d_queries + static_cast<size_t>(q) * a.dim;
    a.out_idx = a.out_idx ? a.out_idx + static_cast<size_t>(q) * b.out_stride : nullptr;
    a.out_score = a.out_score ? a.out_score + static_cast<size_t>(q) * b.out_stride : nullptr;
    a.out_found = a.out_found ? a.out_found + q : nullptr;
    a.out_flags = a.out_flags ? a.out_flags + q : nullptr;
    a.out_cand = a.out_cand ? a.out_cand + static_cast<size_t>(q) * a.cand_slots : nullptr;
    a.scratch.stamps = nullptr;
    for (int i = tid; i < static_cast<int>(b.qpad); i += kScanThreads) qraw[i] = i < a.dim ? a.d_query[i] : 0.f;
    if (tid == 0) {
        mbar_init(reinterpret_cast<uint64_t*>(&misc->stage_bar), 1);
        fence_barrier_init();
        misc->inexact = 0;
    }
    __syncthreads();
    FinalizeCtx f;
    f.ring = ring;
    f.ring_bytes = b.ring_bytes;
    f.row_bytes = b.row_bytes;
    f.lists = lists;
    f.misc = misc;
    f.qraw = qraw;
    f.glists = b.glists + static_cast<size_t>(q) * b.nlists * kBatchKp;
    f.nlists = b.nlists;
    f.list_len = kBatchKp;
    f.bound_floor = b.thr0 ? b.thr0[q] : -CUDART_INF_F;
    finalize_candidates<T>(a, f);
}

// normalised, zero-padded copy of the queries: the tensor-core operand (rows >= nq are zero)
// (T = float for fp32 stores; T = uint16_t: rounded to bf16, nearest-even, for bf16 stores)
template <typename T>
__global__ void __launch_bounds__(256) prep_queries_kernel(const float* __restrict__ q, T* __restrict__ qhat, int nq,
                                                           int nq_pad, int dim, int ld) {
    const int lane = threadIdx.x & 31;
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= nq_pad) return;
    T* dst = qhat + static_cast<size_t>(row) * ld;
    if (row >= nq) {
        for (int i = lane; i < ld; i += 32) dst[i] = T(0);
        return;
    }
    const float* src = q + static_cast<size_t>(row) * dim;
    double acc = 0.0;
    for (int i = lane; i < dim; i += 32) acc += static_cast<double>(src[i]) * static_cast<double>(src[i]);
    acc = warp_sum(acc);
    const float r = acc > 0.0 ? static_cast<float>(1.0 / sqrt(acc)) : 0.f;
    for (int i = lane; i < ld; i += 32) {
        const float x = i < dim ? src[i] * r : 0.f;
        if (sizeof(T) == 4)
            dst[i] = static_cast<T>(x);
        else
            dst[i] = static_cast<T>(f32_to_bf16_bits(x));
    }
}

// ---------------------------------------------------------------- host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled encode_fn() {
    static PFN_encodeTiled fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        cudaGetLastError();
        return reinterpret_cast<PFN_encodeTiled>(p);
    }();
    return fn;
}

// row-major [rows][ld] matrix of fp32 or bf16; a box is 128 bytes of K x box_rows rows
static bool make_map(CUtensorMap* m, const void* base, uint64_t rows, uint64_t ld, uint32_t box_rows, bool bf16) {
    PFN_encodeTiled enc = encode_fn();
    if (!enc) return false;
    const uint32_t esz = bf16 ? 2 : 4;
    const cuuint64_t dims[2] = {ld, rows};
    const cuuint64_t strides[1] = {ld * esz};
    const cuuint32_t box[2] = {128u / esz, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return enc(m, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Shared-memory opt-in of a kernel: per function and per device, so done once for each pair instead of per launch.
template <typename K>
static cudaError_t opt_in_smem(K kernel) {
    static std::mutex mu;
    static std::set<std::pair<const void*, int>> done;
    int dev = 0;
    cudaGetDevice(&dev);
    const std::pair<const void*, int> key(reinterpret_cast<const void*>(kernel), dev);
    std::lock_guard<std::mutex> lk(mu);
    if (done.count(key)) return cudaSuccess;
    const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e == cudaSuccess) done.insert(key);
    return e;
}

// Experiment knobs (environment), read ONCE -- five getenv calls per launch were measurable on the latency path.
// ragstore_debug_reload_env() re-reads them (tests switch kernel variants at run time).
struct BatchKnobs {
    bool no_pair = false, no_twopass = false;
    int stages = 0, debug = 0, stride_max = 32;
    void load() {
        const char* e;
        no_pair = (e = std::getenv("RAGSTORE_BATCH_2SM")) && e[0] == '0';
        no_twopass = (e = std::getenv("RAGSTORE_BATCH_TWOPASS")) && e[0] == '0';
        stages = (e = std::getenv("RAGSTORE_BATCH_STAGES")) ? atoi(e) : 0;
        debug = (e = std::getenv("RAGSTORE_BATCH_DEBUG")) ? atoi(e) : 0;
        const int v = (e = std::getenv("RAGSTORE_BATCH_STRIDE_MAX")) ? atoi(e) : 0;
        stride_max = v >= 1 ? v : 32;
    }
};
static BatchKnobs& knobs() {
    static BatchKnobs k = [] {
        BatchKnobs b;
        b.load();
        return b;
    }();
    return k;
}
void batch_reload_env() { knobs().load(); }

size_t batch_scratch_bytes(int sm_count) {
    return sizeof(uint64_t) * size_t(kBatchMaxQ) * size_t(sm_count) * kBatchKp   // candidate lists
           + sizeof(float) * size_t(kBatchMaxQ) * size_t(kBatchMaxLd)           // normalised query operand
           + sizeof(float) * size_t(kBatchMaxQ);                                 // sampling-pass thresholds
}

int launch_batch_search(const BatchSearchArgs& b, cudaStream_t stream) {
    const ScanArgs& s = b.scan;
    const bool bf = s.dtype == RAGSTORE_BF16 || b.operand_bf16;     // element type of the tensor-core operands
    const bool bf_rows = s.dtype == RAGSTORE_BF16;                   // element type of the rows the finalize step rescoring reads
    const void* op_mat = b.operand_mat ? b.operand_mat : s.mat;
    if ((bf && s.ld % 8 != 0) || b.nq < 1 || b.nq > kBatchMaxQ || s.ld > kBatchMaxLd || s.kp < kBatchKp || s.kp > kMaxKp ||
        (s.kp & (s.kp - 1)) || s.n_rows < 1 || s.n_rows >= 0x7FFF0000ll)
        return static_cast<int>(cudaErrorInvalidValue);
    {   // every TMA load of this path reads memory of the GPU the kernel runs on: the operands are the store's own
        // matrix (or shadow) and its own scratch.  Refuse anything else outright.
        int dev = -1;
        cudaGetDevice(&dev);
        cudaPointerAttributes pa{};
        if (cudaPointerGetAttributes(&pa, b.scratch) != cudaSuccess || pa.type != cudaMemoryTypeDevice || pa.device != dev) {
            cudaGetLastError();
            return static_cast<int>(cudaErrorInvalidDevicePointer);
        }
        if (cudaPointerGetAttributes(&pa, op_mat) != cudaSuccess || pa.type != cudaMemoryTypeDevice || pa.device != dev) {
            cudaGetLastError();
            return static_cast<int>(cudaErrorInvalidDevicePointer);
        }
    }
    const int mblocks = b.nq > kBM ? 2 : 1;
    const int nq_pad = mblocks * kBM;
    uint64_t* glists = reinterpret_cast<uint64_t*>(b.scratch);
    float* qhat = reinterpret_cast<float*>(b.scratch + sizeof(uint64_t) * size_t(kBatchMaxQ) * size_t(s.sm_count) * kBatchKp);
    float* thr0 = qhat + size_t(kBatchMaxQ) * size_t(kBatchMaxLd);

    if (bf)
        prep_queries_kernel<uint16_t><<<(nq_pad * 32 + 255) / 256, 256, 0, stream>>>(b.d_queries, reinterpret_cast<uint16_t*>(qhat), b.nq, nq_pad, s.dim, s.ld);
    else
        prep_queries_kernel<float><<<(nq_pad * 32 + 255) / 256, 256, 0, stream>>>(b.d_queries, qhat, b.nq, nq_pad, s.dim, s.ld);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return static_cast<int>(e);

    CUtensorMap map_q, map_c;
    const BatchKnobs& kn = knobs();
    const bool pair_maps = mblocks == 2 && !kn.no_pair;
    if (!make_map(&map_q, qhat, static_cast<uint64_t>(nq_pad), static_cast<uint64_t>(s.ld), pair_maps ? kBM : nq_pad / kCluster, bf) ||
        !make_map(&map_c, op_mat, static_cast<uint64_t>(s.n_rows), static_cast<uint64_t>(s.ld), pair_maps ? 128 : kBN, bf))
        return static_cast<int>(cudaErrorNotSupported);

    // 129..256 queries: the CTA-pair (cta_group::2) kernel by default -- half the shared-memory fill per corpus row and a
    // double-buffered accumulator; measured equal to 4 % faster than the single-CTA kernel on fp32 stores (both sit at
    // the power-capped tensor rate) and 4-15 % faster on bf16 stores and small corpora.  RAGSTORE_BATCH_2SM=0 opts out.
    const bool pair = pair_maps;
    BatchLayout L;
    L.mblocks = mblocks;
    L.stage_bytes = pair ? 2 * kOperandBytes : (mblocks + 2) * kOperandBytes;
    const uint32_t lists_bytes = sizeof(uint32_t) * kBatchKp * (pair ? kBM : nq_pad);   // row ids of the candidate lists
    const uint32_t keys_bytes = sizeof(uint64_t) * kBatchKp * (pair ? kBM : nq_pad);    // publish staging (reuses the ring)
    const uint32_t budget = 227u * 1024u - 1024u /*alignment slack*/ - lists_bytes - 512u;
    L.n_stages = budget / L.stage_bytes;
    if (L.n_stages > 8) L.n_stages = 8;
    if (kn.stages >= 2 && static_cast<uint32_t>(kn.stages) < L.n_stages) L.n_stages = static_cast<uint32_t>(kn.stages);   // experiments only
    if (L.n_stages < 2 || L.n_stages * L.stage_bytes < keys_bytes) return static_cast<int>(cudaErrorInvalidConfiguration);
    L.off_bars = L.n_stages * L.stage_bytes;
    L.off_tmem = L.off_bars + 8u * (2u * L.n_stages + 4u);
    L.off_lists = (L.off_tmem + 16u + 15u) & ~15u;
    L.total = L.off_lists + lists_bytes + 1024u;

    BatchArgs a;
    a.n_rows = s.n_rows;
    a.nq = b.nq;
    a.mblocks = mblocks;
    a.nkb = bf ? (s.ld + 63) / 64 : (s.ld + 31) / 32;
    a.rnorm = s.rnorm;
    a.glists = glists;
    a.debug = kn.debug;
    const int64_t n_tiles = (s.n_rows + kBN - 1) / kBN;
    const int64_t n_units = pair ? 2 * n_tiles : n_tiles;       // CTAs that have work
    int grid = static_cast<int>(n_units < s.sm_count ? n_units : s.sm_count);
    grid = (grid + kCluster - 1) / kCluster * kCluster;          // whole clusters
    if (grid > s.sm_count) grid = s.sm_count / kCluster * kCluster;
    const int nlists = pair ? grid / 2 : grid;
    a.tile_stride = 1;
    a.thr0 = nullptr;
    a.sample = 0;
    const float* thr_used = nullptr;
    auto launch_as = [&](auto kernel) -> cudaError_t {
        cudaError_t e2 = opt_in_smem(kernel);
        if (e2 != cudaSuccess) return e2;
        kernel<<<grid, kBatchThreads, L.total, stream>>>(map_q, map_c, a, L);
        return cudaGetLastError();
    };
    auto launch_pair = [&](auto kernel) -> cudaError_t {
        cudaError_t e2 = opt_in_smem(kernel);
        if (e2 != cudaSuccess) return e2;
        kernel<<<grid, kPairThreads, L.total, stream>>>(map_q, map_c, a, L);
        return cudaGetLastError();
    };
    auto launch_main = [&]() -> cudaError_t {
        if (pair) return bf ? launch_pair(batch_score_kernel_2sm<true>) : launch_pair(batch_score_kernel_2sm<false>);
        if (mblocks == 2) return bf ? launch_as(batch_score_kernel<2, true>) : launch_as(batch_score_kernel<2, false>);
        return bf ? launch_as(batch_score_kernel<1, true>) : launch_as(batch_score_kernel<1, false>);
    };
    {
        // Two passes on large corpora.  The cost of the fused top-k is the list warm-up: each CTA sees only 1/148 of
        // the rows, so its thresholds rise slowly and ~1000 ln(tiles) insert events per warp stall the accumulator
        // drain.  A sampling pass over every s-th tile (same kernel) gives, per query, the k'-th best score of the
        // sample -- a lower bound of the k'-th best overall -- and the full pass starts from it: only ~32*s rows per
        // query in the WHOLE corpus still beat it.
        const bool two_pass = !kn.no_twopass;
        if (two_pass) {
            // small corpora: stride 1, i.e. the "sample" is a full (cheap: 4 registers per thread, no lists) sweep --
            // still far cheaper than filling cold lists column by column (~200 us per CTA for 256 queries)
            int stride = static_cast<int>(n_tiles / (2ll * nlists));   // every CTA (pair) samples >= 2 tiles when it can
            // ~thr_rank * stride rows per query beat the threshold in the whole corpus; a CTA keeps 32 of its share, so the
            // stride is held where that share (thr_rank * stride / nlists) stays around 16 and lists rarely overflow
            const int thr_rank = b.thr_rank > 0 ? std::min(b.thr_rank, s.kp) : s.kp;
            const int stride_lists = std::max(1, 16 * nlists / thr_rank);
            const int stride_max = std::min(kn.stride_max, stride_lists);   // (knob: experiments)
            stride = stride > stride_max ? stride_max : (stride < 1 ? 1 : stride);
            a.tile_stride = stride;
            a.sample = 1;
            e = launch_main();
            if (e != cudaSuccess) return static_cast<int>(e);
            const int nvals = nlists * 4;
            // threshold = the thr_rank-th largest sampled score: at least that many real rows reach it (fewer samples: no threshold)
            batch_thresholds_kernel<<<b.nq, 256, nvals * sizeof(float), stream>>>(reinterpret_cast<const float*>(glists), nvals, thr_rank, thr0);
            e = cudaGetLastError();
            if (e != cudaSuccess) return static_cast<int>(e);
            a.sample = 0;
            a.tile_stride = 1;
            a.thr0 = thr0;
            thr_used = thr0;
        }
        e = launch_main();
    }
    if (e != cudaSuccess) return static_cast<int>(e);

    BatchFinalizeArgs f;
    f.scan = s;
    f.d_queries = b.d_queries;
    f.glists = glists;
    f.thr0 = thr_used;
    f.nlists = nlists;
    f.out_stride = b.out_stride;
    f.row_bytes = static_cast<uint32_t>(s.ld) * (bf_rows ? 2u : 4u);
    f.qpad = static_cast<uint32_t>(s.ld);
    const uint32_t fixed = f.qpad * 4u + sizeof(uint64_t) * 2u * kConsumerWarps * s.kp + sizeof(ScanMisc) + 256u;
    uint32_t ring = 200u * 1024u - fixed;
    const uint32_t need = static_cast<uint32_t>(s.kp) * (f.row_bytes + 16u);  // room for every candidate row at once
    if (ring > need) ring = need;
    const uint32_t lists_need = static_cast<uint32_t>(nlists) * s.kp * 8u;
    if (ring < lists_need && lists_need + fixed <= 200u * 1024u) ring = lists_need;
    f.ring_bytes = (ring + 127u) & ~127u;
    const size_t fin_smem = f.ring_bytes + fixed;
    if (bf_rows) {
        e = opt_in_smem(batch_finalize_kernel<uint16_t>);
        if (e != cudaSuccess) return static_cast<int>(e);
        batch_finalize_kernel<uint16_t><<<b.nq, kScanThreads, fin_smem, stream>>>(f);
    } else {
        e = opt_in_smem(batch_finalize_kernel<float>);
        if (e != cudaSuccess) return static_cast<int>(e);
        batch_finalize_kernel<float><<<b.nq, kScanThreads, fin_smem, stream>>>(f);
    }
    return static_cast<int>(cudaGetLastError());
}

}  // namespace rag
