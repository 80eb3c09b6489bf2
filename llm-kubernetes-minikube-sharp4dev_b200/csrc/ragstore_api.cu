// ragstore_api.cu -- the C ABI of include/ragstore.h: store life cycle, batched ingest, search.
//
// Host-side replacement for the reference's `class RagIndex` internals
// (Minimal_Agent_RAG/Minimal_RAG/Helpers/RagIndex.cs:8,48,59-67): the `List<RagChunk>` becomes a
// device-resident row-major matrix, `Add` becomes a batched upload + normalise kernel, and the
// LINQ search becomes one fused kernel launch per query.  There is no CPU scoring path in this
// library: without a CUDA device every call fails with RAGSTORE_ERR_NO_DEVICE.
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <limits>
#include <mutex>
#include <cmath>
#include <string>
#include <vector>

#include "dealing.h"
#include "exchange_impl.h"
#include "store_impl.h"

using namespace rag;

namespace {

thread_local std::string g_err;

int fail(int code, const char* what, const char* detail = nullptr) {
    g_err = what;
    if (detail) {
        g_err += ": ";
        g_err += detail;
    }
    return code;
}

#define CU_TRY(expr)                                                              \
    do {                                                                          \
        cudaError_t e__ = (expr);                                                 \
        if (e__ != cudaSuccess) return fail(RAGSTORE_ERR_CUDA, #expr, cudaGetErrorString(e__)); \
    } while (0)

}  // namespace

namespace rag {
int choose_kp(int take) {
    // candidate list = requested hits + a safety margin that the certificate can use
    const int want = take + (take / 4 > 8 ? take / 4 : 8);
    int kp = 32;
    while (kp < want && kp < kMaxKp) kp <<= 1;
    return kp;
}

float scan_eps(int dim_ld, int dtype) {
    // |approx - reference| <= (terms per lane + butterfly + normalisation roundings) * 2^-24
    const int per_chunk = dtype == RAGSTORE_F32 ? 4 : 8;
    const int chunks = (dim_ld + per_chunk - 1) / per_chunk;
    const int dv = (chunks + 31) / 32;
    const int terms = dv * per_chunk;
    return static_cast<float>((terms + 16) * 5.9604644775390625e-08);
}

}  // namespace rag

namespace rag {
int api_fail(int code, const char* what, const char* detail) { return fail(code, what, detail); }
}  // namespace rag


namespace {

// host mirror of global_row_id() (topk_device.cuh)
int64_t host_global_row_id(const ragstore* s, int64_t local) {
    return dealt_global_id(local, s->row_block, s->row_nshards, s->row_shard, s->row_base);
}

int alloc_scratch(ScanScratch& s) {
    uint8_t* p = nullptr;
    CU_TRY(cudaMalloc(&p, kScanScratchBytes));
    CU_TRY(cudaMemset(p, 0, kScanScratchBytes));
    s.ticket = reinterpret_cast<unsigned int*>(p);
    s.stamps = reinterpret_cast<unsigned long long*>(p + 64);
    s.cta_lists = reinterpret_cast<uint64_t*>(p + 256);
    return RAGSTORE_OK;
}

void free_scratch(ScanScratch& s) {
    if (s.ticket) cudaFree(s.ticket);
    s = ScanScratch{};
}

int alloc_ctx(SearchCtx& c, int ld, int sm_count) {
    CU_TRY(cudaStreamCreateWithFlags(&c.st, cudaStreamNonBlocking));
    int rc = alloc_scratch(c.scratch);
    if (rc) return rc;
    CU_TRY(cudaMalloc(&c.batch_scratch, batch_scratch_bytes(sm_count)));
    CU_TRY(cudaMalloc(&c.d_pass, sizeof(ragstore_cand) * kMaxPasses * kMaxKp));
    CU_TRY(cudaMalloc(&c.d_bound, 16));
    const size_t nq = kCtxQueries;
    CU_TRY(cudaMalloc(&c.d_q, nq * ld * sizeof(float)));
    // one block per side: [idx][score][found][flags] packed per call, so the hits come back in ONE copy
    CU_TRY(cudaMalloc(&c.d_idx, nq * kMaxKp * 16 + nq * 8));
    CU_TRY(cudaMallocHost(&c.h_q, nq * ld * sizeof(float)));
    CU_TRY(cudaMallocHost(&c.h_idx, nq * kMaxKp * 16 + nq * 8));
    CU_TRY(cudaMallocHost(&c.h_done, 64));
    *c.h_done = 0;
    return RAGSTORE_OK;
}

void free_ctx(SearchCtx& c) {
    if (c.st) cudaStreamDestroy(c.st);
    free_scratch(c.scratch);
    cudaFree(c.batch_scratch);
    cudaFree(c.d_pass);
    cudaFree(c.d_bound);
    cudaFree(c.d_q);
    cudaFree(c.d_idx);
    cudaFreeHost(c.h_q);
    cudaFreeHost(c.h_idx);
    cudaFreeHost(c.h_done);
    c = SearchCtx{};
}

}  // namespace

namespace rag {

ScanArgs base_args(const ragstore* s, int64_t n, int take, int kp) {
    ScanArgs a{};
    a.mat = s->mat;
    a.rnorm = s->rnorm;
    // the per-row lookups are skipped altogether while every row is stored exactly (the common case)
    a.exact = s->inexact_rows > 0 ? s->exact : nullptr;
    a.raw_slot = s->raw_used > 0 ? s->raw_slot : nullptr;
    a.raw_rows = s->raw_rows;
    a.raw_ld = s->raw_ld;
    a.min_norm_bits = s->d_min_norm;
    a.n_rows = n;
    a.dim = s->dim;
    a.ld = s->ld;
    a.dtype = s->dtype;
    a.row_base = s->row_base;
    a.row_block = s->row_block;
    a.row_nshards = s->row_nshards;
    a.row_shard = s->row_shard;
    a.k_take = take;
    a.kp = kp;
    a.eps = scan_eps(s->ld, s->dtype);
    // bf16 store whose exact scores come from fp32 originals (sidecar): the scanned image of a row differs from the
    // row that is rescored by the bf16 rounding, <= 2^-9 per element -> <= 2^-8 on a cosine (Cauchy-Schwarz + norm)
    if (s->dtype == RAGSTORE_BF16 && s->raw_used > 0) a.eps += 3.90625e-3f;
    a.sm_count = s->sm_count;
    return a;
}

// Single queries on a store with a bf16 shadow: the scan streams the shadow (d * 2 bytes per row instead of d * 4) with the
// fp32 query, so only the rows' rounding to bf16 separates its scores from the fp32 scan's: <= 2^-9 relative per element ->
// <= 2^-9 on a cosine of unit vectors (Cauchy-Schwarz), on top of the bf16 scan's own accumulation bound.  Candidates are
// rescored from the fp32 rows, so hits and scores are those of the fp32 scan whenever the certificate holds; it uses
// candidate lists of at least 64 keys, and callers redo what it cannot certify with the fp32 scan.
int shadow_scan_kp(int take) { return std::max(64, choose_kp(take)); }

// Worth it only when the bytes saved outweigh the longer tail (64-key lists, fp32 rows staged for rescoring): from ~64 MB of
// image up (10k x 768 answers in 44 us from the fp32 rows and in 53 us from the image; 1M x 768 in 0.49 against 0.26 ms).
bool shadow_scan_pays(const ragstore* s, int64_t n) {
    return s->shadow != nullptr && s->dtype == RAGSTORE_F32 && static_cast<double>(n) * s->ld * 2.0 >= 64e6;
}

ScanArgs shadow_scan_args(const ragstore* s, int64_t n, int take, int kp) {
    ScanArgs a = base_args(s, n, take, kp);
    a.mat = s->shadow;
    a.dtype = RAGSTORE_BF16;
    a.rescore_rows = static_cast<const float*>(s->mat);
    a.rescore_ld = s->ld;
    a.eps = scan_eps(s->ld, RAGSTORE_BF16) + 1.953125e-3f + 2e-5f;
    return a;
}

// fp32 store: tf32 products (10-bit mantissas, both operands) + fp32 accumulation vs the reference's score.
// bf16 store: the stored rows enter exactly and the row norm is applied, so only the query's rounding to bf16
// (<= 2^-9 relative per element -> <= 2^-9 on a cosine, Cauchy-Schwarz) and the fp32 accumulation remain; when
// exact scores come from fp32 originals (sidecar) the stored image itself is another 2^-8 away, as in base_args.
float batch_eps(const ragstore* s) {
    // fp32 store nominated from its bf16 shadow: both operands rounded to bf16 (2^-9 relative each, Cauchy-Schwarz on
    // unit vectors -> 2^-8), fp32 accumulation, and the 1e-7 between a sidecar original and its normalised image
    if (s->dtype == RAGSTORE_F32 && s->shadow) return 3.92e-3f + static_cast<float>(s->ld) * 2.4e-7f + 2e-5f;
    if (s->dtype == RAGSTORE_F32) return 2.1e-3f;
    float e = 1.953125e-3f + static_cast<float>(s->ld) * 2.4e-7f + 2e-5f;
    if (s->raw_used > 0) e += 3.90625e-3f;
    return e;
}

// nq scans (one corpus sweep each + ~28 us of fixed cost) against one tensor-core batch (~62 us of fixed cost in its
// five launches, one sweep, and the threshold pass over 1/stride of the corpus -- a whole second sweep on corpora too
// small to sample).  Measured on B200 (tools/probe.py few): 10k x 768: scans 29 us per query, batch 67-77 us for
// 2..15 queries; 4M x 768: 15 queries 27.2 ms as scans, 1.91 ms as one batch.  The batch wins from 2 queries on
// corpora of ~1 GB up and from 3 on small ones.  RAGSTORE_BATCH_MINQ=<n> replaces the model by "at least n queries".
bool batch_eligible(const ragstore* s, int64_t n, int nq, int kp) {
    if (nq < kBatchMinQ || n < kBatchMinRows || kp < 32 || kp > kMaxKp || s->ld > kBatchMaxLd || (s->dtype != RAGSTORE_F32 && s->ld % 8 != 0))
        return false;
    static const int forced_min = [] {
        const char* e = std::getenv("RAGSTORE_BATCH_MINQ");
        return e ? atoi(e) : 0;
    }();
    if (forced_min > 0) return nq >= forced_min;
    const double bytes = static_cast<double>(n) * s->ld * (s->dtype == RAGSTORE_F32 ? 4.0 : 2.0);
    const double t_scan = 28e-6 + bytes / 6.6e12;
    const double tiles_per_cta = static_cast<double>(n) / 256.0 / (s->sm_count > 0 ? s->sm_count : 148);
    const double stride = std::min(16.0, std::max(1.0, std::floor(tiles_per_cta / 2.0)));
    const double t_batch = 62e-6 + (1.0 + 1.0 / stride) * bytes / (nq <= 128 ? 6.2e12 : 4.5e12);
    return nq * t_scan > t_batch;
}

void fill_batch_args(const ragstore* s, int64_t n, int take, int kp, BatchSearchArgs& b) {
    // A bf16-shadow nomination is coarser (eps ~4e-3): it merges the per-CTA lists into at least 128 candidates per query,
    // whose margin (rank k against rank 129 instead of rank 33) the certificate can still prove.
    // Reduced-precision nomination (tf32 / bf16, eps 2e-3 .. 4e-3) needs a wider margin than the fp32 scan: for requests
    // beyond the 32-candidate case the merged set is twice what the scan would keep (k = 100: 256 candidates, so the
    // certificate compares rank 100 with rank 257 instead of rank 129; on 12.5M x 768 bf16 rows that takes the certified
    // share of a 256-query batch from 35 % to ~100 %).
    int kpm = kp >= 64 ? std::min(kMaxKp, 2 * kp) : kp;
    if (s->dtype == RAGSTORE_F32 && s->shadow) kpm = std::max(kpm, 128);
    b.scan = base_args(s, n, take, kpm);
    b.scan.eps = batch_eps(s);
    b.operand_mat = (s->dtype == RAGSTORE_F32 && s->shadow) ? s->shadow : nullptr;
    b.operand_bf16 = b.operand_mat ? 1 : 0;
    b.thr_rank = kp;   // the skip threshold follows what the request needs, not the widened candidate set
}

// Waits until the launch that carries (c.h_done, seq) has made its outputs visible to the host.  A short search (the
// reference's own case is a 30 us kernel) is polled for in page-locked memory: the word arrives when the last CTA is done,
// a few microseconds before the stream reports the kernel retired.  Anything long, or late, takes the ordinary stream wait.
static int wait_done(SearchCtx& c, unsigned int seq, double expected_us) {
    if (expected_us < 150.0) {
        const auto t0 = std::chrono::steady_clock::now();
        volatile unsigned int* w = c.h_done;
        for (int spins = 0;; ++spins) {
            if (*w == seq) return RAGSTORE_OK;
            if ((spins & 255) == 255 && std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(400)) break;
#if defined(__x86_64__)
            __builtin_ia32_pause();
#endif
        }
    }
    CU_TRY(cudaStreamSynchronize(c.st));
    return RAGSTORE_OK;
}

int acquire_ctx(ragstore* s) {
    std::unique_lock<std::mutex> lk(s->ctx_mu);
    s->ctx_cv.wait(lk, [&] { return !s->free_ctx.empty(); });
    const int ci = s->free_ctx.back();
    s->free_ctx.pop_back();
    return ci;
}

void release_ctx(ragstore* s, int ci) {
    {
        std::lock_guard<std::mutex> lk(s->ctx_mu);
        s->free_ctx.push_back(ci);
    }
    s->ctx_cv.notify_one();
}

}  // namespace rag

namespace {

// One tensor-core launch for up to kBatchMaxQ queries (outputs laid out with `stride` hits per query).
int run_batch(const ragstore* s, int64_t n, int take, int kp, const float* d_q, int nq, int64_t* d_idx, double* d_score,
              int32_t* d_found, int32_t* d_flags, int stride, uint8_t* scratch, cudaStream_t st) {
    BatchSearchArgs b{};
    fill_batch_args(s, n, take, kp, b);
    b.scan.out_idx = d_idx;
    b.scan.out_score = d_score;
    b.scan.out_found = d_found;
    b.scan.out_flags = d_flags;
    b.d_queries = d_q;
    b.nq = nq;
    b.out_stride = stride;
    b.scratch = scratch;
    s->n_batch_launches.fetch_add(1, std::memory_order_relaxed);
    return launch_batch_search(b, st);
}

}  // namespace

extern "C" {

int ragstore_version(void) { return 1001; }

int ragstore_debug_reload_env(void) {
    batch_reload_env();
    return RAGSTORE_OK;
}

int ragstore_last_error(char* buf, int buf_len) {
    if (!buf || buf_len <= 0) return 0;
    const int n = static_cast<int>(g_err.size()) < buf_len - 1 ? static_cast<int>(g_err.size()) : buf_len - 1;
    memcpy(buf, g_err.data(), n);
    buf[n] = 0;
    return n;
}

int ragstore_create(int dim, int dtype, int64_t capacity_rows, int device, uint32_t flags, ragstore_t** out) {
    if (!out) return fail(RAGSTORE_ERR_INVALID, "ragstore_create: out is null");
    *out = nullptr;
    if (dim <= 0 || capacity_rows <= 0 || (dtype != RAGSTORE_F32 && dtype != RAGSTORE_BF16))
        return fail(RAGSTORE_ERR_INVALID, "ragstore_create: dim and capacity_rows must be positive, dtype f32|bf16");
    if (capacity_rows >= 0xFFFFFFFFll)
        return fail(RAGSTORE_ERR_UNSUPPORTED, "ragstore_create: at most 2^32-2 rows per store (shard the corpus)");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) {
        cudaGetLastError();
        return fail(RAGSTORE_ERR_NO_DEVICE, "ragstore_create: no usable CUDA device (this library has no CPU path)");
    }
    cudaDeviceProp prop;
    CU_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(RAGSTORE_ERR_NO_DEVICE, "ragstore_create: kernels are built for sm_100a (B200) only");
    DeviceGuard g(device);
    if (!g.ok) return fail(RAGSTORE_ERR_CUDA, "cudaSetDevice failed");

    ragstore* s = new ragstore();
    s->device = device;
    s->dim = dim;
    s->dtype = dtype;
    const int align = dtype == RAGSTORE_F32 ? 4 : 8;  // 16-byte rows
    s->ld = (dim + align - 1) / align * align;
    s->capacity = capacity_rows;
    s->sm_count = prop.multiProcessorCount;
    // probe the scan layout once so that an unsupported dim fails at create, not at first search
    {
        const size_t row_bytes = static_cast<size_t>(s->ld) * s->elem_size();
        if (row_bytes > 24 * 1024) {
            delete s;
            return fail(RAGSTORE_ERR_UNSUPPORTED, "ragstore_create: rows above 24 KB are not covered by the scan kernel");
        }
    }
    auto bail = [&](int rc) {
        ragstore_destroy(s);
        return rc;
    };
    const size_t cap4 = static_cast<size_t>((capacity_rows + 3) & ~3ll) + 64;
    // sidecar of original fp32 rows for inexactly stored ones: on for fp32 stores, off for bf16 stores
    // (lossy by choice) unless the caller says otherwise
    bool sidecar = dtype == RAGSTORE_F32;
    if (flags & RAGSTORE_CREATE_SIDECAR_ON) sidecar = true;
    if (flags & RAGSTORE_CREATE_SIDECAR_OFF) sidecar = false;
    s->raw_ld = (dim + 3) / 4 * 4;
    cudaError_t e;
    int vr;
    if ((vr = s->mat_buf.reserve(static_cast<size_t>(capacity_rows) * s->ld * s->elem_size(), device))) return bail(fail(RAGSTORE_ERR_CUDA, "reserve(matrix)", cudaGetErrorString(static_cast<cudaError_t>(vr))));
    if ((vr = s->rnorm_buf.reserve(cap4 * sizeof(float), device))) return bail(fail(RAGSTORE_ERR_CUDA, "reserve(rnorm)", cudaGetErrorString(static_cast<cudaError_t>(vr))));
    if ((vr = s->exact_buf.reserve(cap4, device))) return bail(fail(RAGSTORE_ERR_CUDA, "reserve(exact)", cudaGetErrorString(static_cast<cudaError_t>(vr))));
    if ((flags & RAGSTORE_CREATE_BF16_SHADOW) && dtype == RAGSTORE_F32 && s->ld % 8 == 0) {
        if ((vr = s->shadow_buf.reserve(static_cast<size_t>(capacity_rows) * s->ld * 2, device))) return bail(fail(RAGSTORE_ERR_CUDA, "reserve(shadow)", cudaGetErrorString(static_cast<cudaError_t>(vr))));
        s->shadow = s->shadow_buf.ptr<uint16_t>();
    }
    s->mat = s->mat_buf.ptr<void>();
    s->rnorm = s->rnorm_buf.ptr<float>();
    s->exact = s->exact_buf.ptr<uint8_t>();
    if (sidecar) {
        if ((vr = s->slot_buf.reserve(cap4 * sizeof(int32_t), device))) return bail(fail(RAGSTORE_ERR_CUDA, "reserve(raw_slot)", cudaGetErrorString(static_cast<cudaError_t>(vr))));
        if ((vr = s->raw_buf.reserve(static_cast<size_t>(capacity_rows) * s->raw_ld * sizeof(float), device))) return bail(fail(RAGSTORE_ERR_CUDA, "reserve(raw_rows)", cudaGetErrorString(static_cast<cudaError_t>(vr))));
        s->raw_slot = s->slot_buf.ptr<int32_t>();
        s->raw_rows = s->raw_buf.ptr<float>();
    }
    if ((e = cudaMalloc(&s->d_inexact, 64)) != cudaSuccess) return bail(fail(RAGSTORE_ERR_CUDA, "cudaMalloc(counters)", cudaGetErrorString(e)));
    if ((e = cudaMemset(s->d_inexact, 0, 64)) != cudaSuccess) return bail(fail(RAGSTORE_ERR_CUDA, "cudaMemset(counters)", cudaGetErrorString(e)));
    s->d_raw_count = reinterpret_cast<unsigned int*>(s->d_inexact + 1);
    s->d_min_norm = reinterpret_cast<uint32_t*>(s->d_inexact + 2);
    {
        const uint32_t inf_bits = 0x7F800000u;
        if ((e = cudaMemcpy(s->d_min_norm, &inf_bits, 4, cudaMemcpyHostToDevice)) != cudaSuccess) return bail(fail(RAGSTORE_ERR_CUDA, "cudaMemcpy(min_norm)", cudaGetErrorString(e)));
    }
    if ((e = cudaStreamCreateWithFlags(&s->ingest_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(fail(RAGSTORE_ERR_CUDA, "cudaStreamCreate", cudaGetErrorString(e)));
    if ((e = cudaStreamCreateWithFlags(&s->dev_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(fail(RAGSTORE_ERR_CUDA, "cudaStreamCreate", cudaGetErrorString(e)));
    int rc = alloc_scratch(s->dev_scratch);
    if (rc) return bail(rc);
    if ((e = cudaMalloc(&s->dev_batch_scratch, batch_scratch_bytes(s->sm_count))) != cudaSuccess) return bail(fail(RAGSTORE_ERR_CUDA, "cudaMalloc(batch scratch)", cudaGetErrorString(e)));
    s->ctxs.resize(kNumCtx);
    for (int i = 0; i < kNumCtx; ++i) {
        rc = alloc_ctx(s->ctxs[i], s->ld, s->sm_count);
        if (rc) return bail(rc);
        s->free_ctx.push_back(i);
    }
    *out = s;
    return RAGSTORE_OK;
}

int ragstore_destroy(ragstore_t* s) {
    if (!s) return RAGSTORE_OK;
    if (s->sharded || !s->shards.empty()) return sharded_destroy(s);
    DeviceGuard g(s->device);
    cudaDeviceSynchronize();
    for (auto& c : s->ctxs) free_ctx(c);
    free_scratch(s->dev_scratch);
    cudaFree(s->dev_batch_scratch);
    if (s->ingest_stream) cudaStreamDestroy(s->ingest_stream);
    if (s->dev_stream) cudaStreamDestroy(s->dev_stream);
    cudaFree(s->d_stage);
    for (int b = 0; b < 2; ++b) {
        if (s->h_stage[b]) cudaFreeHost(s->h_stage[b]);
        if (s->ingest_ev[b]) cudaEventDestroy(s->ingest_ev[b]);
        if (s->k1_ev0[b]) cudaEventDestroy(s->k1_ev0[b]);
        if (s->k1_ev1[b]) cudaEventDestroy(s->k1_ev1[b]);
    }
    if (s->h_counters) cudaFreeHost(s->h_counters);
    cudaFree(s->d_inexact);
    s->shadow_buf.release();
    s->raw_buf.release();
    s->slot_buf.release();
    s->exact_buf.release();
    s->rnorm_buf.release();
    s->mat_buf.release();
    delete s;
    return RAGSTORE_OK;
}

int ragstore_count(ragstore_t* s, int64_t* out_rows) {
    if (!s || !out_rows) return fail(RAGSTORE_ERR_INVALID, "ragstore_count: null argument");
    *out_rows = s->count.load(std::memory_order_acquire);
    return RAGSTORE_OK;
}

int ragstore_set_row_base(ragstore_t* s, int64_t row_base) {
    if (!s || row_base < 0) return fail(RAGSTORE_ERR_INVALID, "ragstore_set_row_base: bad argument");
    if (s && s->sharded) return fail(RAGSTORE_ERR_INVALID, "ragstore_set_row_base: per-device entry point, not available on a multi-device handle (ragstore_create_sharded)");
    s->row_base = row_base;
    return RAGSTORE_OK;
}

int ragstore_set_row_blocking(ragstore_t* s, int64_t block_rows, int n_shards, int shard_index) {
    if (!s || block_rows < 0 || n_shards < 1 || shard_index < 0 || shard_index >= n_shards)
        return fail(RAGSTORE_ERR_INVALID, "ragstore_set_row_blocking: bad argument");
    if (s && s->sharded) return fail(RAGSTORE_ERR_INVALID, "ragstore_set_row_blocking: per-device entry point, not available on a multi-device handle (ragstore_create_sharded)");
    s->row_block = block_rows;
    s->row_nshards = n_shards;
    s->row_shard = shard_index;
    return RAGSTORE_OK;
}

int ragstore_add_many(ragstore_t* s, const float* const* blocks, const int64_t* block_rows, int n_blocks, int row_dim,
                      uint32_t flags, int64_t* out_first_row) {
    if (!s || n_blocks < 0 || (n_blocks > 0 && (!blocks || !block_rows)))
        return fail(RAGSTORE_ERR_INVALID, "ragstore_add_many: null/negative argument");
    for (int b = 0; b < n_blocks; ++b) {
        int64_t first = 0;
        const int rc = ragstore_add(s, blocks[b], block_rows[b], row_dim, flags, &first);   // publishes block b before b + 1 starts
        if (rc) return rc;
        if (b == 0 && out_first_row) *out_first_row = first;
    }
    if (n_blocks == 0 && out_first_row) *out_first_row = s->count.load(std::memory_order_acquire);
    return RAGSTORE_OK;
}

int ragstore_reserve_rows(ragstore_t* s, int64_t n_rows) {
    if (!s || n_rows < 0) return fail(RAGSTORE_ERR_INVALID, "ragstore_reserve_rows: bad argument");
    if (n_rows > s->capacity) return fail(RAGSTORE_ERR_CAPACITY, "ragstore_reserve_rows: above the capacity fixed at create");
    if (s->sharded) return sharded_reserve_rows(s, n_rows);
    std::lock_guard<std::mutex> lk(s->add_mu);
    DeviceGuard g(s->device);
    if (!g.ok) return fail(RAGSTORE_ERR_CUDA, "cudaSetDevice failed");
    const size_t upto = static_cast<size_t>(n_rows);
    const size_t upto4 = ((upto + 3) & ~size_t(3)) + 64;
    int vr = s->mat_buf.ensure(upto * s->ld * s->elem_size(), false, s->ingest_stream);
    if (!vr && s->shadow) vr = s->shadow_buf.ensure(upto * s->ld * 2, false, s->ingest_stream);
    if (!vr) vr = s->rnorm_buf.ensure(upto4 * sizeof(float), true, s->ingest_stream);
    if (!vr) vr = s->exact_buf.ensure(upto4, true, s->ingest_stream);
    if (!vr && s->raw_slot) vr = s->slot_buf.ensure(upto4 * sizeof(int32_t), true, s->ingest_stream);
    if (vr) return fail(RAGSTORE_ERR_CUDA, "ragstore_reserve_rows: mapping device memory", cudaGetErrorString(static_cast<cudaError_t>(vr)));
    CU_TRY(cudaStreamSynchronize(s->ingest_stream));
    return RAGSTORE_OK;
}

int ragstore_add(ragstore_t* s, const float* rows, int64_t n_rows, int row_dim, uint32_t flags, int64_t* out_first_row) {
    if (!s || (!rows && n_rows > 0) || n_rows < 0) return fail(RAGSTORE_ERR_INVALID, "ragstore_add: null/negative argument");
    if (row_dim != s->dim) return fail(RAGSTORE_ERR_INVALID, "ragstore_add: row_dim differs from the store's dim");
    if (s->sharded) return sharded_add(s, rows, n_rows, flags, out_first_row);
    NvtxRange nvtx("ragstore_add: upload + normalise (K1)");
    std::lock_guard<std::mutex> lk(s->add_mu);
    const int64_t first = s->count.load(std::memory_order_acquire);
    if (out_first_row) *out_first_row = first;
    if (n_rows == 0) return RAGSTORE_OK;
    if (first + n_rows > s->capacity) return fail(RAGSTORE_ERR_CAPACITY, "ragstore_add: capacity exceeded");
    DeviceGuard g(s->device);
    if (!g.ok) return fail(RAGSTORE_ERR_CUDA, "cudaSetDevice failed");
    const bool on_device = (flags & RAGSTORE_ADD_DEVICE) != 0;
    const bool in_place = (s->dtype == RAGSTORE_F32);  // fp32: land rows in the matrix, normalise in place
    bool src_pinned = false;  // page-locked caller memory goes over PCIe directly, without the staging copy
    if (!on_device) {
        cudaPointerAttributes pa{};
        if (cudaPointerGetAttributes(&pa, rows) == cudaSuccess) src_pinned = (pa.type == cudaMemoryTypeHost);
        cudaGetLastError();
    }
    const size_t esz = s->elem_size();
    // Two-deep pipeline of ~32 MB chunks: while chunk i is copied over PCIe and normalised, chunk i+1 is being
    // staged into the other pinned buffer (host sources) -- the upload kernel never waits for the CPU copy.
    // host sources: ~32 MB staging chunks; device sources need no staging, so larger chunks (the chunk also bounds
    // how much sidecar memory is mapped ahead of need)
    const int64_t chunk_rows = on_device ? int64_t(1) << 17 : std::max<int64_t>(1, (int64_t(32) << 20) / (int64_t(s->dim) * 4));
    const size_t chunk_floats = static_cast<size_t>(chunk_rows) * s->dim;
    if (!on_device && !src_pinned && !s->h_stage[0]) {
        for (int b = 0; b < 2; ++b) CU_TRY(cudaMallocHost(&s->h_stage[b], chunk_floats * sizeof(float)));
    }
    if (!in_place && !on_device && s->stage_floats < 2 * chunk_floats) {
        cudaFree(s->d_stage);
        s->d_stage = nullptr;
        s->stage_floats = 0;
        CU_TRY(cudaMalloc(&s->d_stage, 2 * chunk_floats * sizeof(float)));
        s->stage_floats = 2 * chunk_floats;
    }
    if (!s->ingest_ev[0]) {
        for (int b = 0; b < 2; ++b) {
            CU_TRY(cudaEventCreateWithFlags(&s->ingest_ev[b], cudaEventDisableTiming));
            CU_TRY(cudaEventCreate(&s->k1_ev0[b]));
            CU_TRY(cudaEventCreate(&s->k1_ev1[b]));
        }
        CU_TRY(cudaMallocHost(&s->h_counters, 2 * 2 * sizeof(unsigned long long)));
    }
    s->last_add_kernel_us = 0.0;
    bool used[2] = {false, false};
    int64_t rows_in_flight[2] = {0, 0};

    auto retire = [&](int b) -> int {  // wait for the chunk that last used buffer b and take over its counters
        if (!used[b]) return RAGSTORE_OK;
        CU_TRY(cudaEventSynchronize(s->ingest_ev[b]));
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, s->k1_ev0[b], s->k1_ev1[b]) == cudaSuccess) s->last_add_kernel_us += 1e3 * ms;
        s->inexact_rows = static_cast<int64_t>(s->h_counters[2 * b]);
        s->raw_used = static_cast<int64_t>(s->h_counters[2 * b + 1] & 0xFFFFFFFFull);
        used[b] = false;
        rows_in_flight[b] = 0;
        return RAGSTORE_OK;
    };

    int64_t chunk = 0;
    for (int64_t r0 = 0; r0 < n_rows; r0 += chunk_rows, ++chunk) {
        const int b = static_cast<int>(chunk & 1);
        const int64_t n = std::min(chunk_rows, n_rows - r0);
        const float* src = rows + r0 * s->dim;
        int rc = retire(b);
        if (rc) return rc;
        {   // map physical memory behind the rows about to land (addresses of existing rows never move)
            const size_t upto = static_cast<size_t>(first + r0 + n);
            const size_t upto4 = ((upto + 3) & ~size_t(3)) + 64;
            int vr = s->mat_buf.ensure(upto * s->ld * esz, false, s->ingest_stream);
            if (!vr && s->shadow) vr = s->shadow_buf.ensure(upto * s->ld * 2, false, s->ingest_stream);
            if (!vr) vr = s->rnorm_buf.ensure(upto4 * sizeof(float), true, s->ingest_stream);
            if (!vr) vr = s->exact_buf.ensure(upto4, true, s->ingest_stream);
            if (!vr && s->raw_slot) vr = s->slot_buf.ensure(upto4 * sizeof(int32_t), true, s->ingest_stream);
            if (!vr && s->raw_slot) {
                // sidecar slots: everything known to be used + every row still in flight + this chunk
                const size_t worst = std::min<size_t>(static_cast<size_t>(s->raw_used + rows_in_flight[b ^ 1] + n),
                                                      static_cast<size_t>(s->capacity));
                vr = s->raw_buf.ensure(worst * s->raw_ld * sizeof(float), false, s->ingest_stream);
            }
            if (vr) return fail(RAGSTORE_ERR_CUDA, "mapping device memory for new rows", cudaGetErrorString(static_cast<cudaError_t>(vr)));
        }
        uint8_t* dst_row = static_cast<uint8_t*>(s->mat) + static_cast<size_t>(first + r0) * s->ld * esz;
        NormaliseArgs a{};
        a.dst = dst_row;
        a.rnorm = s->rnorm + first + r0;
        a.exact = s->exact + first + r0;
        a.raw_slot = s->raw_slot ? s->raw_slot + first + r0 : nullptr;
        a.raw_rows = s->raw_rows;
        a.raw_ld = s->raw_ld;
        a.raw_count = s->d_raw_count;
        a.min_norm_bits = s->d_min_norm;
        a.n_rows = n;
        a.dim = s->dim;
        a.ld = s->ld;
        a.dtype = s->dtype;
        a.prenormalised = (flags & RAGSTORE_ADD_PRENORMALISED) ? 1 : 0;
        a.inexact_counter = s->d_inexact;
        a.shadow = s->shadow ? s->shadow + static_cast<size_t>(first + r0) * s->ld : nullptr;
        const float* h2d_src = src;
        if (!on_device && !src_pinned) {  // stage through pinned memory so that the PCIe copy is truly asynchronous
            memcpy(s->h_stage[b], src, static_cast<size_t>(n) * s->dim * sizeof(float));
            h2d_src = s->h_stage[b];
        }
        if (on_device) {
            // device-resident source: K1 reads it where it lies and writes the normalised rows (one pass, no copy)
            a.src = src;
            a.src_ld = s->dim;
        } else if (in_place) {
            CU_TRY(cudaMemcpy2DAsync(dst_row, static_cast<size_t>(s->ld) * 4, h2d_src, static_cast<size_t>(s->dim) * 4,
                                     static_cast<size_t>(s->dim) * 4, static_cast<size_t>(n),
                                     cudaMemcpyHostToDevice, s->ingest_stream));
            a.src = reinterpret_cast<const float*>(dst_row);
            a.src_ld = s->ld;
        } else {
            float* dstage = s->d_stage + static_cast<size_t>(b) * chunk_floats;
            CU_TRY(cudaMemcpyAsync(dstage, h2d_src, static_cast<size_t>(n) * s->dim * 4, cudaMemcpyHostToDevice, s->ingest_stream));
            a.src = dstage;
            a.src_ld = s->dim;
        }
        CU_TRY(cudaEventRecord(s->k1_ev0[b], s->ingest_stream));
        rc = launch_normalise(a, s->ingest_stream);
        CU_TRY(cudaEventRecord(s->k1_ev1[b], s->ingest_stream));
        if (rc) return fail(RAGSTORE_ERR_CUDA, "normalise kernel launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
        // counters after this chunk ([0] inexact rows, [1] low word: sidecar rows), read back without stalling
        CU_TRY(cudaMemcpyAsync(s->h_counters + 2 * b, s->d_inexact, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s->ingest_stream));
        CU_TRY(cudaEventRecord(s->ingest_ev[b], s->ingest_stream));
        used[b] = true;
        rows_in_flight[b] = n;
    }
    // drain in issue order so that the newest counters win
    const int last = static_cast<int>((chunk - 1) & 1);
    int rc = retire(last ^ 1);
    if (!rc) rc = retire(last);
    if (rc) return rc;
    CU_TRY(cudaStreamSynchronize(s->ingest_stream));
    s->count.store(first + n_rows, std::memory_order_release);  // publish: searches see a prefix
    return RAGSTORE_OK;
}

int ragstore_info(ragstore_t* s, int64_t* out, int n) {
    if (!s || !out) return fail(RAGSTORE_ERR_INVALID, "ragstore_info: null argument");
    if (s->sharded) return sharded_info(s, out, n);
    DeviceGuard g(s->device);
    unsigned long long inexact = 0;
    CU_TRY(cudaMemcpy(&inexact, s->d_inexact, sizeof(inexact), cudaMemcpyDeviceToHost));
    int64_t overlaps = 0;
    if (n > 15) {  // launches that found their scratch in use by another launch (misuse of the device entry points)
        unsigned int w = 0;
        CU_TRY(cudaMemcpy(&w, s->dev_scratch.ticket + 2, sizeof(w), cudaMemcpyDeviceToHost));
        overlaps += w;
        for (auto& c : s->ctxs) {
            if (!c.scratch.ticket) continue;
            CU_TRY(cudaMemcpy(&w, c.scratch.ticket + 2, sizeof(w), cudaMemcpyDeviceToHost));
            overlaps += w;
        }
    }
    const int64_t v[16] = {s->dim, s->ld, s->dtype, s->count.load(), s->capacity, static_cast<int64_t>(inexact), s->device, s->sm_count,
                           s->raw_slot ? 1 : 0, s->raw_used, static_cast<int64_t>(s->mat_buf.mapped()), s->mat_buf.uses_vmm() ? 1 : 0,
                           s->n_scan_launches.load(), s->n_batch_launches.load(), static_cast<int64_t>(s->last_add_kernel_us * 1000.0),
                           overlaps};
    for (int i = 0; i < n && i < 16; ++i) out[i] = v[i];
    return RAGSTORE_OK;
}

int ragstore_debug_stamps(ragstore_t* s, uint64_t* out, int n) {
    if (!s || !out) return fail(RAGSTORE_ERR_INVALID, "ragstore_debug_stamps: null argument");
    if (s && s->sharded) return fail(RAGSTORE_ERR_INVALID, "ragstore_debug_stamps: per-device entry point, not available on a multi-device handle (ragstore_create_sharded)");
    DeviceGuard g(s->device);
    uint64_t v[8] = {0};
    CU_TRY(cudaMemcpy(v, s->dev_scratch.stamps, sizeof(v), cudaMemcpyDeviceToHost));
    for (int i = 0; i < n && i < 8; ++i) out[i] = v[i];
    return RAGSTORE_OK;
}

int ragstore_device_views(ragstore_t* s, void** d_matrix, float** d_rnorm, uint8_t** d_exact) {
    if (!s) return fail(RAGSTORE_ERR_INVALID, "ragstore_device_views: null store");
    if (s && s->sharded) return fail(RAGSTORE_ERR_INVALID, "ragstore_device_views: per-device entry point, not available on a multi-device handle (ragstore_create_sharded)");
    if (d_matrix) *d_matrix = s->mat;
    if (d_rnorm) *d_rnorm = s->rnorm;
    if (d_exact) *d_exact = s->exact;
    return RAGSTORE_OK;
}

int ragstore_search_device(ragstore_t* s, const float* d_queries, int n_queries, int top_k, int64_t* d_idx,
                           double* d_score, int32_t* d_found, int32_t* d_flags, void* stream) {
    if (!s || !d_queries || n_queries < 0 || !d_idx || !d_score) return fail(RAGSTORE_ERR_INVALID, "ragstore_search_device: null argument");
    if (s && s->sharded) return fail(RAGSTORE_ERR_INVALID, "ragstore_search_device: per-device entry point, not available on a multi-device handle (ragstore_create_sharded)");
    NvtxRange nvtx("ragstore_search_device: scan + top-k + rescoring (K2/K3/K4/K5)");
    DeviceGuard g(s->device);
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : s->dev_stream;
    const int64_t n = s->count.load(std::memory_order_acquire);
    const int req = top_k < 1 ? 1 : top_k;
    const int take = static_cast<int>(std::min<int64_t>(n, req));
    if (n == 0) {
        if (d_found) CU_TRY(cudaMemsetAsync(d_found, 0, sizeof(int32_t) * n_queries, st));
        return RAGSTORE_OK;
    }
    if (take > kMaxKp) return fail(RAGSTORE_ERR_UNSUPPORTED, "top_k above 256 is not covered yet");
    const int kp = choose_kp(take);
    if (batch_eligible(s, n, n_queries, kp)) {
        // tensor-core path: the corpus is swept once per 256 queries instead of once per query
        for (int q0 = 0; q0 < n_queries; q0 += kBatchMaxQ) {
            const int nq = std::min(kBatchMaxQ, n_queries - q0);
            const int rc = run_batch(s, n, take, kp, d_queries + static_cast<size_t>(q0) * s->dim, nq,
                                     d_idx + static_cast<size_t>(q0) * req, d_score + static_cast<size_t>(q0) * req,
                                     d_found ? d_found + q0 : nullptr, d_flags ? d_flags + q0 : nullptr, req,
                                     s->dev_batch_scratch, st);
            if (rc) return fail(RAGSTORE_ERR_CUDA, "batched search launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
        }
        return RAGSTORE_OK;
    }
    for (int q = 0; q < n_queries; ++q) {
        ScanArgs a = base_args(s, n, take, kp);
        a.d_query = d_queries + static_cast<size_t>(q) * s->dim;
        a.scratch = s->dev_scratch;
        a.out_idx = d_idx + static_cast<size_t>(q) * req;
        a.out_score = d_score + static_cast<size_t>(q) * req;
        a.out_found = d_found ? d_found + q : nullptr;
        a.out_flags = d_flags ? d_flags + q : nullptr;
        s->n_scan_launches.fetch_add(1, std::memory_order_relaxed);
        const int rc = launch_scan(a, st);
        if (rc) return fail(RAGSTORE_ERR_CUDA, "scan kernel launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
    }
    return RAGSTORE_OK;
}

// one shard's rescored top-k as records; the query comes from device memory (d_query) or, inside the kernel's
// parameter block, from host memory (h_query)
static int search_shard_impl(ragstore_t* s, const float* d_query, const float* h_query, int top_k, ragstore_cand* d_out,
                             int32_t* d_flags, void* stream) {
    if (!s || (!d_query && !h_query) || !d_out) return fail(RAGSTORE_ERR_INVALID, "ragstore_search_shard: null argument");
    if (s && s->sharded) return fail(RAGSTORE_ERR_INVALID, "ragstore_search_shard: per-device entry point, not available on a multi-device handle (ragstore_create_sharded)");
    if (h_query && s->dim > kParamQueryMax)
        return fail(RAGSTORE_ERR_UNSUPPORTED, "ragstore_search_shard_hostq: rows above 1024 elements need the query in device memory");
    DeviceGuard g(s->device);
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : s->dev_stream;
    const int64_t n = s->count.load(std::memory_order_acquire);
    const int req = top_k < 1 ? 1 : top_k;
    if (req > kMaxKp) return fail(RAGSTORE_ERR_UNSUPPORTED, "top_k above 256 is not covered yet");
    const int take = static_cast<int>(std::min<int64_t>(n, req));
    if (n == 0) {
        // an empty shard contributes only padding; 0xFF bytes give id = -1 and a NaN score (ranked last, skipped)
        CU_TRY(cudaMemsetAsync(d_out, 0xFF, sizeof(ragstore_cand) * req, st));
        if (d_flags) {
            const int32_t f = RAGSTORE_HIT_CERTIFIED | RAGSTORE_HIT_EXACT_ROWS;
            CU_TRY(cudaMemcpyAsync(d_flags, &f, sizeof(f), cudaMemcpyHostToDevice, st));
        }
        return RAGSTORE_OK;
    }
    ScanArgs a = base_args(s, n, take, choose_kp(take));
    a.d_query = d_query;
    a.scratch = s->dev_scratch;
    a.out_flags = d_flags;
    a.out_cand = d_out;
    a.cand_slots = req;
    s->n_scan_launches.fetch_add(1, std::memory_order_relaxed);
    const int rc = launch_scan(a, st, h_query);
    if (rc) return fail(RAGSTORE_ERR_CUDA, "scan kernel launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
    return RAGSTORE_OK;
}

int ragstore_search_shard_device(ragstore_t* s, const float* d_query, int top_k, ragstore_cand* d_out,
                                 int32_t* d_flags, void* stream) {
    if (!d_query) return fail(RAGSTORE_ERR_INVALID, "ragstore_search_shard_device: null argument");
    return search_shard_impl(s, d_query, nullptr, top_k, d_out, d_flags, stream);
}

int ragstore_search_shard_hostq(ragstore_t* s, const float* h_query, int top_k, ragstore_cand* d_out,
                                int32_t* d_flags, void* stream) {
    if (!h_query) return fail(RAGSTORE_ERR_INVALID, "ragstore_search_shard_hostq: null argument");
    return search_shard_impl(s, nullptr, h_query, top_k, d_out, d_flags, stream);
}

// Row-sharded search in ONE launch per rank: the shard scan whose last CTA also publishes this shard's hits to every peer,
// waits for theirs and merges (topk_device.cuh, "the exchange step, fused").
int ragstore_search_shard_exchange(ragstore_t* s, ragstore_exchange_t* x, const float* d_query, const float* h_query, int top_k,
                                   ragstore_cand* d_mine, int64_t* d_idx, double* d_score, int32_t* d_found, int32_t* d_flags,
                                   int32_t* d_status, void* stream) {
    if (!s || !x || (!d_query == !h_query) || !d_mine || !d_idx || !d_score || !d_found)
        return fail(RAGSTORE_ERR_INVALID, "ragstore_search_shard_exchange: null argument (give exactly one of d_query / h_query)");
    if (s->sharded) return fail(RAGSTORE_ERR_INVALID, "ragstore_search_shard_exchange: per-device entry point, not available on a multi-device handle");
    if (!x->connected) return fail(RAGSTORE_ERR_INVALID, "ragstore_search_shard_exchange: call ragstore_exchange_connect first");
    if (x->device != s->device) return fail(RAGSTORE_ERR_INVALID, "ragstore_search_shard_exchange: store and exchange live on different devices");
    if (h_query && s->dim > kParamQueryMax)
        return fail(RAGSTORE_ERR_UNSUPPORTED, "ragstore_search_shard_exchange: rows above 1024 elements need the query in device memory");
    const int req = top_k < 1 ? 1 : top_k;
    if (req > kMaxKp) return fail(RAGSTORE_ERR_UNSUPPORTED, "top_k above 256 is not covered by the one-launch form");
    if (req > x->max_records || x->world * req > 4096)
        return fail(RAGSTORE_ERR_INVALID, "ragstore_search_shard_exchange: top_k exceeds what the exchange was created for");
    NvtxRange nvtx("ragstore_search_shard_exchange: shard scan + peer exchange + merge, one launch");
    DeviceGuard g(s->device);
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : s->dev_stream;
    const int64_t n = s->count.load(std::memory_order_acquire);
    if (n == 0) {
        // an empty shard launches no scan: padding records, then the stand-alone exchange kernel (same collective call)
        const float* q = d_query ? d_query : h_query;
        int rc = d_query ? ragstore_search_shard_device(s, q, top_k, d_mine, d_flags, stream) : ragstore_search_shard_hostq(s, q, top_k, d_mine, d_flags, stream);
        if (rc) return rc;
        return ragstore_exchange_merge_flags(x, d_mine, d_flags, 1, req, top_k, d_idx, d_score, d_found, d_flags, d_status, st);
    }
    const int take = static_cast<int>(std::min<int64_t>(n, req));
    ScanArgs a = base_args(s, n, take, choose_kp(take));
    a.d_query = d_query;
    a.scratch = s->dev_scratch;
    a.out_flags = d_flags ? d_flags : reinterpret_cast<int32_t*>(s->dev_scratch.stamps + 7);   // the tail reads this shard's flags back
    a.out_cand = d_mine;
    a.cand_slots = req;
    a.xch = x->d_view;
    a.xseq = ++x->seq;
    a.x_idx = d_idx;
    a.x_score = d_score;
    a.x_found = d_found;
    a.x_flags = d_flags;
    a.x_status = d_status;
    s->n_scan_launches.fetch_add(1, std::memory_order_relaxed);
    const int rc = launch_scan(a, st, h_query);
    if (rc) return fail(RAGSTORE_ERR_CUDA, "scan kernel launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
    return RAGSTORE_OK;
}

int ragstore_search_shard_batch_device(ragstore_t* s, const float* d_queries, int n_queries, int top_k, ragstore_cand* d_out,
                                       int32_t* d_flags, void* stream) {
    if (!s || !d_queries || !d_out || n_queries < 1) return fail(RAGSTORE_ERR_INVALID, "ragstore_search_shard_batch_device: bad argument");
    if (s && s->sharded) return fail(RAGSTORE_ERR_INVALID, "ragstore_search_shard_batch_device: per-device entry point, not available on a multi-device handle (ragstore_create_sharded)");
    DeviceGuard g(s->device);
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : s->dev_stream;
    const int64_t n = s->count.load(std::memory_order_acquire);
    const int req = top_k < 1 ? 1 : top_k;
    if (req > kMaxKp) return fail(RAGSTORE_ERR_UNSUPPORTED, "top_k above 256 is not covered yet");
    const int take = static_cast<int>(std::min<int64_t>(n, req));
    // padding first ({NaN, -1}); shards with rows then overwrite their first `take` records per query
    CU_TRY(cudaMemsetAsync(d_out, 0xFF, sizeof(ragstore_cand) * static_cast<size_t>(req) * n_queries, st));
    if (d_flags) CU_TRY(cudaMemsetAsync(d_flags, 0, sizeof(int32_t) * n_queries, st));
    if (n == 0) return RAGSTORE_OK;
    const int kp = choose_kp(take);
    if (batch_eligible(s, n, n_queries, kp)) {
        for (int q0 = 0; q0 < n_queries; q0 += kBatchMaxQ) {
            const int nq = std::min(kBatchMaxQ, n_queries - q0);
            BatchSearchArgs b{};
            fill_batch_args(s, n, take, kp, b);
            b.scan.out_flags = d_flags ? d_flags + q0 : nullptr;
            b.scan.out_cand = d_out + static_cast<size_t>(q0) * req;
            b.scan.cand_slots = req;
            b.d_queries = d_queries + static_cast<size_t>(q0) * s->dim;
            b.nq = nq;
            b.out_stride = req;
            b.scratch = s->dev_batch_scratch;
            s->n_batch_launches.fetch_add(1, std::memory_order_relaxed);
            const int rc = launch_batch_search(b, st);
            if (rc) return fail(RAGSTORE_ERR_CUDA, "batched search launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
        }
        return RAGSTORE_OK;
    }
    for (int q = 0; q < n_queries; ++q) {
        ScanArgs a = base_args(s, n, take, kp);
        a.d_query = d_queries + static_cast<size_t>(q) * s->dim;
        a.scratch = s->dev_scratch;
        a.out_flags = d_flags ? d_flags + q : nullptr;
        a.out_cand = d_out + static_cast<size_t>(q) * req;
        a.cand_slots = req;
        s->n_scan_launches.fetch_add(1, std::memory_order_relaxed);
        const int rc = launch_scan(a, st);
        if (rc) return fail(RAGSTORE_ERR_CUDA, "scan kernel launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
    }
    return RAGSTORE_OK;
}

int ragstore_merge_shards_batch_device(const ragstore_cand* d_gathered, int n_shards, int n_queries, int per_shard, int top_k,
                                       int64_t* d_idx, double* d_score, int32_t* d_found, int device, void* stream) {
    if (!d_gathered || !d_idx || !d_score || !d_found || n_shards <= 0 || per_shard <= 0 || n_queries <= 0)
        return fail(RAGSTORE_ERR_INVALID, "ragstore_merge_shards_batch_device: bad argument");
    DeviceGuard g(device);
    if (!g.ok) return fail(RAGSTORE_ERR_NO_DEVICE, "ragstore_merge_shards_batch_device: no such device");
    const int req = top_k < 1 ? 1 : top_k;
    const int rc = launch_merge_shards(d_gathered, n_shards, n_queries, per_shard, req, d_idx, d_score, d_found,
                                       static_cast<cudaStream_t>(stream));
    if (rc) return fail(RAGSTORE_ERR_CUDA, "merge kernel launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
    return RAGSTORE_OK;
}

int ragstore_merge_shards_device(const ragstore_cand* d_gathered, int n_shards, int per_shard, int top_k,
                                 int64_t* d_idx, double* d_score, int32_t* d_found, int device, void* stream) {
    if (!d_gathered || !d_idx || !d_score || !d_found || n_shards <= 0 || per_shard <= 0)
        return fail(RAGSTORE_ERR_INVALID, "ragstore_merge_shards_device: bad argument");
    DeviceGuard g(device);
    if (!g.ok) return fail(RAGSTORE_ERR_NO_DEVICE, "ragstore_merge_shards_device: no such device");
    const int req = top_k < 1 ? 1 : top_k;
    const int rc = launch_merge_shards(d_gathered, n_shards, 1, per_shard, req, d_idx, d_score, d_found,
                                       static_cast<cudaStream_t>(stream));
    if (rc) return fail(RAGSTORE_ERR_CUDA, "merge kernel launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
    return RAGSTORE_OK;
}

int ragstore_search(ragstore_t* s, const float* queries, int n_queries, int query_dim, int top_k, int64_t* out_idx,
                    double* out_score, int32_t* out_found, int32_t* out_flags) {
    return rag::search_host_limited(s, -1, queries, n_queries, query_dim, top_k, out_idx, out_score, out_found, out_flags);
}

}  // extern "C"

int rag::search_host_limited(ragstore* s, int64_t n_limit, const float* queries, int n_queries, int query_dim, int top_k, int64_t* out_idx,
                             double* out_score, int32_t* out_found, int32_t* out_flags) {
    if (!s || n_queries < 0 || (n_queries > 0 && (!queries || !out_idx || !out_score || !out_found)))
        return fail(RAGSTORE_ERR_INVALID, "ragstore_search: null argument");
    int64_t n = s->count.load(std::memory_order_acquire);  // snapshot: a prefix of the appended rows
    if (n_limit >= 0 && n_limit < n) n = n_limit;
    const int req = top_k < 1 ? 1 : top_k;                       // Math.Max(1, topK)
    const int take = static_cast<int>(std::min<int64_t>(n, req));
    if (n == 0 || n_queries == 0) {
        for (int q = 0; q < n_queries; ++q) {
            out_found[q] = 0;
            if (out_flags) out_flags[q] = RAGSTORE_HIT_CERTIFIED | RAGSTORE_HIT_EXACT_ROWS;
        }
        return RAGSTORE_OK;
    }
    if (query_dim != s->dim) {
        // RagIndex.cs:126: every Cosine returns -1; the stable sort keeps insertion order
        for (int q = 0; q < n_queries; ++q) {
            for (int j = 0; j < take; ++j) {
                // the shard's first rows in insertion order; under block-cyclic dealing their GLOBAL ids (the merge of
                // all shards' answers then yields the corpus' first rows, as the reference's stable sort does)
                out_idx[static_cast<size_t>(q) * req + j] = host_global_row_id(s, j);
                out_score[static_cast<size_t>(q) * req + j] = -1.0;
            }
            out_found[q] = take;
            if (out_flags) out_flags[q] = RAGSTORE_HIT_CERTIFIED | RAGSTORE_HIT_EXACT_ROWS;
        }
        return RAGSTORE_OK;
    }
    if (take > kMaxTake) return fail(RAGSTORE_ERR_UNSUPPORTED, "more than 8128 hits per query are not covered");
    if (s->sharded) return sharded_search(s, queries, n_queries, top_k, out_idx, out_score, out_found, out_flags);
    NvtxRange nvtx("ragstore_search: host query -> scan/batch + top-k + rescoring -> host hits");
    DeviceGuard g(s->device);
    if (!g.ok) return fail(RAGSTORE_ERR_CUDA, "cudaSetDevice failed");

    const int ci = acquire_ctx(s);
    struct Release {
        ragstore* s;
        int ci;
        ~Release() { release_ctx(s, ci); }
    } rel{s, ci};
    SearchCtx& c = s->ctxs[ci];

    if (take > kMaxKp) {
        // More than 256 hits: the candidate lists hold 256 keys, so the scan is repeated.  Pass p collects (and rescores
        // exactly) the 256 best candidates whose approximate key lies BELOW the worst key of pass p-1; keys are unique
        // (score bits, row), so the passes partition the corpus in approximate order.  The rescored records of all
        // passes are then merged by (fp64 score desc, row asc) like the shards of a sharded corpus.
        const int passes = std::min<int>(kMaxPasses, static_cast<int>((std::min<int64_t>(n, take + 64) + kMaxKp - 1) / kMaxKp));
        for (int qi = 0; qi < n_queries; ++qi) {
            memcpy(c.h_q, queries + static_cast<size_t>(qi) * s->dim, s->dim * sizeof(float));
            CU_TRY(cudaMemcpyAsync(c.d_q, c.h_q, s->dim * sizeof(float), cudaMemcpyHostToDevice, c.st));
            uint64_t key_below = 0;
            int32_t flags_all = RAGSTORE_HIT_EXACT_ROWS;
            double bound_total = 0.0;
            int done_passes = 0;
            bool exhausted = false;
            for (int p = 0; p < passes && !exhausted; ++p) {
                ScanArgs a = base_args(s, n, kMaxKp, kMaxKp);
                a.d_query = c.d_q;
                a.scratch = c.scratch;
                a.key_below = key_below;
                a.out_flags = reinterpret_cast<int32_t*>(c.d_idx);   // scratch word, read back below
                a.out_cand = c.d_pass + static_cast<size_t>(p) * kMaxKp;
                a.cand_slots = kMaxKp;
                a.out_bound = c.d_bound;
                a.out_last_key = reinterpret_cast<uint64_t*>(c.d_bound + 1);
                s->n_scan_launches.fetch_add(1, std::memory_order_relaxed);
                const int rc = launch_scan(a, c.st);
                if (rc) return fail(RAGSTORE_ERR_CUDA, "scan kernel launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
                struct { double bound; uint64_t last; } hb{};
                int32_t hf = 0;
                CU_TRY(cudaMemcpyAsync(&hb, c.d_bound, 16, cudaMemcpyDeviceToHost, c.st));
                CU_TRY(cudaMemcpyAsync(&hf, c.d_idx, 4, cudaMemcpyDeviceToHost, c.st));
                CU_TRY(cudaStreamSynchronize(c.st));
                flags_all &= hf | ~RAGSTORE_HIT_EXACT_ROWS;
                bound_total = hb.bound;
                key_below = hb.last;
                done_passes = p + 1;
                exhausted = hb.last == 0ull || hb.bound == -std::numeric_limits<double>::infinity();  // list not full: nothing left
            }
            // merge the passes' records; outputs land in the context's packed result block
            const size_t nh = static_cast<size_t>(take);
            int64_t* d_idx = c.d_idx;
            double* d_score = reinterpret_cast<double*>(c.d_idx + nh);
            int32_t* d_found = reinterpret_cast<int32_t*>(d_score + nh);
            const int rc = launch_merge_shards(c.d_pass, done_passes, 1, kMaxKp, take, d_idx, d_score, d_found, c.st);
            if (rc) return fail(RAGSTORE_ERR_CUDA, "merge kernel launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
            CU_TRY(cudaMemcpyAsync(c.h_idx, c.d_idx, nh * 16 + 8, cudaMemcpyDeviceToHost, c.st));
            CU_TRY(cudaStreamSynchronize(c.st));
            const int32_t found = *reinterpret_cast<int32_t*>(reinterpret_cast<double*>(c.h_idx + nh) + nh);
            const size_t o = static_cast<size_t>(qi) * req;
            memcpy(out_idx + o, c.h_idx, sizeof(int64_t) * found);
            memcpy(out_score + o, reinterpret_cast<double*>(c.h_idx + nh), sizeof(double) * found);
            out_found[qi] = found;
            if (out_flags) {
                const double ek = found > 0 ? out_score[o + found - 1] : 0.0;
                const bool cert = exhausted || (found == take && ek > bound_total);
                out_flags[qi] = (flags_all & RAGSTORE_HIT_EXACT_ROWS) | (cert ? RAGSTORE_HIT_CERTIFIED : 0);
            }
        }
        return RAGSTORE_OK;
    }
    int kp0 = choose_kp(take);
    const int hs = take;  // hits per query in the context buffers (compact: only what is copied back)
    for (int q0 = 0; q0 < n_queries; q0 += kCtxQueries) {
        const int nq = std::min(kCtxQueries, n_queries - q0);
        {   // pack this call's outputs: [idx nq*hs][score nq*hs][found nq][flags nq]
            const size_t nh = static_cast<size_t>(nq) * hs;
            c.d_score = reinterpret_cast<double*>(c.d_idx + nh);
            c.d_found = reinterpret_cast<int32_t*>(c.d_score + nh);
            c.d_flags = c.d_found + nq;
            c.h_score = reinterpret_cast<double*>(c.h_idx + nh);
            c.h_found = reinterpret_cast<int32_t*>(c.h_score + nh);
            c.h_flags = c.h_found + nq;
        }
        int kp = kp0;
        bool batch_pass = batch_eligible(s, n, nq, kp0);
        // A shadowed fp32 store answers single queries from the bf16 image first (half the bytes); what that cannot certify
        // goes through the fp32 scan below as before.
        bool shadow_pass = !batch_pass && shadow_scan_pays(s, n);
        // Scan launches carry their query inside the kernel's parameter block (rows of up to kParamQueryMax elements):
        // no staging copy, no host->device copy command in front of the kernel.  The tensor-core pass reads all
        // queries from device memory, so a batch is uploaded as before.
        const bool q_in_params = !batch_pass && s->dim <= kParamQueryMax;
        // ... from the context's page-locked block, which the kernels read in place (it is device-addressable): no copy
        // command that would wait behind an ingest chunk on the copy engine while rows are being appended
        if (!q_in_params) memcpy(c.h_q, queries + static_cast<size_t>(q0) * s->dim, static_cast<size_t>(nq) * s->dim * sizeof(float));
        std::vector<int> todo(nq);
        for (int i = 0; i < nq; ++i) todo[i] = i;
        while (!todo.empty()) {
            if (batch_pass) {
                // first pass on the tensor cores; queries whose margin the tf32 scores cannot certify are
                // redone below by the fp32 scan kernel
                const int rc = run_batch(s, n, take, kp0, c.h_q, nq, c.d_idx, c.d_score, c.d_found, c.d_flags, hs,
                                         c.batch_scratch, c.st);
                if (rc) return fail(RAGSTORE_ERR_CUDA, "batched search launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
            }
            // Scan launches write their few hits STRAIGHT into the page-locked host block (device-addressable under
            // unified addressing): k * 16 bytes of posted PCIe writes by the kernel's last CTA instead of a
            // device->host copy command behind the kernel -- one driver call and one DMA round trip less per search.
            const bool direct = !batch_pass;
            unsigned int done_seq = 0;
            for (int i : todo) {
                if (batch_pass) break;
                ScanArgs a = shadow_pass ? shadow_scan_args(s, n, take, shadow_scan_kp(take)) : base_args(s, n, take, kp);
                if (i == todo.back()) {   // launches of one stream finish in order: the last one's word covers them all
                    a.out_done = c.h_done;
                    a.done_seq = done_seq = ++c.done_seq;
                }
                a.d_query = q_in_params ? nullptr : c.h_q + static_cast<size_t>(i) * s->dim;
                a.scratch = c.scratch;
                a.out_idx = c.h_idx + static_cast<size_t>(i) * hs;
                a.out_score = c.h_score + static_cast<size_t>(i) * hs;
                a.out_found = c.h_found + i;
                a.out_flags = c.h_flags + i;
                s->n_scan_launches.fetch_add(1, std::memory_order_relaxed);
                const int rc = launch_scan(a, c.st, q_in_params ? queries + static_cast<size_t>(q0 + i) * s->dim : nullptr);
                if (rc) return fail(RAGSTORE_ERR_CUDA, "scan kernel launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
            }
            if (!direct) {
                CU_TRY(cudaMemcpyAsync(c.h_idx, c.d_idx, static_cast<size_t>(nq) * hs * 16 + static_cast<size_t>(nq) * 8, cudaMemcpyDeviceToHost, c.st));
                CU_TRY(cudaStreamSynchronize(c.st));
            } else {
                const double bytes = static_cast<double>(n) * s->ld * ((s->dtype == RAGSTORE_F32 && !shadow_pass) ? 4.0 : 2.0);
                const int rcw = wait_done(c, done_seq, static_cast<double>(todo.size()) * (30.0 + bytes / 6.5e6));
                if (rcw) return rcw;
            }
            std::vector<int> again;
            for (int i : todo) {
                const bool cert = (c.h_flags[i] & RAGSTORE_HIT_CERTIFIED) != 0;
                if (!cert && (batch_pass || shadow_pass || kp < kMaxKp)) {
                    again.push_back(i);  // rescan this query (fp32 scan after a tf32 pass; else a wider list)
                    continue;
                }
                const size_t o = static_cast<size_t>(q0 + i) * req;
                memcpy(out_idx + o, c.h_idx + static_cast<size_t>(i) * hs, sizeof(int64_t) * take);
                memcpy(out_score + o, c.h_score + static_cast<size_t>(i) * hs, sizeof(double) * take);
                out_found[q0 + i] = c.h_found[i];
                if (out_flags) out_flags[q0 + i] = c.h_flags[i];
            }
            todo.swap(again);
            if (batch_pass)
                batch_pass = false;  // next pass: fp32 scan with the same list length
            else if (shadow_pass)
                shadow_pass = false; // next pass: fp32 scan with the request's own list length
            else
                kp = kMaxKp;
        }
    }
    return RAGSTORE_OK;
}

extern "C" {

int ragstore_hit_filters(const double* scores, int found, double evidence_min_score, double best_factor,
                         double low_score_floor, uint8_t* out_cited, double* out_best, int32_t* out_low_score) {
    if (found < 0 || (found > 0 && !scores)) return fail(RAGSTORE_ERR_INVALID, "ragstore_hit_filters: bad argument");
    if (found == 0) {  // the callers return early on an empty hit list (Program.cs:74,117)
        if (out_best) *out_best = 0.0;
        if (out_low_score) *out_low_score = 0;
        return RAGSTORE_OK;
    }
    double best = scores[0];  // hits.Max(h => h.Score): NaN never wins a comparison, as in Enumerable.Max
    for (int i = 1; i < found; ++i)
        if (scores[i] > best || best != best) best = scores[i];
    const double cut = evidence_min_score > best * best_factor ? evidence_min_score : best * best_factor;  // Math.Max (:126)
    if (out_cited)
        for (int i = 0; i < found; ++i) out_cited[i] = scores[i] >= cut ? 1 : 0;
    if (out_best) *out_best = best;
    if (out_low_score) *out_low_score = best < low_score_floor ? 1 : 0;  // :82
    return RAGSTORE_OK;
}

// ---------------------------------------------------------------- persistence
namespace {
struct SaveHeader {
    char magic[8];  // "RAGSTOR1"
    int32_t version, dim, ld, dtype, has_sidecar, raw_ld;
    uint32_t min_norm_bits, pad;
    int64_t count, inexact_rows, sidecar_rows;
};

int copy_out(FILE* f, const void* d_src, size_t bytes, void* h_pin, size_t pin_bytes) {
    const uint8_t* p = static_cast<const uint8_t*>(d_src);
    for (size_t off = 0; off < bytes; off += pin_bytes) {
        const size_t n = std::min(pin_bytes, bytes - off);
        CU_TRY(cudaMemcpy(h_pin, p + off, n, cudaMemcpyDeviceToHost));
        if (fwrite(h_pin, 1, n, f) != n) return fail(RAGSTORE_ERR_INVALID, "ragstore_save: short write");
    }
    return RAGSTORE_OK;
}
int copy_in(FILE* f, void* d_dst, size_t bytes, void* h_pin, size_t pin_bytes) {
    uint8_t* p = static_cast<uint8_t*>(d_dst);
    for (size_t off = 0; off < bytes; off += pin_bytes) {
        const size_t n = std::min(pin_bytes, bytes - off);
        if (fread(h_pin, 1, n, f) != n) return fail(RAGSTORE_ERR_INVALID, "ragstore_load: file truncated");
        CU_TRY(cudaMemcpy(p + off, h_pin, n, cudaMemcpyHostToDevice));
    }
    return RAGSTORE_OK;
}
}  // namespace

// one single-device store as a self-delimiting image at the current position of f (add_mu of the store is held by the caller)
static int save_image(ragstore* s, FILE* f) {
    DeviceGuard g(s->device);
    const size_t pin_bytes = size_t(32) << 20;
    void* h_pin = nullptr;
    if (cudaMallocHost(&h_pin, pin_bytes) != cudaSuccess) return fail(RAGSTORE_ERR_CUDA, "ragstore_save: cudaMallocHost");
    SaveHeader h{};
    memcpy(h.magic, "RAGSTOR1", 8);
    h.version = 1;
    h.dim = s->dim;
    h.ld = s->ld;
    h.dtype = s->dtype;
    h.has_sidecar = s->raw_slot ? 1 : 0;
    h.raw_ld = s->raw_ld;
    h.pad = s->shadow ? RAGSTORE_CREATE_BF16_SHADOW : 0;   // the shadow itself is not written: it is rebuilt from the rows on load
    h.count = s->count.load();
    h.inexact_rows = s->inexact_rows;
    h.sidecar_rows = s->raw_used;
    int rc = RAGSTORE_OK;
    if (cudaMemcpy(&h.min_norm_bits, s->d_min_norm, 4, cudaMemcpyDeviceToHost) != cudaSuccess) rc = fail(RAGSTORE_ERR_CUDA, "ragstore_save: read min norm");
    if (!rc && fwrite(&h, sizeof(h), 1, f) != 1) rc = fail(RAGSTORE_ERR_INVALID, "ragstore_save: short write");
    const size_t n = static_cast<size_t>(h.count);
    if (!rc) rc = copy_out(f, s->mat, n * s->ld * s->elem_size(), h_pin, pin_bytes);
    if (!rc) rc = copy_out(f, s->rnorm, n * sizeof(float), h_pin, pin_bytes);
    if (!rc) rc = copy_out(f, s->exact, n, h_pin, pin_bytes);
    if (!rc && h.has_sidecar) rc = copy_out(f, s->raw_slot, n * sizeof(int32_t), h_pin, pin_bytes);
    if (!rc && h.has_sidecar) rc = copy_out(f, s->raw_rows, static_cast<size_t>(h.sidecar_rows) * s->raw_ld * sizeof(float), h_pin, pin_bytes);
    cudaFreeHost(h_pin);
    return rc;
}

namespace {
struct ShardedSaveHeader {
    char magic[8];  // "RAGSHRD1"
    int32_t version, n_shards, dim, dtype;
    int64_t block_rows, count;
};
int64_t file_bytes_left(FILE* f) {
    const long here = ftell(f);
    if (here < 0 || fseek(f, 0, SEEK_END) != 0) return -1;
    const long end = ftell(f);
    if (fseek(f, here, SEEK_SET) != 0) return -1;
    return static_cast<int64_t>(end) - here;
}
}  // namespace

int ragstore_save(ragstore_t* s, const char* path) {
    if (!s || !path) return fail(RAGSTORE_ERR_INVALID, "ragstore_save: null argument");
    FILE* f = fopen(path, "wb");
    if (!f) return fail(RAGSTORE_ERR_INVALID, "ragstore_save: cannot open file", path);
    int rc = RAGSTORE_OK;
    if (s->sharded) {
        // a multi-device handle: a small header, then every shard's image in shard order.  No append runs meanwhile, so the
        // shards hold exactly the published prefix.
        int64_t block_rows = 0;
        std::unique_lock<std::mutex> lk(*sharded_add_mutex(s, &block_rows));
        ShardedSaveHeader h{};
        memcpy(h.magic, "RAGSHRD1", 8);
        h.version = 1;
        h.n_shards = static_cast<int32_t>(s->shards.size());
        h.dim = s->dim;
        h.dtype = s->dtype;
        h.block_rows = block_rows;
        h.count = s->count.load();
        if (fwrite(&h, sizeof(h), 1, f) != 1) rc = fail(RAGSTORE_ERR_INVALID, "ragstore_save: short write");
        for (size_t i = 0; !rc && i < s->shards.size(); ++i) {
            std::lock_guard<std::mutex> lk2(s->shards[i]->add_mu);
            rc = save_image(s->shards[i], f);
        }
    } else {
        std::lock_guard<std::mutex> lk(s->add_mu);  // no append while the image is written
        rc = save_image(s, f);
    }
    if (fclose(f) != 0 && !rc) rc = fail(RAGSTORE_ERR_INVALID, "ragstore_save: close failed");
    return rc;
}

// Reads one image from the current position of f into a new store on `device`.  whole_file: the image must end where the
// file ends (a stand-alone image); otherwise it must fit in what is left (an image inside a multi-device file).
static int load_image(FILE* f, bool whole_file, int device, int64_t capacity_rows, ragstore_t** out) {
    *out = nullptr;
    SaveHeader h{};
    if (fread(&h, sizeof(h), 1, f) != 1 || memcmp(h.magic, "RAGSTOR1", 8) != 0 || h.version != 1)
        return fail(RAGSTORE_ERR_INVALID, "ragstore_load: not a ragstore image (version 1)");
    {   // nothing in the header is trusted before it is checked against itself and against the file's size
        const bool dt_ok = h.dtype == RAGSTORE_F32 || h.dtype == RAGSTORE_BF16;
        const int align = h.dtype == RAGSTORE_F32 ? 4 : 8;
        bool ok = dt_ok && h.dim > 0 && h.dim <= (1 << 20) && h.ld == (h.dim + align - 1) / align * align && h.raw_ld == (h.dim + 3) / 4 * 4 &&
                  h.count >= 0 && h.count < 0xFFFFFFFFll && (h.has_sidecar == 0 || h.has_sidecar == 1) && h.sidecar_rows >= 0 &&
                  h.sidecar_rows <= h.count && h.inexact_rows >= 0 && h.inexact_rows <= h.count && (h.has_sidecar || h.sidecar_rows == 0) &&
                  (h.pad == 0 || h.pad == RAGSTORE_CREATE_BF16_SHADOW);
        if (ok) {
            const uint64_t n = static_cast<uint64_t>(h.count), esz = h.dtype == RAGSTORE_F32 ? 4 : 2;
            uint64_t want = n * static_cast<uint64_t>(h.ld) * esz + n * 4 + n;
            if (h.has_sidecar) want += n * 4 + static_cast<uint64_t>(h.sidecar_rows) * static_cast<uint64_t>(h.raw_ld) * 4;
            const int64_t left = file_bytes_left(f);
            ok = left >= 0 && (whole_file ? static_cast<uint64_t>(left) == want : static_cast<uint64_t>(left) >= want);
        }
        if (!ok) return fail(RAGSTORE_ERR_INVALID, "ragstore_load: header is inconsistent (counts, layout or file size do not agree)");
    }
    const int64_t cap = std::max<int64_t>(std::max<int64_t>(capacity_rows, h.count), 1);
    ragstore_t* s = nullptr;
    int rc = ragstore_create(h.dim, h.dtype, cap, device, (h.has_sidecar ? RAGSTORE_CREATE_SIDECAR_ON : RAGSTORE_CREATE_SIDECAR_OFF) | h.pad, &s);
    if (rc) return rc;
    if (s->ld != h.ld || s->raw_ld != h.raw_ld) {
        ragstore_destroy(s);
        return fail(RAGSTORE_ERR_INVALID, "ragstore_load: image layout differs from this build");
    }
    DeviceGuard g(device);
    const size_t pin_bytes = size_t(32) << 20;
    void* h_pin = nullptr;
    if (cudaMallocHost(&h_pin, pin_bytes) != cudaSuccess) rc = fail(RAGSTORE_ERR_CUDA, "ragstore_load: cudaMallocHost");
    const size_t n = static_cast<size_t>(h.count), n4 = ((n + 3) & ~size_t(3)) + 64;
    if (!rc && n) {
        int vr = s->mat_buf.ensure(n * s->ld * s->elem_size(), false, s->ingest_stream);
        if (!vr && s->shadow) vr = s->shadow_buf.ensure(n * s->ld * 2, false, s->ingest_stream);
        if (!vr) vr = s->rnorm_buf.ensure(n4 * sizeof(float), true, s->ingest_stream);
        if (!vr) vr = s->exact_buf.ensure(n4, true, s->ingest_stream);
        if (!vr && h.has_sidecar) vr = s->slot_buf.ensure(n4 * sizeof(int32_t), true, s->ingest_stream);
        if (!vr && h.has_sidecar && h.sidecar_rows) vr = s->raw_buf.ensure(static_cast<size_t>(h.sidecar_rows) * s->raw_ld * sizeof(float), false, s->ingest_stream);
        if (vr) rc = fail(RAGSTORE_ERR_CUDA, "ragstore_load: mapping device memory", cudaGetErrorString(static_cast<cudaError_t>(vr)));
        if (!rc && cudaStreamSynchronize(s->ingest_stream) != cudaSuccess) rc = fail(RAGSTORE_ERR_CUDA, "ragstore_load: sync");
        if (!rc) rc = copy_in(f, s->mat, n * s->ld * s->elem_size(), h_pin, pin_bytes);
        if (!rc) rc = copy_in(f, s->rnorm, n * sizeof(float), h_pin, pin_bytes);
        if (!rc) rc = copy_in(f, s->exact, n, h_pin, pin_bytes);
        if (!rc && h.has_sidecar) rc = copy_in(f, s->raw_slot, n * sizeof(int32_t), h_pin, pin_bytes);
        if (!rc && h.has_sidecar) rc = copy_in(f, s->raw_rows, static_cast<size_t>(h.sidecar_rows) * s->raw_ld * sizeof(float), h_pin, pin_bytes);
        if (!rc && s->shadow) {   // the bf16 image the batched path nominates from, rebuilt from the loaded fp32 rows
            const int lr = launch_shadow_from_rows(static_cast<const float*>(s->mat), s->shadow, static_cast<int64_t>(n) * s->ld, s->ingest_stream);
            if (lr || cudaStreamSynchronize(s->ingest_stream) != cudaSuccess) rc = fail(RAGSTORE_ERR_CUDA, "ragstore_load: rebuilding the bf16 shadow");
        }
    }
    if (!rc) {
        const unsigned long long counters[2] = {static_cast<unsigned long long>(h.inexact_rows), static_cast<unsigned long long>(h.sidecar_rows)};
        if (cudaMemcpy(s->d_inexact, counters, sizeof(counters), cudaMemcpyHostToDevice) != cudaSuccess ||
            cudaMemcpy(s->d_min_norm, &h.min_norm_bits, 4, cudaMemcpyHostToDevice) != cudaSuccess)
            rc = fail(RAGSTORE_ERR_CUDA, "ragstore_load: restoring counters");
    }
    if (h_pin) cudaFreeHost(h_pin);
    if (rc) {
        ragstore_destroy(s);
        return rc;
    }
    s->inexact_rows = h.inexact_rows;
    s->raw_used = h.sidecar_rows;
    s->count.store(h.count, std::memory_order_release);
    *out = s;
    return RAGSTORE_OK;
}

int ragstore_load(const char* path, int device, int64_t capacity_rows, ragstore_t** out) {
    if (!path || !out) return fail(RAGSTORE_ERR_INVALID, "ragstore_load: null argument");
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) return fail(RAGSTORE_ERR_INVALID, "ragstore_load: cannot open file", path);
    const int rc = load_image(f, true, device, capacity_rows, out);
    fclose(f);
    return rc;
}

int ragstore_load_sharded(const char* path, const int* devices, int n_devices, int64_t capacity_rows, ragstore_t** out) {
    if (!path || !out || !devices || n_devices < 1) return fail(RAGSTORE_ERR_INVALID, "ragstore_load_sharded: null argument");
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) return fail(RAGSTORE_ERR_INVALID, "ragstore_load_sharded: cannot open file", path);
    ShardedSaveHeader h{};
    int rc = RAGSTORE_OK;
    if (fread(&h, sizeof(h), 1, f) != 1 || memcmp(h.magic, "RAGSHRD1", 8) != 0 || h.version != 1)
        rc = fail(RAGSTORE_ERR_INVALID, "ragstore_load_sharded: not a multi-device ragstore image (version 1)");
    if (!rc && (h.n_shards != n_devices || h.n_shards < 1 || h.n_shards > 64 || h.block_rows < 1 || h.count < 0 || h.dim < 1))
        rc = fail(RAGSTORE_ERR_INVALID, "ragstore_load_sharded: the image was saved over a different number of shards (rows are dealt by shard count), or its header is damaged");
    std::vector<ragstore_t*> shards;
    const int64_t cap_total = std::max<int64_t>(std::max<int64_t>(capacity_rows, h.count), 1);
    for (int i = 0; !rc && i < n_devices; ++i) {
        ragstore_t* sh = nullptr;
        // per-shard capacity as ragstore_create_sharded sizes it
        const int64_t cap = std::max<int64_t>(shard_rows(cap_total, i, h.block_rows, n_devices), h.block_rows);
        rc = load_image(f, false, devices[i], cap, &sh);
        if (!rc) shards.push_back(sh);
    }
    if (!rc && file_bytes_left(f) != 0) rc = fail(RAGSTORE_ERR_INVALID, "ragstore_load_sharded: trailing bytes after the last shard");
    fclose(f);
    if (!rc) rc = sharded_adopt(shards, devices, h.dim, h.dtype, cap_total, h.block_rows, h.count, out);
    if (rc)
        for (ragstore_t* sh : shards) ragstore_destroy(sh);
    return rc;
}

int ragstore_set_coalescing(ragstore_t* s, int window_us, int max_batch) {
    if (!s || window_us < 0 || max_batch < 1) return fail(RAGSTORE_ERR_INVALID, "ragstore_set_coalescing: bad argument");
    std::lock_guard<std::mutex> lk(s->co.mu);
    s->co.window_us = window_us;
    s->co.max_batch = std::min(max_batch, 1024);
    return RAGSTORE_OK;
}

// Which search path a request of `req` hits takes: requests are only batched with their own kind, so that one caller's
// large top_k cannot push everybody else off the tensor-core path (or hand them its error).
static int coalesce_class(int req) {
    if (req > kMaxKp) return 2;              // repeated scans (top_k above 256)
    return choose_kp(req) == 32 ? 0 : 1;     // 0: candidate lists of 32 (tensor-core path from 2-3 callers up); 1: wider lists
}

int ragstore_search_coalesced(ragstore_t* s, const float* query, int query_dim, int top_k, int64_t* out_idx, double* out_score,
                              int32_t* out_found, int32_t* out_flags) {
    if (!s || !query || !out_idx || !out_score || !out_found) return fail(RAGSTORE_ERR_INVALID, "ragstore_search_coalesced: null argument");
    if (query_dim != s->dim)  // the -1 corner case needs no GPU work: answer it directly
        return ragstore_search(s, query, 1, query_dim, top_k, out_idx, out_score, out_found, out_flags);
    const int req = top_k < 1 ? 1 : top_k;
    // refused before it is queued: an unservable request must not reach (and fail) a batch it shares with others
    if (std::min<int64_t>(s->count.load(std::memory_order_acquire), req) > kMaxTake)
        return fail(RAGSTORE_ERR_UNSUPPORTED, "more than 8128 hits per query are not covered");
    CoPending me{query, top_k, out_idx, out_score, out_found, out_flags};
    // one sweep of the corpus for a whole batch (tensor-core path from 2-3 queries up); runs without the queue's lock
    auto run = [s](const std::vector<CoPending*>& batch) {
        const int nb = static_cast<int>(batch.size());
        int kmax = 1;
        for (auto* p : batch) kmax = std::max(kmax, p->top_k);
        std::vector<float> qbuf(static_cast<size_t>(nb) * s->dim);
        std::vector<int64_t> ibuf(static_cast<size_t>(nb) * kmax);
        std::vector<double> sbuf(static_cast<size_t>(nb) * kmax);
        std::vector<int32_t> fbuf(nb), gbuf(nb);
        for (int i = 0; i < nb; ++i) memcpy(qbuf.data() + static_cast<size_t>(i) * s->dim, batch[i]->q, sizeof(float) * s->dim);
        const int rc = ragstore_search(s, qbuf.data(), nb, s->dim, kmax, ibuf.data(), sbuf.data(), fbuf.data(), gbuf.data());
        const std::string err = rc ? g_err : std::string();
        for (int i = 0; i < nb; ++i) {
            CoPending* p = batch[i];
            if (!rc) {
                // a hit list for k is the prefix of the hit list for any larger k (the order is total)
                const int want = p->top_k < 1 ? 1 : p->top_k;
                const int n = std::min(want, fbuf[i]);
                memcpy(p->idx, ibuf.data() + static_cast<size_t>(i) * kmax, sizeof(int64_t) * n);
                memcpy(p->score, sbuf.data() + static_cast<size_t>(i) * kmax, sizeof(double) * n);
                *p->found = n;
                if (p->flags) *p->flags = gbuf[i];
            }
            p->status = rc;
            p->err = err;
        }
    };
    const int status = s->co.submit(me, [](int k) { return coalesce_class(k < 1 ? 1 : k); }, run);
    if (status) g_err = me.err;
    return status;
}

int ragstore_synth_fill_device(float* d_out, int64_t row0, int64_t n_rows, int dim, int64_t ld, uint64_t seed,
                               int round_bf16, int device, void* stream) {
    if (!d_out || n_rows < 0 || dim <= 0 || ld < dim) return fail(RAGSTORE_ERR_INVALID, "ragstore_synth_fill_device: bad argument");
    DeviceGuard g(device);
    if (!g.ok) return fail(RAGSTORE_ERR_NO_DEVICE, "ragstore_synth_fill_device: no such device");
    const int rc = launch_synth(d_out, row0, n_rows, dim, ld, seed, round_bf16, static_cast<cudaStream_t>(stream));
    if (rc) return fail(RAGSTORE_ERR_CUDA, "synth kernel launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
    return RAGSTORE_OK;
}

}  // extern "C"
