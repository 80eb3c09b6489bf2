// store_impl.h -- the store object behind `ragstore_t` and the host helpers shared by ragstore_api.cu (single-device
// stores) and sharded_store.cu (one handle over several devices).  Not part of the ABI.
#pragma once
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "coalesce.h"
#include "ragstore_internal.h"
#include "vmm_buffer.h"

namespace rag {

struct DeviceGuard {
    int prev = -1;
    bool ok = false;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev == dev) {
            prev = -1;  // already there: nothing to switch, nothing to restore
            ok = true;
        } else {
            ok = cudaSetDevice(dev) == cudaSuccess;
        }
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

// NVTX range around a host-side phase (upload/normalise, scan, batched search, exchange, merge): shows up on an nsys /
// Nsight timeline when a tool is attached; a few nanoseconds of no-op otherwise (NVTX v3 is header-only and resolves the
// tool's injection library lazily).
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};

constexpr int kCtxQueries = 256; // queries per host-search context batch (= one tensor-core launch)
constexpr int kNumCtx = 4;       // concurrent host searches per store
constexpr int kMaxPasses = 32;   // top_k > 256 is served by repeated scans of 256 candidates each
constexpr int kMaxTake = kMaxPasses * kMaxKp - 64;

struct SearchCtx {
    cudaStream_t st = nullptr;
    ScanScratch scratch{};
    uint8_t* batch_scratch = nullptr;  // candidate lists + query operand of the tensor-core path
    ragstore_cand* d_pass = nullptr;   // top_k > 256: the rescored candidates of up to kMaxPasses scans
    double* d_bound = nullptr;         // [1] + uint64 last key [1]
    float* d_q = nullptr;
    int64_t* d_idx = nullptr;
    double* d_score = nullptr;
    int32_t* d_found = nullptr;
    int32_t* d_flags = nullptr;
    float* h_q = nullptr;  // pinned mirrors
    int64_t* h_idx = nullptr;
    double* h_score = nullptr;
    int32_t* h_found = nullptr;
    int32_t* h_flags = nullptr;
    unsigned int* h_done = nullptr;   // page-locked completion word the scan's last CTA writes (see wait_done)
    unsigned int done_seq = 0;
};

int choose_kp(int take);
float scan_eps(int dim_ld, int dtype);

}  // namespace rag

namespace rag {
struct ShardedState;
}
using rag::ScanScratch;
using rag::SearchCtx;
using rag::VmmBuffer;

struct ragstore {
    int device = 0, dim = 0, ld = 0, dtype = 0, sm_count = 0;
    int64_t capacity = 0;
    std::atomic<int64_t> count{0};
    int64_t row_base = 0;
    int64_t row_block = 0;
    int row_nshards = 1, row_shard = 0;
    mutable std::atomic<int64_t> n_scan_launches{0}, n_batch_launches{0};  // kernel launches issued by searches
    VmmBuffer mat_buf, rnorm_buf, exact_buf, slot_buf, raw_buf, shadow_buf;  // address-stable, grown on demand
    uint16_t* shadow = nullptr;    // fp32 store created with RAGSTORE_CREATE_BF16_SHADOW: bf16 image of the stored rows,
                                   // [rows][ld], read by the batched (tensor-core) nomination pass only
    void* mat = nullptr;
    float* rnorm = nullptr;
    uint8_t* exact = nullptr;
    int32_t* raw_slot = nullptr;   // null when the store keeps no sidecar
    float* raw_rows = nullptr;
    int raw_ld = 0;
    int64_t raw_used = 0;          // sidecar rows in use (host copy, refreshed after every add)
    int64_t inexact_rows = 0;      // rows stored inexactly (host copy, refreshed after every add)
    unsigned long long* d_inexact = nullptr;  // [0] inexact rows, then [1] raw_count (u32), [2] min_norm_bits (u32)
    unsigned int* d_raw_count = nullptr;
    uint32_t* d_min_norm = nullptr;
    cudaStream_t ingest_stream = nullptr;
    cudaStream_t dev_stream = nullptr;
    std::mutex add_mu;
    float* d_stage = nullptr;      // device staging (bf16 stores fed from the host), two chunks
    size_t stage_floats = 0;
    float* h_stage[2] = {nullptr, nullptr};   // pinned host staging, one per pipeline slot
    cudaEvent_t ingest_ev[2] = {nullptr, nullptr};
    cudaEvent_t k1_ev0[2] = {nullptr, nullptr}, k1_ev1[2] = {nullptr, nullptr};  // bracket the K1 launches (profiling aid)
    double last_add_kernel_us = 0.0;
    unsigned long long* h_counters = nullptr; // pinned: per slot [inexact rows, sidecar rows]
    rag::Coalescer co;   // coalescing of concurrent single-query callers (ragstore_search_coalesced)
    ScanScratch dev_scratch{};
    uint8_t* dev_batch_scratch = nullptr;
    std::vector<SearchCtx> ctxs;
    std::vector<int> free_ctx;
    std::mutex ctx_mu;
    std::condition_variable ctx_cv;
    // ---- single-process multi-device store (sharded_store.cu): a parent handle owns one child store per shard and no
    // device memory of its own; every public entry point dispatches on `shards`
    std::vector<ragstore*> shards;
    rag::ShardedState* sharded = nullptr;
    size_t elem_size() const { return dtype == RAGSTORE_F32 ? 4 : 2; }
};


namespace rag {

// ---- helpers implemented in ragstore_api.cu
ScanArgs base_args(const ragstore* s, int64_t n, int take, int kp);
// the same search as a scan over the store's bf16 shadow (half the bytes), candidates rescored from the fp32 rows; only for
// stores created with RAGSTORE_CREATE_BF16_SHADOW (s->shadow != null).  Wider error bound: callers retry uncertified
// queries with base_args().
ScanArgs shadow_scan_args(const ragstore* s, int64_t n, int take, int kp);
int shadow_scan_kp(int take);
bool shadow_scan_pays(const ragstore* s, int64_t n);
float batch_eps(const ragstore* s);
bool batch_eligible(const ragstore* s, int64_t n, int nq, int kp);
// store views, candidate-set size, error bound and tensor-core operand of one batched launch (outputs left to the caller)
void fill_batch_args(const ragstore* s, int64_t n, int take, int kp, BatchSearchArgs& b);
int acquire_ctx(ragstore* s);            // blocks until one of the store's search contexts is free
void release_ctx(ragstore* s, int ci);

// ragstore_search on one single-device store, restricted to its first n_limit rows (n_limit < 0: all published rows)
int search_host_limited(ragstore* s, int64_t n_limit, const float* queries, int n_queries, int query_dim, int top_k,
                        int64_t* out_idx, double* out_score, int32_t* out_found, int32_t* out_flags);

// ---- sharded_store.cu: what the public entry points do when the handle is a sharded parent
int sharded_destroy(ragstore* p);
int sharded_add(ragstore* p, const float* rows, int64_t n_rows, uint32_t flags, int64_t* out_first_row);
int sharded_reserve_rows(ragstore* p, int64_t n_rows);
int sharded_search(ragstore* p, const float* queries, int n_queries, int top_k, int64_t* out_idx, double* out_score,
                   int32_t* out_found, int32_t* out_flags);
int sharded_info(ragstore* p, int64_t* out, int n);
std::mutex* sharded_add_mutex(ragstore* p, int64_t* block_rows);   // the handle's append lock (held while it is saved)
// wraps already loaded shard stores (one per device, dealt with block_rows) into a multi-device handle; checks that their
// row counts are those of a corpus of `count` rows.  On failure the shards stay the caller's.
int sharded_adopt(std::vector<ragstore*>& shards, const int* devices, int dim, int dtype, int64_t capacity_rows, int64_t block_rows,
                  int64_t count, ragstore** out);

}  // namespace rag
