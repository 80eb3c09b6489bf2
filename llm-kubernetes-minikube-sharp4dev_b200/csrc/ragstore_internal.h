// ragstore_internal.h -- host-side declarations shared by the .cu files of libragstore.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ragstore.h"

namespace rag {

struct ExchangeView;   // exchange_device.cuh

constexpr int kConsumerWarps = 8;                       // warps that score rows
constexpr int kScanThreads = (kConsumerWarps + 1) * 32; // + one producer warp (bulk-copy issue)
constexpr int kMaxKp = 256;                             // largest candidate list kept per warp/CTA
constexpr int kMaxCtas = 1024;                          // scratch sizing: CTAs of one scan launch

// Scratch one scan launch needs in global memory (beyond inputs/outputs).
struct ScanScratch {
    uint64_t* cta_lists;   // [kMaxCtas * kMaxKp] per-CTA candidate keys
    unsigned int* ticket;  // [1] "last CTA done" counter, self-resetting; ticket[1] = in-use flag, ticket[2] = number of
                           // launches that found the scratch already in use (two searches overlapping on one scratch)
    unsigned long long* stamps;  // [8] %globaltimer at the phase boundaries of the last launch (profiling aid)
};
constexpr size_t kScanScratchBytes = sizeof(uint64_t) * size_t(kMaxCtas) * kMaxKp + 256;

struct ScanArgs {
    // store
    const void* mat;        // row-major [n_rows][ld] of fp32 or bf16
    const float* rnorm;     // [round_up(n_rows,4)] reciprocal norm of each STORED row (0 for a zero row)
    const uint8_t* exact;   // [n_rows] 1 = stored bit-for-bit as added (may be null => treated as 1)
    const int32_t* raw_slot;// [n_rows] slot of the row's original fp32 values in raw_rows, -1 = stored exactly (may be null)
    const float* raw_rows;  // [n_slots][raw_ld] original fp32 rows of the inexactly stored ones (sidecar)
    int raw_ld;
    const float* rescore_rows;      // not null: `mat` is only a bf16 IMAGE of an fp32 store (the shadow); candidates are rescored
    int rescore_ld;                 // from these fp32 rows [n_rows][rescore_ld] (a sidecar original still takes precedence)
    const uint32_t* min_norm_bits;  // fp32 bits of the smallest non-zero original row norm (may be null)
    float min_norm_assumed;         // used when min_norm_bits is null: the caller's promise about the smallest row norm
                                    // (0 = no promise: the certificate then ignores the reference's +1e-9 term)
    int64_t n_rows;
    int dim, ld, dtype;
    int64_t row_base;
    int64_t row_block;      // > 0: block-cyclic global ids (streaming ingest over several shards), see global_row_id()
    int row_nshards, row_shard;
    // query + selection
    const float* d_query;   // raw fp32 [dim]
    int k_take;             // min(n_rows, max(1, top_k))
    int kp;                 // candidate list length: power of two in [32, kMaxKp], >= k_take
    float eps;              // bound on |approximate score - reference score| (DESIGN.md "certificate")
    uint64_t key_below;     // 0 = none; otherwise only candidates with key < key_below are collected (top_k > 256: the
                            // next 256 candidates after those of the previous pass)
    // scratch
    ScanScratch scratch;
    // outputs (device memory)
    int64_t* out_idx;       // [k_take]  (may be null when out_cand is used)
    double* out_score;      // [k_take]
    int32_t* out_found;     // [1]
    int32_t* out_flags;     // [1] RAGSTORE_HIT_*
    ragstore_cand* out_cand;// [cand_slots] shard records, padded with {-inf,-1} (may be null)
    int cand_slots;
    // Row-sharded corpora: when `xch` is set the launch's LAST CTA also does the exchange step of SURVEY 8(e) -- it stores
    // the cand_slots records it has just written to out_cand into every rank's inbox over peer memory, waits for the
    // other ranks' records of call `xseq` and ranks all of them: one launch per query per rank (exchange_device.cuh)
    const ExchangeView* xch;           // device copy of the exchange view (null: no exchange)
    unsigned long long xseq;          // sequence number of this collective call
    int64_t* x_idx;                   // [cand_slots] merged global ids, best first
    double* x_score;                  // [cand_slots]
    int32_t* x_found;                 // [1]
    int32_t* x_flags;                 // [1] RAGSTORE_HIT_* ANDed over all ranks (may be null)
    int32_t* x_status;                // [1] 0, or 1 when a peer's records did not arrive in time (may be null)
    unsigned int* out_done; // [1] page-locked host word: receives done_seq once every output of this launch is visible to the
    unsigned int done_seq;  // host (system-scope fence first).  A host thread may poll it instead of waiting for the stream.
    double* out_bound;      // [1] bound + eps + shrink: no row outside the collected candidates scores above it (may be null)
    uint64_t* out_last_key; // [1] approximate key of the worst collected candidate, 0 if none (may be null)
    int sm_count;
};

// Enqueues the fused scan + top-k + exact-rescore kernel.  Returns cudaError_t as int.
// h_query: when not null (and dim <= kParamQueryMax, a.d_query == nullptr) the raw query travels inside the kernel's
// parameter block instead of device memory: a host caller then needs no host->device copy before the launch.
constexpr int kParamQueryMax = 1024;
int launch_scan(const ScanArgs& a, cudaStream_t stream, const float* h_query = nullptr);

struct NormaliseArgs {
    const float* src;     // [n_rows][src_ld] fp32 (may alias dst for the fp32 store)
    int64_t src_ld;
    void* dst;            // first destination row in the matrix
    float* rnorm;         // destination slice of the per-row reciprocal norms
    uint8_t* exact;       // destination slice of the per-row exact flags
    int32_t* raw_slot;    // destination slice of the sidecar slots (null = no sidecar)
    float* raw_rows;      // sidecar base
    int raw_ld;
    unsigned int* raw_count;              // device counter of sidecar rows in use
    uint32_t* min_norm_bits;              // device scalar: fp32 bits of the smallest non-zero row norm
    int64_t n_rows;
    int dim, ld, dtype;
    int prenormalised;
    unsigned long long* inexact_counter;  // device counter of rows stored inexactly
    uint16_t* shadow;     // fp32 stores with a bf16 shadow: destination slice of the shadow rows (ld elements each), or null
};
int launch_normalise(const NormaliseArgs& a, cudaStream_t stream);
// bf16 image of n fp32 values (a loaded store's shadow is rebuilt from its rows)
int launch_shadow_from_rows(const float* src, uint16_t* dst, int64_t n, cudaStream_t stream);

// ---- batched (tensor-core) path, batch_gemm.cu
constexpr int kBatchMaxQ = 256;    // queries per launch (two 128-row MMA blocks)
constexpr int kBatchMinQ = 2;      // one query always takes the scan kernel; from 2 up a cost model decides (ragstore_api.cu)
constexpr int kBatchMaxLd = 4096;  // padded row length covered by the query operand buffer
constexpr int64_t kBatchMinRows = 4096;
struct BatchSearchArgs {
    ScanArgs scan;           // store views, k_take, kp (size of the MERGED candidate set per query: 32..256; every CTA
                             // keeps 32 keys per query), eps, outputs as per-launch bases
    const void* operand_mat; // rows the tensor cores read: null = scan.mat; else a bf16 shadow of an fp32 store
                             // (same row count and ld): nomination on bf16 operands, rescoring from the fp32 rows
    int operand_bf16;        // 1 when operand_mat is bf16 although scan.dtype is fp32
    int thr_rank;            // rank of the sampled score that becomes the skip threshold (0 = scan.kp).  Smaller than
                             // scan.kp when the candidate set is wider than the request needs (shadow nomination): fewer
                             // rows pass the threshold, so the sampling pass can stay sparse
    const float* d_queries;  // raw queries [nq][dim], device
    int nq;
    int out_stride;          // hits per query in scan.out_idx / scan.out_score
    uint8_t* scratch;        // batch_scratch_bytes(sm_count) bytes of device memory
};
size_t batch_scratch_bytes(int sm_count);
int launch_batch_search(const BatchSearchArgs& b, cudaStream_t stream);
void batch_reload_env();   // re-reads the RAGSTORE_BATCH_* experiment knobs (otherwise read once)

int launch_synth(float* d_out, int64_t row0, int64_t n_rows, int dim, int64_t ld, uint64_t seed,
                 int round_bf16, cudaStream_t stream);

// records `what[: detail]` as the calling thread's ragstore_last_error text and returns `code` (ragstore_api.cu)
int api_fail(int code, const char* what, const char* detail = nullptr);

int launch_merge_shards(const ragstore_cand* d_gathered, int n_shards, int n_queries, int per_shard, int k_take,
                        int64_t* d_idx, double* d_score, int32_t* d_found, cudaStream_t stream);

}  // namespace rag
