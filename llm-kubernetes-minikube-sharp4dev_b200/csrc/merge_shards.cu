// merge_shards.cu -- global top-k over the all-gathered per-shard hit lists.
//
// Each shard contributes its own exactly-rescored top-k as (fp64 score, global row id) records;
// the union contains the global top-k, and (score desc, id asc) is the order a single RagIndex
// over the concatenated corpus would produce (reference Helpers/RagIndex.cs:64-65: stable
// OrderByDescending + Take).  One CTA ranks every record by counting the records that beat it.
#include <math_constants.h>

#include "common.cuh"
#include "ragstore_internal.h"

namespace rag {

// gathered layout: [shard][query][per_shard]; one CTA per query (blockIdx.x)
__global__ void __launch_bounds__(256) merge_shards_kernel(const ragstore_cand* __restrict__ in, int n_shards, int n_queries,
                                                           int per_shard, int k_req, int64_t* out_idx, double* out_score,
                                                           int32_t* out_found) {
    extern __shared__ __align__(16) uint8_t sm[];
    const int n_items = n_shards * per_shard;
    const int qi = blockIdx.x;
    uint64_t* keys = reinterpret_cast<uint64_t*>(sm);
    int64_t* ids = reinterpret_cast<int64_t*>(keys + n_items);
    double* scs = reinterpret_cast<double*>(ids + n_items);
    out_idx += static_cast<size_t>(qi) * k_req;
    out_score += static_cast<size_t>(qi) * k_req;
    out_found += qi;
    __shared__ int n_valid;
    if (threadIdx.x == 0) n_valid = 0;
    __syncthreads();
    int local_valid = 0;
    for (int i = threadIdx.x; i < n_items; i += blockDim.x) {
        const int sh = i / per_shard, j = i % per_shard;
        const ragstore_cand c = in[(static_cast<size_t>(sh) * n_queries + qi) * per_shard + j];
        scs[i] = c.score;
        keys[i] = ord_f64(c.score);
        ids[i] = c.id;
        local_valid += (c.id >= 0);
    }
    if (local_valid) atomicAdd(&n_valid, local_valid);
    __syncthreads();
    const int take = min(n_valid, k_req);
    for (int i = threadIdx.x; i < n_items; i += blockDim.x) {
        const int64_t id = ids[i];
        if (id < 0) continue;
        const uint64_t k = keys[i];
        int rank = 0;
        for (int u = 0; u < n_items; ++u) {
            const int64_t iu = ids[u];
            const uint64_t ku = keys[u];
            rank += (iu >= 0) && ((ku > k) || (ku == k && iu < id));
        }
        if (rank < take) {
            out_idx[rank] = id;
            out_score[rank] = scs[i];
        }
    }
    if (threadIdx.x == 0) *out_found = take;
}

int launch_merge_shards(const ragstore_cand* d_gathered, int n_shards, int n_queries, int per_shard, int k_take,
                        int64_t* d_idx, double* d_score, int32_t* d_found, cudaStream_t stream) {
    const int n_items = n_shards * per_shard;
    if (n_items <= 0 || n_items > 8192 || n_queries <= 0) return static_cast<int>(cudaErrorInvalidValue);
    const size_t smem = static_cast<size_t>(n_items) * 24;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(merge_shards_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem));
        if (e != cudaSuccess) return static_cast<int>(e);
    }
    merge_shards_kernel<<<n_queries, 256, smem, stream>>>(d_gathered, n_shards, n_queries, per_shard, k_take, d_idx, d_score, d_found);
    return static_cast<int>(cudaGetLastError());
}

}  // namespace rag
