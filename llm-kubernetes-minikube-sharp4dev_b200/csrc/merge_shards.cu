// merge_shards.cu -- global top-k over the all-gathered per-shard hit lists; the ranking itself is
// merge_device.cuh (reference order: Helpers/RagIndex.cs:64-65).
#include <math_constants.h>

#include "merge_device.cuh"

namespace rag {

// gathered layout: [shard][query][per_shard]; one CTA per query (blockIdx.x)
__global__ void __launch_bounds__(256) merge_shards_kernel(const ragstore_cand* __restrict__ in, int n_shards, int n_queries,
                                                           int per_shard, int k_req, int64_t* out_idx, double* out_score,
                                                           int32_t* out_found) {
    extern __shared__ __align__(16) uint8_t sm[];
    __shared__ int n_valid;
    merge_query_block(in, static_cast<size_t>(n_queries) * per_shard, n_shards, blockIdx.x, per_shard, k_req, out_idx, out_score,
                      out_found, sm, &n_valid);
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
