// merge_device.cuh -- one CTA ranks the per-shard hit lists of one query (shared by merge_shards.cu and exchange.cu).
//
// Each shard contributes its own exactly-rescored top-k as (fp64 score, global row id) records;
// the union contains the global top-k, and (score desc, id asc) is the order a single RagIndex
// over the concatenated corpus would produce (reference Helpers/RagIndex.cs:64-65: stable
// OrderByDescending + Take).  Every record is ranked by counting the records that beat it.
#pragma once
#include "common.cuh"
#include "ragstore_internal.h"

namespace rag {

// in: shard s's records of query qi start at in[s * shard_stride + qi * per_shard].
// sm: 24 bytes per record (n_shards * per_shard records) of dynamic shared memory; n_valid: one shared int.
// FRESH: read the records with ld.global.cv (never from a possibly stale L1 line) -- for records that a peer GPU wrote
// into this GPU's memory while the kernel was already running (exchange.cu).
template <bool FRESH = false>
__device__ __forceinline__ void merge_query_block(const ragstore_cand* in, size_t shard_stride, int n_shards, int qi, int per_shard,
                                                  int k_req, int64_t* out_idx, double* out_score, int32_t* out_found, uint8_t* sm,
                                                  int* n_valid) {
    const int n_items = n_shards * per_shard;
    uint64_t* keys = reinterpret_cast<uint64_t*>(sm);
    int64_t* ids = reinterpret_cast<int64_t*>(keys + n_items);
    double* scs = reinterpret_cast<double*>(ids + n_items);
    out_idx += static_cast<size_t>(qi) * k_req;
    out_score += static_cast<size_t>(qi) * k_req;
    out_found += qi;
    if (threadIdx.x == 0) *n_valid = 0;
    __syncthreads();
    int local_valid = 0;
    for (int i = threadIdx.x; i < n_items; i += blockDim.x) {
        const int sh = i / per_shard, j = i % per_shard;
        const ragstore_cand* p = in + static_cast<size_t>(sh) * shard_stride + static_cast<size_t>(qi) * per_shard + j;
        double score;
        int64_t id;
        if (FRESH) {
            score = __ldcv(&p->score);
            id = static_cast<int64_t>(__ldcv(reinterpret_cast<const long long*>(&p->id)));
        } else {
            score = p->score;
            id = p->id;
        }
        scs[i] = score;
        keys[i] = ord_f64(score);
        ids[i] = id;
        local_valid += (id >= 0);
    }
    if (local_valid) atomicAdd(n_valid, local_valid);
    __syncthreads();
    const int take = min(*n_valid, k_req);
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
    __syncthreads();  // the caller may reuse sm / n_valid for its next query
}

}  // namespace rag
