// exchange.cu -- the one exchange step of the row-sharded search as ONE kernel over peer memory.
//
// After its shard scan every rank holds n_queries x per_shard exactly-rescored hit records.  Instead of an NCCL
// all-gather followed by a merge launch, one kernel per rank (i) stores its records straight into every peer's inbox
// over NVLink (plain stores to IPC-mapped peer memory) and raises a per-query sequence flag there, (ii) waits until the
// flags of all peers have arrived in its own inbox, (iii) ranks the G x per_shard records of each query
// (merge_device.cuh; order of Helpers/RagIndex.cs:64-65).  Payload is G x k x 16 bytes per query: pure latency, so the
// win is the launches and the collective's fixed cost removed from every query.
//
// Inbox of one rank (one cudaMalloc, shared through a CUDA IPC handle):
//     flags   [2 slots][world][max_queries]  uint64   sequence number of the call that filled the records
//     records [2 slots][world][max_records]  ragstore_cand
// Calls alternate between the two slots.  Two are enough: a rank can only start call s+1 after its own kernel of call s
// finished, i.e. after EVERY rank had published call s, i.e. after every rank's kernel of call s-1 finished -- so nobody
// still reads the slot that call s+1 overwrites.
// Liveness: a block publishes ALL of its queries before it waits for anything and the grid never exceeds the SM count
// (every block is resident), so no wait depends on a block that cannot run; waits are bounded by a timeout that is
// reported in d_status instead of hanging the GPU.
#include <nvtx3/nvToolsExt.h>

#include <cstring>
#include <new>

#include "exchange_impl.h"

namespace rag {

__global__ void __launch_bounds__(256) exchange_merge_kernel(const ExchangeView v, const ragstore_cand* __restrict__ mine, int n_queries,
                                                             int per_shard, int k_req, unsigned long long seq,
                                                             unsigned long long timeout_ns, int64_t* out_idx, double* out_score,
                                                             int32_t* out_found, int32_t* out_status, const int32_t* __restrict__ my_flags,
                                                             int32_t* out_flags) {
    extern __shared__ __align__(16) uint8_t sm[];
    __shared__ int n_valid;
    __shared__ int timed_out;
    __shared__ int all_flags;
    const int tid = threadIdx.x;
    const int slot = static_cast<int>(seq & 1ull);
    const size_t rec_base = (static_cast<size_t>(slot) * v.world + v.rank) * v.max_records;   // my block in every inbox
    const size_t flag_base = (static_cast<size_t>(slot) * v.world + v.rank) * v.max_queries;

    // ---- publish every query this block owns, to every rank (own inbox included)
    for (int q = blockIdx.x; q < n_queries; q += gridDim.x) {
        const uint4* src = reinterpret_cast<const uint4*>(mine + static_cast<size_t>(q) * per_shard);
        for (int i = tid; i < v.world * per_shard; i += blockDim.x) {
            const int r = i / per_shard, j = i % per_shard;
            reinterpret_cast<uint4*>(v.rec[r] + rec_base + static_cast<size_t>(q) * per_shard)[j] = src[j];
        }
    }
    __threadfence_system();
    __syncthreads();
    // The arrival word carries, below the sequence number, this rank's RAGSTORE_HIT_* bits for the query: a result is
    // certified / exact only if every shard's part of it is, so the certificate travels with the hits.
    for (int q = blockIdx.x; q < n_queries; q += gridDim.x)
        if (tid < v.world) st_release_sys(v.flag[tid] + flag_base + q, (seq << 2) | static_cast<unsigned long long>((my_flags ? my_flags[q] : 3) & 3));

    // ---- wait for every rank's records of a query, then rank them
    const ragstore_cand* inbox = v.rec[v.rank] + static_cast<size_t>(slot) * v.world * v.max_records;
    const unsigned long long* inbox_flags = v.flag[v.rank] + static_cast<size_t>(slot) * v.world * v.max_queries;
    for (int q = blockIdx.x; q < n_queries; q += gridDim.x) {
        if (tid == 0) timed_out = 0, all_flags = 3;
        __syncthreads();
        if (tid < v.world) {
            const unsigned long long* f = inbox_flags + static_cast<size_t>(tid) * v.max_queries + q;
            const unsigned long long t0 = globaltimer_ns();
            unsigned long long w;
            while (((w = ld_acquire_sys(f)) >> 2) != seq) {
                if (globaltimer_ns() - t0 > timeout_ns) {
                    timed_out = 1;
                    w = 0;
                    break;
                }
                __nanosleep(64);
            }
            atomicAnd(&all_flags, static_cast<int>(w & 3ull));
        }
        __syncthreads();
        if (tid == 0 && out_status) out_status[q] = timed_out;
        if (tid == 0 && out_flags) out_flags[q] = all_flags;
        merge_query_block<true>(inbox, static_cast<size_t>(v.max_records), v.world, q, per_shard, k_req, out_idx, out_score, out_found, sm,
                          &n_valid);
    }
}

}  // namespace rag

using namespace rag;

namespace {
struct DeviceScope {
    int prev = -1;
    bool ok = false;
    explicit DeviceScope(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev == dev) {
            prev = -1;  // already there: nothing to switch, nothing to restore
            ok = true;
        } else {
            ok = cudaSetDevice(dev) == cudaSuccess;
        }
    }
    ~DeviceScope() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};
}  // namespace

extern "C" {

int ragstore_exchange_create(int device, int world, int rank, int max_records, int max_queries, ragstore_exchange_t** out,
                             void* handle_out) {
    if (!out || !handle_out) return api_fail(RAGSTORE_ERR_INVALID, "ragstore_exchange_create: null argument");
    *out = nullptr;
    if (world < 1 || world > kMaxWorld || rank < 0 || rank >= world || max_records < 1 || max_queries < 1 || max_queries > max_records)
        return api_fail(RAGSTORE_ERR_INVALID, "ragstore_exchange_create: need 1 <= world <= 16, 0 <= rank < world, 1 <= max_queries <= max_records");
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || device < 0 || device >= n_dev) {
        cudaGetLastError();
        return api_fail(RAGSTORE_ERR_NO_DEVICE, "ragstore_exchange_create: no such CUDA device");
    }
    DeviceScope g(device);
    if (!g.ok) return api_fail(RAGSTORE_ERR_CUDA, "ragstore_exchange_create: cudaSetDevice failed");
    ragstore_exchange* x = new (std::nothrow) ragstore_exchange();
    if (!x) return api_fail(RAGSTORE_ERR_CUDA, "ragstore_exchange_create: out of host memory");
    // strides between the per-sender regions are whole 128-byte lines: no two senders ever write the same line
    max_records = (max_records + 7) & ~7;
    max_queries = (max_queries + 15) & ~15;
    x->device = device, x->world = world, x->rank = rank, x->max_records = max_records, x->max_queries = max_queries;
    cudaDeviceGetAttribute(&x->sm_count, cudaDevAttrMultiProcessorCount, device);
    x->flag_bytes = (sizeof(unsigned long long) * 2u * world * max_queries + 255u) & ~size_t(255);
    x->bytes = x->flag_bytes + sizeof(ragstore_cand) * 2u * world * static_cast<size_t>(max_records);
    cudaError_t e = cudaMalloc(&x->base, x->bytes);
    if (e == cudaSuccess) e = cudaMemset(x->base, 0, x->bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, x->base);
    if (e != cudaSuccess) {
        if (x->base) cudaFree(x->base);
        delete x;
        return api_fail(RAGSTORE_ERR_CUDA, "ragstore_exchange_create", cudaGetErrorString(e));
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == RAGSTORE_EXCHANGE_HANDLE_BYTES, "handle size");
    std::memcpy(handle_out, &h, sizeof(h));
    *out = x;
    return RAGSTORE_OK;
}

int ragstore_exchange_connect(ragstore_exchange_t* x, const void* all_handles) {
    if (!x || !all_handles) return api_fail(RAGSTORE_ERR_INVALID, "ragstore_exchange_connect: null argument");
    if (x->connected) return api_fail(RAGSTORE_ERR_INVALID, "ragstore_exchange_connect: already connected");
    DeviceScope g(x->device);
    if (!g.ok) return api_fail(RAGSTORE_ERR_CUDA, "ragstore_exchange_connect: cudaSetDevice failed");
    for (int r = 0; r < x->world; ++r) {
        if (r == x->rank) {
            x->peer[r] = x->base;
            continue;
        }
        cudaIpcMemHandle_t h;
        std::memcpy(&h, static_cast<const uint8_t*>(all_handles) + static_cast<size_t>(r) * sizeof(h), sizeof(h));
        const cudaError_t e = cudaIpcOpenMemHandle(&x->peer[r], h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            for (int u = 0; u < r; ++u)
                if (u != x->rank && x->peer[u]) cudaIpcCloseMemHandle(x->peer[u]), x->peer[u] = nullptr;
            cudaGetLastError();
            return api_fail(RAGSTORE_ERR_CUDA, "ragstore_exchange_connect: cannot map a peer inbox (no peer access between these GPUs?)",
                            cudaGetErrorString(e));
        }
    }
    x->view.world = x->world, x->view.rank = x->rank, x->view.max_records = x->max_records, x->view.max_queries = x->max_queries;
    for (int r = 0; r < x->world; ++r) {
        x->view.flag[r] = static_cast<unsigned long long*>(x->peer[r]);
        x->view.rec[r] = reinterpret_cast<ragstore_cand*>(static_cast<uint8_t*>(x->peer[r]) + x->flag_bytes);
    }
    {
        cudaError_t e = cudaMalloc(&x->d_view, sizeof(ExchangeView));
        if (e == cudaSuccess) e = cudaMemcpy(x->d_view, &x->view, sizeof(ExchangeView), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) return api_fail(RAGSTORE_ERR_CUDA, "ragstore_exchange_connect: device copy of the view", cudaGetErrorString(e));
    }
    x->connected = true;
    return RAGSTORE_OK;
}

int ragstore_exchange_merge(ragstore_exchange_t* x, const ragstore_cand* d_mine, int n_queries, int per_shard, int top_k,
                            int64_t* d_idx, double* d_score, int32_t* d_found, int32_t* d_status, void* stream) {
    return ragstore_exchange_merge_flags(x, d_mine, nullptr, n_queries, per_shard, top_k, d_idx, d_score, d_found, nullptr, d_status, stream);
}

int ragstore_exchange_merge_flags(ragstore_exchange_t* x, const ragstore_cand* d_mine, const int32_t* d_my_flags, int n_queries,
                                  int per_shard, int top_k, int64_t* d_idx, double* d_score, int32_t* d_found, int32_t* d_flags,
                                  int32_t* d_status, void* stream) {
    if (!x || !d_mine || !d_idx || !d_score || !d_found) return api_fail(RAGSTORE_ERR_INVALID, "ragstore_exchange_merge: null argument");
    if (!x->connected) return api_fail(RAGSTORE_ERR_INVALID, "ragstore_exchange_merge: call ragstore_exchange_connect first");
    const int req = top_k < 1 ? 1 : top_k;
    if (n_queries < 1 || per_shard < 1 || n_queries > x->max_queries || static_cast<int64_t>(n_queries) * per_shard > x->max_records ||
        x->world * per_shard > 8192)
        return api_fail(RAGSTORE_ERR_INVALID, "ragstore_exchange_merge: n_queries / per_shard exceed what the exchange was created for");
    nvtxRangePushA("ragstore_exchange_merge: peer stores + flags + merge");
    struct Pop {
        ~Pop() { nvtxRangePop(); }
    } pop;
    DeviceScope g(x->device);
    if (!g.ok) return api_fail(RAGSTORE_ERR_CUDA, "ragstore_exchange_merge: cudaSetDevice failed");
    const size_t smem = static_cast<size_t>(x->world) * per_shard * 24;
    if (smem > 48 * 1024) {
        const cudaError_t e = cudaFuncSetAttribute(exchange_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return api_fail(RAGSTORE_ERR_CUDA, "ragstore_exchange_merge: shared memory", cudaGetErrorString(e));
    }
    const int grid = n_queries < x->sm_count ? n_queries : x->sm_count;   // every block resident: see the liveness note above
    const unsigned long long seq = ++x->seq;
    exchange_merge_kernel<<<grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(x->view, d_mine, n_queries, per_shard, req, seq,
                                                                                  5000000000ull /* 5 s */, d_idx, d_score, d_found,
                                                                                  d_status, d_my_flags, d_flags);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return api_fail(RAGSTORE_ERR_CUDA, "ragstore_exchange_merge: launch", cudaGetErrorString(e));
    return RAGSTORE_OK;
}

int ragstore_exchange_destroy(ragstore_exchange_t* x) {
    if (!x) return RAGSTORE_OK;
    DeviceScope g(x->device);
    if (g.ok) {
        cudaDeviceSynchronize();
        for (int r = 0; r < x->world; ++r)
            if (r != x->rank && x->peer[r]) cudaIpcCloseMemHandle(x->peer[r]);
        if (x->base) cudaFree(x->base);
        if (x->d_view) cudaFree(x->d_view);
        cudaGetLastError();
    }
    delete x;
    return RAGSTORE_OK;
}

}  // extern "C"
