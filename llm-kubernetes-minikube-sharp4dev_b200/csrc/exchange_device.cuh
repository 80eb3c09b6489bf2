// exchange_device.cuh -- device-side view of the peer-memory exchange (csrc/exchange.cu): every rank's inbox as seen from
// this GPU, and the system-scope flag accesses.  Shared by the stand-alone exchange kernel and by the scan kernel's last
// CTA, which can publish, wait and merge itself (one launch per query per rank).
#pragma once
#include "merge_device.cuh"

namespace rag {

constexpr int kMaxWorld = 16;

struct ExchangeView {
    ragstore_cand* rec[kMaxWorld];        // rank r's inbox records
    unsigned long long* flag[kMaxWorld];  // rank r's inbox flags
    int world, rank, max_records, max_queries;
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

}  // namespace rag
