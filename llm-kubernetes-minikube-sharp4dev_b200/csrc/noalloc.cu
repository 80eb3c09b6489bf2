// noalloc.cu -- a search entry point that never allocates: for hosts that must not call cudaMalloc.
//
// Who that is: a serving process that handed every byte of HBM to its own arena at start-up (the
// corpus plus whatever else it runs) and treats any later driver allocation as a bug; and a request
// thread with a latency budget, because cudaMalloc/cudaFree take driver-wide locks and synchronise
// the device, which stalls every search in flight on that GPU.  Such a host builds the matrix itself
// (same layout as the store) and only wants the scan + top-k + rescoring launch.
//
// The scan kernel needs scratch in global memory beyond its inputs and outputs (one candidate list
// per CTA plus the "last CTA" ticket).  Here that scratch is a fixed array in the library's device
// image, so every GPU that loads libragstore.so carries kNoAllocScratchBytes (about 321 KB) until
// the process exits, whether or not this entry point is ever called.
#include <cstdio>
#include <mutex>

#include "ragstore_internal.h"

namespace {

using namespace rag;

constexpr int kNoAllocMaxCtas = 160;   // >= the 148 SMs of a B200; one CTA per SM
constexpr size_t kNoAllocScratchBytes = 256 + sizeof(uint64_t) * size_t(kNoAllocMaxCtas) * kMaxKp;

// In host code this name is only a handle for the runtime, not a device address: it must never be
// handed to a kernel.  The address the kernels use comes from cudaGetSymbolAddress (below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
struct PerDevice {
    bool ready = false;
    unsigned char* scratch = nullptr;  // device address of g_noalloc_scratch on this device
    int sm_count = 0;
};
PerDevice g_dev[kMaxDevices];
std::mutex g_dev_mu;
thread_local char g_noalloc_err[256];

int set_err(int code, const char* msg) {
    snprintf(g_noalloc_err, sizeof(g_noalloc_err), "%s", msg);
    return code;
}

}  // namespace

extern "C" {

const char* ragstore_noalloc_last_error(void) { return g_noalloc_err; }

int64_t ragstore_noalloc_scratch_bytes(void) { return static_cast<int64_t>(kNoAllocScratchBytes); }

int ragstore_search_noalloc(const void* d_matrix, const float* d_rnorm, int64_t n_rows, int dim, int dtype,
                            const float* d_query, int top_k, int64_t* d_idx, double* d_score, int32_t* d_found,
                            int32_t* d_flags, int device, void* stream) {
    g_noalloc_err[0] = 0;
    if (!d_matrix || !d_rnorm || !d_query || !d_idx || !d_score || !d_found)
        return set_err(RAGSTORE_ERR_INVALID, "ragstore_search_noalloc: null pointer argument");
    if (dtype != RAGSTORE_F32 && dtype != RAGSTORE_BF16)
        return set_err(RAGSTORE_ERR_INVALID, "ragstore_search_noalloc: dtype must be RAGSTORE_F32 or RAGSTORE_BF16");
    const int esz = dtype == RAGSTORE_F32 ? 4 : 2;
    if (dim <= 0 || (static_cast<long long>(dim) * esz) % 16 != 0)
        return set_err(RAGSTORE_ERR_INVALID, "ragstore_search_noalloc: rows must be a multiple of 16 bytes (dim % 4 == 0 for fp32, % 8 for bf16)");
    if ((reinterpret_cast<uintptr_t>(d_matrix) & 15) || (reinterpret_cast<uintptr_t>(d_rnorm) & 15))
        return set_err(RAGSTORE_ERR_INVALID, "ragstore_search_noalloc: matrix and rnorm must be 16-byte aligned");
    if (n_rows <= 0 || n_rows >= 0xFFFFFFFFll)
        return set_err(RAGSTORE_ERR_UNSUPPORTED, "ragstore_search_noalloc: n_rows must be in [1, 2^32-2]");
    const int req = top_k < 1 ? 1 : top_k;
    if (req > kMaxKp)
        return set_err(RAGSTORE_ERR_UNSUPPORTED, "ragstore_search_noalloc: top_k above 256 does not fit the fixed scratch");
    if (device < 0 || device >= kMaxDevices)
        return set_err(RAGSTORE_ERR_INVALID, "ragstore_search_noalloc: device ordinal out of range");

    int prev = -1;
    cudaGetDevice(&prev);
    if (cudaSetDevice(device) != cudaSuccess) {
        cudaGetLastError();
        return set_err(RAGSTORE_ERR_NO_DEVICE, "ragstore_search_noalloc: cudaSetDevice failed (no CPU path exists)");
    }
    struct Restore {
        int prev;
        ~Restore() {
            if (prev >= 0) cudaSetDevice(prev);
        }
    } restore{prev};

    PerDevice pd;
    {
        // First call on a device: resolve the scratch array's address ON THAT DEVICE (it is current
        // here) and read the SM count.  Neither allocates.  Do this once before any stream capture.
        std::lock_guard<std::mutex> lk(g_dev_mu);
        if (!g_dev[device].ready) {
            void* p = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess || p == nullptr) {
                cudaGetLastError();
                return set_err(RAGSTORE_ERR_CUDA, "ragstore_search_noalloc: cudaGetSymbolAddress(scratch) failed");
            }
            int sms = 0;
            e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
            if (e != cudaSuccess || sms <= 0) {
                cudaGetLastError();
                return set_err(RAGSTORE_ERR_CUDA, "ragstore_search_noalloc: cannot read the SM count");
            }
            g_dev[device].scratch = static_cast<unsigned char*>(p);
            g_dev[device].sm_count = sms;
            g_dev[device].ready = true;
        }
        pd = g_dev[device];
    }
    if (pd.sm_count > kNoAllocMaxCtas)
        return set_err(RAGSTORE_ERR_UNSUPPORTED, "ragstore_search_noalloc: this GPU has more SMs than the fixed scratch covers (160)");

    const int take = static_cast<int>(n_rows < req ? n_rows : req);
    int kp = 32;
    {
        const int want = take + (take / 4 > 8 ? take / 4 : 8);
        while (kp < want && kp < kMaxKp) kp <<= 1;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // The array is shared by every call and keeps the previous call's bytes.  Only the ticket has to
    // start at zero (the candidate lists are fully rewritten by each launch); clear it in stream order.
    if (cudaMemsetAsync(pd.scratch, 0, 256, st) != cudaSuccess) {
        cudaGetLastError();
        return set_err(RAGSTORE_ERR_CUDA, "ragstore_search_noalloc: cudaMemsetAsync(ticket) failed");
    }
    ScanArgs a{};
    a.mat = d_matrix;
    a.rnorm = d_rnorm;
    a.n_rows = n_rows;
    a.dim = dim;
    a.ld = dim;
    a.dtype = dtype;
    a.d_query = d_query;
    a.k_take = take;
    a.kp = kp;
    {
        const int per_chunk = dtype == RAGSTORE_F32 ? 4 : 8;
        const int dv = ((dim + per_chunk - 1) / per_chunk + 31) / 32;
        a.eps = static_cast<float>((dv * per_chunk + 16) * 5.9604644775390625e-08);
    }
    // Certificate: the store-backed entry points know the smallest row norm they ever ingested and charge the reference's
    // "+1e-9" (RagIndex.cs:134) with it.  Here the rows are the caller's and unit norm by contract, so the term is charged
    // for rows no shorter than 1/2 -- 2e-9 / |q| -- which covers any rounding of a unit row with a wide margin.  A caller
    // that passes shorter rows still gets correctly rescored hits; only RAGSTORE_HIT_CERTIFIED may then promise too much.
    a.min_norm_assumed = 0.5f;
    a.scratch.ticket = reinterpret_cast<unsigned int*>(pd.scratch);
    a.scratch.stamps = reinterpret_cast<unsigned long long*>(pd.scratch + 64);
    a.scratch.cta_lists = reinterpret_cast<uint64_t*>(pd.scratch + 256);
    a.out_idx = d_idx;
    a.out_score = d_score;
    a.out_found = d_found;
    a.out_flags = d_flags;
    a.sm_count = pd.sm_count;
    const int rc = launch_scan(a, st);
    if (rc) {
        snprintf(g_noalloc_err, sizeof(g_noalloc_err), "ragstore_search_noalloc: kernel launch failed: %s",
                 cudaGetErrorString(static_cast<cudaError_t>(rc)));
        return rc == static_cast<int>(cudaErrorInvalidConfiguration) ? RAGSTORE_ERR_UNSUPPORTED : RAGSTORE_ERR_CUDA;
    }
    return RAGSTORE_OK;
}

}  // extern "C"
