// synth.cu -- synthetic benchmark corpus generated on the device; bit-identical
// to oracle/ragindex_oracle.c:rag_synth_rows and oracle/oracle_np.py:synth_rows.  Benchmark
// plumbing, not a reference function: it keeps a 30 GB corpus from crossing PCIe.
#include "common.cuh"
#include "ragstore_internal.h"

namespace rag {

__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ int synth_sample(uint64_t row_key, int64_t col) {
    const uint64_t u = splitmix64(row_key ^ (static_cast<uint64_t>(col) * 0xD1B54A32D192ED03ull + 0x8CB92BA72F3D8DD7ull));
    const int s = static_cast<int>(u & 0xFFFF) + static_cast<int>((u >> 16) & 0xFFFF) +
                  static_cast<int>((u >> 32) & 0xFFFF) + static_cast<int>((u >> 48) & 0xFFFF);
    return s - 131070;
}

__global__ void __launch_bounds__(256) synth_rows_kernel(float* out, int64_t row0, int64_t n_rows, int dim,
                                                         int64_t ld, uint64_t seed, int round_bf16) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp0; r < n_rows; r += nwarps) {
        const uint64_t row_key = splitmix64(seed + 0x632BE59BD9B4E019ull * static_cast<uint64_t>(row0 + r + 1));
        long long n2 = 0;
        for (int c = lane; c < dim; c += 32) {
            const long long s = synth_sample(row_key, c);
            n2 += s * s;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) n2 += __shfl_xor_sync(0xffffffffu, n2, o);
        const double norm = __dsqrt_rn(static_cast<double>(n2));
        float* o = out + r * ld;
        for (int c = lane; c < ld; c += 32) {
            float f = 0.f;
            if (c < dim && n2 != 0) f = static_cast<float>(__ddiv_rn(static_cast<double>(synth_sample(row_key, c)), norm));
            if (round_bf16) f = bf16_bits_to_f32(f32_to_bf16_bits(f));
            o[c] = f;
        }
    }
}

int launch_synth(float* d_out, int64_t row0, int64_t n_rows, int dim, int64_t ld, uint64_t seed, int round_bf16,
                 cudaStream_t stream) {
    if (n_rows <= 0) return 0;
    int64_t blocks = (n_rows + 7) / 8;
    if (blocks > 148 * 32) blocks = 148 * 32;
    synth_rows_kernel<<<static_cast<int>(blocks), 256, 0, stream>>>(d_out, row0, n_rows, dim, ld, seed, round_bf16);
    return static_cast<int>(cudaGetLastError());
}

}  // namespace rag
