// micro-benchmark: DADD dependent-chain latency and F2F.F64.F32 cost on B200 (informs the rescoring chains)
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_dadd(double* out, long long* cyc, double inc, int n, int active) {
    double acc = threadIdx.x;
    if ((int)threadIdx.x >= active) { return; }
    long long t0 = clock64();
    for (int i = 0; i < n; i += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) acc = __dadd_rn(acc, inc);
    }
    long long t1 = clock64();
    out[threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void k_f2f(double* out, long long* cyc, const float* in, int n) {
    double acc = 0;
    float f = in[threadIdx.x];
    long long t0 = clock64();
    for (int i = 0; i < n; i += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { acc = __dadd_rn(acc, (double)f); f = __fmul_rn(f, 1.0001f); }
    }
    long long t1 = clock64();
    out[threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
// the same chain with only `active` lanes of each warp alive: does the F2F cost scale with the lanes in use?
__global__ void k_f2f_lanes(double* out, long long* cyc, const float* in, int n, int active) {
    double acc = 0;
    float f = in[threadIdx.x];
    if ((int)(threadIdx.x & 31) >= active) return;
    long long t0 = clock64();
    for (int i = 0; i < n; i += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { acc = __dadd_rn(acc, (double)f); f = __fmul_rn(f, 1.0001f); }
    }
    long long t1 = clock64();
    out[threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void k_f2f_only(double* out, long long* cyc, const float* in, int n) {
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    float f = in[threadIdx.x];
    long long t0 = clock64();
    for (int i = 0; i < n; i += 4) {
        a0 += (double)f; f = __fmul_rn(f, 1.0001f);
        a1 += (double)f; f = __fmul_rn(f, 1.0001f);
        a2 += (double)f; f = __fmul_rn(f, 1.0001f);
        a3 += (double)f; f = __fmul_rn(f, 1.0001f);
    }
    long long t1 = clock64();
    out[threadIdx.x] = a0 + a1 + a2 + a3;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
    double* out; long long* cyc; float* in;
    cudaMalloc(&out, 8 * 1024); cudaMalloc(&cyc, 8 * 16); cudaMalloc(&in, 4 * 1024);
    cudaMemset(in, 0x3f, 4 * 1024);
    long long h;
    int n = 4096;
    for (int active : {1, 32}) {
        for (int warps : {1, 3, 8}) {
            k_dadd<<<1, 32 * warps>>>(out, cyc, 1e-9, n, active == 1 ? 1 : 32 * warps);
            cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
            printf("DADD chain: warps=%d lanes=%d  %.2f cycles/op\n", warps, active, (double)h / n);
        }
    }
    for (int warps : {1, 3}) {
        k_f2f<<<1, 32 * warps>>>(out, cyc, in, n);
        cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("DADD chain + F2F + FMUL per op: warps=%d  %.2f cycles/op\n", warps, (double)h / n);
        k_f2f_only<<<1, 32 * warps>>>(out, cyc, in, n);
        cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("4 independent (F2F + DADD) chains: warps=%d  %.2f cycles/op\n", warps, (double)h / n);
    }
    for (int warps : {1, 3, 9}) {
        for (int active : {32, 16, 8, 4, 1}) {
            k_f2f_lanes<<<1, 32 * warps>>>(out, cyc, in, n, active);
            cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
            printf("DADD chain + F2F + FMUL per op: warps=%d active lanes per warp=%d  %.2f cycles/op\n", warps, active, (double)h / n);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
