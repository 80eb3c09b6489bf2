// normalise.cu -- K1: batched upload/normalise kernel that fills the device-resident store.
//
// Replaces the store call `_chunks.Add(new RagChunk(id, file, text, emb))`
// (reference Helpers/RagIndex.cs:48) for a whole batch of embeddings, and hoists the per-row norm
// that the reference recomputes inside every Cosine call (RagIndex.cs:132,134) to ingest time.
//
// One warp per row.  Pass 1: sum of squares in fp64.  scale = fl32(1/sqrt(sum)) (1 for a zero or
// non-finite row, or when the caller asserts unit rows).  Pass 2: store x*scale as fp32 or bf16,
// measure the norm of what was actually stored and keep its reciprocal (the scan multiplies by
// it, so the approximate score is a true cosine of the stored row), and record whether the stored
// row is bit-identical to the input ("exact row": scale == 1.0f and, for bf16, a lossless round).
#include <math_constants.h>

#include "common.cuh"
#include "ragstore_internal.h"

namespace rag {

template <typename T>
__global__ void __launch_bounds__(256) normalise_rows_kernel(const NormaliseArgs a) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp0; r < a.n_rows; r += nwarps) {
        const float* src = a.src + r * a.src_ld;
        double acc = 0.0;
        for (int i = lane; i < a.dim; i += 32) {
            const double x = static_cast<double>(src[i]);
            acc += x * x;
        }
        const double n2 = warp_sum(acc);
        float scale = 1.0f;
        if (!a.prenormalised && n2 > 0.0 && n2 < CUDART_INF) scale = static_cast<float>(1.0 / sqrt(n2));
        T* dst = static_cast<T*>(a.dst) + r * static_cast<int64_t>(a.ld);
        double acc2 = 0.0;
        bool same = (scale == 1.0f);
        // fp32 store: "stored == input" is known now; keep the original values of an inexact row in
        // the sidecar BEFORE the in-place write destroys them.
        float* raw = nullptr;
        if (sizeof(T) == 4 && !same && a.raw_slot) {
            unsigned int slot = 0;
            if (lane == 0) slot = atomicAdd(a.raw_count, 1u);
            slot = __shfl_sync(0xffffffffu, slot, 0);
            raw = a.raw_rows + static_cast<size_t>(slot) * a.raw_ld;
            if (lane == 0) a.raw_slot[r] = static_cast<int32_t>(slot);
        }
        for (int i = lane; i < a.ld; i += 32) {
            const float x = i < a.dim ? src[i] : 0.f;  // read before the (possibly aliasing) write
            if (raw && i < a.raw_ld) raw[i] = x;
            const float y = __fmul_rn(x, scale);
            float stored;
            if constexpr (sizeof(T) == 4) {
                dst[i] = y;
                stored = y;
                if (a.shadow) a.shadow[r * static_cast<int64_t>(a.ld) + i] = static_cast<uint16_t>(f32_to_bf16_bits(y));
            } else {
                const uint32_t b = f32_to_bf16_bits(y);
                dst[i] = static_cast<uint16_t>(b);
                stored = bf16_bits_to_f32(b);
                same = same && (__float_as_uint(stored) == __float_as_uint(x));
            }
            acc2 += static_cast<double>(stored) * static_cast<double>(stored);
        }
        const double n2s = warp_sum(acc2);
        same = __all_sync(0xffffffffu, same);
        if (sizeof(T) == 2 && a.raw_slot) {  // bf16 store: exactness is only known after rounding
            if (!same) {
                unsigned int slot = 0;
                if (lane == 0) slot = atomicAdd(a.raw_count, 1u);
                slot = __shfl_sync(0xffffffffu, slot, 0);
                raw = a.raw_rows + static_cast<size_t>(slot) * a.raw_ld;
                for (int i = lane; i < a.raw_ld; i += 32) raw[i] = i < a.dim ? src[i] : 0.f;
                if (lane == 0) a.raw_slot[r] = static_cast<int32_t>(slot);
            }
        }
        if (a.raw_slot && same && lane == 0) a.raw_slot[r] = -1;
        if (lane == 0 && a.min_norm_bits && n2 > 0.0 && n2 < CUDART_INF) {
            // positive floats order like their bit patterns; round the norm DOWN so it stays a lower bound
            const uint32_t nb = __float_as_uint(__double2float_rd(sqrt(n2)));
            if (nb < *reinterpret_cast<volatile const uint32_t*>(a.min_norm_bits)) atomicMin(a.min_norm_bits, nb);
        }
        if (lane == 0) {
            a.rnorm[r] = (n2s > 0.0 && n2s < CUDART_INF) ? static_cast<float>(1.0 / sqrt(n2s)) : 0.f;
            a.exact[r] = same ? 1 : 0;
            if (!same && a.inexact_counter) atomicAdd(a.inexact_counter, 1ull);
        }
    }
}

// Vectorised single-pass variant for 16-byte-aligned rows of up to 1024 elements: every lane keeps its (up to)
// eight float4 of the row in registers between the norm and the store, so each element is read once and
// written once with 128-bit coalesced accesses -- the kernel is a pure HBM stream (2 * d * 4 bytes per fp32 row).
template <typename T>
__global__ void __launch_bounds__(256) normalise_rows_vec_kernel(const NormaliseArgs a) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    const int nch = a.dim >> 2;   // float4 chunks holding data (dim % 4 == 0)
    const int nch_ld = a.ld >> 2; // chunks of the stored fp32 row / half-chunks of the bf16 row (ld % 4 == 0)
    for (int64_t r = warp0; r < a.n_rows; r += nwarps) {
        const float4* src = reinterpret_cast<const float4*>(a.src + r * a.src_ld);
        float4 v[8];
        double acc = 0.0;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const int ch = c * 32 + lane;
            v[c] = ch < nch ? src[ch] : make_float4(0.f, 0.f, 0.f, 0.f);
            acc += static_cast<double>(v[c].x) * v[c].x + static_cast<double>(v[c].y) * v[c].y +
                   static_cast<double>(v[c].z) * v[c].z + static_cast<double>(v[c].w) * v[c].w;
        }
        const double n2 = warp_sum(acc);
        float scale = 1.0f;
        if (!a.prenormalised && n2 > 0.0 && n2 < CUDART_INF) scale = static_cast<float>(1.0 / sqrt(n2));
        bool same = (scale == 1.0f);
        float4* raw = nullptr;
        if (sizeof(T) == 4 && !same && a.raw_slot) {
            unsigned int slot = 0;
            if (lane == 0) slot = atomicAdd(a.raw_count, 1u);
            slot = __shfl_sync(0xffffffffu, slot, 0);
            raw = reinterpret_cast<float4*>(a.raw_rows + static_cast<size_t>(slot) * a.raw_ld);
            if (lane == 0) a.raw_slot[r] = static_cast<int32_t>(slot);
        }
        double acc2 = 0.0;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const int ch = c * 32 + lane;
            if (ch >= nch_ld) continue;
            const float4 x = v[c];
            if (raw && ch < (a.raw_ld >> 2)) raw[ch] = x;
            float4 y = make_float4(__fmul_rn(x.x, scale), __fmul_rn(x.y, scale), __fmul_rn(x.z, scale), __fmul_rn(x.w, scale));
            if constexpr (sizeof(T) == 4) {
                reinterpret_cast<float4*>(static_cast<float*>(a.dst) + r * static_cast<int64_t>(a.ld))[ch] = y;
                if (a.shadow) {   // the same pass also writes the bf16 image the tensor cores nominate from
                    const uint32_t b0 = f32_to_bf16_bits(y.x), b1 = f32_to_bf16_bits(y.y), b2 = f32_to_bf16_bits(y.z), b3 = f32_to_bf16_bits(y.w);
                    reinterpret_cast<uint2*>(a.shadow + r * static_cast<int64_t>(a.ld))[ch] = make_uint2(b0 | (b1 << 16), b2 | (b3 << 16));
                }
            } else {
                const uint32_t b0 = f32_to_bf16_bits(y.x), b1 = f32_to_bf16_bits(y.y), b2 = f32_to_bf16_bits(y.z), b3 = f32_to_bf16_bits(y.w);
                reinterpret_cast<uint2*>(static_cast<uint16_t*>(a.dst) + r * static_cast<int64_t>(a.ld))[ch] = make_uint2(b0 | (b1 << 16), b2 | (b3 << 16));
                y = make_float4(bf16_bits_to_f32(b0), bf16_bits_to_f32(b1), bf16_bits_to_f32(b2), bf16_bits_to_f32(b3));
                same = same && __float_as_uint(y.x) == __float_as_uint(x.x) && __float_as_uint(y.y) == __float_as_uint(x.y) &&
                       __float_as_uint(y.z) == __float_as_uint(x.z) && __float_as_uint(y.w) == __float_as_uint(x.w);
            }
            acc2 += static_cast<double>(y.x) * y.x + static_cast<double>(y.y) * y.y + static_cast<double>(y.z) * y.z +
                    static_cast<double>(y.w) * y.w;
        }
        const double n2s = warp_sum(acc2);
        same = __all_sync(0xffffffffu, same);
        if (sizeof(T) == 2 && a.raw_slot && !same) {  // bf16 store: exactness is only known after rounding
            unsigned int slot = 0;
            if (lane == 0) slot = atomicAdd(a.raw_count, 1u);
            slot = __shfl_sync(0xffffffffu, slot, 0);
            float4* rw = reinterpret_cast<float4*>(a.raw_rows + static_cast<size_t>(slot) * a.raw_ld);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const int ch = c * 32 + lane;
                if (ch < (a.raw_ld >> 2)) rw[ch] = v[c];
            }
            if (lane == 0) a.raw_slot[r] = static_cast<int32_t>(slot);
        }
        if (lane == 0) {
            if (a.raw_slot && same) a.raw_slot[r] = -1;
            if (a.min_norm_bits && n2 > 0.0 && n2 < CUDART_INF) {
                // one shared cell: look before the atomic, so that it fires only when the minimum really drops
                const uint32_t nb = __float_as_uint(__double2float_rd(sqrt(n2)));
                if (nb < *reinterpret_cast<volatile const uint32_t*>(a.min_norm_bits)) atomicMin(a.min_norm_bits, nb);
            }
            a.rnorm[r] = (n2s > 0.0 && n2s < CUDART_INF) ? static_cast<float>(1.0 / sqrt(n2s)) : 0.f;
            a.exact[r] = same ? 1 : 0;
            if (!same && a.inexact_counter) atomicAdd(a.inexact_counter, 1ull);
        }
    }
}

__global__ void __launch_bounds__(256) shadow_from_rows_kernel(const float4* __restrict__ src, uint2* __restrict__ dst, int64_t n4) {
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const float4 v = src[i];
        dst[i] = make_uint2(f32_to_bf16_bits(v.x) | (f32_to_bf16_bits(v.y) << 16), f32_to_bf16_bits(v.z) | (f32_to_bf16_bits(v.w) << 16));
    }
}

int launch_shadow_from_rows(const float* src, uint16_t* dst, int64_t n, cudaStream_t stream) {
    if (n <= 0) return 0;
    if (n % 4) return static_cast<int>(cudaErrorInvalidValue);   // rows are multiples of 8 elements
    shadow_from_rows_kernel<<<148 * 8, 256, 0, stream>>>(reinterpret_cast<const float4*>(src), reinterpret_cast<uint2*>(dst), n / 4);
    return static_cast<int>(cudaGetLastError());
}

int launch_normalise(const NormaliseArgs& a, cudaStream_t stream) {
    if (a.n_rows <= 0) return 0;
    const int threads = 256;
    int64_t blocks = (a.n_rows + 7) / 8;
    if (blocks > 148 * 16) blocks = 148 * 16;
    const bool vec = a.dim % 4 == 0 && a.dim <= 1024 && a.ld == a.dim && a.src_ld % 4 == 0 && a.raw_ld % 4 == 0 &&
                     (reinterpret_cast<uintptr_t>(a.src) & 15) == 0 && (reinterpret_cast<uintptr_t>(a.dst) & 15) == 0 &&
                     (a.dtype == RAGSTORE_F32 || a.ld % 8 == 0);
    if (vec && a.dtype == RAGSTORE_F32)
        normalise_rows_vec_kernel<float><<<static_cast<int>(blocks), threads, 0, stream>>>(a);
    else if (vec)
        normalise_rows_vec_kernel<uint16_t><<<static_cast<int>(blocks), threads, 0, stream>>>(a);
    else if (a.dtype == RAGSTORE_F32)
        normalise_rows_kernel<float><<<static_cast<int>(blocks), threads, 0, stream>>>(a);
    else
        normalise_rows_kernel<uint16_t><<<static_cast<int>(blocks), threads, 0, stream>>>(a);
    return static_cast<int>(cudaGetLastError());
}

}  // namespace rag
