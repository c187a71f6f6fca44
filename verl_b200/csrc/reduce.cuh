// Warp / block reductions and typed mask loads for the memory-bound token kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/vpg_b200.h"

namespace vpg {

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Sum over the block; result valid in thread 0.  `smem` must hold 32 values of T.  Safe to call repeatedly.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* smem) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) smem[w] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    v = (threadIdx.x < nw) ? smem[threadIdx.x] : T(0);
    if (w == 0) v = warp_sum(v);
    return v;
}

// N block sums behind ONE pair of barriers (block_sum pays two per value).  Same summation order as block_sum, value by
// value; results valid in thread 0.  `smem` must hold N * 32 values of T.  Safe to call repeatedly.
template <int N, typename T>
__device__ __forceinline__ void block_sum_n(T (&v)[N], T* smem) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < N; ++k) v[k] = warp_sum(v[k]);
    __syncthreads();
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < N; ++k) smem[k * 32 + w] = v[k];
    }
    __syncthreads();
    if (w == 0) {
        const int nw = (blockDim.x + 31) >> 5;
#pragma unroll
        for (int k = 0; k < N; ++k) v[k] = warp_sum((lane < nw) ? smem[k * 32 + lane] : T(0));
    }
}

// ---- mask element -> float
template <typename MT>
__device__ __forceinline__ float mask_to_float(MT m) {
    return static_cast<float>(m);
}

// load 4 consecutive mask elements starting at idx (idx % 4 == 0 and base 16-byte aligned for the vector paths)
template <typename MT>
__device__ __forceinline__ float4 load_mask4(const MT* __restrict__ p, long long idx);

template <>
__device__ __forceinline__ float4 load_mask4<long long>(const long long* __restrict__ p, long long idx) {
    const longlong2 a = *reinterpret_cast<const longlong2*>(p + idx);
    const longlong2 b = *reinterpret_cast<const longlong2*>(p + idx + 2);
    return make_float4((float)a.x, (float)a.y, (float)b.x, (float)b.y);
}
template <>
__device__ __forceinline__ float4 load_mask4<int>(const int* __restrict__ p, long long idx) {
    const int4 a = *reinterpret_cast<const int4*>(p + idx);
    return make_float4((float)a.x, (float)a.y, (float)a.z, (float)a.w);
}
template <>
__device__ __forceinline__ float4 load_mask4<unsigned char>(const unsigned char* __restrict__ p, long long idx) {
    const uchar4 a = *reinterpret_cast<const uchar4*>(p + idx);
    return make_float4((float)a.x, (float)a.y, (float)a.z, (float)a.w);
}
template <>
__device__ __forceinline__ float4 load_mask4<float>(const float* __restrict__ p, long long idx) {
    return *reinterpret_cast<const float4*>(p + idx);
}

__device__ __forceinline__ float4 load_f4(const float* __restrict__ p, long long idx) {
    return *reinterpret_cast<const float4*>(p + idx);
}
__device__ __forceinline__ void store_f4(float* __restrict__ p, long long idx, float4 v) {
    *reinterpret_cast<float4*>(p + idx) = v;
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static inline size_t mask_elem_size(int dt) {
    return dt == VPG_MASK_I64 ? 8 : dt == VPG_MASK_I32 ? 4 : dt == VPG_MASK_U8 ? 1 : 4;
}
// mask vector loads need 4 elements to be 16-byte (i64: 2x16), 16, 4, 16 byte aligned
static inline bool mask_aligned(const void* p, int dt) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    return dt == VPG_MASK_U8 ? (a & 3) == 0 : (a & 15) == 0;
}

#define VPG_DISPATCH_MASK(dt, MT, ...)                                                   \
    switch (dt) {                                                                        \
        case VPG_MASK_I64: { using MT = long long; __VA_ARGS__; } break;                 \
        case VPG_MASK_I32: { using MT = int; __VA_ARGS__; } break;                       \
        case VPG_MASK_U8: { using MT = unsigned char; __VA_ARGS__; } break;              \
        case VPG_MASK_F32: { using MT = float; __VA_ARGS__; } break;                     \
        default: return ::vpg::set_error(VPG_ERR_INVALID, "unknown mask dtype %d", dt);  \
    }

}  // namespace vpg
