// Shared device / host helpers of the memory-bound token kernels (token_ops.cu, adv_ops.cu): 128-bit packs, grid
// sizing, the "last block done" finalisation pattern.
#pragma once
#include <cuda_runtime.h>

#include <initializer_list>

#include "common.h"
#include "reduce.cuh"

namespace vpg {

constexpr int kThreads = 256;

template <int VEC>
struct Pack {
    float v[VEC];
};

template <int VEC>
__device__ __forceinline__ Pack<VEC> ldv(const float* __restrict__ p, long long i) {
    Pack<VEC> r;
    if constexpr (VEC == 4) {
        const float4 a = load_f4(p, i);
        r.v[0] = a.x, r.v[1] = a.y, r.v[2] = a.z, r.v[3] = a.w;
    } else {
        r.v[0] = p[i];
    }
    return r;
}
template <int VEC, typename MT>
__device__ __forceinline__ Pack<VEC> ldm(const MT* __restrict__ p, long long i) {
    Pack<VEC> r;
    if constexpr (VEC == 4) {
        const float4 a = load_mask4<MT>(p, i);
        r.v[0] = a.x, r.v[1] = a.y, r.v[2] = a.z, r.v[3] = a.w;
    } else {
        r.v[0] = mask_to_float(p[i]);
    }
    return r;
}
template <int VEC>
__device__ __forceinline__ void stv(float* __restrict__ p, long long i, const Pack<VEC>& r) {
    if constexpr (VEC == 4) {
        store_f4(p, i, make_float4(r.v[0], r.v[1], r.v[2], r.v[3]));
    } else {
        p[i] = r.v[0];
    }
}

// Blocks per row.  The grid should be MANY waves of the resident blocks (4-8 per SM) whenever the problem allows it: with
// ~2.5 waves (2048 rows, one block each) the last, partial wave costs 10-25 % of the sweep.  Target: `VPG_TOK_BLOCKS_PER_SM`
// blocks per SM in total, never less than two iterations of work per block.  Default 8: more blocks cost more than the
// partial last wave in every kernel that ends in per-block atomics (A/B at 8..128 per SM: profiles/r2_ab_token_grid.txt).
static inline int blocks_per_row(int64_t B, int64_t L, int vec) {
    static const int per_sm = env_int("VPG_TOK_BLOCKS_PER_SM", 8);
    const int64_t per_block = static_cast<int64_t>(kThreads) * vec * 2;
    int64_t need = ceil_div(L, per_block);
    int64_t want = ceil_div(static_cast<int64_t>(sm_count()) * per_sm, B > 0 ? B : 1);
    int64_t g = need < want ? need : want;
    return static_cast<int>(g < 1 ? 1 : g);
}

// "last block done": returns true in every thread of the block that arrives last
__device__ __forceinline__ bool last_block_done(unsigned int* counter, unsigned int total) {
    __shared__ bool is_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int prev = atomicAdd(counter, 1u);
        is_last = (prev == total - 1);
    }
    __syncthreads();
    if (is_last) __threadfence();
    return is_last;
}

// torch.clamp / torch.maximum / torch.minimum propagate NaN; fminf / fmaxf drop it.  A NaN log-prob must surface as a
// NaN loss exactly like in the reference (it is how a diverged run is noticed), so the NaN-propagating forms are used.
// `max.NaN.f32` / `min.NaN.f32` are ONE instruction each (FMNMX.NAN) where the select-on-isnan form costs four to five;
// the loss kernels are issue-bound, not HBM-bound, without them (ncu: sm throughput 66-72 % at 55 % of DRAM peak).
__device__ __forceinline__ float max_nan(float a, float b) {
    float d;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}
__device__ __forceinline__ float min_nan(float a, float b) {
    float d;
    asm("min.NaN.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return min_nan(max_nan(x, lo), hi); }

static inline int flat_grid(int64_t n, int vec) {
    int64_t g = ceil_div(n, static_cast<int64_t>(kThreads) * vec * 4);
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    if (g > cap) g = cap;
    return static_cast<int>(g < 1 ? 1 : g);
}


// vector path requires 16-byte aligned rows
static inline bool rows_vec_ok(int64_t L, const void* mask, int mask_dtype, std::initializer_list<const void*> ptrs) {
    if (L % 4 != 0) return false;
    if (mask && !mask_aligned(mask, mask_dtype)) return false;
    for (const void* p : ptrs)
        if (p && !aligned16(p)) return false;
    return true;
}


// masked_whiten's error status (torch_functional.py:195-200: the reference raises when the mask sums to 0 or 1),
// published by a one-thread kernel BETWEEN the statistics pass and the normalisation pass: `status` may be a pinned
// (UVA-mapped) host word and `status_event` a cudaEvent_t recorded right behind it, so the host can learn the
// status — and raise like the reference — while the normalisation pass is still running.
static __global__ void whiten_status_kernel(const double* __restrict__ sums, int* status) {
    const double N = sums[2];
    *status = (N == 0.0) ? 1 : (N == 1.0) ? 2 : 0;
    __threadfence_system();
}
static inline cudaError_t publish_whiten_status(const double* sums, int* status, void* status_event,
                                                cudaStream_t stream) {
    whiten_status_kernel<<<1, 1, 0, stream>>>(sums, status);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess && status_event) e = cudaEventRecord(static_cast<cudaEvent_t>(status_event), stream);
    return e;
}

// Warp-shuffle segmented reduction: sum of `v` over the lanes in `peers` (the lane set `__match_any_sync` returned for this
// lane's key), accumulated in lane order so the result does not depend on which lane asks.  All 32 lanes must call.
__device__ __forceinline__ double segment_sum(double v, unsigned peers) {
    double acc = 0.0;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const double o = __shfl_sync(0xffffffffu, v, j);
        if ((peers >> j) & 1u) acc += o;
    }
    return acc;
}

}  // namespace vpg
