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
// ~2.5 waves (2048 rows, one block each) the last, partial wave costs 10-25 % of the sweep — but every block ends in
// barriers and atomics, so how finely a sweep may be cut depends on its epilogue.  `per_sm` is the kernel's own target of
// blocks per SM in total (never less than two iterations of work per block), from the A/B at 8 / 16 / 32 per SM after the
// epilogues were fused (profiles/r2_ab_token_grid_after_epilogue_fusion.txt): 8 for the fused loss (seven block sums), 16 for the sweeps with
// two or three sums, 32 for the ones that end in one atomic or none.  `VPG_TOK_BLOCKS_PER_SM` overrides all of them.
static inline int blocks_per_row(int64_t B, int64_t L, int vec, int per_sm = 8) {
    static const int forced = env_int("VPG_TOK_BLOCKS_PER_SM", 0);
    if (forced > 0) per_sm = forced;
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

// Threads per block of the row-scan kernels (GAE, reverse scans): a block walks its row in chunks of threads * 8 tokens,
// load -> compose -> barrier -> replay -> store, every chunk waiting for the carry of the one to its right.  The phases
// of a block do not overlap each other, so what keeps HBM busy is the NUMBER of independent blocks on an SM, not their
// size.  `VPG_SCAN_THREADS` (32..512, multiple of 32) overrides the choice below.
static inline int scan_block_threads(int64_t rows) {
    static const int forced = [] {
        int v = env_int("VPG_SCAN_THREADS", 0);
        v = (v / 32) * 32;
        return v <= 0 ? 0 : (v > 512 ? 512 : v);
    }();
    if (forced) return forced;
    // many rows: 256 threads, four blocks per SM (A/B at 2048 rows: GAE 0.73 -> 0.81, reverse scans 0.79 -> 0.85 of HBM
    // peak); few rows (16 x 16384 at BASELINE configs[4]): the row is the critical path, the widest block is the fastest
    return rows >= 2 * static_cast<int64_t>(sm_count()) ? 256 : 512;
}

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


// masked_whiten's error status (torch_functional.py:195-200: the reference raises when the mask sums to 0 or 1).  The
// kernel that produces the statistics publishes it from its LAST block (sums[3] is the arrival counter, zeroed with the
// sums): `status` may be a pinned (UVA-mapped) host word and `status_event` a cudaEvent_t recorded right behind that
// kernel, so the host learns the status — and raises like the reference — while the normalisation pass is still
// running.  No launch of its own (the one-thread status kernel of round 1 cost 5.5 us between the two passes).
__device__ __forceinline__ void whiten_publish_status(double* sums, int* status) {
    // Only thread 0 wrote to `sums` (three atomics): it alone fences before it counts the block in.  A block-wide
    // __threadfence here would hold every warp until its own output stores have drained (measured: -3 % on the scans).
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(reinterpret_cast<unsigned int*>(&sums[3]), 1u) == gridDim.x - 1) {
            __threadfence();
            const double N = __ldcg(&sums[2]);
            *status = (N == 0.0) ? 1 : (N == 1.0) ? 2 : 0;
            __threadfence_system();
        }
    }
}
static inline cudaError_t record_whiten_status(void* status_event, cudaStream_t stream) {
    return status_event ? cudaEventRecord(static_cast<cudaEvent_t>(status_event), stream) : cudaSuccess;
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
