// One row-copy kernel behind index_first_axis / pad_input / packed-sequence permutes:
//     out[dst_idx[r], :] = src[src_idx[r], :]      (a null index array is the identity; dst < 0 skips the row)
// Rows are moved in the widest unit (16/8/4/2/1 B) both pointers and the row size allow.  A row is covered by
// 2^k threads (k chosen from the row width), a block by 256 >> k rows, and every thread keeps FOUR rows in flight:
// the four index loads, then the four data loads, are independent, so the idx -> data dependency is paid once per
// four rows.  No integer division anywhere (row / column come from shifts).  HBM-bound byte work.
#pragma once
#include <cuda_runtime.h>
#include <string.h>

#include "common.h"

namespace vpg {

template <typename U>
__global__ void __launch_bounds__(256)
permute_rows_kernel(const U* __restrict__ src, const long long* __restrict__ sidx, const long long* __restrict__ didx,
                    long long n_rows, long long upr, int lg_tpr, U* __restrict__ out) {
    const int tpr = 1 << lg_tpr;
    const int rpb = 256 >> lg_tpr;
    const int tr = threadIdx.x >> lg_tpr, tc = threadIdx.x & (tpr - 1);
    const long long step = static_cast<long long>(gridDim.x) * rpb * 4;
    for (long long r0 = static_cast<long long>(blockIdx.x) * rpb * 4; r0 < n_rows; r0 += step) {
        long long s[4], d[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const long long r = r0 + k * rpb + tr;
            const bool ok = r < n_rows;
            d[k] = ok ? (didx ? didx[r] : r) : -1;
            s[k] = ok ? (sidx ? sidx[r] : r) : 0;
        }
        for (long long c = tc; c < upr; c += tpr) {
            U v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (d[k] >= 0) v[k] = src[s[k] * upr + c];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (d[k] >= 0) out[d[k] * upr + c] = v[k];
        }
    }
}

// pad_input in ONE pass: out has n_out rows, out[didx[r], :] = src[r, :] with didx STRICTLY ASCENDING (what unpad_input
// produces), every other row zero.  The kernel walks the OUTPUT: a block owns a contiguous run of tiles of `tile_rows`
// output rows.  One binary search at the start of the run finds the first source row that lands in it; after that the
// source rows of consecutive tiles are consecutive, so per tile the block reads the next (at most tile_rows) entries of
// didx coalesced, notes in a shared-memory map which output row each one fills, and then sweeps the tile: every output
// row is written exactly once (source row or zeros), coalesced, with four rows in flight per thread.  No memset pass, no
// second write of the valid rows, no serial gap loops.
constexpr int kGapTileMax = 2048;

template <typename U>
__global__ void __launch_bounds__(256)
scatter_rows_gapfill_kernel(const U* __restrict__ src, const long long* __restrict__ didx, long long n_rows, long long n_out,
                            long long upr, int lg_tpr, int tile_rows, long long tiles_per_block, U* __restrict__ out) {
    __shared__ int map[kGapTileMax];
    __shared__ long long s_first;
    __shared__ int s_count;
    const int tpr = 1 << lg_tpr;
    const int rpb = 256 >> lg_tpr;
    const int tr = threadIdx.x >> lg_tpr, tc = threadIdx.x & (tpr - 1);
    const long long n_tiles = (n_out + tile_rows - 1) / tile_rows;
    const long long t_beg = static_cast<long long>(blockIdx.x) * tiles_per_block;
    const long long t_end = t_beg + tiles_per_block < n_tiles ? t_beg + tiles_per_block : n_tiles;
    if (t_beg >= t_end) return;
    if (threadIdx.x == 0) {  // first source row whose destination is >= the first output row of this block
        const long long key = t_beg * tile_rows;
        long long lo = 0, hi = n_rows;
        while (lo < hi) {
            const long long mid = (lo + hi) >> 1;
            if (didx[mid] < key) lo = mid + 1; else hi = mid;
        }
        s_first = lo;
        s_count = 0;
    }
    U zero;
    memset(&zero, 0, sizeof(U));
    __syncthreads();
    long long r_cur = s_first;
    for (long long t = t_beg; t < t_end; ++t) {
        const long long o0 = t * tile_rows;
        const int n_here = static_cast<int>((n_out - o0) < tile_rows ? (n_out - o0) : tile_rows);
        for (int j = threadIdx.x; j < tile_rows; j += 256) map[j] = -1;
        __syncthreads();
        int mine = 0;
        for (int i = threadIdx.x; i < tile_rows; i += 256) {
            const long long r = r_cur + i;
            if (r < n_rows) {
                const long long d = didx[r] - o0;
                if (d >= 0 && d < n_here) map[d] = i, ++mine;
            }
        }
        if (mine) atomicAdd(&s_count, mine);
        __syncthreads();
        const int count = s_count;
        for (int j0 = 0; j0 < n_here; j0 += rpb * 4) {
            int s[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int j = j0 + k * rpb + tr;
                s[k] = j < n_here ? map[j] : -2;
            }
            for (long long c = tc; c < upr; c += tpr) {
                U v[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    v[k] = zero;
                    if (s[k] >= 0) v[k] = src[(r_cur + s[k]) * upr + c];
                }
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (s[k] != -2) out[(o0 + j0 + k * rpb + tr) * upr + c] = v[k];
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) s_count = 0;  // ordered before the next tile's adds by the barrier after the map reset
        r_cur += count;
    }
}

// Narrow rows (the whole row is ONE unit of 1..16 bytes: log-probs [nnz, 1], token ids): the generic kernel above would
// store 4-byte units one by one behind a per-row map.  Here the output tile (tile_rows rows, <= 32 KB) is assembled in
// shared memory — zeroed with 128-bit stores, the source rows of the tile dropped in by destination, then written out as
// full 128-bit vectors — so every global access is coalesced and every output byte is still written exactly once.  The
// start of a block's run is found by a block-wide NT-ary lower bound (3-4 probe rounds at 2^25 rows instead of 25
// dependent loads of one thread); the number of source rows a tile consumed is the sum of per-thread hit counts.
template <int NT>
__device__ __forceinline__ long long block_lower_bound(const long long* __restrict__ a, long long n, long long key) {
    long long lo = 0, hi = n;  // invariant: the answer lies in [lo, hi]; every thread holds the same pair
    while (lo < hi) {
        const long long step = (hi - lo + NT - 1) / NT;
        long long q = lo + (static_cast<long long>(threadIdx.x) + 1) * step - 1;
        if (q > hi - 1) q = hi - 1;
        const int c = __syncthreads_count(a[q] < key);  // probes ascend with threadIdx, so the predicate is a prefix
        if (c == NT) return hi;
        const long long qc = lo + (static_cast<long long>(c) + 1) * step - 1;
        lo += static_cast<long long>(c) * step;
        hi = qc < hi - 1 ? qc : hi - 1;
    }
    return lo;
}

template <typename U, int NT>
__global__ void __launch_bounds__(NT)
scatter_narrow_gapfill_kernel(const U* __restrict__ src, const long long* __restrict__ didx, long long n_rows,
                              long long n_out, int tile_rows, long long tiles_per_block, U* __restrict__ out) {
    extern __shared__ uint4 stage4[];
    U* stage = reinterpret_cast<U*>(stage4);
    __shared__ int s_count[2];
    const long long n_tiles = (n_out + tile_rows - 1) / tile_rows;
    const long long t_beg = static_cast<long long>(blockIdx.x) * tiles_per_block;
    const long long t_end = t_beg + tiles_per_block < n_tiles ? t_beg + tiles_per_block : n_tiles;
    if (t_beg >= t_end) return;
    long long r_cur = block_lower_bound<NT>(didx, n_rows, t_beg * tile_rows);
    if (threadIdx.x < 2) s_count[threadIdx.x] = 0;
    const int tile_vecs = static_cast<int>(static_cast<size_t>(tile_rows) * sizeof(U) / 16);
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    int flip = 0;
    for (long long t = t_beg; t < t_end; ++t, flip ^= 1) {
        const long long o0 = t * tile_rows;
        const int n_here = static_cast<int>((n_out - o0) < tile_rows ? (n_out - o0) : tile_rows);
        for (int j = threadIdx.x; j < tile_vecs; j += NT) stage4[j] = z4;
        __syncthreads();
        int mine = 0;
        for (int i0 = threadIdx.x; i0 < n_here; i0 += 4 * NT) {  // ascending indices: at most n_here source rows land here
            long long d[4];
            U v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const long long r = r_cur + i0 + k * NT;
                const bool ok = r < n_rows;
                d[k] = ok ? didx[r] - o0 : static_cast<long long>(n_here);
                if (ok) v[k] = src[r];
            }
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (d[k] >= 0 && d[k] < n_here) stage[d[k]] = v[k], ++mine;
            if (d[3] >= n_here) break;  // this thread's later rows land beyond the tile as well
        }
        if (mine) atomicAdd(&s_count[flip], mine);
        __syncthreads();
        const int count = s_count[flip];
        if (threadIdx.x == 0) s_count[flip ^ 1] = 0;  // the other slot is idle until the next tile's adds (one barrier away)
        const size_t bytes_here = static_cast<size_t>(n_here) * sizeof(U);
        const int vecs_here = static_cast<int>(bytes_here / 16);
        uint4* o4 = reinterpret_cast<uint4*>(out + o0);
        for (int j = threadIdx.x; j < vecs_here; j += NT) o4[j] = stage4[j];
        for (int j = vecs_here * static_cast<int>(16 / sizeof(U)) + threadIdx.x; j < n_here; j += NT) out[o0 + j] = stage[j];
        __syncthreads();
        r_cur += count;
    }
}

static inline int row_copy_unit(const void* a, const void* b, int64_t row_bytes) {
    const uintptr_t x = reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | static_cast<uintptr_t>(row_bytes);
    if ((x & 15) == 0) return 16;
    if ((x & 7) == 0) return 8;
    if ((x & 3) == 0) return 4;
    if ((x & 1) == 0) return 2;
    return 1;
}

// enqueue the copy; returns a cudaError_t from the launch
static inline cudaError_t launch_permute_rows(const void* src, const long long* sidx, const long long* didx,
                                              int64_t n_rows, int64_t row_bytes, void* out, cudaStream_t stream) {
    const int u = row_copy_unit(src, out, row_bytes);
    const int64_t upr = row_bytes / u;
    int lg = 0;
    while (lg < 8 && (1LL << lg) < upr) ++lg;  // threads per row = min(256, next power of two >= units per row)
    const int64_t rows_per_block = (256 >> lg) * 4;
    int64_t g = ceil_div(n_rows, rows_per_block);
    const int64_t cap = static_cast<int64_t>(sm_count()) * 16;
    if (g > cap) g = cap;
    const unsigned grid = static_cast<unsigned>(g < 1 ? 1 : g);
    switch (u) {
        case 16: permute_rows_kernel<uint4><<<grid, 256, 0, stream>>>((const uint4*)src, sidx, didx, n_rows, upr, lg, (uint4*)out); break;
        case 8: permute_rows_kernel<uint2><<<grid, 256, 0, stream>>>((const uint2*)src, sidx, didx, n_rows, upr, lg, (uint2*)out); break;
        case 4: permute_rows_kernel<uint32_t><<<grid, 256, 0, stream>>>((const uint32_t*)src, sidx, didx, n_rows, upr, lg, (uint32_t*)out); break;
        case 2: permute_rows_kernel<uint16_t><<<grid, 256, 0, stream>>>((const uint16_t*)src, sidx, didx, n_rows, upr, lg, (uint16_t*)out); break;
        default: permute_rows_kernel<uint8_t><<<grid, 256, 0, stream>>>((const uint8_t*)src, sidx, didx, n_rows, upr, lg, (uint8_t*)out); break;
    }
    return cudaGetLastError();
}

static inline cudaError_t launch_scatter_rows_gapfill(const void* src, const long long* didx, int64_t n_rows,
                                                      int64_t n_out, int64_t row_bytes, void* out, cudaStream_t stream) {
    const int u = row_copy_unit(src, out, row_bytes);
    const int64_t upr = row_bytes / u;
    if (upr == 1 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {  // narrow rows: shared-memory staged tiles
        static const int nt = env_int("VPG_PAD_NT", 256), tile_kb = env_int("VPG_PAD_TILE_KB", 16),
                         bps = env_int("VPG_PAD_BPS", 6);  // measured best of the A/B in profiles/r2_ab_pad_input.jsonl
        int64_t tile = (static_cast<int64_t>(tile_kb) * 1024) / u;
        if (tile > 8192) tile = 8192;
        const int64_t n_tiles = ceil_div(n_out, tile);
        int64_t g = n_tiles;
        const int64_t cap = static_cast<int64_t>(sm_count()) * bps;
        if (g > cap) g = cap;
        if (g < 1) g = 1;
        const int64_t tpb = ceil_div(n_tiles, g);
        const int64_t gr = ceil_div(n_tiles, tpb);
        const unsigned grid = static_cast<unsigned>(gr < 1 ? 1 : gr);
        const size_t smem = static_cast<size_t>(tile) * u;
        const int tr = static_cast<int>(tile);
#define VPG_NARROW(U, NT) scatter_narrow_gapfill_kernel<U, NT><<<grid, NT, smem, stream>>>((const U*)src, didx, n_rows, n_out, tr, tpb, (U*)out)
#define VPG_NARROW_NT(U) do { if (nt == 1024) VPG_NARROW(U, 1024); else if (nt == 512) VPG_NARROW(U, 512); else if (nt == 128) VPG_NARROW(U, 128); else VPG_NARROW(U, 256); } while (0)
        switch (u) {
            case 16: VPG_NARROW_NT(uint4); break;
            case 8: VPG_NARROW_NT(uint2); break;
            case 4: VPG_NARROW_NT(uint32_t); break;
            case 2: VPG_NARROW_NT(uint16_t); break;
            default: VPG_NARROW_NT(uint8_t); break;
        }
#undef VPG_NARROW_NT
#undef VPG_NARROW
        return cudaGetLastError();
    }
    int lg = 0;
    while (lg < 8 && (1LL << lg) < upr) ++lg;
    // tile: ~32 KB of output, between one 4-row pass of the block and the shared-memory map
    int64_t tile = (32 * 1024) / (row_bytes > 0 ? row_bytes : 1);
    const int64_t one_pass = (256 >> lg) * 4;
    if (tile < one_pass) tile = one_pass;
    if (tile < 16) tile = 16;
    if (tile > kGapTileMax) tile = kGapTileMax;
    const int64_t n_tiles = ceil_div(n_out, tile);
    int64_t g = n_tiles;
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    const int64_t tpb = ceil_div(n_tiles, g);
    const unsigned grid = static_cast<unsigned>(ceil_div(n_tiles, tpb) < 1 ? 1 : ceil_div(n_tiles, tpb));
    const int tr = static_cast<int>(tile);
    switch (u) {
        case 16: scatter_rows_gapfill_kernel<uint4><<<grid, 256, 0, stream>>>((const uint4*)src, didx, n_rows, n_out, upr, lg, tr, tpb, (uint4*)out); break;
        case 8: scatter_rows_gapfill_kernel<uint2><<<grid, 256, 0, stream>>>((const uint2*)src, didx, n_rows, n_out, upr, lg, tr, tpb, (uint2*)out); break;
        case 4: scatter_rows_gapfill_kernel<uint32_t><<<grid, 256, 0, stream>>>((const uint32_t*)src, didx, n_rows, n_out, upr, lg, tr, tpb, (uint32_t*)out); break;
        case 2: scatter_rows_gapfill_kernel<uint16_t><<<grid, 256, 0, stream>>>((const uint16_t*)src, didx, n_rows, n_out, upr, lg, tr, tpb, (uint16_t*)out); break;
        default: scatter_rows_gapfill_kernel<uint8_t><<<grid, 256, 0, stream>>>((const uint8_t*)src, didx, n_rows, n_out, upr, lg, tr, tpb, (uint8_t*)out); break;
    }
    return cudaGetLastError();
}

}  // namespace vpg
