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
// produces), every other row zero.  The thread group of row r also zeroes the gap (didx[r-1], didx[r]) in front of its
// row (row 0: from 0; the last row additionally the tail up to n_out), so every output byte is written exactly once:
// no memset pass and no second write of the valid rows.  Same row geometry as permute_rows_kernel.
template <typename U>
__global__ void __launch_bounds__(256)
scatter_rows_gapfill_kernel(const U* __restrict__ src, const long long* __restrict__ didx, long long n_rows, long long n_out,
                            long long upr, int lg_tpr, U* __restrict__ out) {
    const int tpr = 1 << lg_tpr;
    const int rpb = 256 >> lg_tpr;
    const int tr = threadIdx.x >> lg_tpr, tc = threadIdx.x & (tpr - 1);
    const long long step = static_cast<long long>(gridDim.x) * rpb * 4;
    U zero;
    memset(&zero, 0, sizeof(U));
    for (long long r0 = static_cast<long long>(blockIdx.x) * rpb * 4; r0 < n_rows; r0 += step) {
        long long d[4], lo[4], hi[4];  // destination row, first gap row in front of it, end of the tail gap behind it
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const long long r = r0 + k * rpb + tr;
            const bool ok = r < n_rows;
            d[k] = ok ? didx[r] : -1;
            lo[k] = (ok && r > 0) ? didx[r - 1] + 1 : 0;
            hi[k] = (ok && r == n_rows - 1) ? n_out : 0;
        }
        for (long long c = tc; c < upr; c += tpr) {
            U v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (d[k] >= 0) v[k] = src[(r0 + k * rpb + tr) * upr + c];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (d[k] >= 0) out[d[k] * upr + c] = v[k];
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (d[k] < 0) continue;
            for (long long g = lo[k]; g < d[k]; ++g)
                for (long long c = tc; c < upr; c += tpr) out[g * upr + c] = zero;
            for (long long g = d[k] + 1; g < hi[k]; ++g)
                for (long long c = tc; c < upr; c += tpr) out[g * upr + c] = zero;
        }
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
    int lg = 0;
    while (lg < 8 && (1LL << lg) < upr) ++lg;
    const int64_t rows_per_block = (256 >> lg) * 4;
    int64_t g = ceil_div(n_rows, rows_per_block);
    const int64_t cap = static_cast<int64_t>(sm_count()) * 16;
    if (g > cap) g = cap;
    const unsigned grid = static_cast<unsigned>(g < 1 ? 1 : g);
    switch (u) {
        case 16: scatter_rows_gapfill_kernel<uint4><<<grid, 256, 0, stream>>>((const uint4*)src, didx, n_rows, n_out, upr, lg, (uint4*)out); break;
        case 8: scatter_rows_gapfill_kernel<uint2><<<grid, 256, 0, stream>>>((const uint2*)src, didx, n_rows, n_out, upr, lg, (uint2*)out); break;
        case 4: scatter_rows_gapfill_kernel<uint32_t><<<grid, 256, 0, stream>>>((const uint32_t*)src, didx, n_rows, n_out, upr, lg, (uint32_t*)out); break;
        case 2: scatter_rows_gapfill_kernel<uint16_t><<<grid, 256, 0, stream>>>((const uint16_t*)src, didx, n_rows, n_out, upr, lg, (uint16_t*)out); break;
        default: scatter_rows_gapfill_kernel<uint8_t><<<grid, 256, 0, stream>>>((const uint8_t*)src, didx, n_rows, n_out, upr, lg, (uint8_t*)out); break;
    }
    return cudaGetLastError();
}

}  // namespace vpg
