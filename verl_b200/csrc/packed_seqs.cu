// Packed ("thd") sequence layout for the Megatron-side callers of the per-token path.
// Reference: verl/models/mcore/util.py:23-142 (preprocess_packed_seqs / postprocess_packed_seqs).
//
// The reference loops over the batch on the host (per sequence: a boolean-mask gather, 2 slice assignments per
// context-parallel chunk, several .item() syncs).  Here the whole layout is index arithmetic on the device:
//   packed_layout_kernel  seqlens -> padded seqlens (multiple of `align`), padded cu_seqlens, {sum, max}   (1 block)
//   packed_map_kernel     valid token n = (sequence i, j-th valid token) -> its slot in the packed stream
//   permute_rows_kernel   (row_copy.cuh) out[dst[n]] = src[src_idx[n]], one pass; used for both directions and for
//                         the backward of the post-processing
// Context parallelism (cp): a padded sequence of Lp tokens is cut into 2*cp chunks of half = Lp/(2*cp); rank r owns
// chunk r and chunk 2*cp-1-r (causal-mask load balance), stored back to back in its local stream at offset
// cu_padded[i]/cp.  The cp-gathered stream is rank-major: position = r * (sum Lp / cp) + local slot.
// HBM-bound integer/byte work; no tensor cores involved.
#include <cuda_runtime.h>

#include "common.h"
#include "row_copy.cuh"

namespace vpg {

__global__ void packed_layout_kernel(const int* __restrict__ seqlens, int B, int align,
                                     int* __restrict__ seqlens_padded, int* __restrict__ cu_padded,
                                     int* __restrict__ totals) {
    __shared__ long long warp_tot[32];
    __shared__ long long carry_s;
    __shared__ int max_s[32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    if (threadIdx.x == 0) {
        carry_s = 0;
        cu_padded[0] = 0;
    }
    int mx = 0;
    __syncthreads();
    for (int base = 0; base < B; base += blockDim.x) {
        const int i = base + threadIdx.x;
        int v = 0;
        if (i < B) {
            const int s = seqlens[i];
            v = s + (align - s % align) % align;
            seqlens_padded[i] = v;
            mx = max(mx, v);
        }
        long long inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const long long o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        if (lane == 31) warp_tot[w] = inc;
        __syncthreads();
        long long pre = 0;
        for (int ww = 0; ww < w; ++ww) pre += warp_tot[ww];
        const long long carry = carry_s;
        if (i < B) cu_padded[i + 1] = static_cast<int>(carry + pre + inc);
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry_s = carry + pre + inc;
        __syncthreads();
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) max_s[w] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        int m = 0;
        for (int ww = 0; ww < nw; ++ww) m = max(m, max_s[ww]);
        totals[0] = static_cast<int>(carry_s);
        totals[1] = m;
    }
}

__global__ void packed_map_kernel(const int* __restrict__ cu_seqlens, const int* __restrict__ cu_padded, int B,
                                  long long T, int cp_size, int cp_rank, long long* __restrict__ gathered_pos,
                                  long long* __restrict__ local_dst) {
    const long long n = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (n >= T) return;
    // sequence of token n: last i with cu_seqlens[i] <= n
    int lo = 0, hi = B;  // invariant: cu[lo] <= n < cu[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (cu_seqlens[mid] <= n) lo = mid; else hi = mid;
    }
    const long long j = n - cu_seqlens[lo];
    const long long p0 = cu_padded[lo];
    if (cp_size <= 1) {
        gathered_pos[n] = p0 + j;
        local_dst[n] = p0 + j;
        return;
    }
    const long long Lp = cu_padded[lo + 1] - p0;
    const long long half = Lp / (2 * cp_size);
    const long long local_total = cu_padded[B] / cp_size;
    long long r, q;
    if (j < half * cp_size) {
        r = j / half;
        q = j - r * half;
    } else {
        r = (Lp - 1 - j) / half;
        q = half + (j - (Lp - half * (r + 1)));
    }
    const long long slot = p0 / cp_size + q;
    gathered_pos[n] = r * local_total + slot;
    local_dst[n] = (r == cp_rank) ? slot : -1;
}

}  // namespace vpg

using namespace vpg;

extern "C" {

int vpg_packed_seq_layout(const int32_t* seqlens, int64_t B, int align, int32_t* seqlens_padded,
                          int32_t* cu_seqlens_padded, int32_t* totals, void* stream) {
    VPG_CHECK_ARG(B >= 0 && align > 0, "invalid packed_seq_layout arguments");
    VPG_CHECK_ARG(cu_seqlens_padded && totals && (B == 0 || (seqlens && seqlens_padded)), "null pointer argument");
    packed_layout_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(seqlens, static_cast<int>(B), align,
                                                                            seqlens_padded, cu_seqlens_padded, totals);
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

int vpg_packed_seq_map(const int32_t* cu_seqlens, const int32_t* cu_seqlens_padded, int64_t B, int64_t T, int cp_size,
                       int cp_rank, int64_t* gathered_pos, int64_t* local_dst, void* stream) {
    VPG_CHECK_ARG(B >= 0 && T >= 0 && cp_size >= 1 && cp_rank >= 0 && cp_rank < cp_size,
                  "invalid packed_seq_map arguments (cp_size=%d cp_rank=%d)", cp_size, cp_rank);
    if (T == 0) return VPG_OK;
    VPG_CHECK_ARG(cu_seqlens && cu_seqlens_padded && gathered_pos && local_dst && B > 0, "null pointer argument");
    packed_map_kernel<<<static_cast<unsigned>(ceil_div(T, 256)), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        cu_seqlens, cu_seqlens_padded, static_cast<int>(B), T, cp_size, cp_rank,
        reinterpret_cast<long long*>(gathered_pos), reinterpret_cast<long long*>(local_dst));
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

int vpg_permute_rows(const void* src, const int64_t* src_idx, const int64_t* dst_idx, int64_t n_rows, int64_t row_bytes,
                     void* out, int64_t n_out, int zero_fill, void* stream_) {
    VPG_CHECK_ARG(n_rows >= 0 && row_bytes > 0 && n_out >= 0, "invalid permute_rows arguments");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (zero_fill && n_out > 0) {
        VPG_CHECK_ARG(out, "null output");
        VPG_CUDA(cudaMemsetAsync(out, 0, static_cast<size_t>(n_out) * row_bytes, stream));
    }
    if (n_rows == 0) return VPG_OK;
    VPG_CHECK_ARG(src && out, "null pointer argument");
    VPG_CUDA(launch_permute_rows(src, reinterpret_cast<const long long*>(src_idx),
                                 reinterpret_cast<const long long*>(dst_idx), n_rows, row_bytes, out, stream));
    count_launch();
    return VPG_OK;
}

}  // extern "C"
