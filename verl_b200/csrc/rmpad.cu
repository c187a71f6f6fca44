// Remove-padding (rmpad) index plumbing: mask -> (indices, seqlens, cu_seqlens, max_seqlen), row gather / scatter.
// Reference: flash_attn.bert_padding (unpad_input / index_first_axis / pad_input; third-party file, called from
// verl/workers/actor/dp_actor.py:110-125,221-232) and verl/utils/torch_functional.py:629-638 (get_unpad_data).
//   indices  = nonzero(mask.flatten())            (int64, ascending)
//   seqlens  = mask.sum(-1, dtype=int32)          (sum of mask VALUES, as the reference computes it)
//   cu_seqlens = pad(cumsum(seqlens, int32), (1,0)),  max_seqlen = seqlens.max()
// Integer outputs are bit-exact.  Stream compaction: tile counts -> single-block scan -> ordered tile write with
// ballot/popc ranks, so the index list is produced in ascending order without a sort.
#include <cuda_runtime.h>

#include "common.h"
#include "row_copy.cuh"
#include "reduce.cuh"

namespace vpg {

constexpr int kTileThreads = 256;
constexpr int kItems = 16;                          // mask elements per thread
constexpr int kTile = kTileThreads * kItems;        // 4096 elements per block

template <typename MT>
__device__ __forceinline__ bool mask_nonzero(MT m) {
    return m != MT(0);
}

// Pass 1 — the ONLY pass over the mask: nonzero count of every tile, the nonzero bitmap (one 32-bit ballot per warp and
// item: 1 bit per element, read back by the write pass instead of the 1-8 byte mask elements), and the per-row sums
// of the mask VALUES (seqlens).  Layout: element = tile0 + j*256 + tid (coalesced); a warp's 32 elements of one item
// almost always sit in one row, so lane 0 keeps a running (row, sum) and issues one integer atomic per row change.
template <typename MT>
__global__ void tile_count_kernel(const MT* __restrict__ mask, long long n, long long S, int* __restrict__ tile_counts,
                                  unsigned int* __restrict__ bits, int* __restrict__ seqlens) {
    __shared__ int red[32];
    const long long tile0 = static_cast<long long>(blockIdx.x) * kTile;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int c = 0;
    long long cur_row = -1;
    long long cur_sum = 0;
    // row arithmetic without per-element 64-bit divisions: ONE division per thread locates the tile; inside the tile an
    // element is < S + kTile past that row's start, so its row offset is a compare (S >= kTile) or a 32-bit division
    const long long tile_row = tile0 / S;
    const long long tile_off = tile0 - tile_row * S;
    const bool wide = S >= kTile;
    const unsigned int S32 = static_cast<unsigned int>(wide ? 1 : S);
    auto row_of = [&](long long loc) -> long long {
        return tile_row + (wide ? (loc >= S ? 1 : 0) : static_cast<long long>(static_cast<unsigned int>(loc) / S32));
    };
#pragma unroll
    for (int j = 0; j < kItems; ++j) {
        const int e = j * kTileThreads + threadIdx.x;  // element inside the tile
        const long long i = tile0 + e;
        const bool in = i < n;
        const MT m = in ? mask[i] : MT(0);
        const bool nz = in && mask_nonzero(m);
        const unsigned bal = __ballot_sync(0xffffffffu, nz);
        if (lane == 0) bits[(static_cast<long long>(blockIdx.x) * kItems + j) * (kTileThreads / 32) + w] = bal;
        c += nz ? 1 : 0;
        // ---- seqlens
        const long long first = i - lane;  // element of lane 0
        if (first < n) {
            const long long last = (first + 31 < n) ? first + 31 : n - 1;
            const long long r0 = row_of(tile_off + (e - lane)), r1 = row_of(tile_off + (e - lane) + (last - first));
            long long v = in ? static_cast<long long>(m) : 0;
            if (r0 == r1) {
                // 0/1 masks (the usual case): the warp's sum is the popcount of the ballot already taken; any other
                // mask value falls back to the shuffle tree
                if (__all_sync(0xffffffffu, v == (nz ? 1 : 0))) {
                    v = __popc(bal);
                } else {
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                }
                if (lane == 0) {
                    if (r0 != cur_row) {
                        if (cur_row >= 0 && cur_sum != 0) atomicAdd(&seqlens[cur_row], static_cast<int>(cur_sum));
                        cur_row = r0, cur_sum = 0;
                    }
                    cur_sum += v;
                }
            } else if (in && v != 0) {  // a row boundary inside the warp's 32 elements (S < 32 or unaligned rows)
                atomicAdd(&seqlens[row_of(tile_off + e)], static_cast<int>(v));
            }
        }
    }
    if (lane == 0 && cur_row >= 0 && cur_sum != 0) atomicAdd(&seqlens[cur_row], static_cast<int>(cur_sum));
    const int tot = block_sum<int>(c, red);
    if (threadIdx.x == 0) tile_counts[blockIdx.x] = tot;
}

// single block: exclusive scan of tile counts -> tile_offsets (long long), inclusive scan of seqlens -> cu_seqlens,
// totals = {nnz, max_seqlen}
__global__ void scan_kernel(const int* __restrict__ tile_counts, int num_tiles, long long* __restrict__ tile_offsets,
                            const int* __restrict__ seqlens, int B, int* __restrict__ cu_seqlens,
                            int* __restrict__ totals) {
    __shared__ long long warp_tot[32];
    __shared__ long long carry_s;
    __shared__ int max_s[32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    // ---- tiles
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < num_tiles; base += blockDim.x) {
        const int i = base + threadIdx.x;
        const long long v = (i < num_tiles) ? tile_counts[i] : 0;
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
        if (i < num_tiles) tile_offsets[i] = carry + pre + inc - v;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry_s = carry + pre + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        tile_offsets[num_tiles] = carry_s;
        totals[0] = static_cast<int>(carry_s);
    }
    __syncthreads();
    // ---- rows: int32 cumsum (wraps like torch's int32 cumsum would)
    if (threadIdx.x == 0) {
        carry_s = 0;
        cu_seqlens[0] = 0;
    }
    int mx = 0;
    __syncthreads();
    for (int base = 0; base < B; base += blockDim.x) {
        const int i = base + threadIdx.x;
        const int v = (i < B) ? seqlens[i] : 0;
        if (i < B) mx = max(mx, v);
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
        if (i < B) cu_seqlens[i + 1] = static_cast<int>(carry + pre + inc);
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
        totals[1] = (B > 0) ? m : 0;
    }
}

// Pass 2: ordered write of the nonzero positions of each tile from the bitmap (the mask is not read again).
// The tile's 128 ballots (item-major, warp-minor == element order) are loaded once, coalesced; their popcounts are
// scanned by four warps; every thread then places its own elements with one shared-memory read per item.
__global__ void tile_write_kernel(const unsigned int* __restrict__ bits, const long long* __restrict__ tile_offsets,
                                  long long* __restrict__ indices) {
    constexpr int kWords = kItems * (kTileThreads / 32);  // 128
    __shared__ unsigned int sbits[kWords];
    __shared__ int excl[kWords];
    __shared__ int wtot[kWords / 32];
    const long long tile0 = static_cast<long long>(blockIdx.x) * kTile;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int inc = 0, cnt = 0;
    if (threadIdx.x < kWords) {
        const unsigned int b = bits[static_cast<long long>(blockIdx.x) * kWords + threadIdx.x];
        sbits[threadIdx.x] = b;
        cnt = __popc(b);
        inc = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        if (lane == 31) wtot[w] = inc;
    }
    __syncthreads();
    if (threadIdx.x < kWords) {
        int pre = 0;
        for (int ww = 0; ww < w; ++ww) pre += wtot[ww];
        excl[threadIdx.x] = pre + inc - cnt;
    }
    __syncthreads();
    const long long off = tile_offsets[blockIdx.x];
    const unsigned int below = (1u << lane) - 1u;
#pragma unroll
    for (int j = 0; j < kItems; ++j) {
        const int q = j * (kTileThreads / 32) + w;
        const unsigned int bal = sbits[q];
        if ((bal >> lane) & 1u)
            indices[off + excl[q] + __popc(bal & below)] = tile0 + static_cast<long long>(j) * kTileThreads + threadIdx.x;
    }
}

// Response-window selection on the rmpad stream (verl/workers/actor/dp_actor.py:135,221-237 folded into one pass):
// token i of the stream sits at padded position indices[i] = b*S + s.  The reference computes log-probs for every
// token, pads back to [B,S] and keeps columns [S-L-1, S-1); only those tokens matter.  For each stream token:
//   keep[i]   = 1 if S-L-1 <= s < S-1
//   dst[i]    = b*L + (s - (S-L-1))        row of the [B*L] response matrix the token lands in (-1 if dropped)
//   labels[i] = ids_rmpad[(i+1) % T]       torch.roll(input_ids_rmpad, -1)
__global__ void response_rows_kernel(const long long* __restrict__ indices, const long long* __restrict__ ids_rmpad,
                                     long long T, long long S, long long L, unsigned char* __restrict__ keep,
                                     long long* __restrict__ dst, long long* __restrict__ labels) {
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= T) return;
    const long long pos = indices[i];
    const long long b = pos / S, s = pos - b * S;
    const long long j = s - (S - L - 1);
    const bool k = (j >= 0) && (j < L);
    keep[i] = k ? 1 : 0;
    dst[i] = k ? b * L + j : -1;
    if (labels) labels[i] = ids_rmpad[(i + 1 == T) ? 0 : i + 1];
}

// Integer row sums of a [B,S] mask -> int64 (attention_mask.sum(dim=1), verl/utils/seqlen_balancing.py:282): one warp
// per row, lanes stride the row (coalesced), exact integer arithmetic.
template <typename MT>
__global__ void row_lengths_kernel(const MT* __restrict__ mask, long long B, long long S, long long* __restrict__ out) {
    const long long row = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= B) return;
    const MT* m = mask + row * S;
    long long c = 0;
    for (long long t = lane; t < S; t += 32) c += static_cast<long long>(m[t]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) out[row] = c;
}

}  // namespace vpg

using namespace vpg;

extern "C" {

int vpg_response_rows(const int64_t* indices, const int64_t* ids_rmpad, int64_t T, int64_t S, int64_t L,
                      uint8_t* keep, int64_t* dst, int64_t* labels, void* stream) {
    VPG_CHECK_ARG(T >= 0 && S > 0 && L >= 0 && L < S, "invalid response window: T=%lld S=%lld L=%lld", (long long)T,
                  (long long)S, (long long)L);
    if (T == 0) return VPG_OK;
    VPG_CHECK_ARG(indices && keep && dst && (!labels || ids_rmpad), "null pointer argument");
    response_rows_kernel<<<static_cast<unsigned>(ceil_div(T, 256)), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const long long*>(indices), reinterpret_cast<const long long*>(ids_rmpad), T, S, L, keep,
        reinterpret_cast<long long*>(dst), reinterpret_cast<long long*>(labels));
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

size_t vpg_unpad_workspace_bytes(int64_t n) {
    const int64_t tiles = ceil_div(n > 0 ? n : 1, kTile);
    // tile offsets (i64) | tile counts (i32) | nonzero bitmap: one u32 per (tile, item, warp) = 1 bit per element
    return static_cast<size_t>(tiles) * sizeof(int) + 16 + static_cast<size_t>(tiles + 1) * sizeof(long long) + 16 +
           static_cast<size_t>(tiles) * kItems * (kTileThreads / 32) * sizeof(unsigned int);
}

int vpg_unpad_indices(const void* mask, int mask_dtype, int64_t B, int64_t S, int64_t* indices, int32_t* seqlens,
                      int32_t* cu_seqlens, int32_t* totals, void* workspace, size_t workspace_bytes, void* stream_) {
    VPG_CHECK_ARG(mask && indices && seqlens && cu_seqlens && totals, "null pointer argument");
    VPG_CHECK_ARG(B >= 0 && S >= 0 && B * S < (1LL << 31), "mask too large (B*S must fit int32)");
    VPG_CHECK_ARG(mask_dtype != VPG_MASK_F32, "float masks are not supported by unpad_indices");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int64_t n = B * S;
    const size_t need = vpg_unpad_workspace_bytes(n);
    if (!workspace || workspace_bytes < need)
        return set_error(VPG_ERR_WORKSPACE, "unpad workspace too small: need %zu, got %zu", need, workspace_bytes);
    VPG_CHECK_ARG((reinterpret_cast<uintptr_t>(workspace) & 7) == 0, "workspace must be 8-byte aligned");
    const int tiles = static_cast<int>(n > 0 ? ceil_div(n, kTile) : 0);
    long long* tile_offsets = static_cast<long long*>(workspace);
    int* tile_counts = reinterpret_cast<int*>(tile_offsets + (ceil_div(n > 0 ? n : 1, kTile) + 1));
    unsigned int* bits = reinterpret_cast<unsigned int*>(
        reinterpret_cast<uint8_t*>(tile_counts) + round_up(static_cast<int64_t>(tiles) * sizeof(int), 16));
    if (B > 0) VPG_CUDA(cudaMemsetAsync(seqlens, 0, static_cast<size_t>(B) * sizeof(int), stream));
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        if (tiles > 0) {
            tile_count_kernel<MT><<<tiles, kTileThreads, 0, stream>>>(static_cast<const MT*>(mask), n, S, tile_counts,
                                                                      bits, seqlens);
            count_launch();
        }
    });
    scan_kernel<<<1, 1024, 0, stream>>>(tile_counts, tiles, tile_offsets, seqlens, static_cast<int>(B), cu_seqlens, totals);
    count_launch();
    if (tiles > 0) {
        tile_write_kernel<<<tiles, kTileThreads, 0, stream>>>(bits, tile_offsets, reinterpret_cast<long long*>(indices));
        count_launch();
    }
    VPG_CUDA(cudaGetLastError());
    return VPG_OK;
}

int vpg_gather_rows(const void* src, const int64_t* indices, int64_t n_idx, int64_t row_bytes, void* out,
                    void* stream_) {
    VPG_CHECK_ARG(n_idx >= 0 && row_bytes > 0, "invalid gather_rows arguments");
    if (n_idx == 0) return VPG_OK;
    VPG_CHECK_ARG(src && indices && out, "null pointer argument");
    VPG_CUDA(launch_permute_rows(src, reinterpret_cast<const long long*>(indices), nullptr, n_idx, row_bytes, out,
                                 static_cast<cudaStream_t>(stream_)));
    count_launch();
    return VPG_OK;
}

int vpg_scatter_rows(const void* src, const int64_t* indices, int64_t n_idx, int64_t row_bytes, void* out,
                     int64_t n_out, int zero_fill, void* stream_) {
    VPG_CHECK_ARG(n_idx >= 0 && row_bytes > 0 && n_out >= 0, "invalid scatter_rows arguments");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (zero_fill && n_out > 0) {
        VPG_CHECK_ARG(out, "null output");
        VPG_CUDA(cudaMemsetAsync(out, 0, static_cast<size_t>(n_out) * row_bytes, stream));
    }
    if (n_idx == 0) return VPG_OK;
    VPG_CHECK_ARG(src && indices && out, "null pointer argument");
    VPG_CUDA(launch_permute_rows(src, nullptr, reinterpret_cast<const long long*>(indices), n_idx, row_bytes, out, stream));
    count_launch();
    return VPG_OK;
}

int vpg_row_lengths(const void* mask, int mask_dtype, int64_t B, int64_t S, int64_t* lengths, void* stream_) {
    VPG_CHECK_ARG(B >= 0 && S >= 0, "invalid row_lengths shape");
    if (B == 0) return VPG_OK;
    VPG_CHECK_ARG(mask && lengths, "null pointer argument");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        row_lengths_kernel<MT><<<static_cast<unsigned>(ceil_div(B * 32, 256)), 256, 0, stream>>>(
            static_cast<const MT*>(mask), B, S, reinterpret_cast<long long*>(lengths));
    });
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

}  // extern "C"
