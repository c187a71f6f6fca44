// Remove-padding (rmpad) index plumbing: mask -> (indices, seqlens, cu_seqlens, max_seqlen), row gather / scatter.
// Reference: flash_attn.bert_padding (unpad_input / index_first_axis / pad_input; third-party file, called from
// verl/workers/actor/dp_actor.py:110-125,221-232) and verl/utils/torch_functional.py:629-638 (get_unpad_data).
//   indices  = nonzero(mask.flatten())            (int64, ascending)
//   seqlens  = mask.sum(-1, dtype=int32)          (sum of mask VALUES, as the reference computes it)
//   cu_seqlens = pad(cumsum(seqlens, int32), (1,0)),  max_seqlen = seqlens.max()
// Integer outputs are bit-exact.  Stream compaction: tile counts -> single-block scan -> ordered tile write with
// ballot/popc ranks, so the index list is produced in ascending order without a sort.
#include <cuda_runtime.h>
#include <limits.h>

#include "common.h"
#include "ptx.cuh"
#include "row_copy.cuh"
#include "reduce.cuh"
#include "token_common.cuh"

namespace vpg {

constexpr int kTileThreads = 256;
constexpr int kItems = 16;                          // mask elements per thread
constexpr int kTile = kTileThreads * kItems;        // 4096 elements per block

template <typename MT>
__device__ __forceinline__ bool mask_nonzero(MT m) {
    return m != MT(0);
}

// Stream compaction in two launches, the mask read ONCE:
//   unpad_count_kernel — streaming pass over the mask.  Warps are autonomous (no block barrier, no tail block): a warp
//     owns a contiguous run of 4096-element tiles and walks them in chunks of 512 elements.  A chunk is staged in shared
//     memory by ONE bulk-copy instruction issued kStages chunks ahead (cp.async.bulk + mbarrier), so the bytes in flight
//     are bounded by shared memory, not by registers or L1 miss tracking.  Per chunk: 16 ballots -> 64 bytes of a flat
//     1-bit-per-element bitmap; per tile: the nonzero count (plain store, the warp owns the tile); per group of 128
//     tiles: one RED on a super-count.  Per-row sums of the mask VALUES (seqlens): for 0/1 masks the popcounts already
//     taken; the row of a chunk costs one 32-bit division per chunk; lane 0 keeps a running (row, sum) and issues one
//     RED per row change — consecutive chunks of a warp are consecutive in memory, so that is one RED per row.
//   unpad_write_kernel — block t writes the ascending int64 indices of tile t from its 512-byte bitmap (the 1-8 byte
//     mask elements are not read again).  Its output offset is a two-level sum (super-counts before its group + tile
//     counts inside the group) taken by one otherwise idle warp under the bitmap load: no scan launch, no chained
//     spinning.  Extra blocks at the end of the grid turn seqlens into cu_seqlens / max_seqlen off the critical path.
constexpr int kWarps = kTileThreads / 32;
constexpr int kWords = kItems * kWarps;      // 128 ballots per tile
constexpr int kStages = 4;                   // bulk copies in flight per warp
constexpr int kSuper = 128;                  // tiles per super-count
constexpr int kRowsPerBlock = 8 * kTileThreads;  // rows scanned by one extra block of the write kernel
constexpr unsigned int kChunk = 32 * kItems;     // 512 elements
constexpr int kChunksPerTile = kTile / kChunk;   // 8

struct UnpadHeader {             // first 16 bytes of the workspace, zeroed before the count pass
    unsigned int extra_done;     // extra blocks of the write kernel that finished
    unsigned int max_biased;     // max row sum, stored as (value ^ 0x80000000) so that 0 is INT_MIN
    unsigned int pad[2];
};

template <typename MT>
__global__ void __launch_bounds__(kTileThreads)
unpad_count_kernel(const MT* __restrict__ mask, unsigned int n, unsigned int S, int num_tiles, int tiles_per_warp,
                   int* __restrict__ tile_counts, int* __restrict__ super_counts, unsigned int* __restrict__ bits,
                   int* __restrict__ seqlens) {
    extern __shared__ __align__(128) unsigned char stage_smem[];
    __shared__ __align__(8) uint64_t bars[kWarps * kStages];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int t_begin = (blockIdx.x * kWarps + w) * tiles_per_warp;
    const int t_end = min(t_begin + tiles_per_warp, num_tiles);
    if (t_begin >= t_end) return;
    const unsigned int ch_begin = static_cast<unsigned int>(t_begin) * kChunksPerTile;
    const unsigned int ch_end = static_cast<unsigned int>(t_end) * kChunksPerTile;  // whole tiles: no bitmap holes
    constexpr unsigned int kChunkBytes = kChunk * sizeof(MT);
    const bool aligned = (reinterpret_cast<uintptr_t>(mask) & 15) == 0;
    MT* const my_stage = reinterpret_cast<MT*>(stage_smem) + static_cast<size_t>(w) * kStages * kChunk;
    uint64_t* const my_bars = bars + w * kStages;
    // a chunk cut by the end of the mask, or a mask whose base is not 16-byte aligned, is read with guarded loads
    auto staged = [&](unsigned int ch) { return aligned && ch * kChunk + kChunk <= n; };
    auto issue = [&](unsigned int ch, int s) {  // lane 0 only
        if (ch < ch_end && staged(ch)) {
            mbar_arrive_expect_tx(&my_bars[s], kChunkBytes);
            tma_load_1d(my_stage + s * kChunk, mask + static_cast<size_t>(ch) * kChunk, kChunkBytes, &my_bars[s]);
        }
    };
    if (lane == 0) {
        for (int s = 0; s < kStages; ++s) mbar_init(&my_bars[s], 1);
        fence_barrier_init();
        fence_proxy_async();
        for (int s = 0; s < kStages; ++s) issue(ch_begin + s, s);
    }
    __syncwarp();
    unsigned int cur_row = 0xffffffffu;
    int cur_sum = 0;
    auto add_row = [&](unsigned int row, int v) {
        if (row != cur_row) {
            if (lane == 0 && cur_sum != 0) atomicAdd(&seqlens[cur_row], cur_sum);
            cur_row = row, cur_sum = 0;
        }
        cur_sum += v;
    };
    int tile_cnt = 0, super_cnt = 0;
    unsigned int k = 0;
    for (unsigned int ch = ch_begin; ch < ch_end; ++ch, ++k) {
        const unsigned int c0 = ch * kChunk;
        const int s = k % kStages;
        MT m[kItems];
        if (staged(ch)) {
            mbar_wait(&my_bars[s], (k / kStages) & 1);
            const MT* buf = my_stage + s * kChunk;
#pragma unroll
            for (int j = 0; j < kItems; ++j) m[j] = buf[j * 32 + lane];
            __syncwarp();  // every lane has read the stage: refill it kStages chunks ahead
            if (lane == 0) issue(ch + kStages, s);
        } else {
#pragma unroll
            for (int j = 0; j < kItems; ++j) {
                const unsigned int i = c0 + j * 32 + lane;
                m[j] = (i < n) ? mask[i] : MT(0);
            }
        }
        unsigned int mine = 0;
        int cnt = 0;
        bool weird = false;
#pragma unroll
        for (int j = 0; j < kItems; ++j) {
            weird |= !(m[j] == MT(0) || m[j] == MT(1));
            const unsigned bal = __ballot_sync(0xffffffffu, mask_nonzero(m[j]));  // elements past n were read as 0
            if (lane == j) mine = bal;  // word j of the chunk's bitmap
            cnt += __popc(bal);
        }
        if (lane < kItems) bits[static_cast<size_t>(ch) * kItems + lane] = mine;  // flat bitmap, element order
        tile_cnt += cnt;
        if ((ch & (kChunksPerTile - 1)) == kChunksPerTile - 1) {  // last chunk of a tile this warp owns
            const int tile = static_cast<int>(ch / kChunksPerTile);
            if (lane == 0) tile_counts[tile] = tile_cnt;
            super_cnt += tile_cnt;
            tile_cnt = 0;
            if ((tile + 1) % kSuper == 0) {
                if (lane == 0 && super_cnt != 0) atomicAdd(&super_counts[tile / kSuper], super_cnt);
                super_cnt = 0;
            }
        }
        if (cnt == 0) continue;
        // ---- seqlens: sum of the mask VALUES per row
        unsigned int row = c0 / S, off = c0 - row * S;
        const bool binary = !__any_sync(0xffffffffu, weird);
        if (binary && off + (kChunk - 1) < S) {  // 0/1 mask, whole chunk inside one row (the usual case)
            add_row(row, cnt);
            continue;
        }
#pragma unroll 1
        for (int j = 0; j < kItems; ++j) {
            const unsigned bal = __shfl_sync(0xffffffffu, mine, j);
            long long mv = 0;  // this lane's element of item j
#pragma unroll
            for (int jj = 0; jj < kItems; ++jj) mv = (jj == j) ? static_cast<long long>(m[jj]) : mv;
            if (off + 31 < S) {  // the item's 32 elements sit in one row
                int v = __popc(bal);
                if (!binary) {
                    v = static_cast<int>(mv);
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                }
                add_row(row, v);
            } else if ((bal >> lane) & 1u) {  // a row boundary inside the 32 elements (S % 32 != 0 or S < 32)
                atomicAdd(&seqlens[row + (off + lane) / S], static_cast<int>(mv));
            }
            off += 32;
            if (S >= 32 && off >= S) off -= S, row += 1;  // S < 32: (off + lane) / S above absorbs the growth
        }
    }
    if (lane == 0 && cur_sum != 0) atomicAdd(&seqlens[cur_row], cur_sum);
    if (lane == 0 && super_cnt != 0) atomicAdd(&super_counts[(t_end - 1) / kSuper], super_cnt);
}

// extra block e of the write kernel: cu_seqlens[r + 1] for rows [e*R, (e+1)*R) (int32 running sums wrap like torch's
// int32 cumsum would), and the max row sum through a biased atomicMax; the extra block that finishes last publishes it
__device__ void unpad_rows_block(int e, int num_extra, const int* __restrict__ seqlens, int B,
                                 int* __restrict__ cu_seqlens, int* __restrict__ totals, UnpadHeader* hdr) {
    __shared__ int red[32];
    __shared__ int s_base;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int r_begin = e * kRowsPerBlock;
    int part = 0;  // sum of the rows before this block's range (nothing to do for the first block)
    for (int r = threadIdx.x; r < r_begin; r += kTileThreads) part += seqlens[r];
    part = block_sum<int>(part, red);
    if (threadIdx.x == 0) s_base = part;
    int v[8], mx = INT_MIN, run = 0;
    const int r0 = r_begin + threadIdx.x * 8;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        v[i] = (r0 + i < B) ? seqlens[r0 + i] : 0;
        if (r0 + i < B) mx = max(mx, v[i]);
        run += v[i];
    }
    int sc = run;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, sc, d);
        if (lane >= d) sc += o;
    }
    __syncthreads();  // s_base visible, red[] free again
    if (lane == 31) red[w] = sc;
    __syncthreads();
    int pre = s_base;
    for (int ww = 0; ww < w; ++ww) pre += red[ww];
    int acc = pre + sc - run;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        acc += v[i];
        if (r0 + i < B) cu_seqlens[r0 + i + 1] = acc;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    __syncthreads();
    if (lane == 0) red[w] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int ww = 1; ww < kWarps; ++ww) mx = max(mx, red[ww]);
        if (e == 0) cu_seqlens[0] = 0;
        atomicMax(&hdr->max_biased, static_cast<unsigned int>(mx) ^ 0x80000000u);
        __threadfence();
        if (atomicAdd(&hdr->extra_done, 1u) == static_cast<unsigned int>(num_extra - 1))
            totals[1] = static_cast<int>(atomicMax(&hdr->max_biased, 0u) ^ 0x80000000u);
    }
}

__global__ void __launch_bounds__(kTileThreads)
unpad_write_kernel(const unsigned int* __restrict__ bits, const int* __restrict__ tile_counts,
                   const int* __restrict__ super_counts, int num_tiles, long long* __restrict__ indices,
                   const int* __restrict__ seqlens, int B, int* __restrict__ cu_seqlens, int* __restrict__ totals,
                   UnpadHeader* hdr) {
    if (static_cast<int>(blockIdx.x) >= num_tiles) {
        unpad_rows_block(static_cast<int>(blockIdx.x) - num_tiles, static_cast<int>(gridDim.x) - num_tiles, seqlens, B,
                         cu_seqlens, totals, hdr);
        return;
    }
    __shared__ unsigned int sbits[kWords];
    __shared__ int excl[kWords];
    __shared__ int wtot[kWords / 32];
    __shared__ int s_prefix;
    const int tile = static_cast<int>(blockIdx.x);
    const long long tile0 = static_cast<long long>(tile) * kTile;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (w == kWarps - 1) {  // nonzeros before this tile: super-counts before its group + tile counts inside the group
        const int sup = tile / kSuper;
        int acc = 0;
        for (int i = lane; i < sup; i += 32) acc += super_counts[i];
        for (int t = sup * kSuper + lane; t < tile; t += 32) acc += tile_counts[t];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) {
            s_prefix = acc;
            if (tile == num_tiles - 1) totals[0] = acc + tile_counts[tile];  // nnz
        }
    }
    int inc = 0, cnt = 0;
    if (threadIdx.x < kWords) {
        const unsigned int b = bits[static_cast<size_t>(tile) * kWords + threadIdx.x];
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
    const long long off = s_prefix;
    const unsigned int below = (1u << lane) - 1u;
#pragma unroll
    for (int j = 0; j < kItems; ++j) {
        const int q = j * kWarps + w;
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
    // header | super-counts (i32) | tile counts (i32) | bitmap: 1 bit per element
    return sizeof(UnpadHeader) + round_up(ceil_div(tiles, kSuper) * sizeof(int), 16) + round_up(tiles * sizeof(int), 16) +
           static_cast<size_t>(tiles) * kWords * sizeof(unsigned int);
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
    if (tiles == 0) {  // empty mask: no valid token, every row length 0
        VPG_CUDA(cudaMemsetAsync(cu_seqlens, 0, static_cast<size_t>(B + 1) * sizeof(int), stream));
        VPG_CUDA(cudaMemsetAsync(totals, 0, 2 * sizeof(int), stream));
        if (B > 0) VPG_CUDA(cudaMemsetAsync(seqlens, 0, static_cast<size_t>(B) * sizeof(int), stream));
        return VPG_OK;
    }
    uint8_t* ws = static_cast<uint8_t*>(workspace);
    const size_t sup_bytes = round_up(ceil_div(static_cast<int64_t>(tiles), kSuper) * sizeof(int), 16);
    const size_t cnt_bytes = round_up(static_cast<int64_t>(tiles) * sizeof(int), 16);
    UnpadHeader* hdr = reinterpret_cast<UnpadHeader*>(ws);
    int* super_counts = reinterpret_cast<int*>(ws + sizeof(UnpadHeader));
    int* tile_counts = reinterpret_cast<int*>(ws + sizeof(UnpadHeader) + sup_bytes);
    unsigned int* bits = reinterpret_cast<unsigned int*>(ws + sizeof(UnpadHeader) + sup_bytes + cnt_bytes);
    VPG_CUDA(cudaMemsetAsync(ws, 0, sizeof(UnpadHeader) + sup_bytes, stream));  // header + super-counts
    VPG_CUDA(cudaMemsetAsync(seqlens, 0, static_cast<size_t>(B) * sizeof(int), stream));
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        // one block per SM ring budget: as many resident blocks as the staging rings allow (<= 4 per SM); every warp
        // owns a contiguous run of tiles
        const size_t smem = static_cast<size_t>(kWarps) * kStages * kChunk * sizeof(MT);
        VPG_CUDA(cudaFuncSetAttribute(unpad_count_kernel<MT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(smem)));  // per device, cheap: set on every call
        int per_sm = static_cast<int>(200 * 1024 / (smem + 1024));
        per_sm = per_sm < 1 ? 1 : per_sm > 4 ? 4 : per_sm;
        const int max_warps = sm_count() * per_sm * kWarps;
        const int tiles_per_warp = static_cast<int>(ceil_div(static_cast<int64_t>(tiles), max_warps));
        const int blocks = static_cast<int>(ceil_div(ceil_div(static_cast<int64_t>(tiles), tiles_per_warp), kWarps));
        unpad_count_kernel<MT><<<blocks, kTileThreads, smem, stream>>>(
            static_cast<const MT*>(mask), static_cast<unsigned int>(n), static_cast<unsigned int>(S), tiles,
            tiles_per_warp, tile_counts, super_counts, bits, seqlens);
        count_launch();
    });
    const int extra = static_cast<int>(ceil_div(B, static_cast<int64_t>(kRowsPerBlock)));
    unpad_write_kernel<<<tiles + extra, kTileThreads, 0, stream>>>(bits, tile_counts, super_counts, tiles,
                                                                  reinterpret_cast<long long*>(indices), seqlens,
                                                                  static_cast<int>(B), cu_seqlens, totals, hdr);
    count_launch();
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
    if (zero_fill == 2 && n_idx > 0 && n_out > 0) {
        // the caller guarantees STRICTLY ASCENDING indices (unpad_input's): one pass writes every output byte once
        VPG_CHECK_ARG(src && indices && out, "null pointer argument");
        VPG_CUDA(launch_scatter_rows_gapfill(src, reinterpret_cast<const long long*>(indices), n_idx, n_out, row_bytes,
                                             out, stream));
        count_launch();
        return VPG_OK;
    }
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
