// Covariance-gated policy losses of the entropy-mechanism recipes — memory-bound token kernels.
// Reference: verl/trainer/ppo/core_algos.py:818-905 (compute_policy_loss_clip_cov) and :908-972
// (compute_policy_loss_kl_cov).  Both gate tokens on cov = (A - mean A) * (log_prob - mean log_prob):
//   clip_cov: tokens with lb < cov < ub that the PPO clip did not already cut are ELIGIBLE; a random subset of
//             max(int(ratio * sum mask), 1) of them gets its loss (and gradient) zeroed.  The reference draws the subset
//             with torch.randperm on the host generator; to stay draw-for-draw identical the eligible set is produced
//             here as flags (ordered compaction = vpg_unpad_indices), the draw stays on the host, and the loss kernel
//             takes the chosen positions back as flags.
//   kl_cov:   the k = max(1, int(N_valid * ratio)) valid tokens with the LARGEST cov get an extra ppo_kl_coef * |lp - old|.
//             The reference copies the valid tokens to the CPU and calls torch.topk; here the k-th largest is found by
//             a 6-pass radix select over 64-bit keys (monotone cov bits << 32 | inverted flat index, so keys are
//             unique and ties go to the lower index), each pass one coalesced sweep + an 11-bit shared-memory
//             histogram, the bin choice made by the last block to finish (no host round trip).
// Every kernel emits the loss MATRIX and its elementwise derivative w.r.t. log_prob; aggregation (all four
// loss_agg_mode values) is vpg_agg_loss on that matrix.  Roofline: HBM (20 B/token per sweep).
#include <cuda_runtime.h>
#include <math.h>

#include "token_common.cuh"

namespace vpg {

// workspace: double sums[10] | u64 kth_key | u64 k_remaining | u32 counters[8] | u32 hist[2048]
struct CovWs {
    double sums[10];  // 8: per-op counter (tokens zeroed / selected); 0 sum m, 1 sum A*m, 2 sum lp*m, 3 #(m>0), 4 sum_{m>0} A, 5 sum_{m>0} lp, 6 sum -(lp-old)*m, 7 sum |lp-old|*m
    unsigned long long kth_key;
    unsigned long long k_remaining;
    unsigned int counters[8];
    unsigned int hist[2048];
};

template <int VEC, typename MT>
__global__ void cov_stats_kernel(const float* __restrict__ old_lp, const float* __restrict__ lp,
                                 const float* __restrict__ adv, const MT* __restrict__ mask, long long n,
                                 CovWs* __restrict__ ws) {
    __shared__ float red[32];
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * VEC;
    for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * VEC; i < n; i += stride) {
        const Pack<VEC> o = ldv<VEC>(old_lp, i), l = ldv<VEC>(lp, i), a = ldv<VEC>(adv, i), m = ldm<VEC, MT>(mask, i);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const bool on = m.v[j] != 0.f, gt = m.v[j] > 0.f;
            const float d = l.v[j] - o.v[j];
            acc[0] += m.v[j];
            acc[1] += (on ? a.v[j] : 0.f) * m.v[j];
            acc[2] += (on ? l.v[j] : 0.f) * m.v[j];
            acc[3] += gt ? 1.f : 0.f;
            acc[4] += gt ? a.v[j] : 0.f;
            acc[5] += gt ? l.v[j] : 0.f;
            acc[6] += (on ? -d : 0.f) * m.v[j];
            acc[7] += (on ? fabsf(d) : 0.f) * m.v[j];
        }
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const float t = block_sum<float>(acc[q], red);
        if (threadIdx.x == 0 && t != 0.f) atomicAdd(&ws->sums[q], static_cast<double>(t));
    }
}

// masked_mean(x, mask) = sum / (sum m + 1e-8) in fp32 (verl/utils/torch_functional.py:171-185)
__device__ __forceinline__ float mm(const CovWs* ws, int q) {
    return static_cast<float>(ws->sums[q]) / (static_cast<float>(ws->sums[0]) + 1e-8f);
}

template <int VEC, typename MT>
__global__ void clip_cov_flags_kernel(const float* __restrict__ old_lp, const float* __restrict__ lp,
                                      const float* __restrict__ adv, const MT* __restrict__ mask, long long n,
                                      float ratio_lo, float ratio_hi, float cov_lb, float cov_ub,
                                      const CovWs* __restrict__ ws, unsigned char* __restrict__ eligible) {
    const float mean_a = mm(ws, 1), mean_l = mm(ws, 2);
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * VEC;
    for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * VEC; i < n; i += stride) {
        const Pack<VEC> o = ldv<VEC>(old_lp, i), l = ldv<VEC>(lp, i), a = ldv<VEC>(adv, i), m = ldm<VEC, MT>(mask, i);
        unsigned char f[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const float ratio = expf(l.v[j] - o.v[j]);
            const float l1 = -a.v[j] * ratio;
            const float l2 = -a.v[j] * clampf(ratio, ratio_lo, ratio_hi);
            const bool valid = m.v[j] > 0.f;
            const bool clipped = (l2 > l1) && valid;
            float cov = (a.v[j] - mean_a) * (l.v[j] - mean_l);
            if (m.v[j] == 0.f || clipped) cov = -INFINITY;
            f[j] = (cov < cov_ub && cov > cov_lb && valid) ? 1 : 0;
        }
        if constexpr (VEC == 4) {
            *reinterpret_cast<uchar4*>(eligible + i) = make_uchar4(f[0], f[1], f[2], f[3]);
        } else {
            eligible[i] = f[0];
        }
    }
}

template <int VEC, typename MT>
__global__ void clip_cov_loss_kernel(const float* __restrict__ old_lp, const float* __restrict__ lp,
                                     const float* __restrict__ adv, const MT* __restrict__ mask, long long n,
                                     float ratio_lo, float ratio_hi, const unsigned char* __restrict__ zeroed,
                                     float* __restrict__ pg_losses, float* __restrict__ dpg, CovWs* __restrict__ ws,
                                     float* __restrict__ out3) {
    __shared__ float red[32];
    float nz = 0.f;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * VEC;
    for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * VEC; i < n; i += stride) {
        const Pack<VEC> o = ldv<VEC>(old_lp, i), l = ldv<VEC>(lp, i), a = ldv<VEC>(adv, i), m = ldm<VEC, MT>(mask, i);
        Pack<VEC> pg, dg;
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const float corr = zeroed[i + j] ? 0.f : 1.f;
            const float ratio = expf(l.v[j] - o.v[j]);
            const float rc = clampf(ratio, ratio_lo, ratio_hi);
            const float l1 = -a.v[j] * ratio, l2 = -a.v[j] * rc;
            pg.v[j] = max_nan(l1, l2) * corr;
            // torch.maximum sends the gradient to the larger branch, half to each on a tie; clamp passes it inside
            // [min, max] (bounds included)
            const float w1 = (l1 > l2) ? 1.f : (l1 == l2 ? 0.5f : 0.f);
            const float inside = (ratio >= ratio_lo && ratio <= ratio_hi) ? 1.f : 0.f;
            dg.v[j] = corr * (w1 * l1 + (1.f - w1) * inside * l1);  // d(-A*ratio)/d lp = -A*ratio = l1
            nz += ((m.v[j] != 0.f) ? (1.f - corr) : 0.f) * m.v[j];
        }
        stv<VEC>(pg_losses, i, pg);
        stv<VEC>(dpg, i, dg);
    }
    const float t = block_sum<float>(nz, red);
    if (threadIdx.x == 0 && t != 0.f) atomicAdd(&ws->sums[8], static_cast<double>(t));  // sum (corr==0)*m
    if (last_block_done(&ws->counters[7], gridDim.x) && threadIdx.x == 0) {
        const float denom = static_cast<float>(ws->sums[0]) + 1e-8f;
        out3[0] = static_cast<float>(ws->sums[6]) / denom;                                       // ppo_kl
        out3[1] = static_cast<float>(*reinterpret_cast<volatile double*>(&ws->sums[8])) / denom;  // pg_clipfrac
        out3[2] = static_cast<float>(ws->sums[0]);
    }
}

// ---- kl_cov: radix select of the k-th largest key
__device__ __forceinline__ unsigned int ordered_bits(float f) {
    const unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ unsigned long long cov_key(float a, float l, float mean_a, float mean_l, long long i) {
    const float cov = (a - mean_a) * (l - mean_l);
    return (static_cast<unsigned long long>(ordered_bits(cov)) << 32) |
           static_cast<unsigned long long>(0xFFFFFFFFu - static_cast<unsigned int>(i));
}
__device__ __forceinline__ float valid_mean(const CovWs* ws, int q) {
    return static_cast<float>(ws->sums[q] / ws->sums[3]);
}

// pass p looks at key bits [shift, shift+bits) of the keys whose higher bits equal the prefix chosen so far
template <int VEC, typename MT>
__global__ void kl_cov_select_kernel(const float* __restrict__ lp, const float* __restrict__ adv,
                                     const MT* __restrict__ mask, long long n, int pass, int shift, int bits,
                                     double ratio, CovWs* __restrict__ ws) {
    __shared__ unsigned int h[2048];
    for (int b = threadIdx.x; b < 2048; b += blockDim.x) h[b] = 0;
    __syncthreads();
    const bool any = ws->sums[3] > 0.0;
    const float mean_a = any ? valid_mean(ws, 4) : 0.f, mean_l = any ? valid_mean(ws, 5) : 0.f;
    const unsigned long long prefix = ws->kth_key;  // bits chosen so far, right-aligned
    const unsigned int bin_mask = (1u << bits) - 1u;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * VEC;
    for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * VEC; i < n; i += stride) {
        const Pack<VEC> l = ldv<VEC>(lp, i), a = ldv<VEC>(adv, i), m = ldm<VEC, MT>(mask, i);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            if (!(m.v[j] > 0.f)) continue;
            const unsigned long long key = cov_key(a.v[j], l.v[j], mean_a, mean_l, i + j);
            if (pass == 0 || (key >> (shift + bits)) == prefix)
                atomicAdd(&h[static_cast<unsigned int>(key >> shift) & bin_mask], 1u);
        }
    }
    __syncthreads();
    for (int b = threadIdx.x; b < 2048; b += blockDim.x)
        if (h[b]) atomicAdd(&ws->hist[b], h[b]);
    if (!last_block_done(&ws->counters[pass], gridDim.x)) return;
    // ---- last block: pick the bin that holds the k-th largest key, from the top bin downwards
    for (int b = threadIdx.x; b < 2048; b += blockDim.x) {
        h[b] = *reinterpret_cast<volatile unsigned int*>(&ws->hist[b]);
        ws->hist[b] = 0;  // ready for the next pass
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        unsigned long long k = ws->k_remaining;
        if (pass == 0) {
            // k = max(1, int(N_valid * ratio))  (core_algos.py:957); no valid token -> nothing is selected
            const long long kk = static_cast<long long>(ws->sums[3] * ratio);
            k = static_cast<unsigned long long>(kk < 1 ? 1 : kk);
        }
        const int lane = threadIdx.x;
        const int top = (1 << bits) - 1;
        // lane owns bins [top - 64*lane - 63, top - 64*lane] (descending); 32 lanes x 64 = 2048 bins
        unsigned long long mine = 0;
        for (int q = 0; q < 64; ++q) {
            const int b = top - 64 * lane - q;
            if (b >= 0) mine += h[b];
        }
        unsigned long long incl = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        const unsigned long long excl = incl - mine;
        const bool owner = (excl < k) && (incl >= k);
        const unsigned int ballot = __ballot_sync(0xffffffffu, owner);
        if (ballot == 0) {  // fewer than k keys in total (only when there is no valid token): select nothing
            if (lane == 0) {
                ws->kth_key = ~0ull;
                ws->k_remaining = 0;
            }
        } else if (owner) {
            unsigned long long run = excl;
            int chosen = 0;
            for (int q = 0; q < 64; ++q) {
                const int b = top - 64 * lane - q;
                if (b < 0) break;
                if (run + h[b] >= k) {
                    chosen = b;
                    break;
                }
                run += h[b];
            }
            ws->kth_key = (pass == 0 ? 0ull : (prefix << bits)) | static_cast<unsigned long long>(chosen);
            ws->k_remaining = k - run;
        }
    }
}

template <int VEC, typename MT>
__global__ void kl_cov_loss_kernel(const float* __restrict__ old_lp, const float* __restrict__ lp,
                                   const float* __restrict__ adv, const MT* __restrict__ mask, long long n,
                                   float kl_coef, float* __restrict__ pg_losses, float* __restrict__ dpg,
                                   CovWs* __restrict__ ws, float* __restrict__ out3) {
    __shared__ float red[32];
    const bool any = ws->sums[3] > 0.0;
    const float mean_a = any ? valid_mean(ws, 4) : 0.f, mean_l = any ? valid_mean(ws, 5) : 0.f;
    const unsigned long long kth = any ? ws->kth_key : ~0ull;
    float nsel = 0.f;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * VEC;
    for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * VEC; i < n; i += stride) {
        const Pack<VEC> o = ldv<VEC>(old_lp, i), l = ldv<VEC>(lp, i), a = ldv<VEC>(adv, i), m = ldm<VEC, MT>(mask, i);
        Pack<VEC> pg, dg;
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const float d = l.v[j] - o.v[j];
            const float l1 = -a.v[j] * expf(d);
            const bool sel = any && (m.v[j] > 0.f) && cov_key(a.v[j], l.v[j], mean_a, mean_l, i + j) >= kth;
            const float sgn = (d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f);
            pg.v[j] = sel ? (l1 + kl_coef * fabsf(d)) : l1;
            dg.v[j] = sel ? (l1 + kl_coef * sgn) : l1;
            nsel += sel ? 1.f : 0.f;
        }
        stv<VEC>(pg_losses, i, pg);
        stv<VEC>(dpg, i, dg);
    }
    const float t = block_sum<float>(nsel, red);
    if (threadIdx.x == 0 && t != 0.f) atomicAdd(&ws->sums[8], static_cast<double>(t));  // #selected
    if (last_block_done(&ws->counters[7], gridDim.x) && threadIdx.x == 0) {
        out3[0] = static_cast<float>(ws->sums[7]) / (static_cast<float>(ws->sums[0]) + 1e-8f);  // ppo_kl_abs
        out3[1] = static_cast<float>(*reinterpret_cast<volatile double*>(&ws->sums[8]));         // tokens selected
        out3[2] = static_cast<float>(ws->sums[0]);
    }
}

static int check_cov_args(const void* a, const void* b, const void* c, const void* m, int64_t n, const void* ws,
                          size_t ws_bytes) {
    VPG_CHECK_ARG(a && b && c && m && ws, "null pointer argument");
    VPG_CHECK_ARG(n > 0 && n < (1LL << 32), "token count must be in [1, 2^32)");
    if (ws_bytes < sizeof(CovWs)) return set_error(VPG_ERR_WORKSPACE, "workspace too small: %zu < %zu", ws_bytes, sizeof(CovWs));
    return VPG_OK;
}

}  // namespace vpg

using namespace vpg;

#define COV_LAUNCH(KERNEL, ...)                                                                    \
    VPG_DISPATCH_MASK(mask_dtype, MT, {                                                            \
        const MT* mk = static_cast<const MT*>(mask);                                               \
        if (vec)                                                                                   \
            KERNEL<4, MT><<<flat_grid(n, 4), kThreads, 0, stream>>>(__VA_ARGS__);                  \
        else                                                                                       \
            KERNEL<1, MT><<<flat_grid(n, 1), kThreads, 0, stream>>>(__VA_ARGS__);                  \
    });                                                                                            \
    VPG_LAUNCH_CHECK()

extern "C" {

size_t vpg_cov_workspace_bytes(void) { return sizeof(CovWs); }

int vpg_clip_cov_flags(const float* old_log_prob, const float* log_prob, const float* advantages, const void* mask,
                       int mask_dtype, int64_t n, float cliprange_low, float cliprange_high, float cov_lb, float cov_ub,
                       uint8_t* eligible, void* workspace, size_t workspace_bytes, void* stream_) {
    if (int rc = check_cov_args(old_log_prob, log_prob, advantages, mask, n, workspace, workspace_bytes)) return rc;
    VPG_CHECK_ARG(eligible, "null pointer argument");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    CovWs* ws = static_cast<CovWs*>(workspace);
    VPG_CUDA(cudaMemsetAsync(ws, 0, sizeof(CovWs), stream));
    const bool vec = rows_vec_ok(n, mask, mask_dtype, {old_log_prob, log_prob, advantages}) &&
                     (reinterpret_cast<uintptr_t>(eligible) & 3) == 0;
    const float lo = 1.f - cliprange_low, hi = 1.f + cliprange_high;
    COV_LAUNCH(cov_stats_kernel, old_log_prob, log_prob, advantages, mk, n, ws);
    COV_LAUNCH(clip_cov_flags_kernel, old_log_prob, log_prob, advantages, mk, n, lo, hi, cov_lb, cov_ub, ws, eligible);
    return VPG_OK;
}

int vpg_clip_cov_loss(const float* old_log_prob, const float* log_prob, const float* advantages, const void* mask,
                      int mask_dtype, int64_t n, float cliprange_low, float cliprange_high, const uint8_t* zeroed,
                      float* pg_losses, float* d_pg_losses, float* out3, void* workspace, size_t workspace_bytes,
                      void* stream_) {
    if (int rc = check_cov_args(old_log_prob, log_prob, advantages, mask, n, workspace, workspace_bytes)) return rc;
    VPG_CHECK_ARG(zeroed && pg_losses && d_pg_losses && out3, "null pointer argument");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    CovWs* ws = static_cast<CovWs*>(workspace);  // sums[0], sums[6] from vpg_clip_cov_flags on the same workspace
    VPG_CUDA(cudaMemsetAsync(&ws->sums[8], 0, sizeof(double), stream));
    VPG_CUDA(cudaMemsetAsync(&ws->counters[7], 0, sizeof(unsigned int), stream));
    const bool vec = rows_vec_ok(n, mask, mask_dtype, {old_log_prob, log_prob, advantages, pg_losses, d_pg_losses});
    const float lo = 1.f - cliprange_low, hi = 1.f + cliprange_high;
    COV_LAUNCH(clip_cov_loss_kernel, old_log_prob, log_prob, advantages, mk, n, lo, hi, zeroed, pg_losses, d_pg_losses,
               ws, out3);
    return VPG_OK;
}

int vpg_kl_cov_loss(const float* old_log_prob, const float* log_prob, const float* advantages, const void* mask,
                    int mask_dtype, int64_t n, double kl_cov_ratio, float ppo_kl_coef, float* pg_losses,
                    float* d_pg_losses, float* out3, void* workspace, size_t workspace_bytes, void* stream_) {
    if (int rc = check_cov_args(old_log_prob, log_prob, advantages, mask, n, workspace, workspace_bytes)) return rc;
    VPG_CHECK_ARG(pg_losses && d_pg_losses && out3, "null pointer argument");
    VPG_CHECK_ARG(kl_cov_ratio > 0, "kl_cov_ratio should be larger than 0.");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    CovWs* ws = static_cast<CovWs*>(workspace);
    VPG_CUDA(cudaMemsetAsync(ws, 0, sizeof(CovWs), stream));
    const bool vec = rows_vec_ok(n, mask, mask_dtype, {old_log_prob, log_prob, advantages, pg_losses, d_pg_losses});
    COV_LAUNCH(cov_stats_kernel, old_log_prob, log_prob, advantages, mk, n, ws);
    // 64-bit keys, most significant bits first: 11 + 11 + 11 + 11 + 11 + 9
    int shift = 64;
    for (int pass = 0; pass < 6; ++pass) {
        const int bits = (pass < 5) ? 11 : 9;
        shift -= bits;
        COV_LAUNCH(kl_cov_select_kernel, log_prob, advantages, mk, n, pass, shift, bits, kl_cov_ratio, ws);
    }
    COV_LAUNCH(kl_cov_loss_kernel, old_log_prob, log_prob, advantages, mk, n, ppo_kl_coef, pg_losses, d_pg_losses, ws,
               out3);
    return VPG_OK;
}

}  // extern "C"
