// Outcome-reward advantage estimators that share GRPO's skeleton (row score -> per-uid group statistics -> per-row
// scalar -> broadcast * mask) and the two reverse-scan estimators (REINFORCE++, ReMax).
// Reference: verl/trainer/ppo/core_algos.py
//   GRPO :245-308, GRPO-PassK :309-369, REINFORCE++-baseline :372-422, RLOO :425-476, OPO :479-530,
//   REINFORCE++ :533-569, ReMax :572-605, GPG :608-667.
// The reference evaluates all of these with Python loops over the batch (and over time steps for REINFORCE++) on
// the driver CPU.  Here: one coalesced sweep for the row statistics (128-bit loads), one single-block kernel for the
// group statistics (B <= a few thousand rows: latency-bound, fp64 atomics in shared/global memory), one coalesced
// sweep that writes the [B,L] advantage.  Compiled with --fmad=false like token_ops.cu.
#include <cuda_runtime.h>

#include "token_common.cuh"

namespace vpg {

// -------------------------------------------------------------------------------------------------------------
// row statistics: score[b] = sum_t rewards[b,t] (NOT masked, core_algos.py:281), len[b] = sum_t mask[b,t]
// -------------------------------------------------------------------------------------------------------------
template <int VEC, typename MT>
__global__ void row_score_len_kernel(const float* __restrict__ rewards, const MT* __restrict__ mask, long long L, int gx,
                                     double* __restrict__ scores, double* __restrict__ lens) {
    __shared__ double red[32];
    const long long b = blockIdx.x / gx;
    const int cx = blockIdx.x % gx;
    const long long base = b * L;
    float s = 0.f, c = 0.f;
    const long long stride = static_cast<long long>(gx) * blockDim.x * VEC;
    for (long long t = (static_cast<long long>(cx) * blockDim.x + threadIdx.x) * VEC; t < L; t += stride) {
        const Pack<VEC> v = ldv<VEC>(rewards, base + t);
        const Pack<VEC> m = ldm<VEC, MT>(mask, base + t);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            s += v.v[j];
            c += m.v[j];
        }
    }
    const double bs = block_sum<double>(static_cast<double>(s), red);
    const double bc = block_sum<double>(static_cast<double>(c), red);
    if (threadIdx.x == 0) {
        atomicAdd(&scores[b], bs);
        atomicAdd(&lens[b], bc);
    }
}

// order-preserving map float -> uint32 (larger float <=> larger key)
__device__ __forceinline__ unsigned int float_key(float f) {
    const unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_float(unsigned int k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

constexpr int kGStat = 5;  // per group: count, sum, M2 (squared deviations), sum(len), sum(len*score)

// single block.  gstat [kGStat*G] doubles and top [2*G] u64 are zeroed by the caller.
__global__ void __launch_bounds__(1024)
outcome_group_kernel(int kind, const double* __restrict__ scores, const double* __restrict__ lens,
                     const int* __restrict__ gid, long long B, long long G, float eps, int norm_by_std, float f_norm,
                     double* __restrict__ gstat, unsigned long long* __restrict__ top, float* __restrict__ adv_row,
                     int* __restrict__ status) {
    __shared__ double red[32];
    __shared__ double bc[4];
    // pass 1: count / sum (+ length-weighted sums, + group arg-max).  Scores are rounded to fp32 first: the reference
    // holds them in an fp32 tensor.
    // Sums per group are warp-shuffle segmented reductions: `__match_any_sync` on the group id yields the lanes of a
    // warp's 32 consecutive rows that share a group, the partial is summed over that set in lane order and its lowest
    // lane issues one atomic per (warp, group) and quantity.
    double nnz_local = 0.0;
    const int lane = threadIdx.x & 31;
    for (long long r0 = threadIdx.x - lane; r0 < B; r0 += blockDim.x) {
        const long long i = r0 + lane;
        const bool ok = i < B;
        const int g = ok ? gid[i] : -1;
        const float sf = ok ? static_cast<float>(scores[i]) : 0.f;
        const double len = (ok && kind == VPG_ADV_OPO) ? lens[i] : 0.0;
        const unsigned peers = __match_any_sync(0xffffffffu, g);
        const bool lead = ok && lane == __ffs(peers) - 1;
        const double sum = segment_sum(static_cast<double>(sf), peers);
        if (lead) {
            atomicAdd(&gstat[kGStat * g + 0], static_cast<double>(__popc(peers)));
            atomicAdd(&gstat[kGStat * g + 1], sum);
        }
        if (kind == VPG_ADV_OPO) {
            const double sl = segment_sum(len, peers);
            // fp32 product as in (len_tensor * score_tensor), core_algos.py:522
            const double slw = segment_sum(static_cast<double>(static_cast<float>(len) * sf), peers);
            if (lead) {
                atomicAdd(&gstat[kGStat * g + 3], sl);
                atomicAdd(&gstat[kGStat * g + 4], slw);
            }
        }
        if (ok && kind == VPG_ADV_GRPO_PASSK) {
            const unsigned long long key = (static_cast<unsigned long long>(float_key(sf)) << 32) |
                                           static_cast<unsigned long long>(0xffffffffu - static_cast<unsigned int>(i));
            atomicMax(&top[g], key);
        }
        if (ok && sf != 0.f) nnz_local += 1.0;
    }
    __syncthreads();
    const double nnz = block_sum<double>(nnz_local, red);
    if (threadIdx.x == 0) bc[0] = nnz;
    // pass 2: squared deviations from the group mean (+ runner-up for pass@k)
    for (long long r0 = threadIdx.x - lane; r0 < B; r0 += blockDim.x) {
        const long long i = r0 + lane;
        const bool ok = i < B;
        const int g = ok ? gid[i] : -1;
        double d = 0.0, n = 0.0;
        float sf = 0.f;
        if (ok) {
            n = gstat[kGStat * g + 0];
            sf = static_cast<float>(scores[i]);
            d = static_cast<double>(sf) - gstat[kGStat * g + 1] / n;
        }
        const unsigned peers = __match_any_sync(0xffffffffu, g);
        const double sq = segment_sum(d * d, peers);
        if (ok && lane == __ffs(peers) - 1) atomicAdd(&gstat[kGStat * g + 2], sq);
        if (ok && kind == VPG_ADV_GRPO_PASSK) {
            const unsigned int arg = 0xffffffffu - static_cast<unsigned int>(top[g] & 0xffffffffull);
            if (static_cast<unsigned int>(i) != arg) {
                const unsigned long long key = (static_cast<unsigned long long>(float_key(sf)) << 32) |
                                               static_cast<unsigned long long>(0xffffffffu - static_cast<unsigned int>(i));
                atomicMax(&top[G + g], key);
            }
            if (n < 2.0) status[0] = 3;  // "Pass@k requires at least 2 samples per group" (core_algos.py:355-358)
        }
    }
    __syncthreads();
    // pass 3: per-row scalar
    double w1 = 0.0, wn = 0.0;  // REINFORCE++-baseline: sum len*s, sum len
    for (long long i = threadIdx.x; i < B; i += blockDim.x) {
        const int g = gid[i];
        const double n = gstat[kGStat * g + 0];
        const float sc = static_cast<float>(scores[i]);
        const bool single = n <= 1.0;
        const float mean = single ? 0.f : static_cast<float>(gstat[kGStat * g + 1] / n);
        const float stdv = single ? 1.f : static_cast<float>(sqrt(gstat[kGStat * g + 2] / (n - 1.0)));  // torch.std
        float a;
        if (kind == VPG_ADV_GRPO) {
            a = norm_by_std ? (sc - mean) / (stdv + eps) : (sc - mean);
        } else if (kind == VPG_ADV_RLOO) {
            // scores[i] * n / (n - 1) - mean * n / (n - 1)  for n > 1, unchanged otherwise (core_algos.py:467-471)
            const float nf = static_cast<float>(n), n1 = static_cast<float>(n - 1.0);
            a = single ? sc : (sc * nf / n1 - mean * nf / n1);
        } else if (kind == VPG_ADV_OPO) {
            const float bsl = single ? 0.f
                                     : static_cast<float>(gstat[kGStat * g + 4]) / static_cast<float>(gstat[kGStat * g + 3]);
            a = sc - bsl;
        } else if (kind == VPG_ADV_GPG) {
            // alpha = bsz / count_nonzero(scores).clamp(min=1)  (core_algos.py:646-647)
            const float alpha = static_cast<float>(B) / static_cast<float>(fmax(bc[0], 1.0));
            a = alpha * (sc - mean) / f_norm;
        } else if (kind == VPG_ADV_RFPP_BASELINE) {
            a = sc - mean;
            w1 += lens[i] * static_cast<double>(a);
            wn += lens[i];
        } else {  // pass@k: only the best response of a group gets r_max - r_second_max (core_algos.py:359-366)
            const unsigned long long k1 = top[g], k2 = top[G + g];
            const unsigned int arg = 0xffffffffu - static_cast<unsigned int>(k1 & 0xffffffffull);
            a = 0.f;
            if (static_cast<unsigned int>(i) == arg && n >= 2.0) {
                a = key_float(static_cast<unsigned int>(k1 >> 32)) - key_float(static_cast<unsigned int>(k2 >> 32));
                if (norm_by_std) a = a / (stdv + eps);
            }
        }
        adv_row[i] = a;
    }
    if (kind != VPG_ADV_RFPP_BASELINE) return;
    // masked_whiten of the broadcast matrix in closed form (binary masks): every valid token of row i holds s_i
    //   mean = sum_i len_i s_i / (N + 1e-8);  var = sum_i len_i (s_i - mean)^2 / (N + 1e-8) * N / (N - 1)
    const double S1 = block_sum<double>(w1, red);
    const double N = block_sum<double>(wn, red);
    if (threadIdx.x == 0) bc[1] = S1, bc[2] = N;
    __syncthreads();
    const double Nn = bc[2];
    if (Nn == 0.0 || Nn == 1.0) {
        if (threadIdx.x == 0) status[0] = (Nn == 0.0) ? 1 : 2;  // masked_var raises (torch_functional.py:195-200)
        return;
    }
    const double mean_w = bc[1] / (Nn + 1e-8);
    double v = 0.0;
    for (long long i = threadIdx.x; i < B; i += blockDim.x) {
        const double d = static_cast<double>(adv_row[i]) - mean_w;
        v += lens[i] * d * d;
    }
    __syncthreads();
    const double V = block_sum<double>(v, red);
    if (threadIdx.x == 0) bc[3] = V;
    __syncthreads();
    const double var = bc[3] / (Nn + 1e-8) * (Nn / (Nn - 1.0));
    const float meanf = static_cast<float>(mean_w);
    const float rstd = rsqrtf(static_cast<float>(var) + 1e-8f);
    for (long long i = threadIdx.x; i < B; i += blockDim.x) adv_row[i] = (adv_row[i] - meanf) * rstd;
}

// out[b,t] = row_val[b] * mask[b,t] (- is never needed); optional second output gets the same values
template <int VEC, typename MT>
__global__ void row_broadcast_kernel(const float* __restrict__ row_val, const MT* __restrict__ mask, long long L, int gx,
                                     float* __restrict__ out) {
    const long long b = blockIdx.x / gx;
    const int cx = blockIdx.x % gx;
    const long long base = b * L;
    const float a = row_val[b];
    const long long stride = static_cast<long long>(gx) * blockDim.x * VEC;
    for (long long t = (static_cast<long long>(cx) * blockDim.x + threadIdx.x) * VEC; t < L; t += stride) {
        const Pack<VEC> m = ldm<VEC, MT>(mask, base + t);
        Pack<VEC> o;
#pragma unroll
        for (int j = 0; j < VEC; ++j) o.v[j] = a * m.v[j];
        stv<VEC>(out, base + t, o);
    }
}

// -------------------------------------------------------------------------------------------------------------
// Reverse scans.  One block per row, each thread owns kScanSeg contiguous tokens in registers; the recurrence on the
// running value is affine per token (run' = a*run + b), so segments compose, a block-wide right-to-left exclusive
// scan gives each segment its incoming value, and the segment is replayed in the reference's operation order.
//   MODE 0  REINFORCE++ (core_algos.py:557-563): R_t = r_t + gamma*run ; run = R_t * m_t          (a = m*gamma, b = m*r)
//   MODE 1  ReMax       (core_algos.py:601):      run = run + r_t*m_t ; R_t = run                   (a = 1,       b = r*m)
// -------------------------------------------------------------------------------------------------------------
constexpr int kScanSeg = 8;

struct Lin {
    float a, b;
};
// h = g after f
__device__ __forceinline__ Lin lin_compose(const Lin& g, const Lin& f) { return Lin{g.a * f.a, g.a * f.b + g.b}; }
__device__ __forceinline__ Lin lin_shfl_down(const Lin& x, int d) {
    return Lin{__shfl_down_sync(0xffffffffu, x.a, d), __shfl_down_sync(0xffffffffu, x.b, d)};
}

template <typename MT, int MODE>
__global__ void __launch_bounds__(512)
reverse_scan_kernel(const float* __restrict__ rewards, const MT* __restrict__ mask, const float* __restrict__ baselines,
                    long long L, float gamma, float* __restrict__ adv_out, float* __restrict__ ret_out,
                    double* __restrict__ sums, int vec, int* status) {
    __shared__ Lin warp_maps[32];
    __shared__ float carry;
    __shared__ double red[32];
    const long long base = static_cast<long long>(blockIdx.x) * L;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const long long span = static_cast<long long>(blockDim.x) * kScanSeg;
    const float bsl = (MODE == 1) ? baselines[blockIdx.x] : 0.f;
    if (threadIdx.x == 0) carry = 0.f;
    float s1 = 0.f, s2 = 0.f, sc = 0.f;
    const long long n_chunks = (L + span - 1) / span;
    for (long long ck = n_chunks - 1; ck >= 0; --ck) {
        __syncthreads();
        const long long t0 = ck * span + static_cast<long long>(threadIdx.x) * kScanSeg;
        float r[kScanSeg], m[kScanSeg];
        const bool full = vec && (t0 + kScanSeg <= L);  // 128-bit path: 32 contiguous, aligned bytes per thread
        if (full) {
#pragma unroll
            for (int q = 0; q < kScanSeg / 4; ++q) {
                const float4 a = load_f4(rewards, base + t0 + 4 * q);
                const float4 c = load_mask4<MT>(mask, base + t0 + 4 * q);
                r[4 * q] = a.x, r[4 * q + 1] = a.y, r[4 * q + 2] = a.z, r[4 * q + 3] = a.w;
                m[4 * q] = c.x, m[4 * q + 1] = c.y, m[4 * q + 2] = c.z, m[4 * q + 3] = c.w;
            }
        } else {
#pragma unroll
            for (int j = 0; j < kScanSeg; ++j) {
                const long long t = t0 + j;
                const bool ok = t < L;
                r[j] = ok ? rewards[base + t] : 0.f;
                m[j] = ok ? mask_to_float(mask[base + t]) : 0.f;  // out-of-range tokens are identity maps (see below)
            }
        }
        Lin seg{1.f, 0.f};
#pragma unroll
        for (int j = kScanSeg - 1; j >= 0; --j) {
            const bool ok = (t0 + j) < L;
            Lin st;
            if (MODE == 0) {
                st.a = ok ? m[j] * gamma : 1.f;
                st.b = ok ? m[j] * r[j] : 0.f;
            } else {
                st.a = 1.f;
                st.b = r[j] * m[j];
            }
            seg = lin_compose(st, seg);
        }
        Lin inc = seg;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const Lin o = lin_shfl_down(inc, d);
            if (lane + d < 32) inc = lin_compose(inc, o);
        }
        if (lane == 0) warp_maps[w] = inc;
        __syncthreads();
        Lin right{1.f, 0.f};
        for (int ww = nw - 1; ww > w; --ww) right = lin_compose(warp_maps[ww], right);
        Lin excl = lin_shfl_down(inc, 1);
        if (lane == 31) excl = Lin{1.f, 0.f};
        excl = lin_compose(excl, right);
        float run = excl.a * carry + excl.b;
        float R[kScanSeg];
#pragma unroll
        for (int j = kScanSeg - 1; j >= 0; --j) {
            if ((t0 + j) < L) {
                if (MODE == 0) {
                    R[j] = r[j] + gamma * run;
                    run = R[j] * m[j];
                } else {
                    run = run + r[j] * m[j];
                    R[j] = run;
                }
            } else {
                R[j] = 0.f;
            }
        }
#pragma unroll
        for (int j = 0; j < kScanSeg; ++j) {
            const long long t = t0 + j;
            if (t < L) {
                if (!full) ret_out[base + t] = R[j];
                if (MODE == 0) {
                    const bool on = m[j] != 0.f;
                    s1 += (on ? R[j] : 0.f) * m[j];
                    s2 += (on ? R[j] * R[j] : 0.f) * m[j];
                    sc += m[j];
                } else if (!full) {
                    adv_out[base + t] = R[j] - bsl * m[j];
                }
            }
        }
        if (full) {
#pragma unroll
            for (int q = 0; q < kScanSeg / 4; ++q) {
                store_f4(ret_out, base + t0 + 4 * q, make_float4(R[4 * q], R[4 * q + 1], R[4 * q + 2], R[4 * q + 3]));
                if (MODE == 1)
                    store_f4(adv_out, base + t0 + 4 * q,
                             make_float4(R[4 * q] - bsl * m[4 * q], R[4 * q + 1] - bsl * m[4 * q + 1],
                                         R[4 * q + 2] - bsl * m[4 * q + 2], R[4 * q + 3] - bsl * m[4 * q + 3]));
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) carry = run;
    }
    if (MODE == 0) {
        const double d1 = block_sum<double>(static_cast<double>(s1), red);
        const double d2 = block_sum<double>(static_cast<double>(s2), red);
        const double dc = block_sum<double>(static_cast<double>(sc), red);
        if (threadIdx.x == 0) {
            atomicAdd(&sums[0], d1);
            atomicAdd(&sums[1], d2);
            atomicAdd(&sums[2], dc);
        }
        whiten_publish_status(sums, status);
    }
}

// advantages = masked_whiten(returns, mask) * mask  (core_algos.py:565-566; torch_functional.py:206-223)
template <int VEC, typename MT>
__global__ void whiten_mask_apply_kernel(const float* __restrict__ values, const MT* __restrict__ mask, long long n,
                                         const double* __restrict__ sums, float* __restrict__ out) {
    const double N = sums[2];
    if (N == 0.0 || N == 1.0) return;  // status word: whiten_publish_status
    const double den = N + 1e-8;
    const double mean_d = sums[0] / den;
    double var_d = (sums[1] - 2.0 * mean_d * sums[0] + mean_d * mean_d * N) / den;
    var_d = var_d * (N / (N - 1.0));
    const float mean = static_cast<float>(mean_d);
    const float rstd = rsqrtf(static_cast<float>(var_d) + 1e-8f);
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * VEC;
    for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * VEC; i < n; i += stride) {
        Pack<VEC> v = ldv<VEC>(values, i);
        const Pack<VEC> m = ldm<VEC, MT>(mask, i);
#pragma unroll
        for (int j = 0; j < VEC; ++j) v.v[j] = ((v.v[j] - mean) * rstd) * m.v[j];
        stv<VEC>(out, i, v);
    }
}

}  // namespace vpg

using namespace vpg;

extern "C" {

size_t vpg_outcome_advantage_workspace_bytes(int64_t B, int64_t G) {
    const size_t b = static_cast<size_t>(B > 0 ? B : 0), g = static_cast<size_t>(G > 0 ? G : 0);
    return (2 * b + kGStat * g + 2 * g) * sizeof(double) + b * sizeof(float) + 16;
}

int vpg_outcome_advantage(int kind, const float* rewards, const void* mask, int mask_dtype, const int32_t* group_ids,
                          int64_t B, int64_t L, int64_t G, float epsilon, int norm_by_std, float f_norm,
                          float* advantages, int32_t* status, void* workspace, size_t workspace_bytes, void* stream_) {
    VPG_CHECK_ARG(rewards && mask && group_ids && advantages && status, "null pointer argument");
    VPG_CHECK_ARG(kind >= VPG_ADV_GRPO && kind <= VPG_ADV_GRPO_PASSK, "unknown outcome advantage kind %d", kind);
    VPG_CHECK_ARG(B > 0 && L > 0 && G > 0 && G <= B, "invalid shape B=%lld L=%lld groups=%lld", (long long)B,
                  (long long)L, (long long)G);
    VPG_CHECK_ARG(B < (1ll << 31), "batch too large");
    const size_t need = vpg_outcome_advantage_workspace_bytes(B, G);
    if (!workspace || workspace_bytes < need)
        return set_error(VPG_ERR_WORKSPACE, "outcome advantage workspace too small: need %zu, got %zu", need,
                         workspace_bytes);
    VPG_CHECK_ARG((reinterpret_cast<uintptr_t>(workspace) & 7) == 0, "workspace must be 8-byte aligned");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    double* scores = static_cast<double*>(workspace);
    double* lens = scores + B;
    double* gstat = lens + B;
    unsigned long long* top = reinterpret_cast<unsigned long long*>(gstat + kGStat * G);
    float* adv_row = reinterpret_cast<float*>(top + 2 * G);
    VPG_CUDA(cudaMemsetAsync(workspace, 0, static_cast<size_t>(2 * B + (kGStat + 2) * G) * sizeof(double), stream));
    VPG_CUDA(cudaMemsetAsync(status, 0, sizeof(int32_t), stream));
    const bool vec = rows_vec_ok(L, mask, mask_dtype, {rewards, advantages});
    const int gx = blocks_per_row(B, L, vec ? 4 : 1, 32);
    const unsigned grid = static_cast<unsigned>(B * gx);
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        const MT* mk = static_cast<const MT*>(mask);
        if (vec)
            row_score_len_kernel<4, MT><<<grid, kThreads, 0, stream>>>(rewards, mk, L, gx, scores, lens);
        else
            row_score_len_kernel<1, MT><<<grid, kThreads, 0, stream>>>(rewards, mk, L, gx, scores, lens);
        count_launch();
        outcome_group_kernel<<<1, 1024, 0, stream>>>(kind, scores, lens, group_ids, B, G, epsilon, norm_by_std, f_norm,
                                                     gstat, top, adv_row, status);
        count_launch();
        if (vec)
            row_broadcast_kernel<4, MT><<<grid, kThreads, 0, stream>>>(adv_row, mk, L, gx, advantages);
        else
            row_broadcast_kernel<1, MT><<<grid, kThreads, 0, stream>>>(adv_row, mk, L, gx, advantages);
    });
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

size_t vpg_reverse_scan_workspace_bytes(int64_t B, int64_t L) {
    (void)B;
    (void)L;
    return 4 * sizeof(double);
}

static inline unsigned scan_threads(int64_t B, int64_t L) {
    int64_t threads = round_up(ceil_div(L, kScanSeg), 32);
    if (threads > scan_block_threads(B)) threads = scan_block_threads(B);
    return static_cast<unsigned>(threads);
}

int vpg_reinforce_plus_plus(const float* rewards, const void* mask, int mask_dtype, int64_t B, int64_t L, float gamma,
                            float* advantages, float* returns, int32_t* status, void* workspace,
                            size_t workspace_bytes, void* stream_, void* status_event) {
    VPG_CHECK_ARG(rewards && mask && advantages && returns && status, "null pointer argument");
    VPG_CHECK_ARG(B > 0 && L > 0, "empty batch");
    if (!workspace || workspace_bytes < vpg_reverse_scan_workspace_bytes(B, L))
        return set_error(VPG_ERR_WORKSPACE, "reinforce++ workspace too small");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    double* sums = static_cast<double*>(workspace);
    VPG_CUDA(cudaMemsetAsync(sums, 0, 4 * sizeof(double), stream));
    const int64_t n = B * L;
    const bool vec = rows_vec_ok(n, mask, mask_dtype, {rewards, returns, advantages});
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        const MT* mk = static_cast<const MT*>(mask);
        reverse_scan_kernel<MT, 0><<<static_cast<unsigned>(B), scan_threads(B, L), 0, stream>>>(
            rewards, mk, nullptr, L, gamma, advantages, returns, sums, (L % kScanSeg == 0) && vec, status);
        count_launch();
        VPG_CUDA(record_whiten_status(status_event, stream));
        if (vec)
            whiten_mask_apply_kernel<4, MT><<<flat_grid(n, 4), kThreads, 0, stream>>>(returns, mk, n, sums, advantages);
        else
            whiten_mask_apply_kernel<1, MT><<<flat_grid(n, 1), kThreads, 0, stream>>>(returns, mk, n, sums, advantages);
    });
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

int vpg_remax(const float* rewards, const float* reward_baselines, const void* mask, int mask_dtype, int64_t B,
              int64_t L, float* advantages, float* returns, void* stream_) {
    VPG_CHECK_ARG(rewards && reward_baselines && mask && advantages && returns, "null pointer argument");
    VPG_CHECK_ARG(B > 0 && L > 0, "empty batch");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int vec = (L % kScanSeg == 0) && rows_vec_ok(L, mask, mask_dtype, {rewards, advantages, returns});
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        reverse_scan_kernel<MT, 1><<<static_cast<unsigned>(B), scan_threads(B, L), 0, stream>>>(
            rewards, static_cast<const MT*>(mask), reward_baselines, L, 1.f, advantages, returns, nullptr, vec, nullptr);
    });
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

}  // extern "C"

// =====================================================================================================
// Driver-side token ops on the batch tensors
// =====================================================================================================
namespace vpg {

__device__ __forceinline__ float kl_value(int kind, float lp, float ref) {
    if (kind == VPG_KL_K1) return lp - ref;
    if (kind == VPG_KL_ABS) return fabsf(lp - ref);
    if (kind == VPG_KL_MSE) {
        const float d = lp - ref;
        return 0.5f * (d * d);
    }
    const float k = clampf(ref - lp, -20.f, 20.f);  // low_var_kl
    return clampf(expf(k) - k - 1.f, -10.f, 10.f);
}

// apply_kl_penalty (verl/trainer/ppo/ray_trainer.py:153-193): kld = kl_penalty(old, ref) * mask;
// token_level_rewards = scores - beta * kld; current_kl = mean_b( masked_mean(kld, mask, axis=-1) ).
// ws (doubles): [0] u32 counter | [1, 1+B) row sums of where(m, kld, 0)*m | [1+B, 1+2B) row mask sums
template <int VEC, typename MT>
__global__ void kl_reward_kernel(const float* __restrict__ scores, const float* __restrict__ old_lp,
                                 const float* __restrict__ ref_lp, const MT* __restrict__ mask, long long B, long long L,
                                 int gx, int kind, float beta, float* __restrict__ rewards, float* __restrict__ out,
                                 double* ws) {
    __shared__ double red[32];
    const long long b = blockIdx.x / gx;
    const int cx = blockIdx.x % gx;
    const long long base = b * L;
    float a = 0.f, c = 0.f;
    const long long stride = static_cast<long long>(gx) * blockDim.x * VEC;
    for (long long t = (static_cast<long long>(cx) * blockDim.x + threadIdx.x) * VEC; t < L; t += stride) {
        const long long i = base + t;
        const Pack<VEC> m = ldm<VEC, MT>(mask, i);
        const Pack<VEC> sc = ldv<VEC>(scores, i);
        const Pack<VEC> lo = ldv<VEC>(old_lp, i);
        const Pack<VEC> lr = ldv<VEC>(ref_lp, i);
        Pack<VEC> o;
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const float kld = kl_value(kind, lo.v[j], lr.v[j]) * m.v[j];
            o.v[j] = sc.v[j] - beta * kld;
            a += (m.v[j] != 0.f ? kld : 0.f) * m.v[j];
            c += m.v[j];
        }
        stv<VEC>(rewards, i, o);
    }
    const double ta = block_sum<double>(static_cast<double>(a), red);
    const double tc = block_sum<double>(static_cast<double>(c), red);
    if (threadIdx.x == 0) {
        atomicAdd(&ws[1 + b], ta);
        atomicAdd(&ws[1 + B + b], tc);
    }
    if (last_block_done(reinterpret_cast<unsigned int*>(ws), gridDim.x)) {
        double s = 0.0;
        for (long long i = threadIdx.x; i < B; i += blockDim.x)
            s += static_cast<double>(static_cast<float>(__ldcg(&ws[1 + i]) / (__ldcg(&ws[1 + B + i]) + 1e-8)));
        s = block_sum<double>(s, red);
        if (threadIdx.x == 0) out[0] = static_cast<float>(s / static_cast<double>(B));
    }
}

// ---- compute_data_metrics (verl/trainer/ppo/metric_utils.py:80-180) ------------------------------------------
// One sweep over the [B,L] tensors.  acc (doubles, zeroed): per masked tensor q in {adv, returns, values}:
//   [5q+0] sum  [5q+1] sum of squares  then max / min keys (u32, order preserving) in keys[2q], keys[2q+1];
//   [15] sum (ret - val)  [16] sum (ret - val)^2  [17] count of masked tokens
// rows (doubles): [0,B) score row sums | [B,2B) reward row sums | [2B,3B) response lengths | [3B,4B) prompt lengths
constexpr int kDmAcc = 18;

__device__ __forceinline__ unsigned int fkey(float f) {
    const unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float fkey_inv(unsigned int k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

template <int VEC, typename MT>
__global__ void data_metrics_kernel(const float* __restrict__ scores, const float* __restrict__ rewards,
                                    const float* __restrict__ adv, const float* __restrict__ rets,
                                    const float* __restrict__ vals, const MT* __restrict__ rmask, long long L, int gx,
                                    long long B, double* __restrict__ acc, unsigned int* __restrict__ keys,
                                    double* __restrict__ rows) {
    const long long b = blockIdx.x / gx;
    const int cx = blockIdx.x % gx;
    const long long base = b * L;
    float s_sc = 0.f, s_rw = 0.f, cnt = 0.f;
    float s1[3] = {0.f, 0.f, 0.f}, s2[3] = {0.f, 0.f, 0.f};
    float mx[3] = {-INFINITY, -INFINITY, -INFINITY}, mn[3] = {INFINITY, INFINITY, INFINITY};
    float d1 = 0.f, d2 = 0.f;
    // every operand is loaded unconditionally (128-bit when VEC == 4): five independent streams in flight per thread,
    // the response_mask.bool() test (metric_utils.py:109) only gates the arithmetic
    const long long stride = static_cast<long long>(gx) * blockDim.x * VEC;
    for (long long t = (static_cast<long long>(cx) * blockDim.x + threadIdx.x) * VEC; t < L; t += stride) {
        const long long i = base + t;
        const Pack<VEC> sc = ldv<VEC>(scores, i), rw = ldv<VEC>(rewards, i), a = ldv<VEC>(adv, i), r = ldv<VEC>(rets, i);
        const Pack<VEC> m = ldm<VEC, MT>(rmask, i);
        Pack<VEC> vl;
        if (vals) {
            vl = ldv<VEC>(vals, i);
        } else {
#pragma unroll
            for (int j = 0; j < VEC; ++j) vl.v[j] = 0.f;
        }
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            s_sc += sc.v[j];
            s_rw += rw.v[j];
            if (m.v[j] != 0.f) {
                cnt += 1.f;
                const float q[3] = {a.v[j], r.v[j], vl.v[j]};
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    s1[k] += q[k];
                    s2[k] += q[k] * q[k];
                    mx[k] = fmaxf(mx[k], q[k]);
                    mn[k] = fminf(mn[k], q[k]);
                }
                const float d = q[1] - q[2];
                d1 += d;
                d2 += d * d;
            }
        }
    }
    // epilogue: the eleven block sums share one barrier pair, the six max / min keys a second one; thread 0 issues every
    // atomic (the per-block cost is what bounds how finely the sweep can be cut: profiles/r2_ab_token_grid_after_epilogue_fusion.txt)
    __shared__ double red_n[11 * 32];
    __shared__ unsigned int red_k[6 * 32];
    double v[11] = {s1[0], s2[0], s1[1], s2[1], s1[2], s2[2], d1, d2, cnt, s_sc, s_rw};
    block_sum_n<11>(v, red_n);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    unsigned int key[6];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        unsigned int kmx = fkey(mx[k]), kmn = fkey(mn[k]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            kmx = max(kmx, __shfl_xor_sync(0xffffffffu, kmx, o));
            kmn = min(kmn, __shfl_xor_sync(0xffffffffu, kmn, o));
        }
        key[2 * k] = kmx, key[2 * k + 1] = kmn;
    }
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 6; ++k) red_k[k * 32 + w] = key[k];
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        const int k = threadIdx.x;
        unsigned int r = red_k[k * 32];
        for (int ww = 1; ww < nw; ++ww) r = (k & 1) ? min(r, red_k[k * 32 + ww]) : max(r, red_k[k * 32 + ww]);
        if (k & 1) atomicMin(&keys[k], r);
        else atomicMax(&keys[k], r);
    }
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) atomicAdd(&acc[5 * k], v[2 * k]), atomicAdd(&acc[5 * k + 1], v[2 * k + 1]);
        atomicAdd(&acc[15], v[6]), atomicAdd(&acc[16], v[7]), atomicAdd(&acc[17], v[8]);
        atomicAdd(&rows[b], v[9]);
        atomicAdd(&rows[B + b], v[10]);
    }
}

// prompt / response lengths from attention_mask [B,S] (metric_utils.py:57-75): one block per row
template <typename MT>
__global__ void seq_lengths_kernel(const MT* __restrict__ attn, long long S, long long L, long long B,
                                   double* __restrict__ rows, int vec) {
    __shared__ double red[32];
    const long long base = static_cast<long long>(blockIdx.x) * S;
    const long long P = S - L;
    float p = 0.f, r = 0.f;
    if (vec) {  // 4 mask elements per load, two loads in flight (rows are 16-byte aligned, S % 4 == 0)
        const long long step = static_cast<long long>(blockDim.x) * 4;
        long long t = static_cast<long long>(threadIdx.x) * 4;
        auto take = [&](const float4& m, long long t0) {
            const float q[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (t0 + j < P) p += q[j];
                else r += q[j];
            }
        };
        for (; t + step < S; t += 2 * step) {
            const float4 m0 = load_mask4<MT>(attn, base + t);
            const float4 m1 = load_mask4<MT>(attn, base + t + step);
            take(m0, t);
            take(m1, t + step);
        }
        if (t < S) take(load_mask4<MT>(attn, base + t), t);
    } else {
        for (long long t = threadIdx.x; t < S; t += blockDim.x) {
            const float m = mask_to_float(attn[base + t]);
            if (t < P) p += m; else r += m;
        }
    }
    const double tp = block_sum<double>(static_cast<double>(p), red);
    const double tr = block_sum<double>(static_cast<double>(r), red);
    if (threadIdx.x == 0) {
        rows[2 * B + blockIdx.x] = tr;
        rows[3 * B + blockIdx.x] = tp;
    }
}

// out[0..25]: score mean/max/min, rewards mean/max/min, adv mean/max/min, returns mean/max/min, values mean/max/min,
// vf_explained_var, response_length mean/max/min/clip_ratio, prompt_length mean/max/min/clip_ratio, n_valid, (pad)
__global__ void data_metrics_finalize_kernel(const double* __restrict__ acc, const unsigned int* __restrict__ keys,
                                             const double* __restrict__ rows, long long B, long long L,
                                             long long P, int use_critic, float* __restrict__ out) {
    __shared__ double red[32];
    __shared__ float fmx[4], fmn[4];
    // sequence-level: mean / max / min over B of the four row quantities (fp32 values like the reference's tensors)
    for (int q = 0; q < 4; ++q) {
        double s = 0.0;
        float mx = -INFINITY, mn = INFINITY;
        double clip = 0.0;
        for (long long i = threadIdx.x; i < B; i += blockDim.x) {
            const float v = static_cast<float>(rows[q * B + i]);
            s += v;
            mx = fmaxf(mx, v);
            mn = fminf(mn, v);
            if (q == 2 && v == static_cast<float>(L)) clip += 1.0;
            if (q == 3 && v == static_cast<float>(P)) clip += 1.0;
        }
        s = block_sum<double>(s, red);
        clip = block_sum<double>(clip, red);
        unsigned int kmx = fkey(mx), kmn = fkey(mn);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            kmx = max(kmx, __shfl_xor_sync(0xffffffffu, kmx, o));
            kmn = min(kmn, __shfl_xor_sync(0xffffffffu, kmn, o));
        }
        __shared__ unsigned int wmx[32], wmn[32];
        if ((threadIdx.x & 31) == 0) wmx[threadIdx.x >> 5] = kmx, wmn[threadIdx.x >> 5] = kmn;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned int a = wmx[0], bmin = wmn[0];
            for (int w = 1; w < (blockDim.x >> 5); ++w) a = max(a, wmx[w]), bmin = min(bmin, wmn[w]);
            fmx[q] = fkey_inv(a), fmn[q] = fkey_inv(bmin);
            const int o = (q < 2) ? 3 * q : (q == 2 ? 16 : 20);
            out[o] = static_cast<float>(s / static_cast<double>(B));
            out[o + 1] = fmx[q];
            out[o + 2] = fmn[q];
            if (q >= 2) out[o + 3] = static_cast<float>(clip / static_cast<double>(B));
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double n = acc[17];
        for (int k = 0; k < 3; ++k) {
            out[6 + 3 * k] = static_cast<float>(acc[5 * k] / n);
            out[7 + 3 * k] = fkey_inv(keys[2 * k]);
            out[8 + 3 * k] = fkey_inv(keys[2 * k + 1]);
        }
        // torch.var (unbiased) of returns and of returns - values over the valid tokens
        const double var_ret = (acc[6] - acc[5] * acc[5] / n) / (n - 1.0);
        const double var_diff = (acc[16] - acc[15] * acc[15] / n) / (n - 1.0);
        out[15] = use_critic ? static_cast<float>(1.0 - var_diff / (var_ret + 1e-5)) : 0.f;
        out[24] = static_cast<float>(n);
        out[25] = 0.f;
    }
}

}  // namespace vpg

extern "C" {

size_t vpg_kl_reward_workspace_bytes(int64_t B) { return static_cast<size_t>(1 + 2 * (B > 0 ? B : 0)) * sizeof(double); }

int vpg_kl_reward(const float* token_level_scores, const float* old_log_probs, const float* ref_log_prob,
                  const void* mask, int mask_dtype, int kind, float beta, int64_t B, int64_t L,
                  float* token_level_rewards, float* current_kl, void* workspace, size_t workspace_bytes,
                  void* stream_) {
    VPG_CHECK_ARG(token_level_scores && old_log_probs && ref_log_prob && mask && token_level_rewards && current_kl,
                  "null pointer argument");
    VPG_CHECK_ARG(B > 0 && L > 0, "empty batch");
    if (kind < 0 || kind > 3) return set_error(VPG_ERR_UNSUPPORTED, "kl_penalty kind %d is not implemented", kind);
    const size_t need = vpg_kl_reward_workspace_bytes(B);
    if (!workspace || workspace_bytes < need) return set_error(VPG_ERR_WORKSPACE, "kl_reward workspace too small");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    VPG_CUDA(cudaMemsetAsync(workspace, 0, need, stream));
    const bool vec = rows_vec_ok(L, mask, mask_dtype, {token_level_scores, old_log_probs, ref_log_prob, token_level_rewards});
    const int gx = blocks_per_row(B, L, vec ? 4 : 1, 32);
    const unsigned grid = static_cast<unsigned>(B * gx);
    double* ws = static_cast<double*>(workspace);
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        const MT* mk = static_cast<const MT*>(mask);
        if (vec)
            kl_reward_kernel<4, MT><<<grid, kThreads, 0, stream>>>(token_level_scores, old_log_probs, ref_log_prob, mk, B,
                                                                   L, gx, kind, beta, token_level_rewards, current_kl, ws);
        else
            kl_reward_kernel<1, MT><<<grid, kThreads, 0, stream>>>(token_level_scores, old_log_probs, ref_log_prob, mk, B,
                                                                   L, gx, kind, beta, token_level_rewards, current_kl, ws);
    });
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

size_t vpg_data_metrics_workspace_bytes(int64_t B) {
    return static_cast<size_t>(kDmAcc + 4 * (B > 0 ? B : 0)) * sizeof(double) + 8 * sizeof(unsigned int);
}

int vpg_data_metrics(const float* token_level_scores, const float* token_level_rewards, const float* advantages,
                     const float* returns, const float* values, const void* response_mask, int rmask_dtype,
                     const void* attention_mask, int amask_dtype, int64_t B, int64_t L, int64_t S, float* out26,
                     void* workspace, size_t workspace_bytes, void* stream_) {
    VPG_CHECK_ARG(token_level_scores && token_level_rewards && advantages && returns && response_mask &&
                      attention_mask && out26,
                  "null pointer argument");
    VPG_CHECK_ARG(B > 0 && L > 0 && S >= L, "invalid shape B=%lld L=%lld S=%lld", (long long)B, (long long)L, (long long)S);
    const size_t need = vpg_data_metrics_workspace_bytes(B);
    if (!workspace || workspace_bytes < need) return set_error(VPG_ERR_WORKSPACE, "data_metrics workspace too small");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    double* acc = static_cast<double*>(workspace);
    double* rows = acc + kDmAcc;
    unsigned int* keys = reinterpret_cast<unsigned int*>(rows + 4 * B);
    VPG_CUDA(cudaMemsetAsync(workspace, 0, need, stream));
    // min keys start at the largest key (memset 0xff on the three min slots is done by initialising all six keys):
    static const unsigned int init_keys[8] = {0u, 0xffffffffu, 0u, 0xffffffffu, 0u, 0xffffffffu, 0u, 0u};
    VPG_CUDA(cudaMemcpyAsync(keys, init_keys, sizeof(init_keys), cudaMemcpyHostToDevice, stream));
    const bool vec = rows_vec_ok(L, response_mask, rmask_dtype,
                                 {token_level_scores, token_level_rewards, advantages, returns, values});
    const int gx = blocks_per_row(B, L, vec ? 4 : 1, 16);
    VPG_DISPATCH_MASK(rmask_dtype, MT, {
        if (vec)
            data_metrics_kernel<4, MT><<<static_cast<unsigned>(B * gx), kThreads, 0, stream>>>(
                token_level_scores, token_level_rewards, advantages, returns, values,
                static_cast<const MT*>(response_mask), L, gx, B, acc, keys, rows);
        else
            data_metrics_kernel<1, MT><<<static_cast<unsigned>(B * gx), kThreads, 0, stream>>>(
                token_level_scores, token_level_rewards, advantages, returns, values,
                static_cast<const MT*>(response_mask), L, gx, B, acc, keys, rows);
    });
    count_launch();
    VPG_DISPATCH_MASK(amask_dtype, MT, {
        seq_lengths_kernel<MT><<<static_cast<unsigned>(B), kThreads, 0, stream>>>(
            static_cast<const MT*>(attention_mask), S, L, B, rows, (S % 4 == 0) && mask_aligned(attention_mask, amask_dtype));
    });
    count_launch();
    data_metrics_finalize_kernel<<<1, 256, 0, stream>>>(acc, keys, rows, B, L, S - L, values != nullptr, out26);
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

}  // extern "C"
