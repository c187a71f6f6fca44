// Memory-bound token kernels: masked reductions, agg_loss, fused PPO actor loss (+ closed-form gradient),
// KL penalty, GRPO group advantage, GAE (+ masked whiten).
// Reference semantics: verl/trainer/ppo/core_algos.py and verl/utils/torch_functional.py (lines cited per kernel).
//
// Design: every [B,L] pass is one coalesced sweep with 128-bit loads (4 tokens / thread / iteration; scalar path for
// unaligned or L % 4 != 0 inputs), per-thread fp32 partials -> warp shuffle -> one double atomic per block and
// quantity.  Scalars that depend on a global sum (token-mean normaliser) are produced by a first, mask-only pass;
// the "last block done" pattern folds the tiny finalisation into the same launch.  Compiled with --fmad=false so
// that elementwise results follow the reference's separately rounded fp32 operations.
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <mutex>
#include <type_traits>
#include <vector>

#include "token_common.cuh"

namespace vpg {

// =====================================================================================================
// masked_sum (verl/utils/torch_functional.py:163-168): out[0] = sum(where(m!=0, v, 0) * m), out[1] = sum(m)
// =====================================================================================================
template <int VEC, typename MT>
__global__ void masked_sum_kernel(const float* __restrict__ values, const MT* __restrict__ mask, long long n,
                                  double* __restrict__ out) {
    __shared__ double red[32];
    float s = 0.f, c = 0.f;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * VEC;
    for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * VEC; i < n; i += stride) {
        const Pack<VEC> v = ldv<VEC>(values, i);
        const Pack<VEC> m = ldm<VEC, MT>(mask, i);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            s += (m.v[j] != 0.f ? v.v[j] : 0.f) * m.v[j];
            c += m.v[j];
        }
    }
    const double bs = block_sum<double>(static_cast<double>(s), red);
    const double bc = block_sum<double>(static_cast<double>(c), red);
    if (threadIdx.x == 0) {
        atomicAdd(&out[0], bs);
        atomicAdd(&out[1], bc);
    }
}

// per-row (axis=-1): out_sum[b] = sum_t where(m,v,0)*m, out_cnt[b] = sum_t m.  One block per row, no atomics.
template <int VEC, typename MT>
__global__ void masked_row_sum_kernel(const float* __restrict__ values, const MT* __restrict__ mask, long long L,
                                      float* __restrict__ row_sum, float* __restrict__ row_cnt) {
    __shared__ double red[32];
    const long long base = static_cast<long long>(blockIdx.x) * L;
    float s = 0.f, c = 0.f;
    for (long long t = static_cast<long long>(threadIdx.x) * VEC; t < L; t += static_cast<long long>(blockDim.x) * VEC) {
        const Pack<VEC> m = ldm<VEC, MT>(mask, base + t);
        Pack<VEC> v;
        if (values) v = ldv<VEC>(values, base + t);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            if (values) s += (m.v[j] != 0.f ? v.v[j] : 0.f) * m.v[j];
            c += m.v[j];
        }
    }
    const double bs = block_sum<double>(static_cast<double>(s), red);
    const double bc = block_sum<double>(static_cast<double>(c), red);
    if (threadIdx.x == 0) {
        if (row_sum) row_sum[blockIdx.x] = static_cast<float>(bs);
        if (row_cnt) row_cnt[blockIdx.x] = static_cast<float>(bc);
    }
}

// =====================================================================================================
// kl_penalty (verl/trainer/ppo/core_algos.py:1034-1069), elementwise; optional d out / d logprob
// =====================================================================================================

__device__ __forceinline__ void kl_term(int kind, float lp, float ref, float& out, float& grad) {
    if (kind == VPG_KL_K1) {
        out = lp - ref;
        grad = 1.f;
    } else if (kind == VPG_KL_ABS) {
        const float d = lp - ref;
        out = fabsf(d);
        grad = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
    } else if (kind == VPG_KL_MSE) {
        const float d = lp - ref;
        out = 0.5f * (d * d);
        grad = d;
    } else {  // low_var_kl
        const float kr = ref - lp;
        const float k = clampf(kr, -20.f, 20.f);
        const float r = expf(k);
        const float raw = r - k - 1.f;
        out = clampf(raw, -10.f, 10.f);
        const bool in_k = (kr >= -20.f) && (kr <= 20.f);
        const bool in_c = (raw >= -10.f) && (raw <= 10.f);
        grad = (in_k && in_c) ? -(r - 1.f) : 0.f;
    }
}

template <int VEC>
__global__ void kl_penalty_kernel(const float* __restrict__ lp, const float* __restrict__ ref, int kind, long long n,
                                  float* __restrict__ out, float* __restrict__ grad) {
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * VEC;
    for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * VEC; i < n; i += stride) {
        const Pack<VEC> a = ldv<VEC>(lp, i);
        const Pack<VEC> b = ldv<VEC>(ref, i);
        Pack<VEC> o, g;
#pragma unroll
        for (int j = 0; j < VEC; ++j) kl_term(kind, a.v[j], b.v[j], o.v[j], g.v[j]);
        stv<VEC>(out, i, o);
        if (grad) stv<VEC>(grad, i, g);
    }
}

// =====================================================================================================
// Fused PPO actor loss.  compute_policy_loss (core_algos.py:722-794) + agg_loss (:686-719) + kl_penalty (:1034-1069)
// composed as in verl/workers/actor/dp_actor.py:422-469, forward and closed-form gradient.
//
// workspace (doubles): [0] two uint32 counters | [1] N = sum(mask) | [2..8) global sums G[6] | [8] a mask value does not
//                      fit the one-byte copy (coop kernel: is not 0 / 1) | [9] a mask value is neither 0 nor 1 |
//                      [10, 10+B) row mask sums | [10+B, 10+4B) row sums of {pg, entropy, kl} * mask
//   G: 0 pg 1 -log_ratio 2 clip 3 clip_lower 4 entropy 5 kl   (all weighted where(m,v,0)*m as masked_mean does)
// =====================================================================================================
struct PpoParams {
    const float* old_lp;
    const float* lp;
    const float* adv;
    const float* entropy;  // may be null
    const float* ref_lp;   // may be null
    long long B, L;
    int gx;
    float ratio_lo, ratio_hi, clip_c;
    int pg_kind;  // VPG_PG_CLIP: compute_policy_loss; VPG_PG_GPG: -log_prob * advantages (core_algos.py:797-815)
    int agg_mode;
    float entropy_coeff, kl_coef;
    int kl_kind;
    float loss_scale;
    float* metrics;
    float* d_lp;
    float* d_ent;
    double* ws;
};

constexpr int kWsHeader = 10;

// pass 1: row mask counts; last block sums them into ws[1].  For 4- and 8-byte integer masks the pass also leaves a ONE-byte
// copy of every mask value in `cmask` (and raises ws[8] when a value does not fit a byte), so that the loss pass — which
// cannot start before the token-mean normaliser is known — re-reads 1 byte per token instead of 8: 38 instead of 44 bytes
// per token for the whole loss (algorithmic floor 36).
template <int VEC, typename MT>
__global__ void mask_count_kernel(const MT* __restrict__ mask, long long B, long long L, int gx, double* ws,
                                  unsigned char* __restrict__ cmask) {
    __shared__ double red[32];
    const long long b = blockIdx.x / gx;
    const int cx = blockIdx.x % gx;
    const long long base = b * L;
    float c = 0.f;
    bool wide = false, odd = false;
    const long long stride = static_cast<long long>(gx) * blockDim.x * VEC;
    auto take = [&](const Pack<VEC>& m, long long t) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            c += m.v[j];
            odd |= !(m.v[j] == 0.f || m.v[j] == 1.f);
        }
        if (cmask) {
            unsigned char q[VEC];
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                wide |= !(m.v[j] >= 0.f && m.v[j] <= 255.f);
                q[j] = static_cast<unsigned char>(m.v[j]);
            }
            if constexpr (VEC == 4) {
                *reinterpret_cast<uchar4*>(cmask + base + t) = make_uchar4(q[0], q[1], q[2], q[3]);
            } else {
                cmask[base + t] = q[0];
            }
        }
    };
    // four packs per iteration, all loads issued before any is consumed (the sweep is latency-bound otherwise: ncu long-
    // scoreboard stall 34 cycles per issue with two)
    long long t = (static_cast<long long>(cx) * blockDim.x + threadIdx.x) * VEC;
    for (; t + 3 * stride < L; t += 4 * stride) {
        const Pack<VEC> m0 = ldm<VEC, MT>(mask, base + t);
        const Pack<VEC> m1 = ldm<VEC, MT>(mask, base + t + stride);
        const Pack<VEC> m2 = ldm<VEC, MT>(mask, base + t + 2 * stride);
        const Pack<VEC> m3 = ldm<VEC, MT>(mask, base + t + 3 * stride);
        take(m0, t);
        take(m1, t + stride);
        take(m2, t + 2 * stride);
        take(m3, t + 3 * stride);
    }
    if (t + stride < L) {  // at most three packs are left
        const Pack<VEC> m0 = ldm<VEC, MT>(mask, base + t);
        const Pack<VEC> m1 = ldm<VEC, MT>(mask, base + t + stride);
        take(m0, t);
        take(m1, t + stride);
        t += 2 * stride;
    }
    if (t < L) take(ldm<VEC, MT>(mask, base + t), t);
    if (cmask && __syncthreads_or(wide ? 1 : 0) && threadIdx.x == 0) ws[8] = 1.0;  // benign race: every writer stores 1
    if (__syncthreads_or(odd ? 1 : 0) && threadIdx.x == 0) ws[9] = 1.0;            // a value that is neither 0 nor 1
    const double bc = block_sum<double>(static_cast<double>(c), red);
    if (threadIdx.x == 0) atomicAdd(&ws[kWsHeader + b], bc);
    unsigned int* counters = reinterpret_cast<unsigned int*>(ws);
    if (last_block_done(&counters[0], gridDim.x)) {
        double s = 0.0;
        for (long long i = threadIdx.x; i < B; i += blockDim.x) s += __ldcg(&ws[kWsHeader + i]);
        s = block_sum<double>(s, red);
        if (threadIdx.x == 0) ws[1] = s;
    }
}

// gradient weight of one token for the chosen aggregation mode (d agg / d loss_mat[b,t])
__device__ __forceinline__ float agg_coef(int mode, double n_total, double row_cnt, long long B, long long L) {
    if (mode == VPG_AGG_TOKEN_MEAN) return static_cast<float>(1.0 / (n_total + 1e-8));
    if (mode == VPG_AGG_SEQ_MEAN_TOKEN_SUM) return static_cast<float>(1.0 / static_cast<double>(B));
    if (mode == VPG_AGG_SEQ_MEAN_TOKEN_MEAN) return static_cast<float>(1.0 / (static_cast<double>(B) * row_cnt));
    return static_cast<float>(1.0 / static_cast<double>(L));
}

__device__ __forceinline__ double agg_finalize(int mode, const double* row_sum, const double* row_cnt, double g_where,
                                               double n_total, long long B, long long L, double* red) {
    // returns the aggregated loss in thread 0
    if (mode == VPG_AGG_TOKEN_MEAN) return g_where / (n_total + 1e-8);
    double s = 0.0;
    for (long long i = threadIdx.x; i < B; i += blockDim.x) {
        const double rs = __ldcg(&row_sum[i]);
        s += (mode == VPG_AGG_SEQ_MEAN_TOKEN_MEAN) ? rs / __ldcg(&row_cnt[i]) : rs;
    }
    s = block_sum<double>(s, red);
    if (mode == VPG_AGG_SEQ_MEAN_TOKEN_SUM_NORM) return s / static_cast<double>(L);
    return s / static_cast<double>(B);
}

// ---- the per-token body of the fused loss, shared by the row-wise kernel and the cooperative single-pass kernel ----
// kSpec: 0 = every option read at run time; 1 / 2 = the two configurations the BASELINE recipes run (clip loss,
// token-mean, 1: no KL term, 2: low_var_kl) with the options folded at compile time — the sweep is ISSUE-bound (ncu of the
// generic body: ~150 instructions per token, sm throughput 66-72 % at 55 % of DRAM peak), so dead accumulators, option
// tests and the select-on-isnan clamps are what the specialisation removes (~105 per token).
struct PpoOpts {
    bool where_w, gpg, has_ent, has_kl, want_grad;
    int kl_kind;
    float ent_g;
};
template <int kSpec>
__device__ __forceinline__ PpoOpts ppo_opts(const PpoParams& p, bool want_grad) {
    PpoOpts o;
    o.where_w = kSpec ? true : (p.agg_mode == VPG_AGG_TOKEN_MEAN);
    o.gpg = kSpec ? false : (p.pg_kind == VPG_PG_GPG);
    o.has_ent = p.entropy != nullptr;
    o.has_kl = kSpec == 1 ? false : kSpec == 2 ? true : (p.ref_lp != nullptr && p.kl_kind >= 0);
    o.kl_kind = kSpec == 2 ? static_cast<int>(VPG_KL_LOW_VAR) : p.kl_kind;
    o.want_grad = want_grad;
    o.ent_g = -p.entropy_coeff;
    return o;
}
struct PpoAcc {
    float a_pg = 0.f, a_lr = 0.f, a_clip = 0.f, a_low = 0.f, a_ent = 0.f, a_kl = 0.f;  // where-weighted (masked_mean)
    float r_pg = 0.f, r_ent = 0.f, r_kl = 0.f;                                           // plain v*m (row sums)
    float a_cnt = 0.f;                                                                   // sum of the mask (kCount)
};

// kBin: the caller knows that every mask value is 0 or 1 (flag from the mask pass), so `where(m, x, 0) * m` is the select
// alone and the gradient weight is `coef` or 0 — bit-identical results, one multiply less per accumulator.
template <int VEC, bool kCount, bool kBin>
__device__ __forceinline__ void ppo_pack(const PpoParams& p, const PpoOpts& o, float coef, const Pack<VEC>& m,
                                         const Pack<VEC>& old, const Pack<VEC>& lp, const Pack<VEC>& adv,
                                         const Pack<VEC>& ent, const Pack<VEC>& ref, Pack<VEC>& dlp, Pack<VEC>& dent,
                                         PpoAcc& a) {
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        const float mj = m.v[j];
        const bool on = mj != 0.f;
        if (kCount) a.a_cnt += mj;
        const float A = adv.v[j];
        const float lr_raw = lp.v[j] - old.v[j];
        const float lr = clampf(lr_raw, -20.f, 20.f);
        const float ratio = expf(lr);
        const float l1 = -A * ratio;
        const float rc = clampf(ratio, p.ratio_lo, p.ratio_hi);
        const float l2 = -A * rc;
        const float c1 = max_nan(l1, l2);
        const float l3 = -A * p.clip_c;
        const float c2 = min_nan(l3, c1);
        const float pg = o.gpg ? (-lp.v[j] * A) : ((A < 0.f) ? c2 : c1);
        const bool clipped = !o.gpg && (l2 > l1);
        const bool low = !o.gpg && (c1 > l3) && (A < 0.f);
        auto wsum = [&](float& accu, bool sel, float x) {  // accu += where(m, x, 0) * m
            if (kBin) accu += sel ? x : 0.f;
            else accu += (sel ? x : 0.f) * mj;
        };
        wsum(a.a_pg, on, pg);
        wsum(a.a_lr, on && !o.gpg, -lr);
        wsum(a.a_clip, on && clipped, 1.f);
        wsum(a.a_low, on && low, 1.f);
        if (!o.where_w) a.r_pg += pg * mj;
        float g = 0.f;
        if (o.want_grad) {
            // gradient of pg wrt log_prob (ties of maximum / min split 1/2, as torch autograd does); d l1 = l1
            const bool in_lr = (lr_raw >= -20.f) && (lr_raw <= 20.f);
            const float dl1 = in_lr ? l1 : 0.f;
            const bool in_clip = (ratio >= p.ratio_lo) && (ratio <= p.ratio_hi);
            const float dl2 = in_clip ? dl1 : 0.f;
            const float g_c1 = (l1 > l2) ? dl1 : ((l1 < l2) ? dl2 : 0.5f * (dl1 + dl2));
            const float g_c2 = (c1 < l3) ? g_c1 : ((c1 == l3) ? 0.5f * g_c1 : 0.f);
            g = o.gpg ? -A : ((A < 0.f) ? g_c2 : g_c1);
        }
        if (o.has_kl) {
            float kv, kg;
            kl_term(o.kl_kind, lp.v[j], ref.v[j], kv, kg);
            wsum(a.a_kl, on, kv);
            if (!o.where_w) a.r_kl += kv * mj;
            g += p.kl_coef * kg;
        }
        const float cw = (kBin && o.where_w) ? (on ? coef : 0.f) : coef * (o.where_w ? (on ? mj : 0.f) : mj);
        dlp.v[j] = cw * g;
        if (o.has_ent) {
            wsum(a.a_ent, on, ent.v[j]);
            if (!o.where_w) a.r_ent += ent.v[j] * mj;
            dent.v[j] = cw * o.ent_g;
        }
    }
}

// block sums + atomics of one block's accumulators: six global sums (+ the mask count) behind one barrier pair, the three
// row sums of the seq-mean modes behind a second one
template <bool kCount>
__device__ __forceinline__ void ppo_block_reduce(const PpoParams& p, const PpoOpts& o, const PpoAcc& a, long long b,
                                                 double* red_n /* 7 * 32 */) {
    double tot[7] = {a.a_pg, a.a_lr, a.a_clip, a.a_low, a.a_ent, a.a_kl, a.a_cnt};
    block_sum_n<7>(tot, red_n);
    double rtot[3] = {a.r_pg, a.r_ent, a.r_kl};
    if (!o.where_w) block_sum_n<3>(rtot, red_n);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < 6; ++k) atomicAdd(&p.ws[2 + k], tot[k]);
        if (kCount) atomicAdd(&p.ws[1], tot[6]);
        if (!o.where_w) {  // row sums feed the seq-mean modes only
            double* rows = p.ws + kWsHeader + p.B;
            atomicAdd(&rows[b], rtot[0]);
            atomicAdd(&rows[p.B + b], rtot[1]);
            atomicAdd(&rows[2 * p.B + b], rtot[2]);
        }
    }
}

// run by every thread of the LAST block: aggregated losses -> metrics[8]
__device__ __forceinline__ void ppo_finalize(const PpoParams& p, const PpoOpts& o, double n_total, double* red) {
    double* ws = p.ws;
    const double* rcnt = ws + kWsHeader;
    const double* rows = ws + kWsHeader + p.B;
    const double G0 = __ldcg(&ws[2]), G1 = __ldcg(&ws[3]), G2 = __ldcg(&ws[4]), G3 = __ldcg(&ws[5]);
    const double G4 = __ldcg(&ws[6]), G5 = __ldcg(&ws[7]);
    const double pg_loss = agg_finalize(p.agg_mode, rows, rcnt, G0, n_total, p.B, p.L, red);
    const double ent_loss = o.has_ent ? agg_finalize(p.agg_mode, rows + p.B, rcnt, G4, n_total, p.B, p.L, red) : 0.0;
    const double kl_loss = o.has_kl ? agg_finalize(p.agg_mode, rows + 2 * p.B, rcnt, G5, n_total, p.B, p.L, red) : 0.0;
    if (threadIdx.x == 0) {
        // fp32 composition as in dp_actor.py:445-469
        const float pgf = static_cast<float>(pg_loss), entf = static_cast<float>(ent_loss);
        const float klf = static_cast<float>(kl_loss);
        float policy = pgf;
        if (o.has_ent && p.entropy_coeff != 0.f) policy = policy - entf * p.entropy_coeff;
        if (o.has_kl) policy = policy + klf * p.kl_coef;
        const double den = n_total + 1e-8;
        p.metrics[0] = policy * p.loss_scale;
        p.metrics[1] = pgf;
        p.metrics[2] = static_cast<float>(G2 / den);
        p.metrics[3] = static_cast<float>(G1 / den);
        p.metrics[4] = static_cast<float>(G3 / den);
        p.metrics[5] = entf;
        p.metrics[6] = klf;
        p.metrics[7] = static_cast<float>(n_total);
    }
}

// kCount: forward only with token-mean aggregation — nothing needs the normaliser before the sweep, so the mask count is
// taken in the same pass (no mask_count_kernel launch, the mask read once).
template <int VEC, typename MT, bool kCount, int kSpec>
__global__ void __launch_bounds__(kThreads, 4) ppo_loss_kernel(const PpoParams p, const MT* __restrict__ mask,
                                                               const unsigned char* __restrict__ cmask) {
    __shared__ double red[32];
    __shared__ double red_n[7 * 32];
    const bool compact = !kCount && cmask != nullptr && p.ws[8] == 0.0;  // the one-byte copy holds every mask value exactly
    const long long b = blockIdx.x / p.gx;
    const int cx = blockIdx.x % p.gx;
    const long long base = b * p.L;
    double* ws = p.ws;
    double n_total = kCount ? 0.0 : ws[1];
    const double row_cnt = kCount ? 1.0 : ws[kWsHeader + b];
    const float coef = kCount ? 0.f : agg_coef(p.agg_mode, n_total, row_cnt, p.B, p.L) * p.loss_scale;
    // kCount: no gradient outputs, the gradient arithmetic is dead code
    const PpoOpts o = ppo_opts<kSpec>(p, !kCount && p.d_lp != nullptr);
    PpoAcc acc;
    const long long stride = static_cast<long long>(p.gx) * blockDim.x * VEC;
    auto sweep = [&](auto bin) {
        constexpr bool kBin = decltype(bin)::value;
        for (long long t = (static_cast<long long>(cx) * blockDim.x + threadIdx.x) * VEC; t < p.L; t += stride) {
            const long long i = base + t;
            const Pack<VEC> m = compact ? ldm<VEC, unsigned char>(cmask, i) : ldm<VEC, MT>(mask, i);
            const Pack<VEC> old = ldv<VEC>(p.old_lp, i);
            const Pack<VEC> lp = ldv<VEC>(p.lp, i);
            const Pack<VEC> adv = ldv<VEC>(p.adv, i);
            Pack<VEC> ent, ref, dlp, dent;
            if (o.has_ent) ent = ldv<VEC>(p.entropy, i);
            if (o.has_kl) ref = ldv<VEC>(p.ref_lp, i);
            ppo_pack<VEC, kCount, kBin>(p, o, coef, m, old, lp, adv, ent, ref, dlp, dent, acc);
            if (o.want_grad) stv<VEC>(p.d_lp, i, dlp);
            if (!kCount && p.d_ent && o.has_ent) stv<VEC>(p.d_ent, i, dent);
        }
    };
    // the mask pass raises ws[9] when a mask value is neither 0 nor 1; the folded bodies then take the select-only sums
    if (!kCount && kSpec != 0 && ws[9] == 0.0) sweep(std::true_type{});
    else sweep(std::false_type{});
    ppo_block_reduce<kCount>(p, o, acc, b, red_n);
    unsigned int* counters = reinterpret_cast<unsigned int*>(ws);
    if (last_block_done(&counters[1], gridDim.x)) {
        if (kCount) n_total = __ldcg(&ws[1]);
        ppo_finalize(p, o, n_total, red);
    }
}

// Token-mean loss WITH gradients in ONE launch and one read of the mask (cooperative launch, every block resident).  The
// gradient of a token-mean loss needs sum(mask) before its first store, which is why the row-wise path runs a mask-only
// pass first and re-reads a one-byte copy (38 bytes per token against the 36 of the inputs and outputs).  Here every
// block owns a fixed set of 1024-token units of the flat [B*L] tokens: phase 1 streams their mask (four 128-bit loads in
// flight), keeps one bit per token in shared memory and adds its count to ws[1]; after the grid barrier phase 2 streams
// the float tensors once and takes the mask from shared memory.  Masks with values other than 0 / 1 (the reference allows
// any weights) raise ws[8] in phase 1 and phase 2 re-reads them from global memory instead.
constexpr int kCoopSliceUnit = kThreads * 4;         // tokens one block covers per step
constexpr long long kCoopMaxSlice = 128 * 1024;      // tokens per block: 32 KB of shared memory (4 mask bits per byte)

template <typename MT, int kSpec>
__global__ void __launch_bounds__(kThreads, 4) ppo_loss_coop_kernel(const PpoParams p, const MT* __restrict__ mask,
                                                                    long long n, long long iters) {
    extern __shared__ unsigned char nib[];  // nib[it * kThreads + tid] = mask bits of the 4 tokens this thread owns in step it
    __shared__ double red[32];
    __shared__ double red_n[7 * 32];
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    double* ws = p.ws;
    // step `it` of block b covers tokens [(it * gridDim.x + b) * kCoopSliceUnit, +kCoopSliceUnit): at any moment the grid
    // streams ONE contiguous window of every tensor (contiguous per-block slices ran 25 % slower: 592 lock-step streams
    // a fixed 224 KB apart)
    const long long step = static_cast<long long>(gridDim.x) * kCoopSliceUnit;
    const long long i0 = static_cast<long long>(blockIdx.x) * kCoopSliceUnit + static_cast<long long>(threadIdx.x) * 4;
    // ---- phase 1: mask -> bits + count
    {
        float c = 0.f;
        bool odd = false;
        auto take = [&](const Pack<4>& m, long long it) {
            unsigned int bits = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                c += m.v[j];
                odd |= !(m.v[j] == 0.f || m.v[j] == 1.f);
                bits |= (m.v[j] != 0.f ? 1u : 0u) << j;
            }
            nib[it * kThreads + threadIdx.x] = static_cast<unsigned char>(bits);
        };
        long long it = 0;
        for (; it + 3 < iters && i0 + (it + 3) * step < n; it += 4) {
            const long long i = i0 + it * step;
            const Pack<4> m0 = ldm<4, MT>(mask, i), m1 = ldm<4, MT>(mask, i + step);
            const Pack<4> m2 = ldm<4, MT>(mask, i + 2 * step), m3 = ldm<4, MT>(mask, i + 3 * step);
            take(m0, it), take(m1, it + 1), take(m2, it + 2), take(m3, it + 3);
        }
        for (; it < iters && i0 + it * step < n; ++it) take(ldm<4, MT>(mask, i0 + it * step), it);
        const int any_odd = __syncthreads_or(odd ? 1 : 0);
        const double bc = block_sum<double>(static_cast<double>(c), red);
        if (threadIdx.x == 0) {
            atomicAdd(&ws[1], bc);
            if (any_odd) ws[8] = 1.0;  // benign race: every writer stores 1
        }
    }
    __threadfence();
    grid.sync();
    // ---- phase 2: the loss sweep
    const double n_total = __ldcg(&ws[1]);
    const bool binary = __ldcg(&ws[8]) == 0.0;
    const float coef = agg_coef(VPG_AGG_TOKEN_MEAN, n_total, 1.0, p.B, p.L) * p.loss_scale;
    const PpoOpts o = ppo_opts<kSpec>(p, p.d_lp != nullptr);
    PpoAcc acc;
    auto sweep = [&](auto bin) {
        constexpr bool kBin = decltype(bin)::value;
        for (long long it = 0; it < iters; ++it) {
            const long long i = i0 + it * step;
            if (i >= n) break;
            Pack<4> m;
            if (kBin) {
                const unsigned int bits = nib[it * kThreads + threadIdx.x];
#pragma unroll
                for (int j = 0; j < 4; ++j) m.v[j] = ((bits >> j) & 1u) ? 1.f : 0.f;
            } else {
                m = ldm<4, MT>(mask, i);
            }
            const Pack<4> old = ldv<4>(p.old_lp, i);
            const Pack<4> lp = ldv<4>(p.lp, i);
            const Pack<4> adv = ldv<4>(p.adv, i);
            Pack<4> ent, ref, dlp, dent;
            if (o.has_ent) ent = ldv<4>(p.entropy, i);
            if (o.has_kl) ref = ldv<4>(p.ref_lp, i);
            ppo_pack<4, false, kBin>(p, o, coef, m, old, lp, adv, ent, ref, dlp, dent, acc);
            if (o.want_grad) stv<4>(p.d_lp, i, dlp);
            if (p.d_ent && o.has_ent) stv<4>(p.d_ent, i, dent);
        }
    };
    if (binary) sweep(std::true_type{});
    else sweep(std::false_type{});
    ppo_block_reduce<false>(p, o, acc, 0, red_n);
    unsigned int* counters = reinterpret_cast<unsigned int*>(ws);
    if (last_block_done(&counters[1], gridDim.x)) ppo_finalize(p, o, n_total, red);
}

// agg_loss for an arbitrary loss matrix: forward sums + gradient d loss / d loss_mat
// kCount: the token-mean FORWARD alone needs no normaliser before its sweep, so the mask count is taken in the same pass
// (no mask_count_kernel launch, the mask read once: 12 B/token).  Otherwise the counts come from mask_count_kernel and,
// for 4- / 8-byte integer masks, the mask is re-read from its one-byte copy (`cmask`, see mask_count_kernel).
template <int VEC, typename MT, bool kCount>
__global__ void agg_loss_kernel(const float* __restrict__ loss_mat, const MT* __restrict__ mask, long long B,
                                long long L, int gx, int mode, float* __restrict__ out, float* __restrict__ d_mat,
                                double* ws, const unsigned char* __restrict__ cmask) {
    __shared__ double red[32];
    const bool compact = !kCount && cmask != nullptr && ws[8] == 0.0;
    const long long b = blockIdx.x / gx;
    const int cx = blockIdx.x % gx;
    const long long base = b * L;
    const double n_known = kCount ? 0.0 : ws[1];
    const double row_cnt = kCount ? 1.0 : ws[kWsHeader + b];
    const float coef = kCount ? 0.f : agg_coef(mode, n_known, row_cnt, B, L);
    const bool where_w = (mode == VPG_AGG_TOKEN_MEAN);
    float a = 0.f, r = 0.f, c = 0.f;
    const long long stride = static_cast<long long>(gx) * blockDim.x * VEC;
    for (long long t = (static_cast<long long>(cx) * blockDim.x + threadIdx.x) * VEC; t < L; t += stride) {
        const long long i = base + t;
        const Pack<VEC> m = compact ? ldm<VEC, unsigned char>(cmask, i) : ldm<VEC, MT>(mask, i);
        const Pack<VEC> v = ldv<VEC>(loss_mat, i);
        Pack<VEC> d;
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const bool on = m.v[j] != 0.f;
            a += (on ? v.v[j] : 0.f) * m.v[j];
            if (kCount) {
                c += m.v[j];
            } else {
                r += v.v[j] * m.v[j];
                d.v[j] = coef * (where_w ? (on ? m.v[j] : 0.f) : m.v[j]);
            }
        }
        if (!kCount && d_mat) stv<VEC>(d_mat, i, d);
    }
    const double ta = block_sum<double>(static_cast<double>(a), red);
    const double tr = block_sum<double>(static_cast<double>(kCount ? c : r), red);
    if (threadIdx.x == 0) {
        atomicAdd(&ws[2], ta);
        if (kCount) atomicAdd(&ws[1], tr);
        else atomicAdd(&ws[kWsHeader + B + b], tr);
    }
    unsigned int* counters = reinterpret_cast<unsigned int*>(ws);
    if (last_block_done(&counters[1], gridDim.x)) {
        const double n_total = kCount ? __ldcg(&ws[1]) : n_known;
        const double v = agg_finalize(mode, ws + kWsHeader + B, ws + kWsHeader, __ldcg(&ws[2]), n_total, B, L, red);
        if (threadIdx.x == 0) out[0] = static_cast<float>(v);
    }
}

// =====================================================================================================
// compute_value_loss (core_algos.py:992-1031): clipped value loss, forward + closed-form gradient wrt vpreds.
//   vpc = max(min(vp, v + c), v - c); l1 = (vp - ret)^2; l2 = (vpc - ret)^2
//   vf_loss = 0.5 * agg(max(l1, l2)); vf_clipfrac = masked_mean(l2 > l1)
// workspace as ppo_loss (mask_count_kernel runs first): G[0] = where-weighted loss sum, G[1] = clip count,
// rows[b] = plain row sums.  out[0] = vf_loss, out[1] = vf_clipfrac.
// =====================================================================================================
template <int VEC, typename MT>
__global__ void value_loss_kernel(const float* __restrict__ vpreds, const float* __restrict__ returns,
                                  const float* __restrict__ values, const MT* __restrict__ mask, long long B,
                                  long long L, int gx, float cliprange, int mode, float* __restrict__ out,
                                  float* __restrict__ d_vpreds, double* ws, const unsigned char* __restrict__ cmask) {
    __shared__ double red[32];
    const bool compact = cmask != nullptr && ws[8] == 0.0;  // the one-byte copy holds every mask value exactly
    const long long b = blockIdx.x / gx;
    const int cx = blockIdx.x % gx;
    const long long base = b * L;
    const double n_total = ws[1];
    const double row_cnt = ws[kWsHeader + b];
    const float coef = agg_coef(mode, n_total, row_cnt, B, L);
    const bool where_w = (mode == VPG_AGG_TOKEN_MEAN);
    float a = 0.f, r = 0.f, cf = 0.f;
    const long long stride = static_cast<long long>(gx) * blockDim.x * VEC;
    for (long long t = (static_cast<long long>(cx) * blockDim.x + threadIdx.x) * VEC; t < L; t += stride) {
        const long long i = base + t;
        const Pack<VEC> m = compact ? ldm<VEC, unsigned char>(cmask, i) : ldm<VEC, MT>(mask, i);
        const Pack<VEC> vp = ldv<VEC>(vpreds, i);
        const Pack<VEC> rt = ldv<VEC>(returns, i);
        const Pack<VEC> vo = ldv<VEC>(values, i);
        Pack<VEC> d;
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const bool on = m.v[j] != 0.f;
            const float hi = vo.v[j] + cliprange, lo = vo.v[j] - cliprange;
            const float y = min_nan(vp.v[j], hi);
            const float vpc = max_nan(y, lo);
            const float e1 = vp.v[j] - rt.v[j], e2 = vpc - rt.v[j];
            const float l1 = e1 * e1, l2 = e2 * e2;
            const float lmax = max_nan(l1, l2);
            a += (on ? lmax : 0.f) * m.v[j];
            if (!where_w) r += lmax * m.v[j];
            cf += ((on && l2 > l1) ? 1.f : 0.f) * m.v[j];
            // d vpc / d vp through torch.min / torch.max (ties split 1/2 like torch autograd)
            const float dmin = (vp.v[j] < hi) ? 1.f : ((vp.v[j] == hi) ? 0.5f : 0.f);
            const float dmax = (y > lo) ? 1.f : ((y == lo) ? 0.5f : 0.f);
            const float g1 = 2.f * e1, g2 = 2.f * e2 * (dmin * dmax);
            const float g = (l1 > l2) ? g1 : ((l1 < l2) ? g2 : 0.5f * (g1 + g2));
            d.v[j] = 0.5f * (coef * (where_w ? (on ? m.v[j] : 0.f) : m.v[j])) * g;
        }
        if (d_vpreds) stv<VEC>(d_vpreds, i, d);
    }
    const double ta = block_sum<double>(static_cast<double>(a), red);
    const double tr = where_w ? 0.0 : block_sum<double>(static_cast<double>(r), red);  // row sums: seq-mean modes only
    const double tc = block_sum<double>(static_cast<double>(cf), red);
    if (threadIdx.x == 0) {
        atomicAdd(&ws[2], ta);
        atomicAdd(&ws[3], tc);
        if (!where_w) atomicAdd(&ws[kWsHeader + B + b], tr);
    }
    unsigned int* counters = reinterpret_cast<unsigned int*>(ws);
    if (last_block_done(&counters[1], gridDim.x)) {
        const double v = agg_finalize(mode, ws + kWsHeader + B, ws + kWsHeader, __ldcg(&ws[2]), n_total, B, L, red);
        if (threadIdx.x == 0) {
            out[0] = 0.5f * static_cast<float>(v);
            out[1] = static_cast<float>(__ldcg(&ws[3]) / (n_total + 1e-8));
        }
    }
}

// =====================================================================================================
// GRPO outcome advantage (core_algos.py:245-308)
// workspace (doubles): [0,B) scores | [B, B+3G) per group {count, sum, sum of squared deviations}
// =====================================================================================================
template <int VEC, typename MT>
__global__ void row_score_kernel(const float* __restrict__ rewards, long long L, int gx, double* __restrict__ scores) {
    __shared__ double red[32];
    const long long b = blockIdx.x / gx;
    const int cx = blockIdx.x % gx;
    const long long base = b * L;
    float s = 0.f;
    const long long stride = static_cast<long long>(gx) * blockDim.x * VEC;
    for (long long t = (static_cast<long long>(cx) * blockDim.x + threadIdx.x) * VEC; t < L; t += stride) {
        const Pack<VEC> v = ldv<VEC>(rewards, base + t);
#pragma unroll
        for (int j = 0; j < VEC; ++j) s += v.v[j];
    }
    const double bs = block_sum<double>(static_cast<double>(s), red);
    if (threadIdx.x == 0) atomicAdd(&scores[b], bs);
}

// Group statistics and the per-row scalar advantage (written as float into adv_row[B]).
// Warp-shuffle SEGMENTED reduction over uid groups: a warp takes 32 consecutive rows, `__match_any_sync` on the group id
// gives every lane the set of lanes of its group (contiguous runs of the n rollouts of a prompt in practice, any
// arrangement in general), the per-group partial is summed over that lane set in lane order with shuffles, and the lowest
// lane of each set issues ONE double atomic per (warp, group) instead of one per row.  Two sweeps like the reference's
// torch.mean / torch.std (sum, then squared deviations from the group mean); the phases are separated by a block barrier
// (one block: up to kGroupRowsPerBlock rows) or a cooperative grid barrier (larger batches: one block per kGroupRowsPerBlock
// rows, at most one per SM, launched cooperatively).
constexpr int kGroupThreads = 1024;
constexpr int kGroupRowsPerBlock = 4096;

template <bool kMultiBlock>
__global__ void __launch_bounds__(kGroupThreads)
grpo_group_kernel(const double* __restrict__ scores, const int* __restrict__ gid, long long B, long long G, float eps,
                  int norm_by_std, double* __restrict__ gstat, float* __restrict__ adv_row) {
    namespace cg = cooperative_groups;
    const int lane = threadIdx.x & 31;
    const long long first = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) - lane;  // row of lane 0
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    auto phase_barrier = [&]() {
        if (kMultiBlock) {
            __threadfence();
            cg::this_grid().sync();
        } else {
            __syncthreads();
        }
    };
    // pass 1: count and sum (scores are rounded to fp32 first: the reference holds them in an fp32 tensor)
    for (long long r0 = first; r0 < B; r0 += stride) {
        const long long i = r0 + lane;
        const bool ok = i < B;
        const int g = ok ? gid[i] : -1;
        const double s = ok ? static_cast<double>(static_cast<float>(scores[i])) : 0.0;
        const unsigned peers = __match_any_sync(0xffffffffu, g);
        const double sum = segment_sum(s, peers);
        if (ok && lane == __ffs(peers) - 1) {
            atomicAdd(&gstat[3 * g + 0], static_cast<double>(__popc(peers)));
            atomicAdd(&gstat[3 * g + 1], sum);
        }
    }
    phase_barrier();
    // pass 2: squared deviations from the group mean
    for (long long r0 = first; r0 < B; r0 += stride) {
        const long long i = r0 + lane;
        const bool ok = i < B;
        const int g = ok ? gid[i] : -1;
        double d = 0.0;
        if (ok) {
            const double n = __ldcg(&gstat[3 * g + 0]);
            d = static_cast<double>(static_cast<float>(scores[i])) - __ldcg(&gstat[3 * g + 1]) / n;
        }
        const unsigned peers = __match_any_sync(0xffffffffu, g);
        const double sq = segment_sum(d * d, peers);
        if (ok && lane == __ffs(peers) - 1) atomicAdd(&gstat[3 * g + 2], sq);
    }
    phase_barrier();
    for (long long r0 = first; r0 < B; r0 += stride) {
        const long long i = r0 + lane;
        if (i >= B) continue;
        const int g = gid[i];
        const double n = __ldcg(&gstat[3 * g + 0]);
        float mean, stdv;
        if (n <= 1.0) {  // singleton group: mean 0, std 1 (core_algos.py:293-295)
            mean = 0.f;
            stdv = 1.f;
        } else {
            mean = static_cast<float>(__ldcg(&gstat[3 * g + 1]) / n);
            stdv = static_cast<float>(sqrt(__ldcg(&gstat[3 * g + 2]) / (n - 1.0)));  // unbiased, torch.std
        }
        const float sc = static_cast<float>(scores[i]);
        adv_row[i] = norm_by_std ? (sc - mean) / (stdv + eps) : (sc - mean);
    }
}

template <int VEC, typename MT>
__global__ void broadcast_mask_kernel(const float* __restrict__ row_val, const MT* __restrict__ mask, long long L,
                                      int gx, float* __restrict__ out) {
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

// =====================================================================================================
// GAE (core_algos.py:193-241).  One block per row; each thread owns a contiguous segment of kSeg tokens held in
// registers.  The reverse recurrence over (nextvalues, lastgaelam) is an affine map per token:
//     nv' = (1-m) nv + m v          lg' = m*gamma*nv + (m*gamma*lam + (1-m)) lg + m (r - v)
// Phase 1 composes each segment's map, phase 2 does a block-wide right-to-left exclusive scan of the segment maps,
// phase 3 replays the segment with its true incoming state using the reference's exact operation order.
// Rows longer than blockDim*kSeg are processed in super-chunks from the right, carrying the state.
// Also accumulates sum(adv*m), sum(adv^2*m), sum(m) (doubles) for the whitening pass.
// =====================================================================================================
constexpr int kSeg = 8;

struct Affine {  // (nv, lg) -> (an*nv + cn, p*nv + q*lg + r)
    float an, cn, p, q, r;
};
__device__ __forceinline__ Affine affine_identity() { return Affine{1.f, 0.f, 0.f, 1.f, 0.f}; }
// h = g after f  (apply f first, then g)
__device__ __forceinline__ Affine affine_compose(const Affine& g, const Affine& f) {
    Affine h;
    h.an = g.an * f.an;
    h.cn = g.an * f.cn + g.cn;
    h.p = g.p * f.an + g.q * f.p;
    h.q = g.q * f.q;
    h.r = g.p * f.cn + g.q * f.r + g.r;
    return h;
}
__device__ __forceinline__ Affine affine_shfl_down(const Affine& a, int delta) {
    Affine o;
    o.an = __shfl_down_sync(0xffffffffu, a.an, delta);
    o.cn = __shfl_down_sync(0xffffffffu, a.cn, delta);
    o.p = __shfl_down_sync(0xffffffffu, a.p, delta);
    o.q = __shfl_down_sync(0xffffffffu, a.q, delta);
    o.r = __shfl_down_sync(0xffffffffu, a.r, delta);
    return o;
}

template <typename MT>
__global__ void __launch_bounds__(512) gae_kernel(const float* __restrict__ rewards, const float* __restrict__ values,
                           const MT* __restrict__ mask, long long L, float gamma, float lam,
                           float* __restrict__ adv_out, float* __restrict__ ret_out, double* __restrict__ sums,
                           int vec, int* status) {
    __shared__ Affine warp_maps[32];
    __shared__ float carry[2];
    __shared__ double red[32];
    const long long base = static_cast<long long>(blockIdx.x) * L;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const float gl = gamma * lam;
    const long long span = static_cast<long long>(blockDim.x) * kSeg;
    if (threadIdx.x == 0) carry[0] = 0.f, carry[1] = 0.f;
    float s1 = 0.f, s2 = 0.f, sc = 0.f;
    // super-chunks from the right end; chunk k covers [start, start+span)
    const long long n_chunks = (L + span - 1) / span;
    for (long long ck = n_chunks - 1; ck >= 0; --ck) {
        __syncthreads();
        const long long t0 = ck * span + static_cast<long long>(threadIdx.x) * kSeg;
        float r[kSeg], v[kSeg], m[kSeg];
        const bool full = vec && (t0 + kSeg <= L);  // 128-bit path: every thread owns 32 contiguous, aligned bytes
        if (full) {
#pragma unroll
            for (int q = 0; q < kSeg / 4; ++q) {
                const float4 a = load_f4(rewards, base + t0 + 4 * q);
                const float4 b = load_f4(values, base + t0 + 4 * q);
                const float4 c = load_mask4<MT>(mask, base + t0 + 4 * q);
                r[4 * q] = a.x, r[4 * q + 1] = a.y, r[4 * q + 2] = a.z, r[4 * q + 3] = a.w;
                v[4 * q] = b.x, v[4 * q + 1] = b.y, v[4 * q + 2] = b.z, v[4 * q + 3] = b.w;
                m[4 * q] = c.x, m[4 * q + 1] = c.y, m[4 * q + 2] = c.z, m[4 * q + 3] = c.w;
            }
        } else {
#pragma unroll
            for (int j = 0; j < kSeg; ++j) {
                const long long t = t0 + j;
                const bool ok = t < L;
                r[j] = ok ? rewards[base + t] : 0.f;
                v[j] = ok ? values[base + t] : 0.f;
                m[j] = ok ? mask_to_float(mask[base + t]) : 0.f;  // out-of-range tokens are identity maps
            }
        }
        // phase 1: compose this segment right-to-left
        Affine seg = affine_identity();
#pragma unroll
        for (int j = kSeg - 1; j >= 0; --j) {
            Affine st;
            st.an = 1.f - m[j];
            st.cn = m[j] * v[j];
            st.p = m[j] * gamma;
            st.q = m[j] * gl + (1.f - m[j]);
            st.r = m[j] * (r[j] - v[j]);
            seg = affine_compose(st, seg);
        }
        // phase 2: exclusive suffix scan over threads (thread i needs the composition of segments i+1.. end)
        Affine inc = seg;  // inclusive suffix: seg_i after (segments to the right)
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const Affine o = affine_shfl_down(inc, d);
            if (lane + d < 32) inc = affine_compose(inc, o);
        }
        if (lane == 0) warp_maps[w] = inc;  // whole-warp map
        __syncthreads();
        // map of everything to the right of this warp
        Affine right = affine_identity();
        for (int ww = nw - 1; ww > w; --ww) right = affine_compose(warp_maps[ww], right);
        // exclusive within the warp: composition of lanes > lane
        Affine excl = affine_shfl_down(inc, 1);
        if (lane == 31) excl = affine_identity();
        excl = affine_compose(excl, right);
        // incoming state for this segment
        const float nv_in = carry[0], lg_in = carry[1];
        float nv = excl.an * nv_in + excl.cn;
        float lg = excl.p * nv_in + excl.q * lg_in + excl.r;
        // phase 3: replay with the reference's operation order
        float a[kSeg];
#pragma unroll
        for (int j = kSeg - 1; j >= 0; --j) {
            const float delta = r[j] + gamma * nv - v[j];
            const float lg_ = delta + gl * lg;
            nv = v[j] * m[j] + (1.f - m[j]) * nv;
            lg = lg_ * m[j] + (1.f - m[j]) * lg;
            a[j] = lg;
        }
        if (full) {
#pragma unroll
            for (int q = 0; q < kSeg / 4; ++q) {
                store_f4(adv_out, base + t0 + 4 * q, make_float4(a[4 * q], a[4 * q + 1], a[4 * q + 2], a[4 * q + 3]));
                store_f4(ret_out, base + t0 + 4 * q,
                         make_float4(a[4 * q] + v[4 * q], a[4 * q + 1] + v[4 * q + 1], a[4 * q + 2] + v[4 * q + 2],
                                     a[4 * q + 3] + v[4 * q + 3]));
            }
        }
#pragma unroll
        for (int j = 0; j < kSeg; ++j) {
            const long long t = t0 + j;
            if (t < L) {
                if (!full) {
                    adv_out[base + t] = a[j];
                    ret_out[base + t] = a[j] + v[j];
                }
                s1 += (m[j] != 0.f ? a[j] : 0.f) * m[j];
                s2 += (m[j] != 0.f ? a[j] * a[j] : 0.f) * m[j];
                sc += m[j];
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {  // state after the leftmost token of this chunk
            carry[0] = nv;
            carry[1] = lg;
        }
    }
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

// sums of v*m, v^2*m, m for masked_whiten
template <int VEC, typename MT>
__global__ void whiten_stats_kernel(const float* __restrict__ values, const MT* __restrict__ mask, long long n,
                                    double* __restrict__ sums, int* status) {
    __shared__ double red[32];
    float s1 = 0.f, s2 = 0.f, sc = 0.f;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * VEC;
    for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * VEC; i < n; i += stride) {
        const Pack<VEC> v = ldv<VEC>(values, i);
        const Pack<VEC> m = ldm<VEC, MT>(mask, i);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const bool on = m.v[j] != 0.f;
            s1 += (on ? v.v[j] : 0.f) * m.v[j];
            s2 += (on ? v.v[j] * v.v[j] : 0.f) * m.v[j];
            sc += m.v[j];
        }
    }
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

// masked_whiten (torch_functional.py:206-223): mean = S1/(N+1e-8); var = E[(x-mean)^2] * N/(N-1);
// out = (x - mean) * rsqrt(var + 1e-8) (+ mean if !shift_mean).  In place allowed (values == out).
template <int VEC>
__global__ void whiten_apply_kernel(const float* values, long long n, const double* __restrict__ sums,
                                    int shift_mean, float* out) {
    const double N = sums[2];
    if (N == 0.0 || N == 1.0) return;  // the reference raises ValueError (torch_functional.py:195-200; status word
                                       // from whiten_publish_status); outputs are left untouched
    const double den = N + 1e-8;
    const double mean_d = sums[0] / den;
    // sum((x-mean)^2 m) = S2 - 2 mean S1 + mean^2 N
    double var_d = (sums[1] - 2.0 * mean_d * sums[0] + mean_d * mean_d * N) / den;
    var_d = var_d * (N / (N - 1.0));
    const float mean = static_cast<float>(mean_d);
    const float rstd = rsqrtf(static_cast<float>(var_d) + 1e-8f);
    // kIlp independent 128-bit loads per thread are issued before the first store: the pass has one input stream only,
    // so a single load in flight per thread leaves the HBM queues half empty (values may alias out: no __restrict__,
    // hence the explicit staging instead of relying on the compiler to hoist loads over stores)
    constexpr int kIlp = 4;
    const long long tile = static_cast<long long>(blockDim.x) * VEC;  // elements one block covers per sub-step
    const long long stride = static_cast<long long>(gridDim.x) * tile * kIlp;
    for (long long i0 = static_cast<long long>(blockIdx.x) * tile * kIlp + static_cast<long long>(threadIdx.x) * VEC;
         i0 < n; i0 += stride) {
        Pack<VEC> v[kIlp];
#pragma unroll
        for (int u = 0; u < kIlp; ++u) {
            const long long i = i0 + u * tile;
            if (i < n) v[u] = ldv<VEC>(values, i);
        }
#pragma unroll
        for (int u = 0; u < kIlp; ++u) {
            const long long i = i0 + u * tile;
            if (i >= n) break;
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                float wv = (v[u].v[j] - mean) * rstd;
                if (!shift_mean) wv += mean;
                v[u].v[j] = wv;
            }
            stv<VEC>(out, i, v[u]);
        }
    }
}

}  // namespace vpg


using namespace vpg;

extern "C" {

int vpg_masked_sum(const float* values, const void* mask, int mask_dtype, int64_t n, double* out, void* stream_) {
    VPG_CHECK_ARG(values && mask && out && n >= 0, "invalid masked_sum arguments");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    VPG_CUDA(cudaMemsetAsync(out, 0, 2 * sizeof(double), stream));
    if (n == 0) return VPG_OK;
    const bool vec = rows_vec_ok(n, mask, mask_dtype, {values});
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        const MT* mk = static_cast<const MT*>(mask);
        if (vec)
            masked_sum_kernel<4, MT><<<flat_grid(n, 4), kThreads, 0, stream>>>(values, mk, n, out);
        else
            masked_sum_kernel<1, MT><<<flat_grid(n, 1), kThreads, 0, stream>>>(values, mk, n, out);
    });
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

int vpg_masked_row_sum(const float* values, const void* mask, int mask_dtype, int64_t B, int64_t L, float* out_sum,
                       float* out_cnt, void* stream_) {
    VPG_CHECK_ARG(mask && B >= 0 && L >= 0 && (out_sum || out_cnt), "invalid masked_row_sum arguments");
    VPG_CHECK_ARG(!out_sum || values, "values required when out_sum is requested");
    if (B == 0) return VPG_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const bool vec = rows_vec_ok(L, mask, mask_dtype, {values});
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        const MT* mk = static_cast<const MT*>(mask);
        if (vec)
            masked_row_sum_kernel<4, MT><<<static_cast<unsigned>(B), kThreads, 0, stream>>>(values, mk, L, out_sum, out_cnt);
        else
            masked_row_sum_kernel<1, MT><<<static_cast<unsigned>(B), kThreads, 0, stream>>>(values, mk, L, out_sum, out_cnt);
    });
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

int vpg_kl_penalty(const float* logprob, const float* ref_logprob, int kind, int64_t n, float* out, float* grad_out,
                   void* stream_) {
    VPG_CHECK_ARG(logprob && ref_logprob && out && n >= 0, "invalid kl_penalty arguments");
    if (kind < 0 || kind > 3) return set_error(VPG_ERR_UNSUPPORTED, "kl_penalty kind %d is not implemented", kind);
    if (n == 0) return VPG_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (rows_vec_ok(n, nullptr, 0, {logprob, ref_logprob, out, grad_out}))
        kl_penalty_kernel<4><<<flat_grid(n, 4), kThreads, 0, stream>>>(logprob, ref_logprob, kind, n, out, grad_out);
    else
        kl_penalty_kernel<1><<<flat_grid(n, 1), kThreads, 0, stream>>>(logprob, ref_logprob, kind, n, out, grad_out);
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

// VPG_TOK_SPEC=0 keeps every loss sweep on its generic (run-time option) body: the A/B switch of the folded bodies
static bool spec_enabled() {
    static const int on = env_int("VPG_TOK_SPEC", 1);
    return on != 0;
}

// VPG_TOK_COOP=0 keeps the token-mean losses with gradients on the two-launch path (mask count, then the sweep)
static bool coop_enabled() {
    static const int on = [] {
        if (env_int("VPG_TOK_COOP", 1) == 0) return 0;
        int dev = 0, ok = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return 0;
        if (cudaDeviceGetAttribute(&ok, cudaDevAttrCooperativeLaunch, dev) != cudaSuccess) return 0;
        return ok;
    }();
    return on != 0;
}

// Resident blocks per SM of a cooperative kernel, asked once per (kernel, KiB of dynamic shared memory) and device: the
// call sits on the host path of every loss evaluation, which is host-bound at the BASELINE micro-batch sizes.
static int coop_occupancy(const void* fn, size_t smem) {
    struct Key {
        const void* fn;
        size_t kib;
        int dev;
        int occ;
    };
    static std::mutex mu;
    static std::vector<Key> cache;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    const size_t kib = (smem + 1023) / 1024;
    std::lock_guard<std::mutex> lk(mu);
    for (const Key& k : cache)
        if (k.fn == fn && k.kib == kib && k.dev == dev) return k.occ;
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, kThreads, kib * 1024) != cudaSuccess) occ = 0;
    cache.push_back(Key{fn, kib, dev, occ});
    return occ;
}

// Grid of the cooperative single-pass kernels: as many blocks as are resident at once (at most 4 per SM), every block
// `iters` units of kCoopSliceUnit tokens (at most kCoopMaxSlice tokens: one mask bit per token has to fit in shared
// memory).  Returns false when the problem does not fit (the caller takes the two-launch path).
static bool coop_plan(const void* fn, int64_t n, long long* iters, unsigned* grid, size_t* smem) {
    // Worth it while the launch it saves matters: at the BASELINE micro-batch sizes (20 k - 262 k tokens) the single launch
    // is 25-35 % faster (C5: 35 -> 24 us).  On long sweeps it LOSES: every warp leaves the grid barrier at the same moment
    // and the sweep is issue-bound, so load and arithmetic phases alternate in lock-step instead of overlapping (ncu at
    // 2^25 tokens: 314 us at 47 % of DRAM peak against 53 + 161 us for the two launches).  Hence at most
    // VPG_TOK_COOP_MAX_ITERS (default 4) units per block, i.e. up to ~2.4 M tokens on 148 SMs.
    static const int max_iters = env_int("VPG_TOK_COOP_MAX_ITERS", 4);
    int64_t blocks = static_cast<int64_t>(sm_count()) * 4;
    const int64_t units = ceil_div(n, kCoopSliceUnit);
    for (int pass = 0; pass < 2; ++pass) {
        const int64_t g = units < blocks ? units : blocks;
        const int64_t it = ceil_div(units, g);
        if (it > max_iters || it * kCoopSliceUnit > kCoopMaxSlice) return false;
        const size_t sm = static_cast<size_t>(it) * kThreads;
        const int occ = coop_occupancy(fn, sm);
        if (occ < 1) return false;
        const int64_t resident = static_cast<int64_t>(occ) * sm_count();
        const int64_t need = ceil_div(units, it);
        if (need <= resident) {
            *iters = it, *grid = static_cast<unsigned>(need), *smem = sm;
            return true;
        }
        blocks = resident;  // fewer blocks fit than assumed: plan for that many
    }
    return false;
}

static size_t ppo_ws_header_bytes(int64_t B) {
    return round_up(static_cast<int64_t>(kWsHeader + 4 * (B > 0 ? B : 0)) * sizeof(double), 16);
}
size_t vpg_ppo_loss_workspace_bytes(int64_t B, int64_t L) {
    // reduction slots + the one-byte mask copy of vpg_ppo_loss ([B*L], see mask_count_kernel)
    return ppo_ws_header_bytes(B) + static_cast<size_t>(round_up((B > 0 ? B : 0) * (L > 0 ? L : 0), 16));
}

int vpg_ppo_loss(const float* old_log_prob, const float* log_prob, const float* advantages, const void* mask,
                 int mask_dtype, const float* entropy, const float* ref_log_prob, int64_t B, int64_t L,
                 float ratio_lo, float ratio_hi, float clip_c, int pg_kind, int agg_mode, float entropy_coeff,
                 float kl_coef, int kl_kind, float loss_scale, float* metrics, float* d_log_prob, float* d_entropy,
                 void* workspace, size_t workspace_bytes, void* stream_) {
    VPG_CHECK_ARG(old_log_prob && log_prob && advantages && mask && metrics, "null pointer argument");
    VPG_CHECK_ARG(pg_kind == VPG_PG_CLIP || pg_kind == VPG_PG_GPG, "unknown policy-loss kind %d", pg_kind);
    VPG_CHECK_ARG(B > 0 && L > 0, "empty batch");
    VPG_CHECK_ARG(agg_mode >= 0 && agg_mode <= 3, "invalid loss_agg_mode %d", agg_mode);
    VPG_CHECK_ARG(clip_c > 1.0f, "clip_ratio_c must be > 1.0 (got %f)", clip_c);  // core_algos.py:761
    if (ref_log_prob && (kl_kind < 0 || kl_kind > 3))
        return set_error(VPG_ERR_UNSUPPORTED, "kl_penalty kind %d is not implemented", kl_kind);
    const size_t need = vpg_ppo_loss_workspace_bytes(B, L);
    if (!workspace || workspace_bytes < need)
        return set_error(VPG_ERR_WORKSPACE, "ppo_loss workspace too small: need %zu, got %zu", need, workspace_bytes);
    VPG_CHECK_ARG((reinterpret_cast<uintptr_t>(workspace) & 7) == 0, "workspace must be 8-byte aligned");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    VPG_CUDA(cudaMemsetAsync(workspace, 0, ppo_ws_header_bytes(B), stream));  // reduction slots only
    const bool vec = rows_vec_ok(L, mask, mask_dtype,
                                 {old_log_prob, log_prob, advantages, entropy, ref_log_prob, d_log_prob, d_entropy});
    // forward only + token-mean: one sweep, the mask count taken on the way
    const bool one_pass = (agg_mode == VPG_AGG_TOKEN_MEAN) && !d_log_prob && !d_entropy;
    const int gx = blocks_per_row(B, L, vec ? 4 : 1, one_pass ? 16 : 8);
    const unsigned grid = static_cast<unsigned>(B * gx);
    // 4- / 8-byte integer masks: the count pass leaves a one-byte copy for the loss pass
    unsigned char* cmask = (mask_dtype == VPG_MASK_I64 || mask_dtype == VPG_MASK_I32)
                               ? static_cast<unsigned char*>(workspace) + ppo_ws_header_bytes(B) : nullptr;
    PpoParams p{};
    p.old_lp = old_log_prob, p.lp = log_prob, p.adv = advantages, p.entropy = entropy, p.ref_lp = ref_log_prob;
    p.B = B, p.L = L, p.gx = gx;
    p.ratio_lo = ratio_lo, p.ratio_hi = ratio_hi, p.clip_c = clip_c, p.pg_kind = pg_kind;
    p.agg_mode = agg_mode, p.entropy_coeff = entropy_coeff, p.kl_coef = kl_coef, p.kl_kind = kl_kind;
    p.loss_scale = loss_scale;
    p.metrics = metrics, p.d_lp = d_log_prob, p.d_ent = d_entropy;
    p.ws = static_cast<double*>(workspace);
    // compile-time folded bodies for the recipes' configurations (vector path only)
    const bool kl_on = ref_log_prob != nullptr && kl_kind >= 0;
    int spec = 0;
    if (vec && pg_kind == VPG_PG_CLIP && agg_mode == VPG_AGG_TOKEN_MEAN && spec_enabled())
        spec = !kl_on ? 1 : (kl_kind == VPG_KL_LOW_VAR ? 2 : 0);
#define VPG_PPO_LAUNCH(V, COUNT, CM)                                                             \
    do {                                                                                         \
        if (V == 4 && spec == 1) ppo_loss_kernel<V, MT, COUNT, (V == 4 ? 1 : 0)><<<grid, kThreads, 0, stream>>>(p, mk, CM); \
        else if (V == 4 && spec == 2) ppo_loss_kernel<V, MT, COUNT, (V == 4 ? 2 : 0)><<<grid, kThreads, 0, stream>>>(p, mk, CM); \
        else ppo_loss_kernel<V, MT, COUNT, 0><<<grid, kThreads, 0, stream>>>(p, mk, CM);          \
    } while (0)
    // token-mean with gradients: one cooperative launch that reads the mask once (see ppo_loss_coop_kernel)
    bool coop_done = false;
    if (spec != 0 && !one_pass && coop_enabled()) {
        VPG_DISPATCH_MASK(mask_dtype, MT, {
            const MT* mk = static_cast<const MT*>(mask);
            const void* fn = spec == 1 ? reinterpret_cast<const void*>(ppo_loss_coop_kernel<MT, 1>)
                                       : reinterpret_cast<const void*>(ppo_loss_coop_kernel<MT, 2>);
            long long n = B * L, iters = 0;
            unsigned cgrid = 0;
            size_t smem = 0;
            if (coop_plan(fn, n, &iters, &cgrid, &smem)) {
                void* kargs[] = {&p, &mk, &n, &iters};
                // a refused cooperative launch (fewer SMs than planned for: MPS share, green context) is not an error:
                // the two-launch path below computes the same result
                if (cudaLaunchCooperativeKernel(fn, dim3(cgrid), dim3(kThreads), kargs, smem, stream) == cudaSuccess)
                    coop_done = true;
                else
                    (void)cudaGetLastError();
            }
        });
    }
    if (coop_done) {
        VPG_LAUNCH_CHECK();
        return VPG_OK;
    }
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        const MT* mk = static_cast<const MT*>(mask);
        if (one_pass) {
            if (vec) VPG_PPO_LAUNCH(4, true, nullptr);
            else VPG_PPO_LAUNCH(1, true, nullptr);
        } else if (vec) {
            mask_count_kernel<4, MT><<<grid, kThreads, 0, stream>>>(mk, B, L, gx, p.ws, cmask);
            count_launch();
            VPG_PPO_LAUNCH(4, false, cmask);
        } else {
            mask_count_kernel<1, MT><<<grid, kThreads, 0, stream>>>(mk, B, L, gx, p.ws, cmask);
            count_launch();
            VPG_PPO_LAUNCH(1, false, cmask);
        }
    });
#undef VPG_PPO_LAUNCH
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

int vpg_agg_loss(const float* loss_mat, const void* mask, int mask_dtype, int64_t B, int64_t L, int agg_mode,
                 float* out, float* d_loss_mat, void* workspace, size_t workspace_bytes, void* stream_) {
    VPG_CHECK_ARG(loss_mat && mask && out, "null pointer argument");
    VPG_CHECK_ARG(B > 0 && L > 0, "empty batch");
    if (agg_mode < 0 || agg_mode > 3) return set_error(VPG_ERR_INVALID, "Invalid loss_agg_mode: %d", agg_mode);
    const size_t need = ppo_ws_header_bytes(B);  // reduction slots only (callers may pass the larger ppo_loss size)
    if (!workspace || workspace_bytes < need)
        return set_error(VPG_ERR_WORKSPACE, "agg_loss workspace too small: need %zu, got %zu", need, workspace_bytes);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    VPG_CUDA(cudaMemsetAsync(workspace, 0, need, stream));
    double* ws = static_cast<double*>(workspace);
    const bool vec = rows_vec_ok(L, mask, mask_dtype, {loss_mat, d_loss_mat});
    const int gx = blocks_per_row(B, L, vec ? 4 : 1, 8);
    const unsigned grid = static_cast<unsigned>(B * gx);
    // one-byte mask copy for the second sweep when the caller's workspace has room for it (the ppo_loss size)
    unsigned char* cmask = ((mask_dtype == VPG_MASK_I64 || mask_dtype == VPG_MASK_I32) &&
                            workspace_bytes >= vpg_ppo_loss_workspace_bytes(B, L))
                               ? static_cast<unsigned char*>(workspace) + need : nullptr;
    const bool one_pass = (agg_mode == VPG_AGG_TOKEN_MEAN) && !d_loss_mat;  // forward only: count in the same sweep
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        const MT* mk = static_cast<const MT*>(mask);
        if (one_pass) {
            if (vec)
                agg_loss_kernel<4, MT, true><<<grid, kThreads, 0, stream>>>(loss_mat, mk, B, L, gx, agg_mode, out, nullptr, ws, nullptr);
            else
                agg_loss_kernel<1, MT, true><<<grid, kThreads, 0, stream>>>(loss_mat, mk, B, L, gx, agg_mode, out, nullptr, ws, nullptr);
        } else if (vec) {
            mask_count_kernel<4, MT><<<grid, kThreads, 0, stream>>>(mk, B, L, gx, ws, cmask);
            count_launch();
            agg_loss_kernel<4, MT, false><<<grid, kThreads, 0, stream>>>(loss_mat, mk, B, L, gx, agg_mode, out, d_loss_mat, ws, cmask);
        } else {
            mask_count_kernel<1, MT><<<grid, kThreads, 0, stream>>>(mk, B, L, gx, ws, cmask);
            count_launch();
            agg_loss_kernel<1, MT, false><<<grid, kThreads, 0, stream>>>(loss_mat, mk, B, L, gx, agg_mode, out, d_loss_mat, ws, cmask);
        }
    });
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

int vpg_value_loss(const float* vpreds, const float* returns, const float* values, const void* mask, int mask_dtype,
                   int64_t B, int64_t L, float cliprange_value, int agg_mode, float* out, float* d_vpreds,
                   void* workspace, size_t workspace_bytes, void* stream_) {
    VPG_CHECK_ARG(vpreds && returns && values && mask && out, "null pointer argument");
    VPG_CHECK_ARG(B > 0 && L > 0, "empty batch");
    if (agg_mode < 0 || agg_mode > 3) return set_error(VPG_ERR_INVALID, "Invalid loss_agg_mode: %d", agg_mode);
    const size_t need = ppo_ws_header_bytes(B);  // reduction slots only (callers may pass the larger ppo_loss size)
    if (!workspace || workspace_bytes < need)
        return set_error(VPG_ERR_WORKSPACE, "value_loss workspace too small: need %zu, got %zu", need, workspace_bytes);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    VPG_CUDA(cudaMemsetAsync(workspace, 0, need, stream));
    double* ws = static_cast<double*>(workspace);
    const bool vec = rows_vec_ok(L, mask, mask_dtype, {vpreds, returns, values, d_vpreds});
    const int gx = blocks_per_row(B, L, vec ? 4 : 1, 16);
    const unsigned grid = static_cast<unsigned>(B * gx);
    // one-byte mask copy for the second sweep when the caller's workspace has room for it (the ppo_loss size)
    unsigned char* cmask = ((mask_dtype == VPG_MASK_I64 || mask_dtype == VPG_MASK_I32) &&
                            workspace_bytes >= vpg_ppo_loss_workspace_bytes(B, L))
                               ? static_cast<unsigned char*>(workspace) + need : nullptr;
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        const MT* mk = static_cast<const MT*>(mask);
        if (vec) {
            mask_count_kernel<4, MT><<<grid, kThreads, 0, stream>>>(mk, B, L, gx, ws, cmask);
            count_launch();
            value_loss_kernel<4, MT><<<grid, kThreads, 0, stream>>>(vpreds, returns, values, mk, B, L, gx,
                                                                    cliprange_value, agg_mode, out, d_vpreds, ws, cmask);
        } else {
            mask_count_kernel<1, MT><<<grid, kThreads, 0, stream>>>(mk, B, L, gx, ws, cmask);
            count_launch();
            value_loss_kernel<1, MT><<<grid, kThreads, 0, stream>>>(vpreds, returns, values, mk, B, L, gx,
                                                                    cliprange_value, agg_mode, out, d_vpreds, ws, cmask);
        }
    });
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

size_t vpg_grpo_workspace_bytes(int64_t B, int64_t G) {
    return static_cast<size_t>((B > 0 ? B : 0) + 3 * (G > 0 ? G : 0)) * sizeof(double) +
           static_cast<size_t>(B > 0 ? B : 0) * sizeof(float) + 16;
}

int vpg_grpo_advantage(const float* rewards, const void* mask, int mask_dtype, const int32_t* group_ids, int64_t B,
                       int64_t L, int64_t G, float epsilon, int norm_by_std, float* advantages, void* workspace,
                       size_t workspace_bytes, void* stream_) {
    VPG_CHECK_ARG(rewards && mask && group_ids && advantages, "null pointer argument");
    VPG_CHECK_ARG(B > 0 && L > 0 && G > 0 && G <= B, "invalid GRPO shape B=%lld L=%lld groups=%lld", (long long)B,
                  (long long)L, (long long)G);
    const size_t need = vpg_grpo_workspace_bytes(B, G);
    if (!workspace || workspace_bytes < need)
        return set_error(VPG_ERR_WORKSPACE, "grpo workspace too small: need %zu, got %zu", need, workspace_bytes);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    double* scores = static_cast<double*>(workspace);
    double* gstat = scores + B;
    float* adv_row = reinterpret_cast<float*>(gstat + 3 * G);
    VPG_CUDA(cudaMemsetAsync(workspace, 0, static_cast<size_t>(B + 3 * G) * sizeof(double), stream));
    const bool vec = rows_vec_ok(L, mask, mask_dtype, {rewards, advantages});
    const int gx = blocks_per_row(B, L, vec ? 4 : 1, 16);
    const unsigned grid = static_cast<unsigned>(B * gx);
    if (vec)
        row_score_kernel<4, float><<<grid, kThreads, 0, stream>>>(rewards, L, gx, scores);
    else
        row_score_kernel<1, float><<<grid, kThreads, 0, stream>>>(rewards, L, gx, scores);
    count_launch();
    {
        int64_t blocks = ceil_div(B, kGroupRowsPerBlock);
        if (blocks > sm_count()) blocks = sm_count();
        if (blocks <= 1) {
            grpo_group_kernel<false><<<1, kGroupThreads, 0, stream>>>(scores, group_ids, B, G, epsilon, norm_by_std, gstat,
                                                                      adv_row);
        } else {  // grid barrier between the sweeps: every block must be resident -> cooperative launch
            const double* a0 = scores;
            const int* a1 = group_ids;
            long long a2 = B, a3 = G;
            float a4 = epsilon;
            int a5 = norm_by_std;
            double* a6 = gstat;
            float* a7 = adv_row;
            void* kargs[] = {&a0, &a1, &a2, &a3, &a4, &a5, &a6, &a7};
            VPG_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(grpo_group_kernel<true>),
                                                 dim3(static_cast<unsigned>(blocks)), dim3(kGroupThreads), kargs, 0, stream));
        }
        count_launch();
    }
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        const MT* mk = static_cast<const MT*>(mask);
        if (vec)
            broadcast_mask_kernel<4, MT><<<grid, kThreads, 0, stream>>>(adv_row, mk, L, gx, advantages);
        else
            broadcast_mask_kernel<1, MT><<<grid, kThreads, 0, stream>>>(adv_row, mk, L, gx, advantages);
    });
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

size_t vpg_gae_workspace_bytes(int64_t B, int64_t L) {
    (void)B;
    (void)L;
    return 4 * sizeof(double);
}

static int whiten_apply(const float* values, int64_t n, const double* sums, int shift_mean, float* out,
                        void* status_event, cudaStream_t stream) {
    VPG_CUDA(record_whiten_status(status_event, stream));  // right behind the kernel whose last block wrote the status
    if (n % 4 == 0 && aligned16(values) && aligned16(out))
        whiten_apply_kernel<4><<<flat_grid(n, 4), kThreads, 0, stream>>>(values, n, sums, shift_mean, out);
    else
        whiten_apply_kernel<1><<<flat_grid(n, 1), kThreads, 0, stream>>>(values, n, sums, shift_mean, out);
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

int vpg_gae(const float* rewards, const float* values, const void* mask, int mask_dtype, int64_t B, int64_t L,
            float gamma, float lam, float* advantages, float* returns, int32_t* status, void* workspace,
            size_t workspace_bytes, void* stream_, void* status_event) {
    VPG_CHECK_ARG(rewards && values && mask && advantages && returns && status, "null pointer argument");
    VPG_CHECK_ARG(B > 0 && L > 0, "empty batch");
    if (!workspace || workspace_bytes < vpg_gae_workspace_bytes(B, L))
        return set_error(VPG_ERR_WORKSPACE, "gae workspace too small");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    double* sums = static_cast<double*>(workspace);
    VPG_CUDA(cudaMemsetAsync(sums, 0, 4 * sizeof(double), stream));
    int64_t threads = round_up(ceil_div(L, kSeg), 32);
    if (threads > scan_block_threads(B)) threads = scan_block_threads(B);
    const int vec = (L % kSeg == 0) && rows_vec_ok(L, mask, mask_dtype, {rewards, values, advantages, returns});
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        gae_kernel<MT><<<static_cast<unsigned>(B), static_cast<unsigned>(threads), 0, stream>>>(
            rewards, values, static_cast<const MT*>(mask), L, gamma, lam, advantages, returns, sums, vec, status);
    });
    VPG_LAUNCH_CHECK();
    return whiten_apply(advantages, B * L, sums, 1, advantages, status_event, stream);
}

int vpg_masked_whiten(const float* values, const void* mask, int mask_dtype, int64_t n, int shift_mean, float* out,
                      int32_t* status, void* workspace, size_t workspace_bytes, void* stream_, void* status_event) {
    VPG_CHECK_ARG(values && mask && out && status && n > 0, "invalid masked_whiten arguments");
    if (!workspace || workspace_bytes < 4 * sizeof(double))
        return set_error(VPG_ERR_WORKSPACE, "masked_whiten workspace too small");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    double* sums = static_cast<double*>(workspace);
    VPG_CUDA(cudaMemsetAsync(sums, 0, 4 * sizeof(double), stream));
    const bool vec = rows_vec_ok(n, mask, mask_dtype, {values});
    VPG_DISPATCH_MASK(mask_dtype, MT, {
        const MT* mk = static_cast<const MT*>(mask);
        if (vec)
            whiten_stats_kernel<4, MT><<<flat_grid(n, 4), kThreads, 0, stream>>>(values, mk, n, sums, status);
        else
            whiten_stats_kernel<1, MT><<<flat_grid(n, 1), kThreads, 0, stream>>>(values, mk, n, sums, status);
    });
    VPG_LAUNCH_CHECK();
    return whiten_apply(values, n, sums, shift_mean, out, status_event, stream);
}

}  // extern "C"

// =====================================================================================================
// GRPO advantage for callers that cannot hand the library a scratch buffer (no counterpart in the reference).
//
// vpg_grpo_advantage needs (B + 3*G) doubles + B floats of scratch between its three launches, and every other entry
// point of this library leaves it to the caller to bring that memory.  Two kinds of hosts cannot: a plain C / ctypes
// host that runs without PyTorch's caching allocator, where the only way to get scratch is cudaMalloc / cudaFree around
// the call, and both of those stall every stream of the device (the rollout engine that shares the GPU included); and a
// trainer colocated with a rollout engine whose memory pool was sized to take all of HBM, where a cudaMalloc in the
// middle of a step simply fails.  For them the scratch lives in the library image itself: one fixed array in device
// global memory per GPU that loads the library, resident from module load until the process exits.
//
// The array is shared by every call on that GPU and keeps whatever the previous call left in it.  The part a call uses
// is cleared on the stream at the start of the call (as vpg_grpo_advantage always does with its workspace), so calls
// that follow one another on a stream are independent; two calls whose kernels can overlap on the GPU (different
// streams, different host threads) would race on the array and must be ordered by the caller.
// =====================================================================================================
namespace vpg {

constexpr int64_t kGrpoStaticMaxRows = VPG_GRPO_STATIC_MAX_ROWS;
// worst case G == B: (B + 3B) doubles + B floats + 16 bytes, rounded up to the array's alignment (576.25 KiB)
constexpr size_t kGrpoStaticBytes =
    ((static_cast<size_t>(kGrpoStaticMaxRows) * (4 * sizeof(double) + sizeof(float)) + 16 + 255) / 256) * 256;

static __device__ __align__(256) unsigned char g_grpo_static_scratch[kGrpoStaticBytes];

// Device address of g_grpo_static_scratch on the CURRENT device.  In host code the array's name is only a handle for the
// runtime's symbol table, not something a kernel can dereference, so the address is looked up once per device with
// cudaGetSymbolAddress (which may also be the call that loads this module onto the device) and cached.
static int grpo_static_scratch(void** out) {
    static std::mutex mu;
    static void* cache[64] = {nullptr};
    int dev = -1;
    VPG_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return set_error(VPG_ERR_UNSUPPORTED, "device ordinal %d outside the address cache (0..63)", dev);
    std::lock_guard<std::mutex> lk(mu);
    if (!cache[dev]) {
        void* p = nullptr;
        VPG_CUDA(cudaGetSymbolAddress(&p, g_grpo_static_scratch));
        if (!p) return set_error(VPG_ERR_CUDA, "cudaGetSymbolAddress returned a null address on device %d", dev);
        cache[dev] = p;
    }
    *out = cache[dev];
    return VPG_OK;
}

}  // namespace vpg

extern "C" {

int vpg_grpo_advantage_static(const float* rewards, const void* mask, int mask_dtype, const int32_t* group_ids,
                              int64_t B, int64_t L, int64_t G, float epsilon, int norm_by_std, float* advantages,
                              void* stream) {
    using namespace vpg;
    if (B > kGrpoStaticMaxRows)
        return set_error(VPG_ERR_WORKSPACE,
                         "vpg_grpo_advantage_static: B=%lld rows exceed the built-in scratch (at most %lld rows); use "
                         "vpg_grpo_advantage with a caller-provided workspace",
                         (long long)B, (long long)kGrpoStaticMaxRows);
    void* scratch = nullptr;
    const int rc = grpo_static_scratch(&scratch);
    if (rc != VPG_OK) return rc;
    // same validation, clearing and launches as the workspace variant; G <= B <= max rows bounds the bytes it touches
    return vpg_grpo_advantage(rewards, mask, mask_dtype, group_ids, B, L, G, epsilon, norm_by_std, advantages, scratch,
                              kGrpoStaticBytes, stream);
}

}  // extern "C"
