// Materialised-logits variants of the per-token statistics.
// Reference: verl/utils/torch_functional.py:64-133 (logprobs_from_logits and variants) and :145-160
// (entropy_from_logits / entropy_from_logits_with_chunking).
// One block per token row; a single sweep over the row keeps an online (max, sum exp, sum x*exp) per thread and the
// label logit, so a [T,V] logits tensor is read from HBM exactly once (the reference needs 3-4 passes).
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "common.h"
#include "reduce.cuh"

namespace vpg {

constexpr int kRowThreads = 512;

struct Osm {  // online softmax state: max, sum exp(x-m), sum x*exp(x-m)
    float m, s, b;
};
__device__ __forceinline__ Osm osm_merge(const Osm& a, const Osm& c) {
    const float m = fmaxf(a.m, c.m);
    const float fa = (a.m == -INFINITY) ? 0.f : expf(a.m - m);
    const float fc = (c.m == -INFINITY) ? 0.f : expf(c.m - m);
    return Osm{m, a.s * fa + c.s * fc, a.b * fa + c.b * fc};
}
template <int N>
__device__ __forceinline__ void osm_update(Osm& st, const float (&x)[N]) {
    float cm = x[0];
#pragma unroll
    for (int j = 1; j < N; ++j) cm = fmaxf(cm, x[j]);
    if (cm == -INFINITY) return;
    const float m = fmaxf(st.m, cm);
    const float f = (st.m == -INFINITY) ? 0.f : expf(st.m - m);
    float s = st.s * f, b = st.b * f;
#pragma unroll
    for (int j = 0; j < N; ++j) {
        const float e = expf(x[j] - m);
        s += e;
        b += (x[j] == -INFINITY) ? 0.f : x[j] * e;
    }
    st.m = m, st.s = s, st.b = b;
}

template <bool kBf16>
__global__ void logits_stats_kernel(const void* __restrict__ logits_, long long V, long long ld,
                                    const long long* __restrict__ labels, float* __restrict__ logprobs,
                                    float* __restrict__ entropy, float* __restrict__ lse_out,
                                    float* __restrict__ part, long long T, long long label_offset) {
    __shared__ Osm sh[kRowThreads / 32];
    const long long row = blockIdx.x;
    Osm st{-INFINITY, 0.f, 0.f};
    if constexpr (kBf16) {
        const __nv_bfloat16* p = reinterpret_cast<const __nv_bfloat16*>(logits_) + row * ld;
        const bool vec = ((reinterpret_cast<uintptr_t>(p) & 15) == 0);
        const long long nvec = vec ? V / 8 : 0;
        for (long long i = threadIdx.x; i < nvec; i += blockDim.x) {
            const uint4 raw = reinterpret_cast<const uint4*>(p)[i];
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
            float x[8];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = __bfloat1622float2(h[j]);
                x[2 * j] = f.x, x[2 * j + 1] = f.y;
            }
            osm_update<8>(st, x);
        }
        for (long long i = nvec * 8 + threadIdx.x; i < V; i += blockDim.x) {
            float x[1] = {__bfloat162float(p[i])};
            osm_update<1>(st, x);
        }
    } else {
        const float* p = reinterpret_cast<const float*>(logits_) + row * ld;
        const bool vec = ((reinterpret_cast<uintptr_t>(p) & 15) == 0);
        const long long nvec = vec ? V / 4 : 0;
        for (long long i = threadIdx.x; i < nvec; i += blockDim.x) {
            const float4 a = reinterpret_cast<const float4*>(p)[i];
            float x[4] = {a.x, a.y, a.z, a.w};
            osm_update<4>(st, x);
        }
        for (long long i = nvec * 4 + threadIdx.x; i < V; i += blockDim.x) {
            float x[1] = {p[i]};
            osm_update<1>(st, x);
        }
    }
    // warp then block merge
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Osm other;
        other.m = __shfl_xor_sync(0xffffffffu, st.m, o);
        other.s = __shfl_xor_sync(0xffffffffu, st.s, o);
        other.b = __shfl_xor_sync(0xffffffffu, st.b, o);
        st = osm_merge(st, other);
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) sh[w] = st;
    __syncthreads();
    if (threadIdx.x == 0) {
        Osm t = sh[0];
        for (int i = 1; i < (blockDim.x >> 5); ++i) t = osm_merge(t, sh[i]);
        if (part) {
            // vocab-shard partials in the wire format of the fused lm-head (vpg_lce_finalize merges them):
            // {max in the log2 domain, sum exp(x-max), sum x*exp(x-max), label logit if the label is in this shard}
            const long long lab = labels ? labels[row] - label_offset : -1;
            float xl = 0.f;
            if (lab >= 0 && lab < V) {
                if constexpr (kBf16)
                    xl = __bfloat162float((reinterpret_cast<const __nv_bfloat16*>(logits_) + row * ld)[lab]);
                else
                    xl = (reinterpret_cast<const float*>(logits_) + row * ld)[lab];
            }
            part[row] = t.m * 1.4426950408889634f;
            part[T + row] = t.s;
            part[2 * T + row] = t.b;
            part[3 * T + row] = xl;
            return;
        }
        const float lse = t.m + logf(t.s);
        if (lse_out) lse_out[row] = lse;
        if (entropy) entropy[row] = lse - t.b / t.s;
        if (logprobs) {
            const long long lab = labels[row];
            float xl = 0.f;
            if (lab >= 0 && lab < V) {
                if constexpr (kBf16)
                    xl = __bfloat162float((reinterpret_cast<const __nv_bfloat16*>(logits_) + row * ld)[lab]);
                else
                    xl = (reinterpret_cast<const float*>(logits_) + row * ld)[lab];
            }
            logprobs[row] = xl - lse;
        }
    }
}

// dlogits[t,v] = g_lp[t]*(1[v==label] - p) - g_ent[t]*p*(x - eb[t]),  p = exp(x - lse[t]),  eb = lse - entropy
template <bool kBf16>
__global__ void logits_stats_bwd_kernel(const void* __restrict__ logits_, long long V, long long ld, int gx,
                                        const long long* __restrict__ labels, const float* __restrict__ lse,
                                        const float* __restrict__ entropy, const float* __restrict__ g_lp,
                                        const float* __restrict__ g_ent, void* __restrict__ dlogits_) {
    const long long row = blockIdx.x / gx;
    const int cx = blockIdx.x % gx;
    const float l = lse[row];
    const float eb = entropy ? (l - entropy[row]) : 0.f;
    const float glp = g_lp ? g_lp[row] : 0.f;
    const float gent = (g_ent && entropy) ? g_ent[row] : 0.f;
    const long long lab = labels ? labels[row] : -1;
    const long long stride = static_cast<long long>(gx) * blockDim.x;
    for (long long v = static_cast<long long>(cx) * blockDim.x + threadIdx.x; v < V; v += stride) {
        float x;
        if constexpr (kBf16)
            x = __bfloat162float((reinterpret_cast<const __nv_bfloat16*>(logits_) + row * ld)[v]);
        else
            x = (reinterpret_cast<const float*>(logits_) + row * ld)[v];
        const float p = expf(x - l);
        float d = glp * ((v == lab ? 1.f : 0.f) - p) - gent * p * (x - eb);
        if constexpr (kBf16)
            (reinterpret_cast<__nv_bfloat16*>(dlogits_) + row * ld)[v] = __float2bfloat16_rn(d);
        else
            (reinterpret_cast<float*>(dlogits_) + row * ld)[v] = d;
    }
}

}  // namespace vpg

using namespace vpg;

extern "C" {

int vpg_logits_stats(const void* logits, int logits_bf16, int64_t T, int64_t V, int64_t ld, const int64_t* labels,
                     float* logprobs, float* entropy, float* lse, void* stream_) {
    VPG_CHECK_ARG(logits && T >= 0 && V > 0 && ld >= V, "invalid logits_stats arguments");
    VPG_CHECK_ARG(!logprobs || labels, "labels required for logprobs");
    if (T == 0) return VPG_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const long long* lab = reinterpret_cast<const long long*>(labels);
    if (logits_bf16)
        logits_stats_kernel<true><<<static_cast<unsigned>(T), kRowThreads, 0, stream>>>(logits, V, ld, lab, logprobs, entropy, lse, nullptr, T, 0);
    else
        logits_stats_kernel<false><<<static_cast<unsigned>(T), kRowThreads, 0, stream>>>(logits, V, ld, lab, logprobs, entropy, lse, nullptr, T, 0);
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

int vpg_logits_partials(const void* logits, int logits_bf16, int64_t T, int64_t V_local, int64_t ld,
                         const int64_t* labels, int64_t label_offset, float* partials, void* stream_) {
    VPG_CHECK_ARG(logits && partials && T >= 0 && V_local > 0 && ld >= V_local, "invalid logits_partials arguments");
    if (T == 0) return VPG_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const long long* lab = reinterpret_cast<const long long*>(labels);
    if (logits_bf16)
        logits_stats_kernel<true><<<static_cast<unsigned>(T), kRowThreads, 0, stream>>>(logits, V_local, ld, lab, nullptr, nullptr, nullptr, partials, T, label_offset);
    else
        logits_stats_kernel<false><<<static_cast<unsigned>(T), kRowThreads, 0, stream>>>(logits, V_local, ld, lab, nullptr, nullptr, nullptr, partials, T, label_offset);
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

int vpg_logits_stats_bwd(const void* logits, int logits_bf16, int64_t T, int64_t V, int64_t ld,
                         const int64_t* labels, const float* lse, const float* entropy, const float* g_lp,
                         const float* g_ent, void* dlogits, void* stream_) {
    VPG_CHECK_ARG(logits && dlogits && lse && T >= 0 && V > 0 && ld >= V, "invalid logits_stats_bwd arguments");
    if (T == 0) return VPG_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    int64_t gx = ceil_div(V, 256 * 8);
    const int64_t want = ceil_div(static_cast<int64_t>(sm_count()) * 8, T);
    if (gx > want) gx = want;
    if (gx < 1) gx = 1;
    const unsigned grid = static_cast<unsigned>(T * gx);
    const long long* lab = reinterpret_cast<const long long*>(labels);
    if (logits_bf16)
        logits_stats_bwd_kernel<true><<<grid, 256, 0, stream>>>(logits, V, ld, static_cast<int>(gx), lab, lse, entropy, g_lp, g_ent, dlogits);
    else
        logits_stats_bwd_kernel<false><<<grid, 256, 0, stream>>>(logits, V, ld, static_cast<int>(gx), lab, lse, entropy, g_lp, g_ent, dlogits);
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

}  // extern "C"
