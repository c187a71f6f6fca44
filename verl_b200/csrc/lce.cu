// Fused lm-head -> log-prob / entropy (forward partials, finalize, backward) and the plain tcgen05 GEMM entry.
// Replaces verl/utils/kernel/kernels.py:474-663 (forward) and :1345-1553 (backward) of the reference.
#include <cuda.h>
#include <cuda_runtime.h>

#include <stdlib.h>

#include <atomic>
#include <mutex>
#include <vector>

#include "common.h"
#include "tc_gemm.cuh"

namespace vpg {

// --------------------------------------------------------------------------------------- tensor maps (host)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
    });
    return fn;
}

// bf16 2-D tensor map over a row-major [outer][inner] matrix with row stride ld (elements), 128B swizzle.
// cuTensorMapEncodeTiled is a DRIVER entry point: it fails with CUDA_ERROR_INVALID_CONTEXT on a thread that has not
// bound the device's primary context yet (e.g. autograd's worker thread of device 0 before its first runtime call).
// cudaFree(nullptr) binds it; once per (thread, device).
static void bind_primary_context() {
    static thread_local int bound_dev = -1;
    int dev = -1;
    if (cudaGetDevice(&dev) == cudaSuccess && dev != bound_dev) {
        (void)cudaFree(nullptr);
        bound_dev = dev;
    }
}

// Tuning knobs: read from the environment ONCE, at the first launch of the process (never per launch).
struct Tuning {
    int cta_group = 2;       // VPG_CTA_GROUP=1 selects single-CTA tiles (A/B and odd SM counts)
    int sync_slack = 2;      // VPG_SYNC_SLACK: pacing slack in checkpoints (0 disables the soft lock-step)
    int ds_stcs = 0;         // VPG_DS_STCS=1: streaming stores for dlogits
    int worker_perm = 0;     // VPG_WORKER_PERM=1: neighbour workers stream the same vocab split
    int cluster1 = 1;        // VPG_CLUSTER1=2: launch single-CTA tiles as clusters of two
    int reserve_pairs = -1;  // VPG_BWD_RESERVE_PAIRS: override of the caller's reserve_cta_pairs
    int bwd_ws_mb = 3072;    // VPG_BWD_WS_MB: default dlogits workspace budget (MiB)
    int splitk = 1;          // VPG_SPLITK=0 disables the split-K tail of the gradient GEMMs
    int wide = 3;            // VPG_GEMM_WIDE bit mask: 256x512 tiles of CTA pairs in d_weight / plain GEMMs (1), d_hidden (2)
    int ds_tma = 1;          // VPG_DS_TMA=0: dlogits by per-lane global stores instead of smem + TMA tensor stores
    int hint_a[4] = {0, 0, 0, 0}, hint_b[4] = {0, 0, 0, 0};  // L2 eviction hints per kernel {FWD, DS, DH, DW}
};
static const Tuning& tuning() {
    static Tuning t = [] {
        Tuning t;
        const char* cg = getenv("VPG_CTA_GROUP");
        t.cta_group = (cg && cg[0] == '1') ? 1 : 2;
        t.sync_slack = env_int("VPG_SYNC_SLACK", 2);
        t.ds_stcs = env_int("VPG_DS_STCS", 0);
        t.worker_perm = env_int("VPG_WORKER_PERM", 0);
        t.cluster1 = env_int("VPG_CLUSTER1", 1);
        t.reserve_pairs = env_int("VPG_BWD_RESERVE_PAIRS", -1);
        t.bwd_ws_mb = env_int("VPG_BWD_WS_MB", 3072);
        t.splitk = env_int("VPG_SPLITK", 1);
        t.ds_tma = env_int("VPG_DS_TMA", 1);
        t.wide = env_int("VPG_GEMM_WIDE", 3);
        const char* tags[4] = {"FWD", "DS", "DH", "DW"};
        for (int i = 0; i < 4; ++i) {
            char name[32];
            t.hint_a[i] = env_int("VPG_HINT_A", 0);
            t.hint_b[i] = env_int("VPG_HINT_B", 0);
            snprintf(name, sizeof(name), "VPG_HINT_A_%s", tags[i]);
            t.hint_a[i] = env_int(name, t.hint_a[i]);
            snprintf(name, sizeof(name), "VPG_HINT_B_%s", tags[i]);
            t.hint_b[i] = env_int(name, t.hint_b[i]);
        }
        return t;
    }();
    return t;
}

// ------------------------------------------------------------------------------------ per-launch timing (bench)
// vpg_timing_enable(1): every tcgen05 launch (and split-K reduce) is bracketed by two CUDA events on ITS stream;
// vpg_timing_collect() synchronises them and returns (kind, milliseconds) per launch.  Off by default: no events.
enum : int { TK_FWD = 0, TK_DS = 1, TK_DW = 2, TK_DH = 3, TK_GEMM = 4, TK_SPLITK = 5 };
struct TimedLaunch {
    int kind;
    cudaEvent_t e0, e1;
};
static std::atomic<int> g_timing_on{0};
static std::mutex g_timing_mu;
static std::vector<TimedLaunch> g_timed;
static std::vector<cudaEvent_t> g_event_pool;

static cudaEvent_t timing_event() {
    std::lock_guard<std::mutex> lk(g_timing_mu);
    if (!g_event_pool.empty()) {
        cudaEvent_t e = g_event_pool.back();
        g_event_pool.pop_back();
        return e;
    }
    cudaEvent_t e = nullptr;
    if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
    return e;
}
struct TimingScope {
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int kind;
    cudaStream_t stream;
    TimingScope(int kind_, cudaStream_t s) : kind(kind_), stream(s) {
        if (!g_timing_on.load(std::memory_order_relaxed)) return;
        e0 = timing_event(), e1 = timing_event();
        if (e0 && e1) (void)cudaEventRecord(e0, stream);
    }
    ~TimingScope() {
        if (!e0 || !e1) return;
        (void)cudaEventRecord(e1, stream);
        std::lock_guard<std::mutex> lk(g_timing_mu);
        g_timed.push_back({kind, e0, e1});
    }
};

static int make_tmap(CUtensorMap* tm, const void* base, int64_t inner, int64_t outer, int64_t ld, int box_inner,
                     int box_outer, int dt, CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
    const int esz = (dt == VPG_DT_F32) ? 4 : 2;
    EncodeTiledFn fn = get_encode_fn();
    bind_primary_context();
    if (!fn) return set_error(VPG_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    if ((reinterpret_cast<uintptr_t>(base) & 15) != 0 || (ld * esz) % 16 != 0)
        return set_error(VPG_ERR_INVALID, "TMA operand must be 16-byte aligned with a 16-byte-multiple row stride");
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(outer)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * esz};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_outer)};
    cuuint32_t estr[2] = {1, 1};
    const CUtensorMapDataType ty = dt == VPG_DT_F32   ? CU_TENSOR_MAP_DATA_TYPE_TFLOAT32  // rounds to TF32 on the way in (no truncation bias)
                                   : dt == VPG_DT_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                                      : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    CUresult r = fn(tm, ty, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(VPG_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return VPG_OK;
}

// K-major operand stored [rows][K]; MN-major operand stored [K][rows]
static int make_operand_tmap(CUtensorMap* tm, const void* base, int64_t rows, int64_t K, int64_t ld, bool mn_major,
                             int tile_rows, int dt = VPG_DT_BF16) {
    if (!mn_major) return make_tmap(tm, base, K, rows, ld, dt == VPG_DT_F32 ? kBlockK / 2 : kBlockK, tile_rows, dt);
    if (dt == VPG_DT_F32) return set_error(VPG_ERR_UNSUPPORTED, "fp32 (TF32) operands must be K-major");
    return make_tmap(tm, base, rows, K, ld, 64, kBlockK, dt);
}

static unsigned int next_launch_id() {  // one counter for all tc_gemm_kernel instantiations
    static std::atomic<unsigned int> counter{1};
    return counter.fetch_add(1);
}

// cta_group used by every tcgen05 launch: 2 (CTA pairs, 256x256 tiles) unless VPG_CTA_GROUP=1 is set
static int cta_group() { return tuning().cta_group; }
static int tile_m() { return kBlockM * cta_group(); }
static int num_workers() { return cta_group() == 2 ? sm_count() / 2 : sm_count(); }
static int b_tile_rows() { return kBlockN / cta_group(); }
// fp32 scratch of the split-K tail of one gradient GEMM: at most one partial tile per worker
// Output-tile width of a MODE_GEMM launch with N columns: 512 (wide tiles, CTA pairs only) when that pads N no further
// than 256-column tiles would, else 256.
enum : int { WIDE_DW = 1, WIDE_DH = 2 };
static int gemm_tile_n(int64_t N, int which = WIDE_DW) {
    if (cta_group() == 2 && (tuning().wide & which) && round_up(N, 2 * kBlockN) == round_up(N, kBlockN)) return 2 * kBlockN;
    return kBlockN;
}
// tiling fields of a plain GEMM with an M x N output (one tile per unit)
static void set_gemm_tiling(TcParams& p, int64_t M, int64_t N, int which = WIDE_DW) {
    const int tn = gemm_tile_n(N, which);
    p.wide = tn > kBlockN ? 1 : 0;
    p.num_m_tiles = static_cast<int>(ceil_div(M, tile_m()));
    p.num_n_tiles = static_cast<int>(ceil_div(N, tn));
    p.num_splits = p.num_n_tiles;
    p.tiles_per_split = 1;
    p.M = static_cast<int>(M);
    p.N = static_cast<int>(N);
}
static size_t splitk_scratch_bytes() { return static_cast<size_t>(num_workers()) * tile_m() * (2 * kBlockN) * 4; }

// Split-K plan of a MODE_GEMM launch on `workers` persistent workers: the tiles of the last, partial scheduling round
// (tiles % workers of them) are cut into ksplit K ranges, so that round costs 1/ksplit of a tile time.
static void plan_splitk(TcParams& p, int workers, float* scratch) {
    const int64_t tiles = static_cast<int64_t>(p.num_m_tiles) * p.num_n_tiles;
    p.tiles_full = static_cast<int>(tiles);
    p.ksplit = 1;
    p.splitk_ws = nullptr;
    if (!scratch || !tuning().splitk || workers <= 0 || tiles <= 0) return;
    const int64_t r = tiles % workers;
    if (r == 0) return;
    int ks = static_cast<int>(workers / r);
    if (ks > 8) ks = 8;
    while (ks > 1 && p.num_k_blocks / ks < 16) --ks;  // every part keeps a mainloop worth starting
    if (ks <= 1) return;
    p.tiles_full = static_cast<int>(tiles - r);
    p.ksplit = ks;
    p.splitk_ws = scratch;
}

template <int kMode, bool kAMn, bool kBMn, int kCG, bool kTf32 = false, bool kWide = false>
static int launch_tc_cg(const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p_in, cudaStream_t stream,
                        int timing_kind, const CUtensorMap* tc = nullptr) {
    auto kern = tc_gemm_kernel<kMode, kAMn, kBMn, kCG, kTf32, kWide>;
    constexpr int kTileN = kWide ? 2 * kBlockN : kBlockN;
    const Tuning& tn = tuning();
    TcParams p = p_in;
    const unsigned int id = next_launch_id();
    p.sync_slack = tn.sync_slack;
    p.sync_epoch = id & 0x7FFu;
    p.pace_slot = static_cast<int>(id % kPaceSlots);
    p.ds_stcs = tn.ds_stcs;
    p.ds_tma = (kMode == MODE_DS && tc && tn.ds_tma) ? 1 : 0;
    const int hk = kMode == MODE_FWD ? 0 : kMode == MODE_DS ? 1 : kAMn ? 3 : 2;
    p.hint_a = tn.hint_a[hk];
    p.hint_b = tn.hint_b[hk];
    if (kMode != MODE_GEMM) p.tiles_full = p.num_m_tiles * p.num_splits, p.ksplit = 1, p.splitk_ws = nullptr;
    constexpr int smem = kMode == MODE_DS ? TileCfg<kCG>::kSmemBytesDs
                                          : TileCfg<kCG>::kSmemBytes + (kMode == MODE_GEMM ? kRedScratchBytes : 0);
    // per-device attribute: set every call (cheap) so multi-device processes are handled
    cudaError_t attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (attr_err != cudaSuccess)
        return set_error(VPG_ERR_CUDA, "cudaFuncSetAttribute(smem=%d) failed: %s", smem, cudaGetErrorString(attr_err));
    const int64_t tiles = static_cast<int64_t>(p.num_m_tiles) * p.num_splits;
    const int64_t units = (kMode == MODE_GEMM) ? p.tiles_full + (tiles - p.tiles_full) * p.ksplit : tiles;
    if (units <= 0) return VPG_OK;
    int workers = static_cast<int>(units < num_workers() ? units : num_workers());
    if (p.max_workers > 0 && workers > p.max_workers) workers = p.max_workers;
    if (p.ar_world > 1) workers = p.max_workers > 0 ? p.max_workers : num_workers();  // all-reduce warps of every CTA
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(static_cast<unsigned>(workers * kCG));
    cfg.blockDim = dim3(kNumThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (kCG == 1 && tn.cluster1 == 2 && (workers % 2) == 0) ? 2 : kCG;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e;
    {
        TimingScope ts(timing_kind, stream);
        e = cudaLaunchKernelEx(&cfg, kern, ta, tb, tc ? *tc : ta, p);
    }
    count_launch();
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return set_error(VPG_ERR_CUDA, "tc_gemm_kernel launch failed: %s", cudaGetErrorString(e));
    }
    if (kMode == MODE_GEMM && p.ksplit > 1) {
        const int tail_tiles = static_cast<int>(tiles) - p.tiles_full;
        const int64_t threads = static_cast<int64_t>(tail_tiles) * (kBlockM * kCG) * kTileN / 4;
        TimingScope ts(TK_SPLITK, stream);
        splitk_reduce_kernel<<<static_cast<unsigned>(ceil_div(threads, 256)), 256, 0, stream>>>(
            p.splitk_ws, tail_tiles, p.ksplit, p.tiles_full, p.num_n_tiles, kBlockM * kCG, kTileN, p.M, p.N, p.out, p.ldc,
            p.out_mode, p.accum, p.ld_accum, p.f16);
        VPG_LAUNCH_CHECK();
    }
    return VPG_OK;
}

template <int kMode, bool kAMn, bool kBMn>
static int launch_tc(const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, cudaStream_t stream,
                     int timing_kind, const CUtensorMap* tc = nullptr) {
    if constexpr (kMode == MODE_GEMM) {
        if (p.wide) {
            if (cta_group() != 2) return set_error(VPG_ERR_INVALID, "wide GEMM tiles need CTA pairs");
            return launch_tc_cg<kMode, kAMn, kBMn, 2, false, true>(ta, tb, p, stream, timing_kind, tc);
        }
    }
    if (cta_group() == 2) return launch_tc_cg<kMode, kAMn, kBMn, 2>(ta, tb, p, stream, timing_kind, tc);
    return launch_tc_cg<kMode, kAMn, kBMn, 1>(ta, tb, p, stream, timing_kind, tc);
}
// lm-head statistics / dlogits launches in the operand type `dt` of the call: bf16 / fp16 (kind::f16) or fp32 (TF32)
template <int kMode>
static int launch_tc_lmhead(const CUtensorMap& ta, const CUtensorMap& tb, TcParams p, int dt, cudaStream_t stream,
                            const CUtensorMap* tc = nullptr) {
    p.f16 = (dt == VPG_DT_F16);
    const int tk = kMode == MODE_FWD ? TK_FWD : TK_DS;
    if (dt != VPG_DT_F32) return launch_tc<kMode, false, false>(ta, tb, p, stream, tk, tc);
    if (cta_group() == 2) return launch_tc_cg<kMode, false, false, 2, true>(ta, tb, p, stream, tk, tc);
    return launch_tc_cg<kMode, false, false, 1, true>(ta, tb, p, stream, tk, tc);
}

// Choice of the number of vocab splits per token tile (a unit = token tile x one split's run of vocab tiles, dealt
// round-robin to the persistent workers).  Cost of a split count s, in milliseconds:
//   * the static schedule itself, simulated: the most loaded worker's vocab tiles x the time of one tile step;
//   * DRAM traffic, which is time because the step is power-capped (~0.1 ms per GB, from the split sweeps in
//     profiles/r2_sweep_fwd_splits.jsonl): the weight shard streams through L2 once per round of grid/s token tiles
//     (traffic grows with s), and the grid/s token tiles in flight (tile_m x H each) are re-fetched for every vocab tile
//     once they no longer fit beside those streams (~48 MB; traffic grows as s falls).  Qwen2.5-7B (H = 3584): 4 splits,
//     17 GB; Llama-3-70B (H = 8192): 6-8 splits (4 splits: 111.1 ms, 8 splits: 106.7 ms for the forward).
// `row_bytes` = bytes of one operand row (H x element size); 0 keeps the balance-only choice.
static int choose_splits(int num_m_tiles, int num_n_tiles, int grid, int64_t row_bytes = 0) {
    int best_s = 1;
    double best_cost = 1e30;
    const int max_s = num_n_tiles < 64 ? num_n_tiles : 64;
    // one 256-column tile step of one worker at the sustained rate (2 * tile_m * 256 * K flops; K = row_bytes / 2)
    const double tile_ms = row_bytes > 0 ? 2.0 * tile_m() * kBlockN * (row_bytes / 2.0) / (1.4e15 / num_workers()) * 1e3 : 1.0;
    const double w_gb = static_cast<double>(num_n_tiles) * kBlockN * row_bytes / 1e9;
    const double a_tile_gb = static_cast<double>(tile_m()) * row_bytes / 1e9;
    for (int s = 1; s <= max_s; ++s) {
        const int tps = (num_n_tiles + s - 1) / s;
        const int s_eff = (num_n_tiles + tps - 1) / tps;  // non-empty splits
        if (s_eff != s) continue;
        const int64_t units = static_cast<int64_t>(num_m_tiles) * s;
        const int last = num_n_tiles - (s - 1) * tps;
        // worst CTA load under u = blockIdx + i*grid
        int64_t worst = 0;
        const int g = static_cast<int>(units < grid ? units : grid);
        for (int cta = 0; cta < g; ++cta) {
            int64_t load = 0;
            for (int64_t u = cta; u < units; u += g) load += ((u % s) == s - 1) ? last : tps;
            if (load > worst) worst = load;
        }
        double cost;
        if (row_bytes > 0) {
            const double in_flight = static_cast<double>(ceil_div(g, s) < num_m_tiles ? ceil_div(g, s) : num_m_tiles);
            const double rounds = static_cast<double>(num_m_tiles) / in_flight;       // sweeps of the weight shard
            const double foot_mb = in_flight * a_tile_gb * 1e9 / 1048576.0;
            const double thrash = foot_mb > 48.0 ? 1.0 - 48.0 / foot_mb : 0.0;          // share of token-tile reads that miss
            const double dram_gb = w_gb * rounds + num_m_tiles * a_tile_gb * (1.0 + thrash * (tps - 1));
            cost = static_cast<double>(worst) * tile_ms * (1.0 + 0.002 * s) + 0.1 * dram_gb;
        } else {
            // mild preference for >= 4 concurrent vocab streams (L2 reuse of the token tile) and fewer partials
            cost = static_cast<double>(worst) * (1.0 + 0.002 * s) * (s < 4 && num_m_tiles * s >= grid ? 1.03 : 1.0);
        }
        if (cost < best_cost) {
            best_cost = cost;
            best_s = s;
        }
    }
    return best_s;
}

// --------------------------------------------------------------------------------------- finalize kernel
// partials [4][P][T] -> logprobs, entropy, lse, eb.  One thread per token; reads are coalesced along T.
__global__ void lce_finalize_kernel(const float* __restrict__ part, int P, long long T, float inv_temp,
                                    float* __restrict__ logprobs, float* __restrict__ entropy,
                                    float* __restrict__ lse_out, float* __restrict__ eb_out) {
    const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= T) return;
    const long long plane = static_cast<long long>(P) * T;
    float m = -INFINITY;
    for (int i = 0; i < P; ++i) m = fmaxf(m, part[i * T + t]);
    float s = 0.f, b = 0.f, tg = 0.f;
    for (int i = 0; i < P; ++i) {
        const float mi = part[i * T + t];
        const float f = (mi == -INFINITY) ? 0.f : exp2f(mi - m);
        s = fmaf(part[plane + i * T + t], f, s);
        b = fmaf(part[2 * plane + i * T + t], f, b);
        tg += part[3 * plane + i * T + t];
    }
    const float ln2 = 0.6931471805599453f;
    const float lse = (m + log2f(s)) * ln2;
    const float eb = (b / s) * inv_temp;
    if (logprobs) logprobs[t] = tg * inv_temp - lse;
    if (entropy) entropy[t] = lse - eb;
    if (lse_out) lse_out[t] = lse;
    if (eb_out) eb_out[t] = eb;
}

// fp32 -> bf16 (round to nearest even), 8 elements per thread; used after the fused d_hidden all-reduce
__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * 8;
    for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 8; i < n; i += stride) {
        if (i + 8 <= n) {
            const float4 a = *reinterpret_cast<const float4*>(src + i);
            const float4 b = *reinterpret_cast<const float4*>(src + i + 4);
            uint4 v;
            v.x = pack_bf16x2(a.x, a.y), v.y = pack_bf16x2(a.z, a.w);
            v.z = pack_bf16x2(b.x, b.y), v.w = pack_bf16x2(b.z, b.w);
            *reinterpret_cast<uint4*>(dst + i) = v;
        } else {
            for (long long j = i; j < n; ++j) dst[j] = __float2bfloat16_rn(src[j]);
        }
    }
}

static int check_lce_shapes(int64_t T, int64_t H, int64_t V) {
    VPG_CHECK_ARG(T >= 0 && H > 0 && V > 0, "invalid shape T=%lld H=%lld V=%lld", (long long)T, (long long)H,
                  (long long)V);
    VPG_CHECK_ARG(H % 64 == 0, "hidden size must be a multiple of 64 (got %lld)", (long long)H);
    VPG_CHECK_ARG(T < (1LL << 31) - 256 && V < (1LL << 31) - 256, "T and V must fit in int32");
    return VPG_OK;
}

}  // namespace vpg

using namespace vpg;

extern "C" {

int vpg_timing_enable(int on) {
    g_timing_on.store(on ? 1 : 0);
    if (!on) {  // drop whatever was recorded and not collected
        std::lock_guard<std::mutex> lk(g_timing_mu);
        for (auto& t : g_timed) g_event_pool.push_back(t.e0), g_event_pool.push_back(t.e1);
        g_timed.clear();
    }
    return VPG_OK;
}

int vpg_timing_collect(int* kinds, float* ms, int max_n) {
    std::vector<TimedLaunch> taken;
    {
        std::lock_guard<std::mutex> lk(g_timing_mu);
        taken.swap(g_timed);
    }
    int n = 0;
    for (auto& t : taken) {
        float dt = -1.f;
        if (cudaEventSynchronize(t.e1) == cudaSuccess) (void)cudaEventElapsedTime(&dt, t.e0, t.e1);
        if (n < max_n && kinds && ms) kinds[n] = t.kind, ms[n] = dt, ++n;
        std::lock_guard<std::mutex> lk(g_timing_mu);
        g_event_pool.push_back(t.e0), g_event_pool.push_back(t.e1);
    }
    (void)cudaGetLastError();
    return n;
}

int vpg_lce_num_splits(int64_t T, int64_t H, int64_t V, int sm) {
    (void)H;
    if (T <= 0 || V <= 0) return 1;
    if (sm <= 0) sm = sm_count();
    const int workers = cta_group() == 2 ? sm / 2 : sm;
    return choose_splits(static_cast<int>(ceil_div(T, tile_m())), static_cast<int>(ceil_div(V, kBlockN)), workers, H * 2);
}

int vpg_lce_fwd_partials(const void* hidden, const void* weight, const int64_t* labels, int64_t T, int64_t H,
                         int64_t V, float temperature, int64_t label_offset, int num_splits, float* partials,
                         void* stream) {
    return vpg_lce_fwd_partials_ex(VPG_DT_BF16, hidden, weight, labels, T, H, V, temperature, label_offset, num_splits,
                                   partials, 0, num_splits, 0, nullptr, 0, stream);
}

int vpg_lce_fwd_partials_dt(int dtype, const void* hidden, const void* weight, const int64_t* labels, int64_t T,
                            int64_t H, int64_t V, float temperature, int64_t label_offset, int num_splits,
                            float* partials, void* stream) {
    return vpg_lce_fwd_partials_ex(dtype, hidden, weight, labels, T, H, V, temperature, label_offset, num_splits,
                                   partials, 0, num_splits, 0, nullptr, 0, stream);
}

int vpg_lce_fwd_partials_push(const void* hidden, const void* weight, const int64_t* labels, int64_t T, int64_t H,
                              int64_t V, float temperature, int64_t label_offset, int num_splits, float* partials,
                              int part_split0, int part_total, int partials_is_multicast, void* const* push_ptrs,
                              int n_push, void* stream) {
    return vpg_lce_fwd_partials_ex(VPG_DT_BF16, hidden, weight, labels, T, H, V, temperature, label_offset, num_splits,
                                   partials, part_split0, part_total, partials_is_multicast, push_ptrs, n_push, stream);
}

int vpg_lce_fwd_partials_ex(int dt, const void* hidden, const void* weight, const int64_t* labels, int64_t T,
                            int64_t H, int64_t V, float temperature, int64_t label_offset, int num_splits,
                            float* partials, int part_split0, int part_total, int partials_is_multicast,
                            void* const* push_ptrs, int n_push, void* stream) {
    int rc = check_lce_shapes(T, H, V);
    if (rc) return rc;
    VPG_CHECK_ARG(dt == VPG_DT_BF16 || dt == VPG_DT_F16 || dt == VPG_DT_F32, "unknown operand dtype %d", dt);
    VPG_CHECK_ARG(hidden && weight && labels && (partials || n_push > 0), "null pointer argument");
    VPG_CHECK_ARG(part_split0 >= 0 && part_total >= part_split0 + num_splits, "partials slot [%d, %d) exceeds %d parts",
                  part_split0, part_split0 + num_splits, part_total);
    VPG_CHECK_ARG(n_push >= 0 && n_push <= kMaxPeers && (n_push == 0 || push_ptrs), "n_push must be in [0, %d]",
                  kMaxPeers);
    VPG_CHECK_ARG(!(partials_is_multicast && n_push > 0), "choose multicast or unicast push, not both");
    VPG_CHECK_ARG(temperature > 0.f, "temperature must be > 0 (got %f)", temperature);
    if (T == 0) return VPG_OK;
    TcParams p{};
    p.num_m_tiles = static_cast<int>(ceil_div(T, tile_m()));
    p.num_n_tiles = static_cast<int>(ceil_div(V, kBlockN));
    VPG_CHECK_ARG(dt == VPG_DT_BF16 || (n_push == 0 && !partials_is_multicast),
                  "the push variants take bf16 operands only");
    p.num_k_blocks = static_cast<int>(H / (dt == VPG_DT_F32 ? kBlockK / 2 : kBlockK));
    VPG_CHECK_ARG(num_splits >= 1 && num_splits <= p.num_n_tiles, "num_splits %d out of range [1,%d]", num_splits,
                  p.num_n_tiles);
    p.num_splits = num_splits;
    p.tiles_per_split = static_cast<int>(ceil_div(p.num_n_tiles, num_splits));
    p.M = static_cast<int>(T);
    p.N = static_cast<int>(V);
    p.scale_log2 = 1.4426950408889634f / temperature;
    p.inv_temp = 1.0f / temperature;
    p.labels = reinterpret_cast<const long long*>(labels);
    p.label_offset = label_offset;
    p.partials = partials;
    p.part_split0 = part_split0;
    p.part_total = part_total;
    p.partials_mc = partials_is_multicast ? 1 : 0;
    p.n_push = n_push;
    for (int i = 0; i < n_push; ++i) {
        VPG_CHECK_ARG(push_ptrs[i], "push_ptrs[%d] is null", i);
        p.push_ptrs[i] = static_cast<float*>(push_ptrs[i]);
    }
    p.worker_perm = tuning().worker_perm ? num_splits : 0;
    CUtensorMap ta, tb;
    if ((rc = make_operand_tmap(&ta, hidden, T, H, H, false, kBlockM, dt))) return rc;
    if ((rc = make_operand_tmap(&tb, weight, V, H, H, false, b_tile_rows(), dt))) return rc;
    return launch_tc_lmhead<MODE_FWD>(ta, tb, p, dt, static_cast<cudaStream_t>(stream));
}

int vpg_lce_finalize(const float* partials, int num_parts, int64_t T, float temperature, float* logprobs,
                     float* entropy, float* lse, float* eb, void* stream) {
    VPG_CHECK_ARG(partials && num_parts >= 1 && T >= 0, "invalid finalize arguments");
    VPG_CHECK_ARG(temperature > 0.f, "temperature must be > 0");
    if (T == 0) return VPG_OK;
    const int threads = 256;
    lce_finalize_kernel<<<static_cast<unsigned>(ceil_div(T, threads)), threads, 0, static_cast<cudaStream_t>(stream)>>>(
        partials, num_parts, T, 1.0f / temperature, logprobs, entropy, lse, eb);
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

int64_t vpg_lce_bwd_default_chunk(int64_t T, int64_t H, int64_t V) {
    return vpg_lce_bwd_chunk_for_budget(T, H, V, static_cast<int64_t>(tuning().bwd_ws_mb) << 20);
}

int64_t vpg_lce_bwd_chunk_for_budget(int64_t T, int64_t H, int64_t V, int64_t budget) {
    // Largest chunk whose bf16 dlogits buffer fits the budget, then the chunk count (up to +8) that wastes the fewest
    // waves of the persistent grid in the three kernels of a chunk (cost in units of tile x k-block; the gradient
    // GEMMs cut the tiles of their last partial round along K, so that round costs 1/ksplit of a tile).
    const int64_t t_pad = round_up(T > 0 ? T : 1, 2 * kBlockM);
    const int64_t v_tiles = ceil_div(V, kBlockN);
    if (budget < (1 << 20)) budget = 1 << 20;  // bytes of the 16-bit dlogits workspace
    int64_t max_tiles = budget / (t_pad * 2) / kBlockN;
    if (max_tiles < 1) max_tiles = 1;
    if (max_tiles > v_tiles) max_tiles = v_tiles;
    const int64_t W = num_workers(), tm = tile_m();
    // a wide tile step costs `scale` 256-column steps
    const int64_t tn_w = gemm_tile_n(H, WIDE_DW), tn_h = gemm_tile_n(H, WIDE_DH), sc_w = tn_w / kBlockN, sc_h = tn_h / kBlockN;
    const int64_t m_t = ceil_div(t_pad, tm), n_w = ceil_div(H, tn_w), n_h = ceil_div(H, tn_h);
    const int64_t min_chunks = ceil_div(v_tiles, max_tiles);
    int64_t best_c = max_tiles;
    double best_cost = 1e300;
    auto rounds = [&](int64_t tiles) -> double {  // scheduling rounds of a gradient GEMM with the split-K tail
        const int64_t r = tiles % W;
        if (r == 0 || !tuning().splitk) return static_cast<double>(ceil_div(tiles, W));
        int64_t ks = W / r;
        if (ks > 8) ks = 8;
        return static_cast<double>(tiles / W) + 1.0 / static_cast<double>(ks < 1 ? 1 : ks);
    };
    // Fixed cost of one more chunk, in the same unit (one 256x256x64 tile step of one worker ~ 0.44 us at the sustained
    // rate): three launches with their ramps (~0.12 ms) plus the read-modify-write of the fp32 [T, H] accumulator in the
    // d_hidden epilogue (8 B per element at HBM rate).  Measured at configs[1]: 8 chunks cost 4.2 ms more than one block
    // (profiles/r2_bench_c2_n1*.json), i.e. ~0.6 ms per chunk.
    const double chunk_overhead = 300.0 + 3.0e-6 * static_cast<double>(t_pad) * static_cast<double>(H);
    for (int64_t nc = min_chunks; nc <= min_chunks + 8 && nc <= v_tiles; ++nc) {
        const int64_t c = ceil_div(v_tiles, nc);
        double cost = 0;
        for (int64_t done = 0; done < v_tiles; done += c) {
            const int64_t ct = (v_tiles - done) < c ? (v_tiles - done) : c;  // tiles in this chunk
            cost += static_cast<double>(ceil_div(m_t * ct, W)) * (H / kBlockK);                  // dlogits
            cost += rounds(ceil_div(ct * kBlockN, tm) * n_w) * (t_pad / kBlockK) * sc_w;          // d_weight
            cost += rounds(m_t * n_h) * (ct * kBlockN / kBlockK) * sc_h;                          // d_hidden
            cost += chunk_overhead;                                                               // launches, fp32 RMW
        }
        if (cost < best_cost) {
            best_cost = cost;
            best_c = c;
        }
    }
    return best_c * kBlockN;
}

// How the backward cuts its dlogits workspace.  SINGLE: the whole [T, V] block fits the budget.  VOCAB: column chunks
// (the reference's default _Split_Dlogits_N, kernels.py:1492-1548): d_weight rows are final per chunk, d_hidden accumulates
// in an fp32 [T, H] buffer.  TOKEN: row chunks (the reference's _Split_Dlogits_M): d_hidden rows are final per chunk,
// d_weight accumulates in an fp32 [V, H] buffer.  Auto picks the orientation with the SMALLER accumulator: a vocab shard
// of a long micro-batch (BASELINE configs[4]: T = 262144, V/8 = 19008) accumulates 0.39 GB instead of 5.4 GB per chunk.
enum : int { BWD_SINGLE = 0, BWD_VOCAB = 1, BWD_TOKEN = 2 };
struct BwdPlan {
    int mode;
    int64_t chunk;  // columns (VOCAB / SINGLE) or rows (TOKEN) of the dlogits block
    int n_chunks;
    size_t off_acc, off_splitk, bytes;
};
static BwdPlan plan_bwd(int64_t T, int64_t H, int64_t V, int64_t chunk_v, int64_t chunk_t) {
    BwdPlan pl{};
    const int64_t t_pad = round_up(T > 0 ? T : 1, 2 * kBlockM);
    const int64_t v_pad = round_up(V, kBlockN);
    const int64_t budget = static_cast<int64_t>(tuning().bwd_ws_mb) << 20;
    if (chunk_v > 0) {
        pl.mode = chunk_v >= v_pad ? BWD_SINGLE : BWD_VOCAB;
        pl.chunk = chunk_v >= v_pad ? v_pad : chunk_v;
    } else if (chunk_t > 0) {
        pl.mode = chunk_t >= t_pad ? BWD_SINGLE : BWD_TOKEN;
        pl.chunk = chunk_t >= t_pad ? v_pad : chunk_t;
    } else if (t_pad * v_pad * 2 <= budget) {
        pl.mode = BWD_SINGLE, pl.chunk = v_pad;
    } else if (T > V) {
        int64_t rows = budget / (v_pad * 2) / (2 * kBlockM) * (2 * kBlockM);
        if (rows < 2 * kBlockM) rows = 2 * kBlockM;
        const int64_t n = ceil_div(t_pad, rows);
        pl.mode = BWD_TOKEN, pl.chunk = round_up(ceil_div(t_pad, n), 2 * kBlockM);
    } else {
        pl.chunk = vpg_lce_bwd_chunk_for_budget(T, H, V, budget);
        pl.mode = pl.chunk >= v_pad ? BWD_SINGLE : BWD_VOCAB;
        if (pl.chunk > v_pad) pl.chunk = v_pad;
    }
    const int64_t ds_rows = pl.mode == BWD_TOKEN ? pl.chunk : t_pad;
    const int64_t ds_cols = pl.mode == BWD_TOKEN ? v_pad : pl.chunk;
    pl.n_chunks = static_cast<int>(pl.mode == BWD_TOKEN ? ceil_div(t_pad, pl.chunk) : ceil_div(v_pad, pl.chunk));
    size_t bytes = round_up(ds_rows * ds_cols * 2, 256);
    pl.off_acc = bytes;
    if (pl.mode == BWD_VOCAB) bytes += round_up(T * H * 4, 256);  // fp32 d_hidden accumulator
    if (pl.mode == BWD_TOKEN) bytes += round_up(V * H * 4, 256);  // fp32 d_weight accumulator
    pl.off_splitk = bytes;
    bytes += splitk_scratch_bytes();  // split-K tail partial tiles
    pl.bytes = bytes + 256;
    return pl;
}

size_t vpg_lce_bwd_workspace_bytes(int64_t T, int64_t H, int64_t V, int64_t chunk_v) {
    if (T <= 0) return 0;
    return plan_bwd(T, H, V, chunk_v, 0).bytes;
}

int vpg_lce_bwd_plan(int64_t T, int64_t H, int64_t V, int64_t chunk_v, int64_t chunk_t, int64_t* out4) {
    VPG_CHECK_ARG(out4 && T > 0 && H > 0 && V > 0, "invalid arguments");
    const BwdPlan pl = plan_bwd(T, H, V, chunk_v, chunk_t);
    out4[0] = pl.mode, out4[1] = pl.chunk, out4[2] = pl.n_chunks, out4[3] = static_cast<int64_t>(pl.bytes);
    return VPG_OK;
}

int vpg_lce_bwd(const void* hidden, const void* weight, const int64_t* labels, const float* lse, const float* eb,
                const float* g_lp, const float* g_ent, int64_t T, int64_t H, int64_t V, float temperature,
                int64_t label_offset, int64_t chunk_v, void* d_hidden, void* d_weight, void* workspace,
                size_t workspace_bytes, void* stream_) {
    return vpg_lce_bwd_dt(VPG_DT_BF16, hidden, weight, labels, lse, eb, g_lp, g_ent, T, H, V, temperature, label_offset,
                          chunk_v, d_hidden, d_weight, workspace, workspace_bytes, stream_);
}

int vpg_lce_bwd_dt(int dtype, const void* hidden, const void* weight, const int64_t* labels, const float* lse,
                   const float* eb, const float* g_lp, const float* g_ent, int64_t T, int64_t H, int64_t V,
                   float temperature, int64_t label_offset, int64_t chunk_v, void* d_hidden, void* d_weight,
                   void* workspace, size_t workspace_bytes, void* stream) {
    VPG_CHECK_ARG(dtype == VPG_DT_BF16 || dtype == VPG_DT_F16, "operand dtype must be VPG_DT_BF16 or VPG_DT_F16");
    vpg_lce_bwd_args a{};
    a.struct_size = sizeof(a);
    a.dtype = dtype, a.hidden = hidden, a.weight = weight, a.labels = labels, a.lse = lse, a.eb = eb;
    a.g_lp = g_lp, a.g_ent = g_ent, a.T = T, a.H = H, a.V = V, a.temperature = temperature;
    a.label_offset = label_offset, a.chunk_v = chunk_v, a.d_hidden = d_hidden, a.d_weight = d_weight;
    a.workspace = workspace, a.workspace_bytes = workspace_bytes, a.stream = stream;
    return vpg_lce_bwd_ex(&a);
}

int vpg_lce_bwd_f32(const void* hidden_f32, const void* weight_f32, const void* hidden_bf16, const void* weight_bf16,
                    const int64_t* labels, const float* lse, const float* eb, const float* g_lp, const float* g_ent,
                    int64_t T, int64_t H, int64_t V, float temperature, int64_t label_offset, int64_t chunk_v,
                    void* d_hidden_f32, void* d_weight_f32, void* workspace, size_t workspace_bytes, void* stream) {
    vpg_lce_bwd_args a{};
    a.struct_size = sizeof(a);
    a.dtype = VPG_DT_F32, a.hidden = hidden_f32, a.weight = weight_f32, a.hidden16 = hidden_bf16;
    a.weight16 = weight_bf16, a.labels = labels, a.lse = lse, a.eb = eb, a.g_lp = g_lp, a.g_ent = g_ent;
    a.T = T, a.H = H, a.V = V, a.temperature = temperature, a.label_offset = label_offset, a.chunk_v = chunk_v;
    a.d_hidden = d_hidden_f32, a.d_weight = d_weight_f32, a.workspace = workspace, a.workspace_bytes = workspace_bytes;
    a.stream = stream;
    return vpg_lce_bwd_ex(&a);
}

int vpg_lce_bwd_mcred(const void* hidden, const void* weight, const int64_t* labels, const float* lse, const float* eb,
                      const float* g_lp, const float* g_ent, int64_t T, int64_t H, int64_t V, float temperature,
                      int64_t label_offset, int64_t chunk_v, void* d_hidden, void* d_weight, void* workspace,
                      size_t workspace_bytes, float* d_hidden_mc_f32, void* stream) {
    vpg_lce_bwd_args a{};
    a.struct_size = sizeof(a);
    a.dtype = VPG_DT_BF16, a.hidden = hidden, a.weight = weight, a.labels = labels, a.lse = lse, a.eb = eb;
    a.g_lp = g_lp, a.g_ent = g_ent, a.T = T, a.H = H, a.V = V, a.temperature = temperature;
    a.label_offset = label_offset, a.chunk_v = chunk_v, a.d_hidden = d_hidden, a.d_weight = d_weight;
    a.workspace = workspace, a.workspace_bytes = workspace_bytes, a.d_hidden_mc_f32 = d_hidden_mc_f32, a.stream = stream;
    return vpg_lce_bwd_ex(&a);
}

int vpg_lce_bwd_ex(const vpg_lce_bwd_args* args) {
    VPG_CHECK_ARG(args && args->struct_size == sizeof(vpg_lce_bwd_args), "vpg_lce_bwd_args: struct_size mismatch");
    const vpg_lce_bwd_args& a = *args;
    const int64_t T = a.T, H = a.H, V = a.V;
    int rc = check_lce_shapes(T, H, V);
    if (rc) return rc;
    const int dt = a.dtype;
    VPG_CHECK_ARG(dt == VPG_DT_BF16 || dt == VPG_DT_F16 || dt == VPG_DT_F32, "unknown operand dtype %d", dt);
    VPG_CHECK_ARG(a.hidden && a.weight && a.labels && a.lse && a.eb && (a.d_hidden || a.d_hidden_mc_f32) && a.d_weight,
                  "null pointer argument");
    VPG_CHECK_ARG(a.temperature > 0.f, "temperature must be > 0");
    cudaStream_t stream = static_cast<cudaStream_t>(a.stream);
    float* d_hidden_mc_f32 = a.d_hidden_mc_f32;
    void* dh_ready_event = d_hidden_mc_f32 ? nullptr : a.d_hidden_ready_event;  // the multimem.red variant has its own schedule
    const bool fused_ar = a.ar_world > 1;
    // operand type of the call.  fp32: the recompute runs as TF32 on the fp32 operands like the forward, dlogits are
    // bf16, the two gradient GEMMs read the caller's bf16 copies and write fp32 gradients.
    const bool f32 = (dt == VPG_DT_F32);
    const int dt16 = f32 ? VPG_DT_BF16 : dt;
    const void* hidden16 = f32 ? a.hidden16 : a.hidden;
    const void* weight16 = f32 ? a.weight16 : a.weight;
    const int64_t gsz = f32 ? 4 : 2;  // bytes per gradient element
    VPG_CHECK_ARG(!f32 || (hidden16 && weight16 && !d_hidden_mc_f32), "fp32 operands need the bf16 copies (hidden16, weight16)");
    VPG_CHECK_ARG(dt == VPG_DT_BF16 || !d_hidden_mc_f32, "the fused d_hidden reduction takes bf16 operands only");
    if (fused_ar) {
        VPG_CHECK_ARG(dt == VPG_DT_BF16 && !d_hidden_mc_f32 && !dh_ready_event,
                      "the in-kernel all-reduce takes bf16 operands and excludes the other d_hidden reduction modes");
        VPG_CHECK_ARG(a.ar_world <= kMaxPeers && a.ar_rank >= 0 && a.ar_rank < a.ar_world && a.ar_d_hidden_mc &&
                          a.ar_flags && a.ar_peer_flags && a.d_hidden,
                      "in-kernel all-reduce: world in [2,%d], rank, multicast address, flags and peer flags required",
                      kMaxPeers);
        VPG_CHECK_ARG((reinterpret_cast<uintptr_t>(a.ar_d_hidden_mc) & 15) == 0 && (T * H) % 8 == 0,
                      "in-kernel all-reduce: the buffer must be a whole number of aligned 16-byte vectors");
    }
    if (T == 0) {
        VPG_CUDA(cudaMemsetAsync(a.d_weight, 0, static_cast<size_t>(V) * H * gsz, stream));
        return VPG_OK;
    }
    VPG_CHECK_ARG(a.chunk_v % kBlockN == 0 && a.chunk_t % (2 * kBlockM) == 0,
                  "chunk_v must be a multiple of %d and chunk_t of %d", kBlockN, 2 * kBlockM);
    const BwdPlan pl = plan_bwd(T, H, V, a.chunk_v, a.chunk_t);
    const int64_t v_pad = round_up(V, kBlockN);
    if (!a.workspace || a.workspace_bytes < pl.bytes)
        return set_error(VPG_ERR_WORKSPACE, "lce_bwd workspace too small: need %zu bytes, got %zu", pl.bytes,
                         a.workspace_bytes);
    VPG_CHECK_ARG((reinterpret_cast<uintptr_t>(a.workspace) & 255) == 0, "workspace must be 256-byte aligned");
    uint8_t* ws = reinterpret_cast<uint8_t*>(a.workspace);
    __nv_bfloat16* ds = reinterpret_cast<__nv_bfloat16*>(ws);
    float* acc_buf = reinterpret_cast<float*>(ws + pl.off_acc);  // fp32 d_hidden (VOCAB) or d_weight (TOKEN) accumulator
    float* splitk_scratch = reinterpret_cast<float*>(ws + pl.off_splitk);
    const bool by_token = pl.mode == BWD_TOKEN;
    const int num_chunks = pl.n_chunks;
    const int64_t ldds = by_token ? v_pad : pl.chunk;
    const int64_t esz = f32 ? 4 : 2;

    for (int ci = 0; ci < num_chunks; ++ci) {
        // this chunk's block of the [T, V] dlogits: rows [t0, t0 + tn), columns [v0, v0 + vc)
        const int64_t t0 = by_token ? static_cast<int64_t>(ci) * pl.chunk : 0;
        const int64_t tn = by_token ? ((T - t0) < pl.chunk ? (T - t0) : pl.chunk) : T;
        const int64_t v0 = by_token ? 0 : static_cast<int64_t>(ci) * pl.chunk;
        const int64_t vc = by_token ? V : ((V - v0) < pl.chunk ? (V - v0) : pl.chunk);
        const int64_t t_pad = round_up(tn, 2 * kBlockM);  // dlogits rows: multiple of the largest m tile (256)
        const int m_tiles_t = static_cast<int>(t_pad / tile_m());
        const uint8_t* h_chunk = reinterpret_cast<const uint8_t*>(a.hidden) + t0 * H * esz;
        const __nv_bfloat16* h16_chunk = reinterpret_cast<const __nv_bfloat16*>(hidden16) + t0 * H;
        const uint8_t* w_chunk = reinterpret_cast<const uint8_t*>(a.weight) + v0 * H * esz;
        const __nv_bfloat16* w16_chunk = reinterpret_cast<const __nv_bfloat16*>(weight16) + v0 * H;
        uint8_t* d_hidden = a.d_hidden ? reinterpret_cast<uint8_t*>(a.d_hidden) + t0 * H * gsz : nullptr;
        const int n_tiles_c = static_cast<int>(ceil_div(vc, kBlockN));
        const bool first = ci == 0, last = ci == num_chunks - 1;
        CUtensorMap ta, tb;
        {  // 1. recompute logits tile -> dlogits block (16-bit)
            TcParams p{};
            p.num_m_tiles = m_tiles_t;
            p.num_n_tiles = n_tiles_c;
            p.num_k_blocks = static_cast<int>(H / (f32 ? kBlockK / 2 : kBlockK));
            p.num_splits = choose_splits(m_tiles_t, n_tiles_c, num_workers(), H * esz);
            p.tiles_per_split = static_cast<int>(ceil_div(n_tiles_c, p.num_splits));
            p.M = static_cast<int>(tn);
            p.N = static_cast<int>(vc);
            p.scale_log2 = 1.4426950408889634f / a.temperature;
            p.inv_temp = 1.0f / a.temperature;
            p.labels = reinterpret_cast<const long long*>(a.labels) + t0;
            p.label_offset = a.label_offset + v0;
            p.lse = a.lse + t0;
            p.eb = a.eb + t0;
            p.g_lp = a.g_lp ? a.g_lp + t0 : nullptr;
            p.g_ent = a.g_ent ? a.g_ent + t0 : nullptr;
            p.ds = ds;
            p.ldds = ldds;
            if ((rc = make_operand_tmap(&ta, h_chunk, tn, H, H, false, kBlockM, dt))) return rc;
            if ((rc = make_operand_tmap(&tb, w_chunk, vc, H, H, false, b_tile_rows(), dt))) return rc;
            // dlogits store map: [t_pad][ldds] 16-bit, 32 x 32 boxes in the 64-byte-swizzle layout of the staging tiles
            CUtensorMap tc;
            if ((rc = make_tmap(&tc, ds, ldds, t_pad, ldds, 32, 32, dt16, CU_TENSOR_MAP_SWIZZLE_64B))) return rc;
            if ((rc = launch_tc_lmhead<MODE_DS>(ta, tb, p, dt, stream, &tc))) return rc;
        }
        const int64_t vc_cols = static_cast<int64_t>(n_tiles_c) * kBlockN;  // dlogits columns actually written
        // 2. d_weight[v0:v0+vc, :] (+)= dlogits^T . hidden[t0:t0+tn]      (M = vc, N = H, K = t_pad)
        auto run_d_weight = [&](int max_workers, bool with_allreduce) -> int {
            int rc = 0;
            CUtensorMap ta, tb;
            TcParams p{};
            set_gemm_tiling(p, vc, H);
            p.num_k_blocks = static_cast<int>(t_pad / kBlockK);
            uint8_t* dw = reinterpret_cast<uint8_t*>(a.d_weight) + v0 * H * gsz;
            p.ldc = H;
            if (!by_token || num_chunks == 1) {  // rows of this vocab chunk are final
                p.out = dw, p.out_mode = f32 ? OUT_F32 : OUT_BF16;
            } else if (f32) {  // fp32 gradients accumulate in place over the token chunks
                p.out = dw, p.out_mode = first ? OUT_F32 : OUT_F32_ACC;
                p.accum = reinterpret_cast<float*>(dw), p.ld_accum = H;
            } else if (first) {
                p.out = acc_buf, p.out_mode = OUT_F32;
            } else if (!last) {
                p.out = acc_buf, p.out_mode = OUT_F32_ACC;
            } else {
                p.out = dw, p.out_mode = OUT_BF16_FROM_ACC;
                p.accum = acc_buf, p.ld_accum = H;
            }
            p.f16 = (dt16 == VPG_DT_F16);
            p.max_workers = max_workers;
            plan_splitk(p, max_workers > 0 && max_workers < num_workers() ? max_workers : num_workers(), splitk_scratch);
            if (with_allreduce) {
                p.ar_world = a.ar_world, p.ar_rank = a.ar_rank, p.ar_epoch = a.ar_epoch;
                p.ar_mc = a.ar_d_hidden_mc;
                p.ar_nvec = T * H / 8;
                p.ar_flags = a.ar_flags;
                // budget of each wait for a peer inside the kernel (a peer that raised or diverged never sets its flag)
                static const int ar_timeout_ms = env_int("VPG_AR_TIMEOUT_MS", 60000);
                p.ar_timeout_ns = ar_timeout_ms > 0 ? static_cast<unsigned long long>(ar_timeout_ms) * 1000000ull : 0ull;
                for (int i = 0; i < a.ar_world; ++i) {
                    VPG_CHECK_ARG(a.ar_peer_flags[i], "ar_peer_flags[%d] is null", i);
                    p.ar_peer_flags[i] = a.ar_peer_flags[i];
                }
            }
            if ((rc = make_operand_tmap(&ta, ds, vc_cols, t_pad, ldds, true, kBlockM, dt16))) return rc;
            if ((rc = make_operand_tmap(&tb, h16_chunk, H, tn, H, true, kBlockN, dt16))) return rc;
            if ((rc = launch_tc<MODE_GEMM, true, true>(ta, tb, p, stream, TK_DW))) return rc;
            return VPG_OK;
        };
        // 3. d_hidden[t0:t0+tn] (+)= dlogits . W[v0:v0+vc, :]          (M = tn, N = H, K = vc)
        auto run_d_hidden = [&]() -> int {
            int rc = 0;
            CUtensorMap ta, tb;
            TcParams p{};
            set_gemm_tiling(p, tn, H, WIDE_DH);
            p.num_m_tiles = m_tiles_t;
            p.num_k_blocks = static_cast<int>(ceil_div(vc, kBlockK));
            p.ldc = H;
            bool splitk_ok = true;
            if (d_hidden_mc_f32 && (by_token || last)) {
                VPG_CHECK_ARG(H % 32 == 0, "fused d_hidden all-reduce needs H %% 32 == 0");
                // multimem.red variant: the epilogue adds this rank's fp32 sum (over all its vocab chunks) into every
                // rank's fp32 buffer through the multicast address, tile by tile
                p.out = d_hidden_mc_f32 + t0 * H, p.out_mode = OUT_F32_MCRED;
                p.accum = (!by_token && num_chunks > 1) ? acc_buf : nullptr, p.ld_accum = H;
                splitk_ok = false;
            } else if (by_token || num_chunks == 1) {  // rows of this token chunk are final
                p.out = d_hidden, p.out_mode = f32 ? OUT_F32 : OUT_BF16;
            } else if (f32) {  // fp32 gradients accumulate in place over the vocab chunks
                p.out = d_hidden, p.out_mode = first ? OUT_F32 : OUT_F32_ACC;
                p.accum = reinterpret_cast<float*>(d_hidden), p.ld_accum = H;
            } else if (first) {
                p.out = acc_buf, p.out_mode = OUT_F32;
            } else if (!last) {
                p.out = acc_buf, p.out_mode = OUT_F32_ACC;
            } else {
                p.out = d_hidden, p.out_mode = OUT_BF16_FROM_ACC;
                p.accum = acc_buf, p.ld_accum = H;
            }
            p.f16 = (dt16 == VPG_DT_F16);
            plan_splitk(p, num_workers(), splitk_ok ? splitk_scratch : nullptr);
            if ((rc = make_operand_tmap(&ta, ds, t_pad, vc_cols, ldds, false, kBlockM, dt16))) return rc;
            if ((rc = make_operand_tmap(&tb, w16_chunk, H, vc, H, true, kBlockN, dt16))) return rc;
            if ((rc = launch_tc<MODE_GEMM, false, true>(ta, tb, p, stream, TK_DH))) return rc;
            return VPG_OK;
        };
        if (last && fused_ar) {
            // d_hidden first (it is complete on every rank before that rank enters the next kernel); the d_weight GEMM
            // then carries the all-reduce of d_hidden in its spare warps, on all its SMs
            if ((rc = run_d_hidden())) return rc;
            if ((rc = run_d_weight(0, true))) return rc;
        } else if (last && dh_ready_event) {
            // d_hidden first; the event releases the caller's all-reduce, which then runs under the d_weight GEMM on
            // the SMs that GEMM can spare for free: the fewest workers with the same number of scheduling rounds
            if ((rc = run_d_hidden())) return rc;
            VPG_CUDA(cudaEventRecord(static_cast<cudaEvent_t>(dh_ready_event), stream));
            const int64_t tiles = ceil_div(vc, static_cast<int64_t>(tile_m())) * ceil_div(H, static_cast<int64_t>(gemm_tile_n(H)));
            const int64_t rounds = ceil_div(tiles, static_cast<int64_t>(num_workers()));
            int spare_cap = static_cast<int>(ceil_div(tiles, rounds));
            // at least this many CTA pairs are left to the collective (caller's choice; env override for tuning)
            const int reserve = tuning().reserve_pairs >= 0 ? tuning().reserve_pairs : a.reserve_cta_pairs;
            if (reserve > 0 && spare_cap > num_workers() - reserve) spare_cap = num_workers() - reserve;
            if ((rc = run_d_weight(spare_cap < 1 ? 1 : spare_cap, false))) return rc;
        } else {
            if ((rc = run_d_weight(0, false))) return rc;
            if ((rc = run_d_hidden())) return rc;
        }
    }
    return VPG_OK;
}

int vpg_cast_f32_bf16(const float* src, void* dst_bf16, int64_t n, void* stream) {
    VPG_CHECK_ARG(src && dst_bf16 && n >= 0, "invalid cast arguments");
    VPG_CHECK_ARG(((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst_bf16)) & 15) == 0,
                  "cast operands must be 16-byte aligned");
    if (n == 0) return VPG_OK;
    int64_t g = ceil_div(n, 256 * 8 * 4);
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    if (g > cap) g = cap;
    cast_f32_bf16_kernel<<<static_cast<unsigned>(g), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        src, static_cast<__nv_bfloat16*>(dst_bf16), n);
    VPG_LAUNCH_CHECK();
    return VPG_OK;
}

int vpg_gemm_bf16(const void* A, const void* B, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb,
                  int a_kmajor, int b_kmajor, void* C, int64_t ldc, int out_mode, float* accum_f32,
                  int64_t ld_accum, void* stream_) {
    return vpg_gemm_bf16_ws(A, B, M, N, K, lda, ldb, a_kmajor, b_kmajor, C, ldc, out_mode, accum_f32, ld_accum,
                            nullptr, 0, stream_);
}

size_t vpg_gemm_splitk_workspace_bytes(void) { return splitk_scratch_bytes(); }

int vpg_gemm_bf16_ws(const void* A, const void* B, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb,
                     int a_kmajor, int b_kmajor, void* C, int64_t ldc, int out_mode, float* accum_f32,
                     int64_t ld_accum, void* workspace, size_t workspace_bytes, void* stream_) {
    VPG_CHECK_ARG(A && B && C, "null pointer argument");
    VPG_CHECK_ARG(M > 0 && N > 0 && K > 0, "invalid GEMM shape");
    VPG_CHECK_ARG(out_mode >= 0 && out_mode <= 3, "invalid out_mode %d", out_mode);
    VPG_CHECK_ARG(ldc % 8 == 0, "ldc must be a multiple of 8");
    VPG_CHECK_ARG(out_mode != OUT_BF16_FROM_ACC || (accum_f32 && ld_accum % 4 == 0), "accum buffer required");
    VPG_CHECK_ARG(!workspace || (workspace_bytes >= splitk_scratch_bytes() &&
                                 (reinterpret_cast<uintptr_t>(workspace) & 15) == 0),
                  "split-K workspace too small or misaligned (vpg_gemm_splitk_workspace_bytes)");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    TcParams p{};
    set_gemm_tiling(p, M, N);
    p.num_k_blocks = static_cast<int>(ceil_div(K, kBlockK));
    p.out = C;
    p.ldc = ldc;
    p.out_mode = out_mode;
    p.accum = accum_f32;
    p.ld_accum = ld_accum;
    plan_splitk(p, num_workers(), static_cast<float*>(workspace));
    CUtensorMap ta, tb;
    int rc;
    if ((rc = make_operand_tmap(&ta, A, M, K, lda, !a_kmajor, kBlockM))) return rc;
    if ((rc = make_operand_tmap(&tb, B, N, K, ldb, !b_kmajor, b_tile_rows()))) return rc;
    if (a_kmajor && b_kmajor) return launch_tc<MODE_GEMM, false, false>(ta, tb, p, stream, TK_GEMM);
    if (a_kmajor && !b_kmajor) return launch_tc<MODE_GEMM, false, true>(ta, tb, p, stream, TK_GEMM);
    if (!a_kmajor && b_kmajor) return launch_tc<MODE_GEMM, true, false>(ta, tb, p, stream, TK_GEMM);
    return launch_tc<MODE_GEMM, true, true>(ta, tb, p, stream, TK_GEMM);
}

}  // extern "C"
