/*
 * vpg_b200.h — C ABI of libvpg_b200.so: the B200-native (sm_100a) per-token policy-gradient hot path of verl.
 *
 * The reference (rfahrn/verl) is pure Python and has no FFI; its boundary for this path is a set of Python
 * functions.  Each entry point below names the reference function it replaces (file:line under
 * /root/reference).  The Python mirror in verl_b200/ binds these with ctypes and re-exposes the reference names.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer owned by the caller (PyTorch's allocator); the library never allocates
 *    or frees device memory, never synchronises, and enqueues all work on `stream` (a cudaStream_t);
 *  - return value 0 = success, negative = error (vpg_last_error() gives the message); no exceptions, no exit;
 *  - "f32" buffers are float, "bf16" are __nv_bfloat16 bit patterns, "i64" int64_t, "i32" int32_t;
 *  - masks are int64 / int32 / uint8(bool) / float32 selected by `mask_dtype` (VPG_MASK_*): the reference passes
 *    int64 attention/response masks (verl/protocol.py), bool and float also occur in its tests.
 */
#ifndef VPG_B200_H_
#define VPG_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VPG_OK 0
#define VPG_ERR_INVALID (-1)
#define VPG_ERR_CUDA (-2)
#define VPG_ERR_WORKSPACE (-3)
#define VPG_ERR_UNSUPPORTED (-4)

/* mask / value element types */
/* 16-bit operand type of the lm-head kernels (*_dt entry points) */
#define VPG_DT_BF16 0
#define VPG_DT_F16 1
#define VPG_DT_F32 2 /* fp32 operands multiplied as TF32 (the reference's tl.dot on fp32 inputs does the same) */

#define VPG_MASK_I64 0
#define VPG_MASK_I32 1
#define VPG_MASK_U8 2
#define VPG_MASK_F32 3

/* agg_loss modes — verl/trainer/ppo/core_algos.py:686-719 */
#define VPG_AGG_TOKEN_MEAN 0
#define VPG_AGG_SEQ_MEAN_TOKEN_SUM 1
#define VPG_AGG_SEQ_MEAN_TOKEN_MEAN 2
#define VPG_AGG_SEQ_MEAN_TOKEN_SUM_NORM 3

/* kl_penalty kinds — verl/trainer/ppo/core_algos.py:1034-1069 */
#define VPG_KL_K1 0      /* "kl" / "k1"          */
#define VPG_KL_ABS 1     /* "abs"                */
#define VPG_KL_MSE 2     /* "mse" / "k2"         */
#define VPG_KL_LOW_VAR 3 /* "low_var_kl" / "k3"  */
#define VPG_KL_NONE (-1)

/* policy-gradient term of vpg_ppo_loss */
#define VPG_PG_CLIP 0 /* compute_policy_loss, core_algos.py:722-794 (clip + dual-clip) */
#define VPG_PG_GPG 1  /* compute_policy_loss_gpg, core_algos.py:797-815: -log_prob * advantages */

/* outcome-reward advantage estimators sharing GRPO's skeleton — verl/trainer/ppo/core_algos.py */
#define VPG_ADV_GRPO 0          /* :245-308 */
#define VPG_ADV_RLOO 1          /* :425-476 */
#define VPG_ADV_OPO 2           /* :479-530 */
#define VPG_ADV_GPG 3           /* :608-667 */
#define VPG_ADV_RFPP_BASELINE 4 /* :372-422 */
#define VPG_ADV_GRPO_PASSK 5    /* :309-369 */

const char* vpg_version(void);
/* last error message of the calling thread ("" if none) */
const char* vpg_last_error(void);
/* number of kernels this library has launched in the calling process (monotonic; for bench.py's gpu_launches) */
int64_t vpg_launch_count(void);
/* cudaStreamSynchronize(stream): the wait behind the host reads of token counts (unpad_input's nnz / max_seqlen — the
 * reference syncs at the same points with nonzero() and .item(), flash_attn bert_padding / torch_functional.py:629-638) */
int vpg_stream_synchronize(void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Fused lm-head -> per-token log-prob + entropy.
 * Replaces verl/utils/kernel/kernels.py:474-663 (efficient_entropy_forward: Triton mainloop :120-254 + epilogue
 * :257-346) behind verl/utils/kernel/linear_cross_entropy.py:42-87.
 *
 * Stage 1 (vpg_lce_fwd_partials): tcgen05/TMEM GEMM hidden[T,H] x weight[V,H]^T fed by TMA with the online
 *   softmax statistics fused in the epilogue; the [T,V] logits are never written.  Per token and per vocab split
 *   s it writes 4 floats into `partials` laid out [4][num_splits][T]:
 *     plane 0: m  = max_v  (x_v * log2(e)/temperature)           (log2 domain)
 *     plane 1: s  = sum_v  2^(x_v*log2e/temp - m)
 *     plane 2: b  = sum_v  x_v * 2^(x_v*log2e/temp - m)           (x = raw dot product)
 *     plane 3: t  = x_{label}  if the label falls in this split (after subtracting label_offset) else 0
 * Stage 2 (vpg_lce_finalize): merges `num_parts` partial sets (splits, and for vocab-parallel runs the gathered
 *   partials of all ranks) into logprobs[T], entropy[T] and the two per-token scalars the backward needs:
 *   lse[T] (natural-log log-sum-exp of logits/temperature) and eb[T] (= sum_v p_v * logit_v).
 *
 * hidden: bf16 [T,H] row-major, weight: bf16 [V,H] row-major (nn.Linear layout), labels: i64 [T].
 * H % 64 == 0 is required (the reference requires H % 128 == 0, kernels.py:504).
 * label_offset = rank * V_local for vocab-parallel shards (kernels.py:232), 0 otherwise.
 * ------------------------------------------------------------------------------------------------------------ */
int vpg_lce_num_splits(int64_t T, int64_t H, int64_t V, int sm_count);
int vpg_lce_fwd_partials(const void* hidden_bf16, const void* weight_bf16, const int64_t* labels, int64_t T,
                         int64_t H, int64_t V, float temperature, int64_t label_offset, int num_splits,
                         float* partials /* [4][num_splits][T] */, void* stream);
int vpg_lce_finalize(const float* partials /* [4][num_parts][T] */, int num_parts, int64_t T, float temperature,
                     float* logprobs, float* entropy, float* lse, float* eb, void* stream);
/* Same as vpg_lce_fwd_partials for operands of type `dtype` (VPG_DT_BF16 | VPG_DT_F16 | VPG_DT_F32): the reference
 * kernel takes hidden / weight in any float dtype (kernels.py:485-490); bf16 and fp16 run on tcgen05 kind::f16, fp32 on
 * kind::tf32 (half the tensor rate). */
int vpg_lce_fwd_partials_dt(int dtype, const void* hidden, const void* weight, const int64_t* labels, int64_t T,
                            int64_t H, int64_t V, float temperature, int64_t label_offset, int num_splits,
                            float* partials /* [4][num_splits][T] */, void* stream);

/* Vocab-parallel "push" variant of stage 1 — replaces the exchange of verl/utils/kernel/kernels.py:609-661 (all_reduce
 * MAX + SUM + SUM of per-token partials).  The GEMM epilogue writes this rank's partials straight into the GATHERED
 * layout [4][part_total][T] at parts [part_split0, part_split0 + num_splits) (part_split0 = rank * num_splits,
 * part_total = world * num_splits) of EVERY rank, tile by tile, over NVLink:
 *   partials_is_multicast != 0: `partials` is an NVSwitch multicast address bound to all ranks' buffers (multimem.st);
 *   n_push > 0: `push_ptrs[0..n_push)` (HOST array of device pointers) are the peer-mapped buffers of all ranks
 *               (this rank's own included); plain NVLink stores;
 *   neither: plain local store (the single-GPU entry point above is this with part_split0 = 0).
 * The caller owns the cross-rank barrier between this launch and vpg_lce_finalize(num_parts = part_total). */
int vpg_lce_fwd_partials_push(const void* hidden_bf16, const void* weight_bf16, const int64_t* labels, int64_t T,
                              int64_t H, int64_t V, float temperature, int64_t label_offset, int num_splits,
                              float* partials, int part_split0, int part_total, int partials_is_multicast,
                              void* const* push_ptrs, int n_push, void* stream);

/* Backward.  Replaces verl/utils/kernel/kernels.py:1345-1553 (efficient_entropy_backward; default strategy
 * _Split_Dlogits_N :1230-1342 + two cuBLAS GEMMs :1538-1548).
 *   dlogits[t,v] = ( g_lp[t]*(1[v==label_t] - p) - g_ent[t]*p*(logit - eb[t]) ) / temperature
 *   d_hidden = dlogits . W        d_weight = dlogits^T . hidden
 * The vocab range is processed in chunks of `chunk_v` columns (multiple of 256): recompute tile -> dlogits chunk
 * (bf16 workspace, never the logits) -> two tcgen05 GEMMs.  d_hidden is accumulated in fp32 across chunks.
 * Workspace size: vpg_lce_bwd_workspace_bytes().  g_lp / g_ent may be NULL (treated as zeros).
 * d_hidden: bf16 [T,H]; d_weight: bf16 [V,H].  For vocab-parallel shards the caller all-reduces d_hidden. */
size_t vpg_lce_bwd_workspace_bytes(int64_t T, int64_t H, int64_t V, int64_t chunk_v);
int64_t vpg_lce_bwd_default_chunk(int64_t T, int64_t H, int64_t V); /* budget: $VPG_BWD_WS_MB (default 3072) MiB */
/* Largest chunk whose dlogits buffer fits `budget_bytes`, then the chunk count (up to +8) that wastes the fewest waves
 * of the persistent grid.  Fewer chunks = fewer launches and fp32 d_hidden round trips: one 20 GB chunk at configs[1]
 * is 2.5 % faster than seven 2.85 GB chunks. */
int64_t vpg_lce_bwd_chunk_for_budget(int64_t T, int64_t H, int64_t V, int64_t budget_bytes);
int vpg_lce_bwd(const void* hidden_bf16, const void* weight_bf16, const int64_t* labels, const float* lse,
                const float* eb, const float* g_lp, const float* g_ent, int64_t T, int64_t H, int64_t V,
                float temperature, int64_t label_offset, int64_t chunk_v, void* d_hidden_bf16, void* d_weight_bf16,
                void* workspace, size_t workspace_bytes, void* stream);
/* Same as vpg_lce_bwd for 16-bit operands of type `dtype`: the dlogits workspace, d_hidden and d_weight are written in
 * that type (the reference returns gradients in hidden.dtype, kernels.py:1393-1394). */
int vpg_lce_bwd_dt(int dtype, const void* hidden, const void* weight, const int64_t* labels, const float* lse,
                   const float* eb, const float* g_lp, const float* g_ent, int64_t T, int64_t H, int64_t V,
                   float temperature, int64_t label_offset, int64_t chunk_v, void* d_hidden, void* d_weight,
                   void* workspace, size_t workspace_bytes, void* stream);

/* fp32 operands: the logits recompute runs as TF32 on the fp32 operands (consistent with the forward), dlogits are
 * bf16, and the two gradient GEMMs read bf16 copies of hidden / weight supplied by the caller; d_hidden [T,H] and
 * d_weight [V,H] are written in fp32. */
int vpg_lce_bwd_f32(const void* hidden_f32, const void* weight_f32, const void* hidden_bf16, const void* weight_bf16,
                    const int64_t* labels, const float* lse, const float* eb, const float* g_lp, const float* g_ent,
                    int64_t T, int64_t H, int64_t V, float temperature, int64_t label_offset, int64_t chunk_v,
                    void* d_hidden_f32, void* d_weight_f32, void* workspace, size_t workspace_bytes, void* stream);

/* Backward with every option as an explicit argument (the entry points above are thin wrappers over it).
 * Replaces verl/utils/kernel/kernels.py:1345-1553 plus, for vocab-parallel shards, the d_hidden all-reduce the
 * reference leaves to its callers (tests/utils/test_linear_cross_entropy_tp.py:313-314).  Zero-initialise the struct,
 * set struct_size = sizeof(vpg_lce_bwd_args), fill what applies.  Three mutually exclusive ways to reduce d_hidden
 * across the vocab-parallel group, all optional:
 *  (a) ar_world > 1 — IN-KERNEL NVLS all-reduce: d_hidden must be this rank's address of a symmetric (peer-mapped,
 *      multicast-bound) bf16 [T,H] buffer and ar_d_hidden_mc the multicast address of the same buffer.  The backward
 *      computes d_hidden BEFORE its last d_weight GEMM; that GEMM's otherwise idle warp (one per CTA) then pulls the
 *      switch-reduced sum of this rank's 1/world slice (multimem.ld_reduce) and broadcasts it into every rank's buffer
 *      (multimem.st) while the tensor cores work on d_weight.  ar_flags: this rank's 18 uint32 flags (zeroed once at
 *      allocation) in symmetric memory; ar_peer_flags[r]: rank r's flag array as mapped into this process (HOST array
 *      of ar_world device pointers, own rank included); ar_epoch: call counter, identical on all ranks, in
 *      [1, 2^31) (the top bit marks a poisoned flag).
 *      When the call's kernels have completed on a rank and ar_flags[17] is still 0, its d_hidden holds the sum over
 *      all ranks.  Every wait for a peer inside the kernel has a budget ($VPG_AR_TIMEOUT_MS, default 60000, 0 = none):
 *      a rank that gives up on a peer (it never launched the matching call) skips its share of the reduction, stores
 *      ar_epoch in ar_flags[17] and poisons its done flag so that late peers fail too instead of returning a partial
 *      sum.  The caller reads ar_flags[17] after the call (non-zero = that call's d_hidden is not the sum).
 *  (b) d_hidden_ready_event (cudaEvent_t) — recorded as soon as d_hidden is complete; d_hidden is computed before the
 *      last d_weight GEMM, which runs on the fewest persistent workers that keep its number of scheduling rounds
 *      (at least reserve_cta_pairs CTA pairs stay free), so the caller's own all-reduce (NCCL on a second stream)
 *      overlaps it.
 *  (c) d_hidden_mc_f32 — multimem.red.add.f32 in the epilogue of the d_hidden GEMM (see vpg_lce_bwd_mcred). */
typedef struct vpg_lce_bwd_args {
    size_t struct_size;
    int dtype;               /* VPG_DT_BF16 | VPG_DT_F16 | VPG_DT_F32 */
    const void* hidden;      /* [T,H] in `dtype` */
    const void* weight;      /* [V,H] in `dtype` */
    const void* hidden16;    /* VPG_DT_F32 only: bf16 copies read by the two gradient GEMMs */
    const void* weight16;
    const int64_t* labels;
    const float* lse;
    const float* eb;
    const float* g_lp;       /* may be NULL */
    const float* g_ent;      /* may be NULL */
    int64_t T, H, V;
    float temperature;
    int64_t label_offset;
    int64_t chunk_v;         /* > 0: dlogits in column chunks of this width (multiple of 256) */
    int64_t chunk_t;         /* > 0 (and chunk_v == 0): dlogits in row chunks of this height (multiple of 256);
                                both 0: the library plans (vpg_lce_bwd_plan) */
    void* d_hidden;          /* [T,H] in `dtype` (fp32 for VPG_DT_F32) */
    void* d_weight;          /* [V,H] likewise */
    void* workspace;
    size_t workspace_bytes;  /* >= vpg_lce_bwd_workspace_bytes() */
    void* stream;
    float* d_hidden_mc_f32;        /* (c) */
    void* d_hidden_ready_event;    /* (b) */
    int reserve_cta_pairs;         /* (b) */
    int ar_world, ar_rank;         /* (a) */
    uint32_t ar_epoch;
    void* ar_d_hidden_mc;
    uint32_t* ar_flags;
    uint32_t* const* ar_peer_flags;
} vpg_lce_bwd_args;
int vpg_lce_bwd_ex(const vpg_lce_bwd_args* args);
/* The chunking the backward will use for (chunk_v, chunk_t) as passed to vpg_lce_bwd_ex:
 * out4 = {mode (0 one block, 1 column chunks with an fp32 [T,H] d_hidden accumulator — the reference's _Split_Dlogits_N,
 * 2 row chunks with an fp32 [V,H] d_weight accumulator — its _Split_Dlogits_M), chunk width / height, number of chunks,
 * workspace bytes}.  With both 0 the dlogits block is bounded by $VPG_BWD_WS_MB (default 3072 MiB) and the orientation
 * with the smaller accumulator is taken (row chunks when T > V, e.g. a vocab shard of a long micro-batch). */
int vpg_lce_bwd_plan(int64_t T, int64_t H, int64_t V, int64_t chunk_v, int64_t chunk_t, int64_t* out4);

/* Forward stage 1 with every option explicit: operand dtype + the push arguments of vpg_lce_fwd_partials_push. */
int vpg_lce_fwd_partials_ex(int dtype, const void* hidden, const void* weight, const int64_t* labels, int64_t T,
                            int64_t H, int64_t V, float temperature, int64_t label_offset, int num_splits,
                            float* partials, int part_split0, int part_total, int partials_is_multicast,
                            void* const* push_ptrs, int n_push, void* stream);

/* Per-launch device timing of the tcgen05 kernels (bench.py's per-kernel roofline): while enabled, every launch is
 * bracketed by two CUDA events on the stream it is launched on.  vpg_timing_collect synchronises those events and
 * returns up to max_n (kind, milliseconds) pairs in launch order, clearing the record.
 * kind: 0 lm-head forward, 1 dlogits recompute, 2 d_weight GEMM, 3 d_hidden GEMM, 4 plain GEMM, 5 split-K reduce. */
int vpg_timing_enable(int on);
int vpg_timing_collect(int* kinds, float* ms, int max_n);

/* Vocab-parallel backward with the d_hidden all-reduce fused into the d_hidden GEMM (the reference leaves this
 * all-reduce to the caller: tests/utils/test_linear_cross_entropy_tp.py:313-314).  d_hidden_mc_f32 is an NVSwitch
 * multicast address bound to an fp32 [T,H] buffer on every rank, zeroed on all ranks (barrier) before the call: the
 * epilogue of the last vocab chunk's d_hidden GEMM issues multimem.red.add.f32 per tile, so the sum over ranks builds
 * up in every rank's buffer while the GEMM runs.  After a barrier each rank converts its copy with
 * vpg_cast_f32_bf16.  d_hidden_mc_f32 == NULL behaves exactly like vpg_lce_bwd (d_hidden may then not be NULL). */
int vpg_lce_bwd_mcred(const void* hidden_bf16, const void* weight_bf16, const int64_t* labels, const float* lse,
                      const float* eb, const float* g_lp, const float* g_ent, int64_t T, int64_t H, int64_t V,
                      float temperature, int64_t label_offset, int64_t chunk_v, void* d_hidden_bf16,
                      void* d_weight_bf16, void* workspace, size_t workspace_bytes, float* d_hidden_mc_f32,
                      void* stream);
int vpg_cast_f32_bf16(const float* src, void* dst_bf16, int64_t n, void* stream);

/* Plain bf16 GEMM on the same tcgen05 mainloop (used by the backward; exported for tests):
 *   C[M,N] (bf16 or f32) = A . B with
 *   a_kmajor=1: A is [M,K] row-major, else A is [K,M] row-major (i.e. A^T stored);
 *   b_kmajor=1: B is [N,K] row-major, else B is [K,N] row-major.
 *   out_mode 0: store bf16, 1: store f32, 2: f32 += , 3: bf16 = f32_accum_in + acc (accum_in given separately) */
int vpg_gemm_bf16(const void* A, const void* B, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb,
                  int a_kmajor, int b_kmajor, void* C, int64_t ldc, int out_mode, float* accum_f32,
                  int64_t ld_accum, void* stream);
/* Same with a scratch buffer (vpg_gemm_splitk_workspace_bytes(), 16-byte aligned) that enables the split-K tail: the
 * tiles of the last, partial scheduling round of the persistent grid are cut along K into fp32 partial tiles, summed in
 * a fixed order by a second small launch, so that round costs 1/ksplit of a tile time. */
size_t vpg_gemm_splitk_workspace_bytes(void);
int vpg_gemm_bf16_ws(const void* A, const void* B, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb,
                     int a_kmajor, int b_kmajor, void* C, int64_t ldc, int out_mode, float* accum_f32,
                     int64_t ld_accum, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Materialised-logits variants — verl/utils/torch_functional.py:64-133 (logprobs_from_logits*) and :145-160
 * (entropy_from_logits*).  logits: [T,V] f32 or bf16 (logits_bf16 != 0), row stride ld (elements).
 * One pass over each row: online log-sum-exp, label gather and sum p*logit.  Either output may be NULL.
 * ------------------------------------------------------------------------------------------------------------ */
int vpg_logits_stats(const void* logits, int logits_bf16, int64_t T, int64_t V, int64_t ld, const int64_t* labels,
                     float* logprobs, float* entropy, float* lse, void* stream);
/* Vocab-sharded materialised logits — verl/utils/megatron/tensor_parallel.py:109-158 (vocab_parallel_entropy,
 * vocab_parallel_log_probs_from_logits).  logits: this rank's [T, V_local] shard (f32 or bf16); a label hits the shard
 * when 0 <= label - label_offset < V_local (label_offset = rank * V_local; labels may be NULL).  Writes this rank's
 * per-token partials [4][T] f32 in the fused lm-head's wire format {max * log2(e), sum exp(x-max), sum x*exp(x-max),
 * label logit or 0}: ONE all-gather of them followed by vpg_lce_finalize(temperature 1) replaces the reference's
 * three dependent all-reduces (MAX, SUM, SUM). */
int vpg_logits_partials(const void* logits, int logits_bf16, int64_t T, int64_t V_local, int64_t ld,
                        const int64_t* labels, int64_t label_offset, float* partials, void* stream);
/* backward of the above: dlogits = g_lp*(onehot - p) - g_ent*p*(logit - eb), with eb = lse - entropy */
int vpg_logits_stats_bwd(const void* logits, int logits_bf16, int64_t T, int64_t V, int64_t ld,
                         const int64_t* labels, const float* lse, const float* entropy, const float* g_lp,
                         const float* g_ent, void* dlogits, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * rmpad index plumbing — flash_attn.bert_padding (unpad_input / index_first_axis / pad_input), third-party file
 * used at verl/workers/actor/dp_actor.py:110-125,221-232, and verl/utils/torch_functional.py:629-638
 * (get_unpad_data).  Integer outputs are bit-exact.
 *   mask [B,S] -> indices i64 [nnz] (ascending flat positions of non-zero mask), seqlens i32 [B],
 *   cu_seqlens i32 [B+1], totals i32 [2] = {nnz, max_seqlen} (write-only for the kernels: device memory or a
 *   pinned, UVA-mapped host pair the caller reads after a stream synchronize).
 *   workspace: vpg_unpad_workspace_bytes(B*S).
 * ------------------------------------------------------------------------------------------------------------ */
size_t vpg_unpad_workspace_bytes(int64_t n);
int vpg_unpad_indices(const void* mask, int mask_dtype, int64_t B, int64_t S, int64_t* indices, int32_t* seqlens,
                      int32_t* cu_seqlens, int32_t* totals, void* workspace, size_t workspace_bytes, void* stream);
/* Response-window selection on the rmpad token stream — folds verl/workers/actor/dp_actor.py:135 (roll of the ids),
 * :221-232 (pad back) and :236-237 (slice [:, -L-1:-1]) into index arithmetic, so the lm-head only has to run on the
 * rows whose output survives the slice.  indices [T] = padded flat positions (vpg_unpad_indices); per stream token:
 * keep[i] = 1 if its column s satisfies S-L-1 <= s < S-1; dst[i] = b*L + s-(S-L-1) (row of the [B*L] response
 * matrix, -1 if dropped); labels[i] = ids_rmpad[(i+1) % T] (labels / ids_rmpad may be NULL). */
int vpg_response_rows(const int64_t* indices, const int64_t* ids_rmpad, int64_t T, int64_t S, int64_t L,
                      uint8_t* keep, int64_t* dst, int64_t* labels, void* stream);
/* out[i, :] = src[indices[i], :]  rows of row_bytes bytes (index_first_axis) */
int vpg_gather_rows(const void* src, const int64_t* indices, int64_t n_idx, int64_t row_bytes, void* out,
                    void* stream);
/* out[indices[i], :] = src[i, :]; `out` has n_out rows (pad_input / index_put_first_axis, flash-attn bert_padding).
 * zero_fill 0: the other rows are left as they are; 1: `out` is zero-filled first (memset + scatter, any index order);
 * 2: the caller guarantees STRICTLY ASCENDING indices (what vpg_unpad_indices produces) and ONE pass writes every output
 * row exactly once — scattered rows and the zero gaps between them — with no memset and no second write. */
int vpg_scatter_rows(const void* src, const int64_t* indices, int64_t n_idx, int64_t row_bytes, void* out,
                     int64_t n_out, int zero_fill, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Token ops — verl/trainer/ppo/core_algos.py and verl/utils/torch_functional.py.
 * All value tensors are f32 [B,L] contiguous; mask as described above.
 * ------------------------------------------------------------------------------------------------------------ */

/* masked_sum/masked_mean building block (torch_functional.py:163-185): out[0] = sum(where(m,v,0)*m),
 * out[1] = sum(m).  `out` (2 doubles) is zeroed by the call.  axis=None only. */
int vpg_masked_sum(const float* values, const void* mask, int mask_dtype, int64_t n, double* out, void* stream);
/* per-row variant: out_sum[b], out_cnt[b] (float) — used for axis=-1 and the seq-mean agg modes */
int vpg_masked_row_sum(const float* values, const void* mask, int mask_dtype, int64_t B, int64_t L, float* out_sum,
                       float* out_cnt, void* stream);

/* kl_penalty elementwise forward (core_algos.py:1034-1069); grad_out (may be NULL) receives d kl / d logprob */
int vpg_kl_penalty(const float* logprob, const float* ref_logprob, int kind, int64_t n, float* out, float* grad_out,
                   void* stream);

/* Fused actor loss, forward + closed-form gradient in one pass:
 *   compute_policy_loss (core_algos.py:722-794) + entropy bonus + KL loss + agg_loss as composed in
 *   verl/workers/actor/dp_actor.py:422-469:
 *     policy_loss = agg(pg) - entropy_coeff*agg(entropy) + kl_coef*agg(kl_penalty(lp, ref))   , times loss_scale
 *   ratio_lo = (float)(1 - cliprange_low), ratio_hi = (float)(1 + cliprange_high): the clamp bounds, rounded to
 *   fp32 by the caller exactly as torch.clamp rounds its Python-float bounds.
 *   entropy / ref_log_prob may be NULL (coeffs ignored).
 *   metrics[8] (f32, device): {policy_loss, pg_loss, pg_clipfrac, ppo_kl, pg_clipfrac_lower, entropy_agg, kl_agg,
 *                             mask_sum}
 *   d_log_prob / d_entropy (may be NULL): gradient of policy_loss (incl. loss_scale) wrt log_prob / entropy.
 *   pg_kind = VPG_PG_GPG replaces the clipped term by -log_prob*advantages (compute_policy_loss_gpg,
 *   core_algos.py:797-815; the three clip statistics are 0 then, as in the reference).
 *   workspace: vpg_ppo_loss_workspace_bytes(B, L). */
size_t vpg_ppo_loss_workspace_bytes(int64_t B, int64_t L);
int vpg_ppo_loss(const float* old_log_prob, const float* log_prob, const float* advantages, const void* mask,
                 int mask_dtype, const float* entropy, const float* ref_log_prob, int64_t B, int64_t L,
                 float ratio_lo, float ratio_hi, float clip_c, int pg_kind, int agg_mode, float entropy_coeff,
                 float kl_coef, int kl_kind, float loss_scale, float* metrics, float* d_log_prob, float* d_entropy,
                 void* workspace, size_t workspace_bytes, void* stream);

/* agg_loss forward for an arbitrary loss matrix (core_algos.py:686-719); out[0] = loss; d_loss_mat (may be NULL)
 * receives d loss / d loss_mat.  workspace as vpg_ppo_loss_workspace_bytes. */
int vpg_agg_loss(const float* loss_mat, const void* mask, int mask_dtype, int64_t B, int64_t L, int agg_mode,
                 float* out, float* d_loss_mat, void* workspace, size_t workspace_bytes, void* stream);

/* compute_value_loss (core_algos.py:992-1031): out[0] = vf_loss = 0.5*agg(max((vp-ret)^2, (clip(vp)-ret)^2)),
 * out[1] = vf_clipfrac; d_vpreds (may be NULL) receives d vf_loss / d vpreds.  workspace as vpg_ppo_loss. */
int vpg_value_loss(const float* vpreds, const float* returns, const float* values, const void* mask, int mask_dtype,
                   int64_t B, int64_t L, float cliprange_value, int agg_mode, float* out, float* d_vpreds,
                   void* workspace, size_t workspace_bytes, void* stream);

/* compute_grpo_outcome_advantage (core_algos.py:245-308): scores = row-sum(rewards); per group (group_ids i32 [B],
 * values in [0,num_groups), built on the host from the uid strings in first-seen order) mean and unbiased std,
 * singleton -> (0,1); adv = (score-mean)/(std+eps) (or score-mean) broadcast * mask.
 * workspace: vpg_grpo_workspace_bytes(B, num_groups). */
size_t vpg_grpo_workspace_bytes(int64_t B, int64_t num_groups);
int vpg_grpo_advantage(const float* token_level_rewards, const void* mask, int mask_dtype, const int32_t* group_ids,
                       int64_t B, int64_t L, int64_t num_groups, float epsilon, int norm_by_std, float* advantages,
                       void* workspace, size_t workspace_bytes, void* stream);

/* vpg_grpo_advantage without a workspace argument.  No counterpart in the reference: it exists for hosts that cannot
 * bring scratch memory to the call — a C / ctypes host without PyTorch's caching allocator (its only source of scratch
 * would be cudaMalloc / cudaFree around the call, which stall every stream on the device, the colocated rollout engine's
 * too), or a trainer whose GPU memory was handed wholesale to another pool so that an allocation in mid-step fails.
 * Inputs and the `advantages` output are still the caller's; results are those of vpg_grpo_advantage.
 *  - The scratch is a fixed array inside the library image (global memory, 576.25 KiB, 256-byte aligned).  Every GPU on
 *    which the library gets loaded carries it until the process exits, whether or not this function is ever called.
 *  - Limit: B <= VPG_GRPO_STATIC_MAX_ROWS (any L, any G <= B).  A larger B returns VPG_ERR_WORKSPACE and launches
 *    nothing; use vpg_grpo_advantage for such batches.
 *  - The first call on each device (made with that device current) looks the array's device address up through
 *    cudaGetSymbolAddress and caches it.  Make that first call during start-up, before any stream capture is begun.
 *  - The call itself allocates and frees nothing: a memset of the bytes it uses and three kernel launches on `stream`.
 *  - One array per GPU, shared by all calls, holding whatever the last call left behind.  Each call clears the part it
 *    needs first, so back-to-back calls on one stream are independent.  Calls that may overlap on the GPU (two streams,
 *    two host threads) race on the array: order them yourself (events / one stream). */
#define VPG_GRPO_STATIC_MAX_ROWS 16384
int vpg_grpo_advantage_static(const float* token_level_rewards, const void* mask, int mask_dtype,
                              const int32_t* group_ids, int64_t B, int64_t L, int64_t num_groups, float epsilon,
                              int norm_by_std, float* advantages, void* stream);

/* The other outcome-reward estimators with GRPO's skeleton (kind = VPG_ADV_*): RLOO, OPO, GPG,
 * REINFORCE++-baseline (followed by masked_whiten over the broadcast matrix, binary masks), GRPO-pass@k.
 * f_norm is GPG's normaliser.  status[0] (i32, device): 0 ok, 1 / 2 = mask sum 0 / 1 (REINFORCE++-baseline's
 * masked_whiten raises ValueError), 3 = a pass@k group with fewer than 2 samples (ValueError in the reference).
 * workspace: vpg_outcome_advantage_workspace_bytes(B, num_groups). */
size_t vpg_outcome_advantage_workspace_bytes(int64_t B, int64_t num_groups);
int vpg_outcome_advantage(int kind, const float* token_level_rewards, const void* mask, int mask_dtype,
                          const int32_t* group_ids, int64_t B, int64_t L, int64_t num_groups, float epsilon,
                          int norm_by_std, float f_norm, float* advantages, int32_t* status, void* workspace,
                          size_t workspace_bytes, void* stream);

/* compute_reinforce_plus_plus_outcome_advantage (core_algos.py:533-569): discounted reverse return with reset after
 * masked-out tokens, advantages = masked_whiten(returns) * mask.  status / status_event as vpg_gae.
 * compute_remax_outcome_advantage (core_algos.py:572-605): returns = reverse cumsum of rewards*mask,
 * advantages = returns - reward_baselines[b] * mask. */
size_t vpg_reverse_scan_workspace_bytes(int64_t B, int64_t L);
int vpg_reinforce_plus_plus(const float* token_level_rewards, const void* mask, int mask_dtype, int64_t B, int64_t L,
                            float gamma, float* advantages, float* returns, int32_t* status, void* workspace,
                            size_t workspace_bytes, void* stream, void* status_event);
int vpg_remax(const float* token_level_rewards, const float* reward_baselines, const void* mask, int mask_dtype,
              int64_t B, int64_t L, float* advantages, float* returns, void* stream);

/* apply_kl_penalty (verl/trainer/ppo/ray_trainer.py:153-193): kld = kl_penalty(old_log_probs, ref_log_prob) * mask;
 * token_level_rewards = token_level_scores - beta * kld; current_kl[0] = mean over rows of masked_mean(kld, mask, -1).
 * workspace: vpg_kl_reward_workspace_bytes(B). */
size_t vpg_kl_reward_workspace_bytes(int64_t B);
int vpg_kl_reward(const float* token_level_scores, const float* old_log_probs, const float* ref_log_prob,
                  const void* mask, int mask_dtype, int kind, float beta, int64_t B, int64_t L,
                  float* token_level_rewards, float* current_kl, void* workspace, size_t workspace_bytes, void* stream);

/* compute_data_metrics (verl/trainer/ppo/metric_utils.py:80-180) in one sweep.  values may be NULL (use_critic=False).
 * attention_mask is [B,S] (prompt columns [0,S-L), response columns [S-L,S)); response_mask [B,L] is used as a
 * boolean.  out26 (f32, device): score mean/max/min [0..2], rewards [3..5], advantages [6..8], returns [9..11],
 * values [12..14], vf_explained_var [15], response_length mean/max/min/clip_ratio [16..19], prompt_length
 * mean/max/min/clip_ratio [20..23], number of valid tokens [24], reserved [25]. */
size_t vpg_data_metrics_workspace_bytes(int64_t B);
int vpg_data_metrics(const float* token_level_scores, const float* token_level_rewards, const float* advantages,
                     const float* returns, const float* values, const void* response_mask, int rmask_dtype,
                     const void* attention_mask, int amask_dtype, int64_t B, int64_t L, int64_t S, float* out26,
                     void* workspace, size_t workspace_bytes, void* stream);

/* compute_gae_advantage_return (core_algos.py:193-241): reverse recurrence per row with masked carry-through,
 * returns = adv + values, then masked_whiten (torch_functional.py:206-223) over the whole batch.
 * status[0] (i32) = 0 ok, 1 mask sum == 0, 2 mask sum == 1 (the reference raises ValueError there).  The word is
 * written by the last block of the recurrence / statistics kernel, i.e. BEFORE the normalisation pass starts; it may
 * live in device memory or in pinned host memory (UVA-mapped).  status_event (cudaEvent_t, may be NULL) is recorded
 * right behind that kernel, so a host that waits on it learns the status while the normalisation pass is still running.
 * workspace: vpg_gae_workspace_bytes(B, L). */
size_t vpg_gae_workspace_bytes(int64_t B, int64_t L);
int vpg_gae(const float* token_level_rewards, const float* values, const void* mask, int mask_dtype, int64_t B,
            int64_t L, float gamma, float lam, float* advantages, float* returns, int32_t* status, void* workspace,
            size_t workspace_bytes, void* stream, void* status_event);

/* masked_whiten (torch_functional.py:206-223) over all elements. status / status_event as above. */
int vpg_masked_whiten(const float* values, const void* mask, int mask_dtype, int64_t n, int shift_mean, float* out,
                      int32_t* status, void* workspace, size_t workspace_bytes, void* stream, void* status_event);

/* ---- dynamic micro-batches (verl/utils/seqlen_balancing.py) ------------------------------------------------------
 * attention_mask.sum(dim=1) (:282) as exact integers: lengths[b] (i64, device) = sum_s mask[b,s]. */
int vpg_row_lengths(const void* mask, int mask_dtype, int64_t B, int64_t S, int64_t* lengths, void* stream);
/* get_seqlen_balanced_partitions / karmarkar_karp (seqlen_balancing.py:26-190) — HOST scheduler, the one entry point
 * whose pointers are HOST memory (no device work, no stream): partitions the n sequence lengths into k sets with
 * the largest-differencing method and the reference's exact tie-breaking (heapq order included), so that the
 * partition lists are bit-identical.  part_indices [n] holds the partitions back to back, each sorted ascending;
 * part_offsets [k+1] their bounds.  Errors mirror the reference's assertions (n < k, n % k != 0 with equal_size,
 * an empty partition) as VPG_ERR_INVALID. */
int vpg_seqlen_balanced_partitions(const int64_t* seqlens_host, int64_t n, int64_t k, int equal_size,
                                   int64_t* part_indices_host, int64_t* part_offsets_host);

/* ---- packed ("thd") sequences for the Megatron-side callers (verl/models/mcore/util.py:23-142) --------------------
 * preprocess_packed_seqs / postprocess_packed_seqs as index arithmetic on the device.
 * layout: seqlens [B] i32 -> seqlens_padded[b] = seqlens[b] rounded up to a multiple of `align` (= tp_size, or
 *   tp_size*cp_size*2 with context parallelism, util.py:39-42), cu_seqlens_padded [B+1] i32, totals = {sum, max}.
 * map: for every valid token n of the rmpad stream (cu_seqlens [B+1] from vpg_unpad_indices): gathered_pos[n] = its
 *   row in the cp-gathered packed stream (rank-major: rank * (sum/cp) + local slot; chunk r and chunk 2cp-1-r of each
 *   padded sequence live on rank r, util.py:53-72,124-140), local_dst[n] = its slot in THIS rank's packed stream or
 *   -1 when another cp rank owns it.  cp_size 1: both = cu_seqlens_padded[b] + j.
 * permute_rows: out[dst_idx[n], :] = src[src_idx[n], :] (rows of row_bytes bytes; NULL index = identity; dst < 0
 *   skips the row); `out` (n_out rows) is zero-filled first when zero_fill != 0. */
int vpg_packed_seq_layout(const int32_t* seqlens, int64_t B, int align, int32_t* seqlens_padded,
                          int32_t* cu_seqlens_padded, int32_t* totals, void* stream);
int vpg_packed_seq_map(const int32_t* cu_seqlens, const int32_t* cu_seqlens_padded, int64_t B, int64_t T, int cp_size,
                       int cp_rank, int64_t* gathered_pos, int64_t* local_dst, void* stream);
int vpg_permute_rows(const void* src, const int64_t* src_idx, const int64_t* dst_idx, int64_t n_rows, int64_t row_bytes,
                     void* out, int64_t n_out, int zero_fill, void* stream);

/* ---- covariance-gated policy losses (verl/trainer/ppo/core_algos.py:818-972) --------------------------------------
 * Both produce the loss MATRIX pg_losses [n] and its elementwise derivative w.r.t. log_prob d_pg_losses [n]
 * (n = B*L flat); aggregate with vpg_agg_loss.  workspace: vpg_cov_workspace_bytes() bytes, 8-byte aligned; the first
 * double of it is sum(mask) after vpg_clip_cov_flags (the host needs it for clip_num, core_algos.py:879).
 * clip_cov (:818-905), two calls around the host's random draw (torch.randperm on the host generator, :884-886):
 *   vpg_clip_cov_flags: eligible[i] = 1 where cov_lb < cov < cov_ub, mask > 0 and the PPO clip did not cut the token,
 *     cov = (A - masked_mean(A)) * (lp - masked_mean(lp));
 *   vpg_clip_cov_loss (same workspace): zeroed[i] != 0 marks the drawn tokens (corr = 0);
 *     pg_losses = max(-A*ratio, -A*clamp(ratio, 1-low, 1+high)) * corr;  out3 = {ppo_kl, pg_clipfrac, sum(mask)}.
 * kl_cov (:908-972), one call: the k = max(1, int(N_valid * kl_cov_ratio)) valid tokens with the largest
 *   cov = (A - mean_valid A) * (lp - mean_valid lp) get pg_losses = -A*ratio + ppo_kl_coef*|lp - old|, the others
 *   -A*ratio (ties at the threshold go to the lower flat index; the reference's CPU torch.topk leaves them
 *   unspecified).  out3 = {ppo_kl_abs, number of tokens selected, sum(mask)}. */
size_t vpg_cov_workspace_bytes(void);
int vpg_clip_cov_flags(const float* old_log_prob, const float* log_prob, const float* advantages, const void* mask,
                       int mask_dtype, int64_t n, float cliprange_low, float cliprange_high, float cov_lb, float cov_ub,
                       uint8_t* eligible, void* workspace, size_t workspace_bytes, void* stream);
int vpg_clip_cov_loss(const float* old_log_prob, const float* log_prob, const float* advantages, const void* mask,
                      int mask_dtype, int64_t n, float cliprange_low, float cliprange_high, const uint8_t* zeroed,
                      float* pg_losses, float* d_pg_losses, float* out3, void* workspace, size_t workspace_bytes,
                      void* stream);
int vpg_kl_cov_loss(const float* old_log_prob, const float* log_prob, const float* advantages, const void* mask,
                    int mask_dtype, int64_t n, double kl_cov_ratio, float ppo_kl_coef, float* pg_losses,
                    float* d_pg_losses, float* out3, void* workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VPG_B200_H_ */
