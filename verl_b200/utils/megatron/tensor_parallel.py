"""Vocab-parallel statistics of MATERIALISED logit shards — the Megatron-side twin of the fused lm-head
(mirror of verl/utils/megatron/tensor_parallel.py:109-186).

    vocab_parallel_entropy(vocab_parallel_logits)                          :142-151  (autograd fn :109-139)
    vocab_parallel_log_probs_from_logits(logits, labels)                    :154-158  (= -megatron.core.tensor_parallel.
                                                                                        vocab_parallel_cross_entropy)
    vocab_parallel_log_probs_from_logits_response_rmpad(...)                :161-186

Each rank reads its [T, V/tp] shard ONCE (vpg_logits_partials: online max / sum-exp / sum x*exp / label logit per
row), the ranks exchange 16 bytes per token with ONE all-gather, and vpg_lce_finalize merges the partials — where the
reference makes 4-5 passes over the shard, saves a second [T, V/tp] tensor (the soft-max) for the backward and issues
three dependent all-reduces (MAX, SUM, SUM).  The backward recomputes p = exp(x - lse) from the saved per-token
scalars in one pass.

The tensor-parallel group: pass `group=`, or register one with set_tensor_model_parallel_group(); if neither is given
and megatron.core is importable its `parallel_state` group is used (as in the reference); otherwise a single rank.
megatron-core's in-place `logits -= max` side effect on the caller's tensor is NOT reproduced (inputs are read-only).
"""

from __future__ import annotations

import torch
import torch.distributed as dist

from ... import ops
from ..kernel.kernels import exchange_partials

_TP_GROUP = None


def set_tensor_model_parallel_group(group):
    global _TP_GROUP
    _TP_GROUP = group


def _resolve_group(group):
    if group is not None:
        return group
    if _TP_GROUP is not None:
        return _TP_GROUP
    try:
        from megatron.core import parallel_state as mpu  # absent in this image; the reference's only source
    except Exception:
        return None
    return mpu.get_tensor_model_parallel_group()


def _rank_world(group):
    if group is None or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def _merged_stats(logits2d, labels, group):
    rank, world = _rank_world(group)
    part = ops.logits_partials(logits2d, labels, label_offset=rank * logits2d.shape[1])
    if world > 1:
        part = exchange_partials(part, group)
    return ops.lce_finalize(part, 1.0)  # logprobs, entropy, lse, eb


class _VocabParallelEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, vocab_parallel_logits: torch.Tensor, group=None) -> torch.Tensor:
        lead = vocab_parallel_logits.shape[:-1]
        x = vocab_parallel_logits.reshape(-1, vocab_parallel_logits.shape[-1])
        if x.dtype not in (torch.float32, torch.bfloat16):  # the kernels read fp32 / bf16; save what they will read
            x = x.float()
        _, entropy, lse, _ = _merged_stats(x, None, group)
        ctx.save_for_backward(x, lse, entropy)
        ctx.in_shape, ctx.in_dtype = vocab_parallel_logits.shape, vocab_parallel_logits.dtype
        return entropy.view(lead)

    @staticmethod
    def backward(ctx, grad_output: torch.Tensor):
        x, lse, entropy = ctx.saved_tensors
        g = grad_output.reshape(-1).to(torch.float32).contiguous()
        d = ops.logits_stats_backward(x.contiguous(), None, lse, entropy, None, g)  # -g * p * (x - sum p*x)
        return d.view(ctx.in_shape).to(ctx.in_dtype), None


class _VocabParallelLogProbs(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits: torch.Tensor, labels: torch.Tensor, group=None) -> torch.Tensor:
        x = logits.reshape(-1, logits.shape[-1])
        if x.dtype not in (torch.float32, torch.bfloat16):
            x = x.float()
        lab = labels.reshape(-1).to(torch.int64).contiguous()
        logprobs, _, lse, _ = _merged_stats(x, lab, group)
        rank, _ = _rank_world(group)
        ctx.save_for_backward(x, lab, lse)
        ctx.offset = rank * x.shape[1]
        ctx.in_shape, ctx.in_dtype = logits.shape, logits.dtype
        return logprobs.view(labels.shape)

    @staticmethod
    def backward(ctx, grad_output: torch.Tensor):
        x, lab, lse = ctx.saved_tensors
        g = grad_output.reshape(-1).to(torch.float32).contiguous()
        d = ops.logits_stats_backward(x.contiguous(), lab - ctx.offset, lse, None, g, None)  # g * (onehot - p)
        return d.view(ctx.in_shape).to(ctx.in_dtype), None, None


def vocab_parallel_entropy(vocab_parallel_logits: torch.Tensor, group=None) -> torch.Tensor:
    """(total_nnz, vocab_size // tp_size) -> (total_nnz,) entropy of the full-vocab soft-max."""
    return _VocabParallelEntropy.apply(vocab_parallel_logits, _resolve_group(group))


def vocab_parallel_log_probs_from_logits(logits, labels, group=None):
    """log-softmax over the full vocab at `labels` (global ids) from this rank's shard; labels.shape out."""
    return _VocabParallelLogProbs.apply(logits, labels, _resolve_group(group))


def vocab_parallel_log_probs_from_logits_response_rmpad(input_ids, attention_mask, logits_rmpad, response_length,
                                                        group=None):
    """tensor_parallel.py:161-186: labels = the rmpad id stream rolled by one; per-token log-probs padded back to
    [batch, seqlen] and cut to the response window [:, -response_length-1:-1]."""
    from ...bert_padding import pad_input, unpad_input

    batch_size, seqlen = input_ids.shape
    input_ids_rmpad, indices, *_ = unpad_input(input_ids.unsqueeze(-1), attention_mask=attention_mask)
    rolled = torch.roll(input_ids_rmpad.squeeze(-1), shifts=-1, dims=0)
    full_log_probs_rmpad = vocab_parallel_log_probs_from_logits(logits=logits_rmpad, labels=rolled, group=group)
    full_output = pad_input(hidden_states=full_log_probs_rmpad.unsqueeze(-1), indices=indices, batch=batch_size,
                            seqlen=seqlen)
    return full_output.squeeze(-1)[:, -response_length - 1:-1]
