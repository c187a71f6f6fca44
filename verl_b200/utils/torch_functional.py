"""Hot-path subset of verl/utils/torch_functional.py with the same names and signatures, running on libvpg_b200.

Covered (reference line ranges in each docstring): gather_from_labels, logprobs_from_logits (+ _v2 / _naive /
_flash_attn aliases), entropy_from_logits (+ _with_chunking), masked_sum / masked_mean / masked_var /
masked_whiten, get_response_mask, get_unpad_data, log_probs_from_logits_response_rmpad,
log_probs_from_logits_all_rmpad.  LR schedulers, tokenizer padding and the distributed statistics helpers of the
reference file are outside the hot path and are not provided here.

All functions require CUDA tensors; there is no CPU fallback.
"""

from __future__ import annotations

import torch

from .. import bert_padding, ops
from ..staging import staged

__all__ = [
    "gather_from_labels",
    "logprobs_from_logits",
    "logprobs_from_logits_v2",
    "logprobs_from_logits_naive",
    "logprobs_from_logits_flash_attn",
    "entropy_from_logits",
    "entropy_from_logits_with_chunking",
    "masked_sum",
    "masked_mean",
    "masked_var",
    "masked_whiten",
    "get_response_mask",
    "get_unpad_data",
    "log_probs_from_logits_response_rmpad",
    "log_probs_from_logits_all_rmpad",
    "clip_by_value",
]


def _no_autograd(name, t):
    """These three are forward-only kernels here; the reference versions are differentiable torch compositions whose
    in-tree callers all run under no_grad.  Fail loudly instead of returning a silently detached result."""
    if torch.is_grad_enabled() and t.requires_grad:
        raise NotImplementedError(f"verl_b200.{name} has no backward: call it under torch.no_grad() (masked_mean, "
                                  "agg_loss and logprobs_from_logits are the differentiable entry points)")


@staged
def gather_from_labels(data, label):
    """verl/utils/torch_functional.py:49-61 — data (..., V), label (...) -> (...)."""
    _no_autograd("gather_from_labels", data)
    flat = data.reshape(-1, data.shape[-1]).contiguous()
    idx = label.reshape(-1).to(torch.int64) + torch.arange(flat.shape[0], device=flat.device) * flat.shape[1]
    return ops.gather_rows(flat.view(-1, 1), idx).view(label.shape)


class _LogitsStats(torch.autograd.Function):
    """One-pass log-prob / entropy over materialised logits with a fused backward."""

    @staticmethod
    def forward(ctx, logits2d, labels, want_lp, want_ent):
        lp, ent, lse = ops.logits_stats(logits2d, labels, want_logprobs=want_lp, want_entropy=want_ent)
        ctx.save_for_backward(logits2d, labels if labels is not None else torch.empty(0), lse,
                              ent if ent is not None else torch.empty(0))
        ctx.has_labels = labels is not None
        ctx.has_ent = ent is not None
        ctx.set_materialize_grads(False)
        outs = (lp if lp is not None else lse.new_empty(0), ent if ent is not None else lse.new_empty(0))
        return outs

    @staticmethod
    def backward(ctx, g_lp, g_ent):
        logits2d, labels, lse, ent = ctx.saved_tensors
        x = logits2d.contiguous()
        d = ops.logits_stats_backward(
            x,
            labels if ctx.has_labels else None,
            lse,
            ent if ctx.has_ent else None,
            g_lp.float().contiguous() if (g_lp is not None and ctx.has_labels) else None,
            g_ent.float().contiguous() if (g_ent is not None and ctx.has_ent) else None,
        )
        return d, None, None, None


def _stats(logits, labels, want_lp, want_ent):
    if not logits.is_cuda:
        raise RuntimeError("verl_b200 ops require CUDA tensors (there is no CPU fallback)")
    x = logits if logits.dtype in (torch.float32, torch.bfloat16) else logits.float()
    x2 = x.reshape(-1, x.shape[-1])
    lab = labels.reshape(-1).to(torch.int64).contiguous() if labels is not None else None
    lp, ent = _LogitsStats.apply(x2, lab, want_lp, want_ent)
    return lp, ent


@staged
def logprobs_from_logits(logits, labels, inplace_backward=True):
    """verl/utils/torch_functional.py:64-92.  Output has logits' dtype and shape logits.shape[:-1]."""
    lp, _ = _stats(logits, labels, True, False)
    return lp.view(logits.shape[:-1]).to(logits.dtype)


def logprobs_from_logits_flash_attn(logits, labels, inplace_backward=True):
    """verl/utils/torch_functional.py:95-100 (flash-attn Triton cross entropy in the reference)."""
    return logprobs_from_logits(logits, labels, inplace_backward)


def logprobs_from_logits_naive(logits, labels):
    """verl/utils/torch_functional.py:110-113"""
    return logprobs_from_logits(logits, labels)


def logprobs_from_logits_v2(logits, labels):
    """verl/utils/torch_functional.py:116-133"""
    return logprobs_from_logits(logits, labels)


def clip_by_value(x, tensor_min, tensor_max):
    """verl/utils/torch_functional.py:136-142"""
    return torch.max(torch.min(x, tensor_max), tensor_min)


@staged
def entropy_from_logits(logits: torch.Tensor):
    """verl/utils/torch_functional.py:145-149.  Output has logits' dtype."""
    _, ent = _stats(logits, None, False, True)
    return ent.view(logits.shape[:-1]).to(logits.dtype)


@staged
def entropy_from_logits_with_chunking(logits: torch.Tensor, chunk_size: int = 2048):
    """verl/utils/torch_functional.py:152-160 — returns fp32; chunking is unnecessary here (one pass, no temporaries)."""
    _, ent = _stats(logits, None, False, True)
    return ent.view(logits.shape[:-1])


# ------------------------------------------------------------------------------------------------- masked ops
class _MaskedMean(torch.autograd.Function):
    """masked_mean(axis=None) == agg_loss('token-mean'); backward comes out of the same kernel."""

    @staticmethod
    def forward(ctx, values, mask):
        v2 = values.reshape(1, -1) if values.dim() < 2 else values
        m2 = mask.reshape(1, -1) if mask.dim() < 2 else mask
        out, d = ops.agg_loss_fwd(v2, m2, "token-mean", want_grad=values.requires_grad)
        ctx.shape = values.shape
        ctx.in_dtype = values.dtype
        if d is not None:
            ctx.save_for_backward(d)
        return out

    @staticmethod
    def backward(ctx, g):
        (d,) = ctx.saved_tensors
        return (d * g).view(ctx.shape).to(ctx.in_dtype), None


def _axis_sums(values, mask, axis):
    """(sum, count) over an arbitrary axis / tuple of axes: the reduced axes are moved last (a strided copy when they
    are not already) and the row-sum kernel reduces them."""
    nd = values.dim()
    axes = tuple(sorted({a % nd for a in (axis if isinstance(axis, (tuple, list)) else (axis,))}))
    keep = [d for d in range(nd) if d not in axes]
    out_shape = [values.shape[d] for d in keep]
    if axes != tuple(range(nd - len(axes), nd)):
        values, mask = values.permute(*keep, *axes), mask.permute(*keep, *axes)
    red = int(torch.Size([values.shape[i] for i in range(len(keep), nd)]).numel())
    s, c = ops.masked_row_sums(values.reshape(-1, red).contiguous(), mask.reshape(-1, red).contiguous())
    return s.view(out_shape), c.view(out_shape)


@staged
def masked_sum(values, mask, axis=None):
    """verl/utils/torch_functional.py:163-168: (where(mask.bool(), values, 0) * mask).sum(axis)."""
    _no_autograd("masked_sum", values)
    if axis is None:
        return ops.masked_sums(values, mask.expand_as(values))[0].to(torch.float32)
    return _axis_sums(values, mask.expand_as(values), axis)[0]


@staged
def masked_mean(values, mask, axis=None):
    """verl/utils/torch_functional.py:171-185: masked_sum / (mask.sum(axis) + 1e-8)."""
    mask = mask.expand_as(values)
    if axis is None:
        return _MaskedMean.apply(values, mask)
    s, c = _axis_sums(values, mask, axis)
    return s / (c + 1e-8)


@staged
def masked_var(values, mask, unbiased=True):
    """verl/utils/torch_functional.py:188-203"""
    mean = masked_mean(values, mask)
    centered_values = values - mean
    variance = masked_mean(centered_values**2, mask)
    if unbiased:
        mask_sum = mask.sum()
        if mask_sum == 0:
            raise ValueError("At least one element in the mask has to be 1.")
        if mask_sum == 1:
            raise ValueError("The sum of the mask is one, which can cause a division by zero.")
        bessel_correction = mask_sum / (mask_sum - 1)
        variance = variance * bessel_correction
    return variance


@staged
def masked_whiten(values, mask, shift_mean=True):
    """verl/utils/torch_functional.py:206-223 (two sweeps: statistics, then normalise)."""
    _no_autograd("masked_whiten", values)
    return ops.masked_whiten(values, mask.expand_as(values), shift_mean).view(values.shape)


def get_response_mask(response_id: torch.Tensor, eos_token: int | list[int] = 2, dtype=torch.int64):
    """verl/utils/torch_functional.py:226-247 (index bookkeeping, torch ops)."""
    eos_mask = torch.isin(response_id, torch.tensor(eos_token, device=response_id.device)).int()
    return (eos_mask.cumsum(dim=1) - eos_mask).eq(0).to(dtype)


# ------------------------------------------------------------------------------------------------- rmpad
@staged
def get_unpad_data(attention_mask):
    """verl/utils/torch_functional.py:629-638 -> (indices, cu_seqlens, max_seqlen_in_batch)."""
    indices, _, cu_seqlens, max_seqlen = ops.unpad_indices(attention_mask)
    return indices, cu_seqlens, max_seqlen


@staged
def log_probs_from_logits_response_rmpad(input_ids, attention_mask, logits_rmpad, response_length):
    """verl/utils/torch_functional.py:438-462"""
    batch_size, seqlen = input_ids.shape
    input_ids_rmpad, indices, *_ = bert_padding.unpad_input(input_ids.unsqueeze(-1), attention_mask=attention_mask)
    input_ids_rmpad = input_ids_rmpad.squeeze(-1)
    input_ids_rmpad_rolled = torch.roll(input_ids_rmpad, shifts=-1, dims=0)
    full_log_probs_rmpad = logprobs_from_logits(logits=logits_rmpad, labels=input_ids_rmpad_rolled)
    full_output = bert_padding.pad_input(
        hidden_states=full_log_probs_rmpad.unsqueeze(-1), indices=indices, batch=batch_size, seqlen=seqlen
    )
    return full_output.squeeze(-1)[:, -response_length - 1 : -1]


@staged
def log_probs_from_logits_all_rmpad(input_ids_rmpad, logits_rmpad, indices, batch_size, seqlen, response_length):
    """verl/utils/torch_functional.py:465-490"""
    input_ids_rmpad = input_ids_rmpad.transpose(0, 1)  # transpose back to [total_nnz, 1]
    input_ids_rmpad = input_ids_rmpad.squeeze(-1)
    input_ids_rmpad_rolled = torch.roll(input_ids_rmpad, shifts=-1, dims=0)
    full_log_probs_rmpad = logprobs_from_logits(logits=logits_rmpad, labels=input_ids_rmpad_rolled)
    full_output = bert_padding.pad_input(
        hidden_states=full_log_probs_rmpad.unsqueeze(-1), indices=indices, batch=batch_size, seqlen=seqlen
    )
    return full_output.squeeze(-1)[:, -response_length - 1 : -1]


@staged
def log_probs_from_logits_response(input_ids, logits, response_length):
    """verl/utils/torch_functional.py:421-435: logits [B,S,V] = model(input_ids) -> log-probs of the last
    ``response_length`` tokens, [B, response_length]."""
    response_logits = logits[:, -response_length - 1 : -1]
    response = input_ids[:, -response_length:]
    return logprobs_from_logits(logits=response_logits, labels=response)


# ------------------------------------------------------------------------------------------------- distributed stats
def distributed_masked_mean(local_tensor, local_mask):
    """verl/utils/torch_functional.py:752-771: global mean of the valid elements across the default process group.
    One masked-sum pass gives {sum, count}; ONE all-reduce of the pair replaces the reference's two."""
    import torch.distributed as dist

    pair = ops.masked_sums(local_tensor, local_mask)  # float64 [2]
    dist.all_reduce(pair, op=dist.ReduceOp.SUM)
    return (pair[0] / pair[1]).to(torch.float32)


def distributed_mean_max_min_std(local_tensor, compute_max=True, compute_min=True, compute_std=True):
    """verl/utils/torch_functional.py:709-749 -> (mean, max, min, std), None for disabled entries.  Metrics-side helper
    (a handful of scalars per call): torch reductions locally, the reference's collectives packed into at most three."""
    import torch.distributed as dist

    x = local_tensor.float()
    sums = torch.stack([x.sum(), x.new_tensor(float(x.numel()))])
    dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    global_mean = sums[0] / sums[1]
    local_max = local_min = global_std = None
    if compute_max or compute_min:
        ext = torch.stack([x.max(), -x.min()])  # one MAX all-reduce serves both
        dist.all_reduce(ext, op=dist.ReduceOp.MAX)
        local_max = ext[0] if compute_max else None
        local_min = -ext[1] if compute_min else None
    if compute_std:
        square_diff = torch.sum(torch.pow(x - global_mean, 2))
        dist.all_reduce(square_diff, op=dist.ReduceOp.SUM)
        global_std = torch.sqrt(square_diff / (sums[1] - 1))
    return global_mean, local_max, local_min, global_std
