"""rmpad index plumbing with the interface of flash_attn.bert_padding (third-party, pure-torch file; flash-attn
2.8.3 is what the reference environment pins closest: docker/verl0.5-cu126-torch2.7.1-fa2.8.0/Dockerfile.base:50-53),
as called from verl/workers/actor/dp_actor.py:40,110-125,221-232 and verl/utils/torch_functional.py:451-462,480-490.

    unpad_input(hidden_states, attention_mask, unused_mask=None)
        -> (hidden_states_rmpad, indices, cu_seqlens, max_seqlen_in_batch, seqused)      (5 values, flash-attn >= 2.7)
    pad_input(hidden_states, indices, batch, seqlen)
    index_first_axis(x, indices) / index_put_first_axis(values, indices, first_axis_dim)   (autograd-aware)

indices are int64 and ascending, cu_seqlens int32 — bit-exact with the reference; gathers/scatters are exact copies.
"""

from __future__ import annotations

import torch

from . import ops


class IndexFirstAxis(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input, indices):
        ctx.save_for_backward(indices)
        assert input.ndim >= 2
        ctx.first_axis_dim = input.shape[0]
        return ops.gather_rows(input, indices)

    @staticmethod
    def backward(ctx, grad_output):
        (indices,) = ctx.saved_tensors
        return ops.scatter_rows(grad_output, indices, ctx.first_axis_dim), None


index_first_axis = IndexFirstAxis.apply


class IndexPutFirstAxis(torch.autograd.Function):
    @staticmethod
    def forward(ctx, values, indices, first_axis_dim):
        ctx.save_for_backward(indices)
        assert indices.ndim == 1
        assert values.ndim >= 2
        return ops.scatter_rows(values, indices, first_axis_dim)

    @staticmethod
    def backward(ctx, grad_output):
        (indices,) = ctx.saved_tensors
        return ops.gather_rows(grad_output, indices), None, None


index_put_first_axis = IndexPutFirstAxis.apply


def unpad_input(hidden_states, attention_mask, unused_mask=None):
    """hidden_states (batch, seqlen, ...), attention_mask (batch, seqlen) with 1 = valid.

    Returns (hidden_states [total_nnz, ...], indices [total_nnz] int64, cu_seqlens [batch+1] int32,
             max_seqlen_in_batch int, seqused [batch] int32).
    """
    all_masks = (attention_mask + unused_mask) if unused_mask is not None else attention_mask
    indices, seqlens_all, cu_seqlens, max_seqlen = ops.unpad_indices(all_masks)
    if unused_mask is not None:
        _, used_seqlens, _, _ = ops.unpad_indices(attention_mask)
    else:
        used_seqlens = seqlens_all
    flat = hidden_states.reshape(hidden_states.shape[0] * hidden_states.shape[1], *hidden_states.shape[2:])
    if flat.ndim == 1:
        out = index_first_axis(flat.unsqueeze(-1), indices).squeeze(-1)
    else:
        out = index_first_axis(flat, indices)
    return out, indices, cu_seqlens, max_seqlen, used_seqlens


def pad_input(hidden_states, indices, batch, seqlen):
    """hidden_states (total_nnz, ...) -> (batch, seqlen, ...) with zeros at padded positions."""
    output = index_put_first_axis(hidden_states, indices, batch * seqlen)
    return output.view(batch, seqlen, *hidden_states.shape[1:])
