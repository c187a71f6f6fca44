"""Drop-in for verl/utils/experimental/torch_functional.py (FusedLinearForPPOFunction :71-198, FusedLinearForPPO
:201-216) — the reference's DEFAULT fused backend (`fused_kernels_backend: torch`, verl/trainer/config/
ppo_trainer.yaml:90-94), reached from verl/models/transformers/dense_common.py:88-121.

Same call and same output contract as the reference: hidden_states [T,H] or [B,S,H], vocab_weights [V,H], input_ids
[T] or [B,S]; log_probs / entropy come back shaped like input_ids in hidden_states' dtype (the reference allocates
them with hidden_states.new_zeros, :101-102), gradients in the operands' dtypes.  The work is the sm_100a lm-head
path (tcgen05 GEMM + fused online soft-max, recompute in the backward): there is no 512-token chunk loop and
`chunk_size` is accepted for signature compatibility only.  The reference rounds its logits to the operand dtype
before the soft-max (bf16 matmul output, :33-37); here they stay in the fp32 accumulator.
"""

from __future__ import annotations

from typing import Optional

import torch

from ..kernel import kernels

__all__ = ["FusedLinearForPPOFunction", "FusedLinearForPPO"]


class FusedLinearForPPOFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, hidden_states, vocab_weights, input_ids, temperature: float = 1.0, chunk_size: int = 512):
        ctx.set_materialize_grads(False)
        orig_ndim = hidden_states.ndim
        assert orig_ndim in (2, 3), f"Invalid hidden_states shape, received {hidden_states.shape}"
        orig_batch_size = -1
        if orig_ndim == 3:
            assert input_ids.ndim == 2, f"input_ids shape doesn't match, {hidden_states.shape} {input_ids.shape}"
            orig_batch_size = hidden_states.shape[0]
            hidden_states = hidden_states.flatten(0, 1)
            input_ids = input_ids.flatten(0, 1)
        hidden2d = hidden_states.contiguous()
        labels = input_ids.contiguous().to(torch.int64)
        weight = vocab_weights.contiguous()
        reduction = kernels.get_entropy_reduction_enum_number("none")
        log_probs, entropy, lse, _acc, eb = kernels.efficient_entropy_forward(hidden2d, weight, labels, reduction,
                                                                             float(temperature), None)
        log_probs = log_probs.to(hidden2d.dtype)
        entropy = entropy.to(hidden2d.dtype)
        if orig_ndim == 3:
            log_probs = log_probs.view(orig_batch_size, -1)
            entropy = entropy.view(orig_batch_size, -1)
        ctx.save_for_backward(hidden2d, weight, labels, lse, eb)
        ctx.orig_batch_size = orig_batch_size
        ctx.orig_ndim = orig_ndim
        ctx.temperature = float(temperature)
        ctx.weight_dtype = vocab_weights.dtype
        return log_probs, entropy

    @staticmethod
    def backward(ctx, dlog_probs: Optional[torch.Tensor], dentropy: Optional[torch.Tensor]):
        assert dlog_probs is not None or dentropy is not None
        hidden2d, weight, labels, lse, eb = ctx.saved_tensors
        if dlog_probs is not None:
            dlog_probs = dlog_probs.reshape(-1).contiguous()
        if dentropy is not None:
            dentropy = dentropy.reshape(-1).contiguous()
        d_hidden, d_weight = kernels.efficient_entropy_backward(
            dlog_probs, dentropy, hidden2d, weight, labels, lse, None, eb,
            kernels.get_entropy_reduction_enum_number("none"), False, ctx.temperature, None)
        if ctx.orig_ndim == 3:
            d_hidden = d_hidden.view(ctx.orig_batch_size, -1, hidden2d.shape[-1])
        need_h, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        return (d_hidden if need_h else None, d_weight if need_w else None, None, None, None)


class FusedLinearForPPO(torch.nn.Module):
    def __init__(self, chunk_size: int = 512):
        super().__init__()
        self.chunk_size = chunk_size

    def forward(self, hidden_states, vocab_weights, input_ids, temperature: float = 1.0):
        input_ids = input_ids.to(torch.int64)
        return FusedLinearForPPOFunction.apply(hidden_states, vocab_weights, input_ids, temperature, self.chunk_size)
