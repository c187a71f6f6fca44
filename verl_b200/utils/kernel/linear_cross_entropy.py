"""Drop-in for verl/utils/kernel/linear_cross_entropy.py (LinearCrossEntropy :40-114, linear_cross_entropy :117).

    from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy
    logprobs, entropy = linear_cross_entropy(hidden, weight, labels, temperature, "none", dist_process_group)

hidden [T,H] or [B,S,H] bf16, weight [V,H] bf16 (nn.Linear layout), labels int64.  Outputs are fp32.
Shapes: 2-D input and [1,T,H] return flat [T] exactly like the reference; [B,S,H] with B>1 returns [B,S] (the
shape the reference's torch backend returns and dp_actor.py:255-256 needs) — a superset of the Triton op's
behaviour.
"""

from __future__ import annotations

import typing

import torch
import torch.distributed as dist

from . import kernels


class LinearCrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(
        ctx,
        hidden: torch.Tensor,
        weight: torch.Tensor,
        labels: torch.Tensor,
        temperature: typing.Optional[float] = 1.0,
        reduction: typing.Optional[str] = "none",
        dist_process_group: typing.Optional[dist.ProcessGroup] = None,
    ) -> list[torch.Tensor]:
        assert isinstance(temperature, float), f"temperature must be a float, but got {type(temperature)}"
        assert isinstance(reduction, str), f"reduction must be a str, but got {type(reduction)}"
        with torch.cuda.nvtx.range("LinearCrossEntropy-forward"):
            REDUCTION = kernels.get_entropy_reduction_enum_number(reduction.lower())

            original_hidden_shape = hidden.shape
            if len(hidden.shape) != 2:
                hidden = hidden.reshape(-1, hidden.shape[-1])
            if len(labels.shape) != 1:
                labels = labels.reshape(-1)
            hidden = hidden.contiguous()
            labels = labels.contiguous()
            if labels.dtype != torch.int64:
                labels = labels.to(torch.int64)

            logprobs, entropy, _maximum, _accumulate, _entropy_b = kernels.efficient_entropy_forward(
                hidden, weight, labels, REDUCTION, temperature, dist_process_group
            )

            ctx.save_for_backward(hidden, weight, labels, _maximum, _entropy_b)
            ctx.original_hidden_shape = original_hidden_shape
            ctx.REDUCTION = REDUCTION
            ctx.dist_process_group = dist_process_group
            ctx.should_return_fp32_grad = False
            ctx.temperature = temperature
            ctx.set_materialize_grads(False)

            ctx.out_shape = None
            if len(original_hidden_shape) == 3 and original_hidden_shape[0] > 1:
                ctx.out_shape = tuple(original_hidden_shape[:2])
                entropy = entropy.view(ctx.out_shape)
                if REDUCTION == kernels.EntropyReductionEnum._None:
                    logprobs = logprobs.view(ctx.out_shape)
        return logprobs, entropy

    @staticmethod
    def backward(ctx, dlogprobs: torch.Tensor, dentropy: torch.Tensor) -> list[torch.Tensor]:
        with torch.cuda.nvtx.range("LinearCrossEntropy-backward"):
            (hidden, weight, labels, _maximum, _entropy_b) = ctx.saved_tensors
            if dlogprobs is not None and dlogprobs.dim() > 1:
                dlogprobs = dlogprobs.reshape(-1)
            if dentropy is not None and dentropy.dim() > 1:
                dentropy = dentropy.reshape(-1)
            d_hidden, d_weight = kernels.efficient_entropy_backward(
                dlogprobs.contiguous() if dlogprobs is not None else None,
                dentropy.contiguous() if dentropy is not None else None,
                hidden,
                weight,
                labels,
                _maximum,
                None,
                _entropy_b,
                ctx.REDUCTION,
                ctx.should_return_fp32_grad,
                ctx.temperature,
                ctx.dist_process_group,
            )
            d_hidden = d_hidden.view(ctx.original_hidden_shape)

        return (d_hidden, d_weight, None, None, None, None)


linear_cross_entropy = LinearCrossEntropy.apply
