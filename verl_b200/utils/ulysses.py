"""Ulysses sequence-parallel plumbing on either side of the per-token path (mirror of verl/utils/ulysses.py,
SURVEY.md §8(f).4): the rmpad token stream is padded to a multiple of sp_size and sliced per rank before the model
(`ulysses_pad_and_slice_inputs`, :278-321, called dp_actor.py:139-160), and the per-token log-probs / entropies of
the ranks are gathered back and un-padded after the lm-head (`gather_outputs_and_unpad`, :243-275, called
dp_actor.py:206-219).

This is collective plumbing, not a kernel: one `all_gather_into_tensor` (NCCL over NVLink on the GPU box, gloo in the
CPU tests) and views.  The head/sequence all-to-all of the attention layers (`gather_seq_scatter_heads`, :62-101)
belongs to the transformer body and is out of scope.
"""

from __future__ import annotations

from typing import Any, Optional

import torch
import torch.distributed as dist
from torch import Tensor
from torch.distributed import ProcessGroup

_ULYSSES_SEQUENCE_PARALLEL_GROUP = None


def set_ulysses_sequence_parallel_group(group: dist.ProcessGroup):
    global _ULYSSES_SEQUENCE_PARALLEL_GROUP
    _ULYSSES_SEQUENCE_PARALLEL_GROUP = group


def get_ulysses_sequence_parallel_group() -> Optional[dist.ProcessGroup]:
    return _ULYSSES_SEQUENCE_PARALLEL_GROUP


def get_ulysses_sequence_parallel_world_size(group: ProcessGroup = None) -> int:
    group = get_ulysses_sequence_parallel_group() if group is None else group
    return dist.get_world_size(group) if group else 1


def get_ulysses_sequence_parallel_rank(group: ProcessGroup = None) -> int:
    group = get_ulysses_sequence_parallel_group() if group is None else group
    return dist.get_rank(group) if group else 0


def _pad_tensor(x: Tensor, dim: int, padding_size: int) -> Tensor:
    """zero-extend `dim` by padding_size (ulysses.py:104-108)"""
    tail = list(x.shape)
    tail[dim] = padding_size
    return torch.cat((x, x.new_zeros(tail)), dim=dim)


def _unpad_tensor(x: Tensor, dim: int, padding_size: int) -> Tensor:
    """drop the last padding_size entries of `dim` (ulysses.py:111-114); a view"""
    return x.narrow(dim, 0, x.size(dim) - padding_size)


def slice_input_tensor(x: Tensor, dim: int, padding: bool = True, group: ProcessGroup = None) -> Tensor:
    """ulysses.py:117-130: this rank's contiguous 1/sp slice along `dim` (zero-padded first when it does not divide)."""
    group = get_ulysses_sequence_parallel_group() if group is None else group
    world, rank = dist.get_world_size(group), get_ulysses_sequence_parallel_rank(group)
    rem = x.size(dim) % world
    if padding and rem:
        x = _pad_tensor(x, dim, world - rem)
    per_rank = x.size(dim) // world
    return x.narrow(dim, rank * per_rank, per_rank).contiguous()


def all_gather_tensor(local_tensor: Tensor, group: Optional[dist.ProcessGroup] = None, async_op: bool = False):
    """ulysses.py:155-162: concatenation of every rank's tensor along axis 0."""
    group = get_ulysses_sequence_parallel_group() if group is None else group
    sp_world_size = dist.get_world_size(group=group)
    output_shape = list(local_tensor.shape)
    output_shape[0] = output_shape[0] * sp_world_size
    output = torch.empty(output_shape, dtype=local_tensor.dtype, device=local_tensor.device)
    dist.all_gather_into_tensor(output, local_tensor.contiguous(), group=group, async_op=async_op)
    return output


class Gather(torch.autograd.Function):
    """ulysses.py:194-234: all-gather along `gather_dim`; the backward keeps this rank's slice of the gradient,
    multiplied by the group size when grad_scaler is set (the loss is averaged over the sp ranks' replicas)."""

    @staticmethod
    def forward(ctx: Any, group: dist.ProcessGroup, local_tensor: Tensor, gather_dim: int, grad_scaler: bool = True,
                async_op=False) -> Tensor:
        ctx.group = group
        ctx.gather_dim = gather_dim
        ctx.grad_scaler = grad_scaler
        ctx.sp_world_size = dist.get_world_size(group=group)
        ctx.sp_rank = dist.get_rank(group=group)
        ctx.part_size = local_tensor.size(gather_dim)
        output = all_gather_tensor(local_tensor, group, async_op)
        if gather_dim % max(local_tensor.dim(), 1) == 0:
            return output  # the all-gather already concatenates along axis 0: no second copy
        return torch.cat(output.split(local_tensor.size(0), dim=0), dim=gather_dim)

    @staticmethod
    def backward(ctx: Any, grad_output: Tensor) -> Any:
        if ctx.grad_scaler:
            grad_output = grad_output * ctx.sp_world_size
        return (None, grad_output.split(ctx.part_size, dim=ctx.gather_dim)[ctx.sp_rank].contiguous(), None, None, None)


def gather_outpus_and_unpad(*args, **kwargs):
    raise RuntimeError(
        "please use verl.utils.ulysses.gather_outputs_and_unpad instead of verl.utils.ulysses.gather_outpus_and_unpad"
    )


def gather_outputs_and_unpad(x: Tensor, gather_dim: int, unpad_dim: int = None, padding_size: int = 0,
                             grad_scaler: bool = True, group: Optional[dist.ProcessGroup] = None):
    """ulysses.py:243-275: gather `x` across the sp group and drop `padding_size` trailing elements of `unpad_dim`
    (the ulysses_pad padding).  Without a group `x` is returned unchanged."""
    group = get_ulysses_sequence_parallel_group() if group is None else group
    if group is None:
        return x
    x = Gather.apply(group, x, gather_dim, grad_scaler)
    if unpad_dim is not None:
        assert isinstance(padding_size, int), "padding size is not given or is not an integer"
        if padding_size == 0:
            return x
        x = _unpad_tensor(x, unpad_dim, padding_size)
    return x


def ulysses_pad(input_ids_rmpad: torch.Tensor, position_ids_rmpad: Optional[torch.Tensor] = None, sp_size: int = 1):
    """ulysses.py:278-293: right-pad the [1,T] id stream with 0 and the position ids with 0..pad-1 ([1,T] or mrope
    [3,1,T]) so that T divides by sp_size.  -> (ids, position_ids, pad_size)."""
    if position_ids_rmpad is not None:
        assert position_ids_rmpad.size(-2) == 1
        assert input_ids_rmpad.size(-1) == position_ids_rmpad.size(-1)
    if sp_size <= 1:
        return input_ids_rmpad, position_ids_rmpad, 0
    pad_size = -input_ids_rmpad.shape[1] % sp_size
    if pad_size == 0:
        return input_ids_rmpad, position_ids_rmpad, 0
    input_ids_rmpad = _pad_tensor(input_ids_rmpad, 1, pad_size)
    if position_ids_rmpad is not None:
        ramp = torch.arange(pad_size, device=position_ids_rmpad.device, dtype=torch.int64)
        ramp = ramp.expand(*position_ids_rmpad.shape[:-1], pad_size)  # [1,pad] or [3,1,pad]
        position_ids_rmpad = torch.cat((position_ids_rmpad, ramp), dim=-1)
    return input_ids_rmpad, position_ids_rmpad, pad_size


def ulysses_pad_and_slice_inputs(input_ids_rmpad: torch.Tensor, position_ids_rmpad: Optional[torch.Tensor] = None,
                                 sp_size: int = 1):
    """ulysses.py:296-321: pad, then keep this rank's slice of the stream (ids along dim 1; position ids along the
    reference's `dim=1` as well, i.e. the token axis of [1,T])."""
    input_ids_rmpad, position_ids_rmpad, pad_size = ulysses_pad(input_ids_rmpad, position_ids_rmpad, sp_size)
    input_ids_rmpad = slice_input_tensor(input_ids_rmpad, dim=1, padding=False)
    if position_ids_rmpad is not None:
        position_ids_rmpad = slice_input_tensor(position_ids_rmpad, dim=1, padding=False)
    return input_ids_rmpad, position_ids_rmpad, pad_size


def validate_ulysses_config(num_heads, ulysses_sequence_size):
    if ulysses_sequence_size > 1:
        assert num_heads % ulysses_sequence_size == 0, (
            f"num_heads ({num_heads}) must be divisible by ulysses sequence size({ulysses_sequence_size})"
        )
