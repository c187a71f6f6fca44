"""Sequence-parallel (Ulysses) plumbing on either side of the per-token path — SURVEY.md §8(f).4.

The reference (verl/utils/ulysses.py) pads the rmpad token stream to a multiple of the sequence-parallel size and hands
every rank one contiguous slice before the model (`ulysses_pad_and_slice_inputs` :296-321, called from dp_actor.py:
139-160), then gathers the per-token log-probs / entropies of the ranks and drops the padding after the lm-head
(`gather_outputs_and_unpad` :243-275, called from dp_actor.py:206-219).  Only those two call sites are on the path, so
this module provides exactly the names they (and `compat.install`) need, with the reference's argument meaning.  The
head/sequence all-to-all inside attention belongs to the transformer body and is out of scope.

Implementation notes (nothing here is a kernel — one `all_gather_into_tensor` over NCCL/NVLink, gloo in the CPU tests):
* the gather is ONE collective into a `[world, ...]` buffer; gathering along axis 0 is then a free reshape, any other
  axis one `movedim` + `reshape` (a single strided copy), never a per-rank `cat`;
* padding uses `torch.nn.functional.pad`; slices are `narrow` views made contiguous once.
"""

from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F

__all__ = [
    "set_ulysses_sequence_parallel_group", "get_ulysses_sequence_parallel_group",
    "get_ulysses_sequence_parallel_world_size", "get_ulysses_sequence_parallel_rank", "slice_input_tensor", "Gather",
    "gather_outputs_and_unpad", "ulysses_pad", "ulysses_pad_and_slice_inputs", "validate_ulysses_config",
    "all_gather_tensor", "gather_outpus_and_unpad",
]

_state = {"group": None}


def set_ulysses_sequence_parallel_group(group):
    _state["group"] = group


def get_ulysses_sequence_parallel_group():
    return _state["group"]


def _resolve(group):
    return _state["group"] if group is None else group


def get_ulysses_sequence_parallel_world_size(group=None) -> int:
    g = _resolve(group)
    return 1 if g is None else dist.get_world_size(g)


def get_ulysses_sequence_parallel_rank(group=None) -> int:
    g = _resolve(group)
    return 0 if g is None else dist.get_rank(g)


def _zero_extend(x: torch.Tensor, dim: int, extra: int) -> torch.Tensor:
    """`extra` zeros appended along `dim`"""
    if extra <= 0:
        return x
    dim %= x.dim()
    spec = [0, 0] * (x.dim() - 1 - dim) + [0, extra]  # F.pad counts axes from the last one
    return F.pad(x, spec)


def slice_input_tensor(x: torch.Tensor, dim: int, padding: bool = True, group=None) -> torch.Tensor:
    """This rank's contiguous 1/world share of `x` along `dim`; with ``padding`` the axis is zero-extended to a multiple
    of the group size first (verl/utils/ulysses.py:117-130)."""
    g = _resolve(group)
    world, rank = dist.get_world_size(g), dist.get_rank(g)
    if padding:
        x = _zero_extend(x, dim, -x.size(dim) % world)
    share = x.size(dim) // world
    return x.narrow(dim, rank * share, share).contiguous()


def _stacked_all_gather(src: torch.Tensor, group, async_op: bool = False):
    """ONE `all_gather_into_tensor` of a contiguous `src` into a fresh `[world, *src.shape]` buffer -> (buffer, work)."""
    world = dist.get_world_size(group)
    stacked = src.new_empty((world, *src.shape))
    work = dist.all_gather_into_tensor(stacked.view(world * src.shape[0], *src.shape[1:]) if src.dim() else stacked, src,
                                       group=group, async_op=async_op)
    return stacked, work


def all_gather_tensor(local_tensor: torch.Tensor, group=None, async_op: bool = False) -> torch.Tensor:
    """Plain (non-differentiable) gather along axis 0: rank r's rows land at `[r*n, (r+1)*n)` of the result
    (verl/utils/ulysses.py:155-162).  As there, ``async_op=True`` returns the buffer right away with the collective still
    in flight on the backend's stream: the caller orders its own use of the result."""
    g = _resolve(group)
    src = local_tensor.contiguous()
    stacked, _ = _stacked_all_gather(src, g, async_op)
    return stacked.view(stacked.shape[0] * src.shape[0], *src.shape[1:]) if src.dim() else stacked


def gather_outpus_and_unpad(*args, **kwargs):
    """The misspelt name the reference keeps only to point callers at the right one (verl/utils/ulysses.py:237-240)."""
    raise RuntimeError(
        "please use verl.utils.ulysses.gather_outputs_and_unpad instead of verl.utils.ulysses.gather_outpus_and_unpad")


class Gather(torch.autograd.Function):
    """Differentiable all-gather along `gather_dim` (verl/utils/ulysses.py:194-234).  Backward: this rank's slice of the
    incoming gradient, times the group size when ``grad_scaler`` (every sp rank holds a replica of the loss)."""

    @staticmethod
    def forward(ctx, group, local_tensor, gather_dim, grad_scaler=True, async_op=False):
        world = dist.get_world_size(group)
        ctx.meta = (dist.get_rank(group), world, gather_dim % max(local_tensor.dim(), 1), local_tensor.size(gather_dim),
                    bool(grad_scaler))
        src = local_tensor.contiguous()
        stacked, _ = _stacked_all_gather(src, group)
        axis = ctx.meta[2]
        if src.dim() == 0:
            return stacked
        if axis == 0:
            return stacked.view(world * src.shape[0], *src.shape[1:])
        # [world, d0, .., d_axis, ..] -> [d0, .., world * d_axis, ..]: rank-major along the gathered axis
        out_shape = list(src.shape)
        out_shape[axis] *= world
        return stacked.movedim(0, axis).reshape(out_shape)

    @staticmethod
    def backward(ctx, grad_output):
        rank, world, axis, width, scale = ctx.meta
        mine = grad_output.narrow(axis, rank * width, width)
        mine = mine * world if scale else mine.contiguous()
        return None, mine.contiguous(), None, None, None


def gather_outputs_and_unpad(x: torch.Tensor, gather_dim: int, unpad_dim: Optional[int] = None, padding_size: int = 0,
                             grad_scaler: bool = True, group=None) -> torch.Tensor:
    """Gather `x` over the sequence-parallel group along `gather_dim`, then cut the `padding_size` trailing entries that
    `ulysses_pad` added along `unpad_dim` (verl/utils/ulysses.py:243-275).  No group configured: `x` itself."""
    g = _resolve(group)
    if g is None:
        return x
    full = Gather.apply(g, x, gather_dim, grad_scaler)
    if unpad_dim is None:
        return full
    if not isinstance(padding_size, int):
        raise AssertionError("padding size is not given or is not an integer")
    return full if padding_size == 0 else full.narrow(unpad_dim, 0, full.size(unpad_dim) - padding_size)


def ulysses_pad(input_ids_rmpad: torch.Tensor, position_ids_rmpad: Optional[torch.Tensor] = None, sp_size: int = 1):
    """Make the token axis of the `[1, T]` id stream divisible by `sp_size` (verl/utils/ulysses.py:278-293): ids get
    zeros, position ids (`[1, T]`, or mrope `[3, 1, T]`) continue with 0, 1, .., pad-1.  -> (ids, position_ids, pad)."""
    if position_ids_rmpad is not None:
        assert position_ids_rmpad.size(-2) == 1
        assert input_ids_rmpad.size(-1) == position_ids_rmpad.size(-1)
    pad = -input_ids_rmpad.shape[1] % sp_size if sp_size > 1 else 0
    if pad == 0:
        return input_ids_rmpad, position_ids_rmpad, 0
    ids = _zero_extend(input_ids_rmpad, 1, pad)
    pos = position_ids_rmpad
    if pos is not None:
        pos = _zero_extend(pos, -1, pad)
        pos[..., -pad:] = torch.arange(pad, device=pos.device, dtype=pos.dtype)
    return ids, pos, pad


def ulysses_pad_and_slice_inputs(input_ids_rmpad: torch.Tensor, position_ids_rmpad: Optional[torch.Tensor] = None,
                                 sp_size: int = 1):
    """`ulysses_pad`, then each rank keeps its slice of dim 1 of both tensors (verl/utils/ulysses.py:296-321; dim 1 is
    the token axis of `[1, T]` position ids — the reference slices mrope position ids along the same dim)."""
    ids, pos, pad = ulysses_pad(input_ids_rmpad, position_ids_rmpad, sp_size)
    ids = slice_input_tensor(ids, dim=1, padding=False)
    if pos is not None:
        pos = slice_input_tensor(pos, dim=1, padding=False)
    return ids, pos, pad


def validate_ulysses_config(num_heads, ulysses_sequence_size):
    """the head axis must split evenly over the sequence-parallel ranks (verl/utils/ulysses.py:324-328)"""
    if ulysses_sequence_size > 1 and num_heads % ulysses_sequence_size != 0:
        raise AssertionError(f"num_heads ({num_heads}) must be divisible by ulysses sequence size({ulysses_sequence_size})")
