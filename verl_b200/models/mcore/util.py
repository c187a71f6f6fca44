"""Packed ("thd") sequence plumbing for the Megatron-side callers of the per-token path
(mirror of verl/models/mcore/util.py:23-142).

    preprocess_packed_seqs(input_ids, attention_mask, pre_process=True)    -> (packed [1, sum_padded/cp, ...], params)
    postprocess_packed_seqs(output, packed_seq_params, attention_mask, batch_size, seq_len, post_process=True)

Each sequence's valid tokens are packed back to back, every sequence padded to a multiple of tp_size (or
tp_size*cp_size*2 under context parallelism, where rank r keeps chunk r and chunk 2cp-1-r of each sequence).  The
reference does this with a host loop over the batch; here the whole layout is three small kernels (vpg_unpad_indices,
vpg_packed_seq_layout, vpg_packed_seq_map) and the data moves in ONE vpg_permute_rows pass per direction.  The
post-processing is autograd-aware (like the reference, only this rank's slice of the gathered outputs carries a
gradient).

The parallel sizes come from keyword arguments; when they are omitted and megatron.core is importable its
parallel_state is asked, as the reference does; otherwise tp = cp = 1.  `PackedSeqParams` has the fields of
megatron.core.packed_seq_params.PackedSeqParams (third-party, absent here) that the reference fills.
"""

from __future__ import annotations

from dataclasses import dataclass, field

import torch

from ... import ops


@dataclass
class PackedSeqParams:
    qkv_format: str = "thd"
    cu_seqlens_q: torch.Tensor = None
    cu_seqlens_kv: torch.Tensor = None
    cu_seqlens_q_padded: torch.Tensor = None
    cu_seqlens_kv_padded: torch.Tensor = None
    max_seqlen_q: int = None
    max_seqlen_kv: int = None
    # layout maps reused by postprocess_packed_seqs (not part of the Megatron dataclass)
    _vpg: dict = field(default=None, repr=False, compare=False)


def _parallel_sizes(tp_size, cp_size, cp_rank, cp_group):
    if tp_size is None or cp_size is None or cp_rank is None:
        try:
            from megatron.core import parallel_state as mpu
        except Exception:
            mpu = None
        if tp_size is None:
            tp_size = mpu.get_tensor_model_parallel_world_size() if mpu else 1
        if cp_size is None:
            cp_size = mpu.get_context_parallel_world_size() if mpu else 1
        if cp_rank is None:
            cp_rank = mpu.get_context_parallel_rank() if mpu else 0
        if cp_group is None and mpu is not None and cp_size > 1:
            cp_group = mpu.get_context_parallel_group()
    return int(tp_size), int(cp_size), int(cp_rank), cp_group


def _layout(attention_mask, tp_size, cp_size, cp_rank):
    mask = attention_mask if attention_mask.dtype != torch.bool else attention_mask.view(torch.uint8)
    indices, seqlens, cu_seqlens, _ = ops.unpad_indices(mask.reshape(mask.shape[0], -1))
    align = tp_size * cp_size * 2 if cp_size > 1 else tp_size
    _, cu_padded, total_padded, max_padded = ops.packed_seq_layout(seqlens, align)
    gathered_pos, local_dst = ops.packed_seq_map(cu_seqlens, cu_padded, indices.numel(), cp_size, cp_rank)
    return dict(indices=indices, cu_padded=cu_padded, total_padded=total_padded, max_padded=max_padded,
                gathered_pos=gathered_pos, local_dst=local_dst, cp_size=cp_size, cp_rank=cp_rank)


def preprocess_packed_seqs(input_ids: torch.Tensor, attention_mask: torch.Tensor, pre_process: bool = True, *,
                           tp_size: int = None, cp_size: int = None, cp_rank: int = None, cp_group=None):
    """mcore/util.py:23-87.  input_ids [B, S, ...]; attention_mask [B, S] (nonzero = valid)."""
    tp_size, cp_size, cp_rank, cp_group = _parallel_sizes(tp_size, cp_size, cp_rank, cp_group)
    lay = _layout(attention_mask, tp_size, cp_size, cp_rank)
    lay["cp_group"] = cp_group
    cu = lay["cu_padded"]
    params = PackedSeqParams(qkv_format="thd", cu_seqlens_q=cu, max_seqlen_q=lay["max_padded"], cu_seqlens_kv=cu,
                             max_seqlen_kv=lay["max_padded"], cu_seqlens_q_padded=cu, cu_seqlens_kv_padded=cu, _vpg=lay)
    if not pre_process:
        return input_ids, params
    B, S = input_ids.shape[:2]
    flat = input_ids.reshape(B * S, *input_ids.shape[2:])
    packed = ops.permute_rows(flat, lay["indices"], lay["local_dst"], lay["total_padded"] // cp_size)
    return packed.unsqueeze(0), params


class _Unpack(torch.autograd.Function):
    """packed local stream [1, P, ...] -> padded [B, S, ...] (zeros at masked positions)"""

    @staticmethod
    def forward(ctx, output, lay, batch_size, seq_len):
        cp_size = lay["cp_size"]
        local = output[0].contiguous()
        if cp_size > 1:
            # every rank's packed stream, rank-major; only data moves — the gradient flows to the local slice alone
            gathered = torch.empty((cp_size * local.shape[0], *local.shape[1:]), dtype=local.dtype, device=local.device)
            torch.distributed.all_gather_into_tensor(gathered, local, group=lay.get("cp_group"))
        else:
            gathered = local
        ctx.lay, ctx.local_rows = lay, local.shape[0]
        out = ops.permute_rows(gathered, lay["gathered_pos"], lay["indices"], batch_size * seq_len)
        return out.view(batch_size, seq_len, *local.shape[1:])

    @staticmethod
    def backward(ctx, grad):
        lay = ctx.lay
        g = grad.reshape(-1, *grad.shape[2:]).contiguous()
        d_local = ops.permute_rows(g, lay["indices"], lay["local_dst"], ctx.local_rows)
        return d_local.unsqueeze(0), None, None, None


def postprocess_packed_seqs(output: torch.Tensor, packed_seq_params, attention_mask: torch.Tensor, batch_size: int,
                            seq_len: int, post_process: bool = True, *, tp_size: int = None, cp_size: int = None,
                            cp_rank: int = None, cp_group=None) -> torch.Tensor:
    """mcore/util.py:90-142.  output [1, packed_len, ...] -> [batch_size, seq_len, ...]."""
    if not post_process:
        return output
    lay = getattr(packed_seq_params, "_vpg", None)
    if lay is None:  # params built elsewhere (e.g. Megatron's own dataclass): rebuild the maps from the mask
        tp_size, cp_size, cp_rank, cp_group = _parallel_sizes(tp_size, cp_size, cp_rank, cp_group)
        lay = _layout(attention_mask, tp_size, cp_size, cp_rank)
        lay["cp_group"] = cp_group
    return _Unpack.apply(output, lay, batch_size, seq_len)
