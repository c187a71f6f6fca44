"""Host side of the fused lm-head op — mirror of verl/utils/kernel/kernels.py (enums :46-117, host functions
efficient_entropy_forward :474-663 and efficient_entropy_backward :1345-1553), backed by libvpg_b200's tcgen05
kernels instead of Triton.

Differences a maintainer should know (all outputs match the reference within its own test tolerances):
* the three saved per-token tensors are (lse, 1, eb) in the reference's (maximum, accumulate, entropy_b) slots:
  the reference identity ``logprob = target - maximum - log(accumulate)`` holds with maximum=lse, accumulate=1;
* vocab-parallel runs exchange ONE all-gather of the packed [4,S,T] partials instead of the reference's three
  dependent all-reduces (kernels.py:612,617,644) — or, with ``set_vocab_parallel_exchange("nvlink")``, no separate
  collective at all: the GEMM epilogue stores the partials into every rank's gathered buffer over NVLink
  (multimem.st / peer stores, see peer.py), and with ``set_reduce_d_hidden(True)`` the d_hidden GEMM reduces into
  every rank's buffer through the NVSwitch (multimem.red) while it runs;
* ``BackwardEnum`` values are accepted by ``set_backward_method`` for API compatibility; the backward always uses
  the split-along-vocab strategy (the reference default, ``_Split_Dlogits_N``) on tcgen05 GEMMs.
"""

from __future__ import annotations

import os
import typing
from dataclasses import dataclass

import torch
import torch.distributed as dist

from ... import ops


@dataclass
class EntropyReductionEnum:
    """Enum for the reduction method of cross entropy (verl/utils/kernel/kernels.py:46-54)."""

    _None = 0
    _Sum = 1
    _Mean = 2


def get_entropy_reduction_enum_number(reduction: str) -> int:
    """verl/utils/kernel/kernels.py:57-70"""
    if reduction == "none":
        return EntropyReductionEnum._None
    if reduction == "sum":
        return EntropyReductionEnum._Sum
    if reduction == "mean":
        return EntropyReductionEnum._Mean
    raise ValueError(f"Invalid reduction: {reduction}")


def get_entropy_reduction_enum(ce_reduction: int) -> int:
    """verl/utils/kernel/kernels.py:73-87"""
    if ce_reduction in (0, 1, 2):
        return ce_reduction
    raise ValueError(f"Invalid ce_reduction: {ce_reduction}")


@dataclass
class BackwardEnum:
    """verl/utils/kernel/kernels.py:90-101"""

    _Total_Fuse_MN = 0
    _Total_Separate = 1
    _Split_Dlogits_N = 2
    _Split_Dlogits_M = 3


@dataclass
class Config:
    _backward: int = BackwardEnum._Split_Dlogits_N
    _use_triton: bool = False  # there is no Triton in this build
    _chunk_v: int = 0  # 0 = library default (vpg_lce_bwd_default_chunk)
    # vocab-parallel exchange: "nccl" (one all-gather of the partials) or "nvlink" (fused into the GEMM epilogue)
    _tp_exchange: str = os.environ.get("VPG_TP_EXCHANGE", "nccl")
    # False = reference semantics (caller all-reduces d_hidden); True = the op returns the summed d_hidden
    _tp_reduce_d_hidden: bool = False
    # how the op sums d_hidden when asked to: "fused" (in-kernel NVLS all-reduce inside the d_weight GEMM), "nccl" (one
    # all-reduce after the GEMMs), "overlap" (NCCL on a second stream under a d_weight GEMM that spares SMs) or
    # "multimem" (multimem.red from the d_hidden GEMM epilogue; each GPU's NVLink ingress is world x |d_hidden|:
    # ~3 ms slower than NCCL at 8 GPUs for configs[1], hence opt-in); "auto" = fused when NVSwitch multicast is there
    _tp_reduce_mode: str = os.environ.get("VPG_TP_REDUCE", "nccl")
    # "fused" mode only: True returns the symmetric slot itself as d_hidden (no copy; valid until the backward after next)
    _tp_alias_slot: bool = os.environ.get("VPG_TP_ALIAS_SLOT", "0") == "1"


_config = Config()


def set_backward_method(backward_method: int):
    """verl/utils/kernel/kernels.py:112-117.  Recorded only; see the module docstring."""
    global _config
    _config._backward = backward_method


def set_backward_chunk(chunk_v: int):
    """Vocab columns of dlogits materialised at a time in the backward (multiple of 256; 0 = default)."""
    global _config
    _config._chunk_v = int(chunk_v)


def set_vocab_parallel_exchange(mode: str):
    """"nccl": all-gather of the packed partials after the GEMM.  "nvlink": the GEMM epilogue writes the partials into
    every rank's symmetric buffer itself (NVSwitch multicast or peer-mapped stores); only a barrier follows."""
    if mode not in ("nccl", "nvlink"):
        raise ValueError(f"Invalid vocab-parallel exchange: {mode}")
    _config._tp_exchange = mode


def set_reduce_d_hidden(enabled: bool, mode: str = "nccl", alias_slot: bool | None = None):
    """True: the backward returns d_hidden already summed over the vocab-parallel group (the reference leaves that
    all-reduce to its callers).  mode "nccl": one all-reduce after the GEMMs.  mode "multimem" (needs the "nvlink"
    exchange and NVSwitch multicast): the d_hidden GEMM's epilogue reduces into every rank's fp32 buffer through the
    switch (multimem.red) while it runs.  mode "overlap": d_hidden is computed before the last d_weight GEMM, which
    spares the SMs it does not need for its number of scheduling rounds; the NCCL all-reduce runs on a second stream
    under that GEMM.  mode "fused" (needs the "nvlink" exchange, NVSwitch multicast and bf16 operands): d_hidden is written
    into a symmetric slot and all-reduced INSIDE the d_weight GEMM kernel by its idle warps (two-shot NVLS:
    multimem.ld_reduce + multimem.st), no SM and no launch of its own; the sum lands in one of two alternating symmetric
    slots and is returned as a copy (T*H bf16, small next to the GEMMs) — ``alias_slot=True`` returns the slot itself,
    valid only until the backward after next, for callers that consume d_hidden at once.  The kernel's waits for its
    peers are bounded ($VPG_AR_TIMEOUT_MS); a failed exchange raises from the next backward or from
    ``peer.workspace(group, device).check_allreduce()``.  mode "auto": "fused" when available, else
    "overlap" when `_overlap_plan` says the hidden all-reduce is worth the extra scheduling rounds of the d_weight GEMM,
    else "nccl"."""
    if mode not in ("nccl", "multimem", "overlap", "auto", "fused"):
        raise ValueError(f"Invalid d_hidden reduction mode: {mode}")
    _config._tp_reduce_d_hidden = bool(enabled)
    _config._tp_reduce_mode = mode
    if alias_slot is not None:
        _config._tp_alias_slot = bool(alias_slot)


_overlap_streams: dict = {}


def _overlap_stream(device):
    """(second stream, event, raw cudaEvent_t) of `device` for the overlapped d_hidden all-reduce."""
    idx = torch.device(device).index
    idx = torch.cuda.current_device() if idx is None else idx
    cur = _overlap_streams.get(idx)
    if cur is None:
        with torch.cuda.device(idx):
            side = torch.cuda.Stream(device=idx)
            ready = torch.cuda.Event()
            ready.record()  # torch creates the cudaEvent_t at the first record
            ready.synchronize()
        cur = _overlap_streams[idx] = (side, ready, int(ready.cuda_event))
    return cur


# NCCL's all-reduce kernel only makes progress under a persistent one-CTA-per-SM GEMM when ALL its CTAs (16 channels
# on an NVSwitch box) are resident at once; measured at 8 GPUs / configs[1]: 8 free SMs -> slower than the serial
# all-reduce, 16 free SMs -> the all-reduce disappears under the d_weight GEMM (backward 22.3 -> 20.4 ms).
_OVERLAP_PAIRS = int(os.environ.get("VPG_TP_OVERLAP_PAIRS", "8"))
_SM_TFLOPS = 9.3          # sustained bf16 TFLOP/s of one SM inside a long step (1380 / 148)
_ALLREDUCE_BUSBW = 400e9  # bytes/s, NCCL NVLS all-reduce of a few hundred MB on one NVSwitch box


def _overlap_plan(T: int, H: int, v_local: int, world: int, device, workers: int = 0):
    """-> (pays, reserve_cta_pairs): does hiding the d_hidden all-reduce under the last d_weight GEMM pay for the extra
    scheduling rounds that GEMM needs when `_OVERLAP_PAIRS` CTA pairs are kept free?  (single-chunk backward assumed:
    with several vocab chunks the last chunk's GEMM is shorter and the estimate is conservative)"""
    workers = workers or torch.cuda.get_device_properties(device).multi_processor_count // 2  # CTA pairs
    tiles = -(-v_local // 256) * -(-H // 256)
    free_rounds = -(-tiles // workers)
    left = workers - _OVERLAP_PAIRS
    if left < 1:
        return False, 0
    reserved_rounds = -(-tiles // min(left, -(-tiles // free_rounds)))
    round_ms = 2.0 * 256 * 256 * T / (2 * _SM_TFLOPS * 1e12) * 1e3
    cost_ms = (reserved_rounds - free_rounds) * round_ms
    allreduce_ms = T * H * 2 * 2.0 * (world - 1) / world / _ALLREDUCE_BUSBW * 1e3
    return allreduce_ms > 1.5 * cost_ms + 0.2, _OVERLAP_PAIRS


def _peer_workspace(group, device):
    from . import peer

    return peer.workspace(group, device)


def init_vocab_parallel(group, device, max_tokens: int, hidden_size: int, vocab_local: int = 0):
    """COLLECTIVE, explicit initialisation of the NVLink-fused vocab-parallel path: allocates (and rendezvouses) every
    symmetric buffer for up to ``max_tokens`` tokens now, so that no forward / backward allocates later.  Without it
    the buffers are created on first use (also collectively; a size disagreement between ranks raises)."""
    S = ops.lce_num_splits(max_tokens, hidden_size, vocab_local) if vocab_local else 16
    return _peer_workspace(group, device).reserve(int(max_tokens), int(hidden_size), max(int(S), 1))


def _fused_allreduce_ok(hidden, world) -> bool:
    return (world > 1 and _config._tp_reduce_d_hidden and _config._tp_reduce_mode in ("fused", "auto")
            and _config._tp_exchange == "nvlink" and hidden.dtype == torch.bfloat16)


def _check_inputs(hidden, weight, labels):
    # same preconditions as verl/utils/kernel/kernels.py:485-490,504
    assert hidden.is_cuda and weight.is_cuda and labels.is_cuda
    assert weight.device == hidden.device and labels.device == hidden.device
    assert hidden.dim() == 2 and weight.dim() == 2 and labels.dim() == 1
    assert hidden.is_contiguous() and weight.is_contiguous() and labels.is_contiguous()
    assert hidden.shape[0] == labels.shape[0] and hidden.shape[1] == weight.shape[1]
    assert hidden.shape[1] % 64 == 0  # the reference asks for % 128 (kernels.py:504); one 128-byte TMA row here
    assert hidden.dtype == weight.dtype and hidden.dtype in (torch.bfloat16, torch.float16, torch.float32), (
        "the sm_100a kernels take bf16 / fp16 (tcgen05 kind::f16) or fp32 (kind::tf32) operands, both alike")
    assert labels.dtype == torch.int64


def exchange_partials(partials: torch.Tensor, group) -> torch.Tensor:
    """Vocab-parallel exchange: ONE all-gather of the packed per-token partials {m, s, b, t} ([4,S,T] per rank)
    -> [4, world*S, T].  The merge done by vpg_lce_finalize is associative and commutative, so rank order is free.
    Replaces the reference's all_reduce(MAX) + all_reduce(SUM) + all_reduce(SUM) (kernels.py:612,617,644)."""
    world = dist.get_world_size(group)
    four, S, T = partials.shape
    gathered = torch.empty((world * four, S, T), dtype=partials.dtype, device=partials.device)
    dist.all_gather_into_tensor(gathered, partials.contiguous(), group=group)
    return gathered.view(world, four, S, T).permute(1, 0, 2, 3).reshape(four, world * S, T).contiguous()


def efficient_entropy_forward(
    hidden: torch.Tensor,
    weight: torch.Tensor,
    labels: torch.Tensor,
    reduction: typing.Optional[int] = 2,
    temperature: typing.Optional[float] = 1.0,
    dist_process_group: typing.Optional[dist.ProcessGroup] = None,
) -> list[torch.Tensor]:
    """forward host function — verl/utils/kernel/kernels.py:474-663.

    Returns (logprobs, entropy, maximum, accumulate, entropy_b) like the reference.
    """
    _check_inputs(hidden, weight, labels)
    REDUCTION = get_entropy_reduction_enum(reduction)
    rank = 0 if dist_process_group is None else dist.get_rank(dist_process_group)
    world = 1 if dist_process_group is None else dist.get_world_size(dist_process_group)
    vocab_local = weight.shape[0]

    if world > 1 and _config._tp_exchange == "nvlink" and hidden.dtype == torch.bfloat16:  # push kernels: bf16 only
        # fused GEMM + all-gather: every rank's epilogue pushes its partials into all gathered buffers over NVLink
        T, H = hidden.shape
        S = ops.lce_num_splits(T, H, vocab_local)
        px = _peer_workspace(dist_process_group, hidden.device)
        gathered, mc_ptr, peer_ptrs = px.partials(S, T)
        ops.lce_forward_partials_push(hidden, weight, labels, temperature, rank * vocab_local, S, gathered,
                                      part_split0=rank * S, multicast_ptr=mc_ptr, peer_ptrs=peer_ptrs)
        px.barrier()
        partials = gathered
    else:
        partials = ops.lce_forward_partials(hidden, weight, labels, temperature, label_offset=rank * vocab_local)
        if world > 1:
            partials = exchange_partials(partials, dist_process_group)
    logprobs, entropy, lse, eb = ops.lce_finalize(partials, temperature)

    if REDUCTION == EntropyReductionEnum._Sum:
        logprobs = logprobs.sum()
    elif REDUCTION == EntropyReductionEnum._Mean:
        logprobs = logprobs.sum() / logprobs.numel()
    accumulate = torch.ones_like(lse)
    return (logprobs, entropy, lse, accumulate, eb)


def efficient_entropy_backward(
    dlogprobs: torch.Tensor,
    dentropy: torch.Tensor,
    hidden: torch.Tensor,
    weight: torch.Tensor,
    labels: torch.Tensor,
    maximum: torch.Tensor,
    acc: torch.Tensor,
    entropy_b: torch.Tensor,
    reduction: typing.Optional[int] = 2,
    should_return_fp32_grad: bool = False,
    temperature: typing.Optional[float] = 1.0,
    dist_process_group: typing.Optional[dist.ProcessGroup] = None,
) -> list[torch.Tensor]:
    """backward host function — verl/utils/kernel/kernels.py:1345-1553.

    d_hidden is NOT all-reduced across a vocab-parallel group, exactly like the reference (its callers do it:
    tests/utils/test_linear_cross_entropy_tp.py:313-314).
    """
    _check_inputs(hidden, weight, labels)
    REDUCTION = get_entropy_reduction_enum(reduction)
    num_tokens = hidden.shape[0]
    rank = 0 if dist_process_group is None else dist.get_rank(dist_process_group)

    g_lp = None
    if dlogprobs is not None:
        if REDUCTION == EntropyReductionEnum._None:
            assert dlogprobs.shape == (num_tokens,)
            g_lp = dlogprobs.to(torch.float32).contiguous()
        else:
            assert dlogprobs.dim() == 0
            g = dlogprobs.to(torch.float32)
            if REDUCTION == EntropyReductionEnum._Mean:
                g = g / num_tokens
            g_lp = g.expand(num_tokens).contiguous()
    g_ent = None
    if dentropy is not None:
        assert dentropy.shape == (num_tokens,)
        g_ent = dentropy.to(torch.float32).contiguous()

    # maximum + log(acc) is the log-sum-exp for both this module's and the reference's saved triple
    lse = maximum if acc is None else maximum + torch.log(acc)
    world = 1 if dist_process_group is None else dist.get_world_size(dist_process_group)
    fused_reduce = False
    if (world > 1 and _config._tp_reduce_d_hidden and _config._tp_reduce_mode == "multimem"
            and _config._tp_exchange == "nvlink" and hidden.dtype == torch.bfloat16):
        px = _peer_workspace(dist_process_group, hidden.device)
        accum, accum_mc = px.d_hidden_accum(hidden.shape[0], hidden.shape[1])
        fused_reduce = accum_mc != 0
    ar = None
    if not fused_reduce and _fused_allreduce_ok(hidden, world):
        px = _peer_workspace(dist_process_group, hidden.device)
        if px.has_multicast:
            ar = px.d_hidden_slot(hidden.shape[0], hidden.shape[1])
        elif _config._tp_reduce_mode == "fused":
            raise RuntimeError("set_reduce_d_hidden(mode='fused') needs NVSwitch multicast on this group")
    if ar is not None:
        # GEMM + all-reduce in ONE kernel: d_hidden lands in the symmetric slot, the d_weight GEMM's idle warps sum it
        # over the ranks through the switch while the tensor cores run; kernel completion = reduction complete
        d_hidden, d_weight = ops.lce_backward(
            hidden, weight, labels, lse.contiguous(), entropy_b.contiguous(), g_lp, g_ent, temperature,
            label_offset=rank * weight.shape[0], chunk_v=_config._chunk_v, allreduce=ar)
        px.commit_d_hidden_slot(ar)  # epoch / slot parity advance only once the launch went through
        if not _config._tp_alias_slot:
            # the slot is overwritten by the backward after next; autograd may adopt the returned tensor as hidden.grad
            # without a copy, and torch.autograd.grad hands it to the caller: give them memory of their own
            d_hidden = d_hidden.clone()
    elif fused_reduce:
        # fused GEMM + all-reduce: zero every rank's fp32 buffer, barrier, the d_hidden GEMM's epilogue reduces into all
        # of them through the switch (multimem.red), barrier, convert this rank's copy
        accum.zero_()
        px.barrier()
        d_weight = ops.lce_backward_mcred(
            hidden, weight, labels, lse.contiguous(), entropy_b.contiguous(), g_lp, g_ent, temperature,
            rank * weight.shape[0], accum, accum_mc, chunk_v=_config._chunk_v,
        )
        px.barrier()
        d_hidden = ops.cast_f32_bf16(accum)
    else:
        reduce = world > 1 and _config._tp_reduce_d_hidden
        overlap, reserve = False, 0
        if reduce and _config._tp_reduce_mode in ("overlap", "auto") and hidden.dtype != torch.float32:
            pays, reserve = _overlap_plan(hidden.shape[0], hidden.shape[1], weight.shape[0], world, hidden.device)
            overlap = pays or _config._tp_reduce_mode == "overlap"
        side = ready = None
        if overlap:
            side, ready, ready_ptr = _overlap_stream(hidden.device)
        d_hidden, d_weight = ops.lce_backward(
            hidden, weight, labels, lse.contiguous(), entropy_b.contiguous(), g_lp, g_ent, temperature,
            label_offset=rank * weight.shape[0], chunk_v=_config._chunk_v,
            d_hidden_ready_event=ready_ptr if overlap else 0, reserve_cta_pairs=reserve if overlap else 0,
        )
        if overlap:
            # the kernels recorded `ready` right behind the d_hidden GEMM: the all-reduce starts there, on the SMs the
            # d_weight GEMM spared, and the caller's stream joins after both
            main = torch.cuda.current_stream(hidden.device)
            side.wait_event(ready)
            with torch.cuda.stream(side):
                dist.all_reduce(d_hidden, group=dist_process_group)
            main.wait_stream(side)
        elif reduce:
            dist.all_reduce(d_hidden, group=dist_process_group)
    if should_return_fp32_grad:
        d_hidden, d_weight = d_hidden.float(), d_weight.float()
    return d_hidden, d_weight
