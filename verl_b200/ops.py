"""Thin tensor-level wrappers over the C ABI (include/vpg_b200.h).

Everything here takes/returns torch CUDA tensors, allocates outputs and workspaces with torch's caching allocator
and enqueues on the current stream.  No op has a CPU or PyTorch fallback: a non-CUDA input raises.
The reference-named Python API (verl_b200.utils.*, verl_b200.trainer.*) is built on these.
"""

from __future__ import annotations

import ctypes
import math

import threading

import torch

from . import _lib
from ._lib import ADV_KINDS, AGG_MODES, KL_KINDS, PG_KINDS, check, mask_arg, on_device, ptr, stream_ptr

_F32 = torch.float32


def _lib_():
    return _lib.load()


def _require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("verl_b200 ops require CUDA tensors (there is no CPU fallback; the reference-named "
                               "driver-side functions stage host tensors through verl_b200.staging)")


def _f32c(t: torch.Tensor | None) -> torch.Tensor | None:
    if t is None:
        return None
    if t.dtype != _F32:
        t = t.to(_F32)
    return t.contiguous()


def _ws(nbytes: int, device) -> torch.Tensor:
    # float64 storage keeps the 8-byte alignment the kernels ask for (torch allocations are 512-byte aligned)
    return torch.empty((max(int(nbytes), 8) + 7) // 8, dtype=torch.float64, device=device)


# --------------------------------------------------------------------------------------------------- lm-head
def lce_num_splits(T: int, H: int, V: int) -> int:
    return int(_lib_().vpg_lce_num_splits(T, H, V, 0))


def lce_forward_partials(hidden, weight, labels, temperature: float, label_offset: int = 0, num_splits: int = 0):
    """-> partials [4, S, T] fp32 (see vpg_lce_fwd_partials)."""
    _require_cuda(hidden, weight, labels)
    T, H = hidden.shape
    V = weight.shape[0]
    S = num_splits or lce_num_splits(T, H, V)
    part = torch.empty((4, S, T), dtype=_F32, device=hidden.device)
    with on_device(hidden.device):
        check(_lib_().vpg_lce_fwd_partials_dt(operand_dtype(hidden, weight), ptr(hidden), ptr(weight), ptr(labels), T,
                                              H, V, float(temperature), int(label_offset), S, ptr(part),
                                              stream_ptr(hidden.device)))
    return part


_BWD_MODES = {0: "single block", 1: "vocab chunks (fp32 d_hidden accumulator)", 2: "token chunks (fp32 d_weight accumulator)"}


def backward_plan(T: int, H: int, V: int, chunk_v: int = 0, chunk_t: int = 0) -> dict:
    """How the backward will cut its dlogits workspace (vpg_lce_bwd_plan): mode, chunk, chunks, workspace_bytes."""
    import ctypes

    out = (ctypes.c_int64 * 4)()
    check(_lib_().vpg_lce_bwd_plan(int(T), int(H), int(V), int(chunk_v), int(chunk_t), out))
    return {"mode": _BWD_MODES[int(out[0])], "chunk": int(out[1]), "chunks": int(out[2]), "workspace_bytes": int(out[3])}


def backward_chunk(T: int, H: int, V: int, device, chunk_v: int = 0) -> int:
    """Explicit column-chunk width to pass to the backward, or 0 = let the library plan.  An explicit ``chunk_v`` wins.
    Otherwise the library bounds the 16-bit dlogits block by $VPG_BWD_WS_MB (default 3072 MiB — what the
    reference's own default backward peaks at for 64k tokens: 9 504-column dlogits splits, kernels.py:1492-1501, plus its
    outputs = 3.1 GB) and chunks along the vocabulary or along
    the tokens, whichever needs the smaller fp32 accumulator.  Only under memory pressure (the planned workspace exceeds
    half of what is free right now) this falls back to narrower vocabulary chunks.  The whole-vocabulary single block
    (20 GB at BASELINE configs[1]) is opt-in: ``kernels.set_backward_chunk`` or the environment variable."""
    if chunk_v and chunk_v > 0:
        return int(chunk_v)
    lib = _lib_()
    # what this process can still get: free on the device + what torch's caching allocator holds but is not using
    free, _total = torch.cuda.mem_get_info(device)
    avail = free + max(0, torch.cuda.memory_reserved(device) - torch.cuda.memory_allocated(device))
    if int(lib.vpg_lce_bwd_workspace_bytes(T, H, V, 0)) <= avail // 2:
        return 0
    chunk = int(lib.vpg_lce_bwd_default_chunk(T, H, V))
    while chunk > 256 and int(lib.vpg_lce_bwd_workspace_bytes(T, H, V, chunk)) > avail // 2:
        chunk = max(256, (chunk // 2) // 256 * 256)
    need = int(lib.vpg_lce_bwd_workspace_bytes(T, H, V, chunk))
    if need > avail:
        # narrower chunks only shrink the dlogits block; the fp32 [T,H] accumulator of column chunking stays
        raise RuntimeError(
            f"lm-head backward: even {chunk}-column dlogits chunks need a {need / 2**20:.0f} MiB workspace (of which "
            f"{T * H * 4 / 2**20:.0f} MiB is the fp32 [T,H] d_hidden accumulator, which chunking cannot shrink) but only "
            f"{avail / 2**20:.0f} MiB are available on {device}; split the micro-batch (fewer tokens per call)")
    return chunk


def operand_dtype(hidden, weight) -> int:
    """VPG_DT_* of the lm-head operands, both tensors alike: bf16 / fp16 (tcgen05 kind::f16) or fp32 (kind::tf32)."""
    if hidden.dtype != weight.dtype:
        raise TypeError(f"hidden ({hidden.dtype}) and weight ({weight.dtype}) must have the same dtype")
    if hidden.dtype == torch.bfloat16:
        return 0
    if hidden.dtype == torch.float16:
        return 1
    if hidden.dtype == torch.float32:
        return 2
    raise TypeError(f"the sm_100a lm-head kernels take bf16, fp16 or fp32 operands, got {hidden.dtype}")


def lce_forward_partials_push(hidden, weight, labels, temperature: float, label_offset: int, num_splits: int,
                              gathered, part_split0: int, multicast_ptr: int = 0, peer_ptrs=None):
    """Vocab-parallel stage 1 with the all-gather fused into the GEMM epilogue (vpg_lce_fwd_partials_push).

    ``gathered`` is this rank's [4, P, T] fp32 view of the symmetric buffer (P = world * num_splits); the kernel writes
    parts [part_split0, part_split0 + num_splits) of EVERY rank's buffer through ``multicast_ptr`` (NVSwitch
    multicast) or, when that is 0, through ``peer_ptrs`` (peer-mapped addresses of the same view on all ranks)."""
    import ctypes

    _require_cuda(hidden, weight, labels, gathered)
    T, H = hidden.shape
    V = weight.shape[0]
    four, P, T2 = gathered.shape
    assert four == 4 and T2 == T and gathered.is_contiguous() and gathered.dtype == _F32
    if multicast_ptr:
        dst, is_mc, arr, n_push = int(multicast_ptr), 1, None, 0
    else:
        ptrs = [int(x) for x in (peer_ptrs or [])]
        if not ptrs:
            raise ValueError("lce_forward_partials_push needs a multicast address or the peer addresses")
        arr = (ctypes.c_void_p * len(ptrs))(*ptrs)
        dst, is_mc, n_push = gathered.data_ptr(), 0, len(ptrs)
    with on_device(hidden.device):
        check(_lib_().vpg_lce_fwd_partials_push(ptr(hidden), ptr(weight), ptr(labels), T, H, V, float(temperature),
                                                int(label_offset), int(num_splits), dst, int(part_split0), int(P),
                                                is_mc, arr, n_push, stream_ptr(hidden.device)))
    return gathered


def _bwd_workspace(T, H, V, device, chunk_v):
    lib = _lib_()
    chunk_v = backward_chunk(T, H, V, device, chunk_v)
    nbytes = int(lib.vpg_lce_bwd_workspace_bytes(T, H, V, int(chunk_v)))
    return chunk_v, nbytes, torch.empty(max(nbytes, 256), dtype=torch.uint8, device=device)


def lce_backward_workspace_info(T: int, H: int, V: int, device, chunk_v: int = 0) -> dict:
    """{mode, chunk, chunks, workspace_bytes} the backward would use right now (bench.py reports it)."""
    return backward_plan(T, H, V, backward_chunk(T, H, V, device, chunk_v))


def lce_backward_mcred(hidden, weight, labels, lse, eb, g_lp, g_ent, temperature: float, label_offset: int,
                       accum_local, accum_multicast_ptr: int, chunk_v: int = 0):
    """Backward whose d_hidden GEMM reduces into the multicast fp32 [T,H] buffer of every rank (vpg_lce_bwd_mcred).
    -> d_weight bf16 [V,H]; the summed d_hidden is in ``accum_local`` once all ranks have passed a barrier."""
    _require_cuda(hidden, weight, labels, lse, eb, accum_local)
    T, H = hidden.shape
    V = weight.shape[0]
    assert accum_local.shape == (T, H) and accum_local.dtype == _F32 and accum_multicast_ptr
    lib = _lib_()
    chunk_v, nbytes, ws = _bwd_workspace(T, H, V, hidden.device, chunk_v)
    d_weight = torch.empty_like(weight)
    with on_device(hidden.device):
        check(lib.vpg_lce_bwd_mcred(ptr(hidden), ptr(weight), ptr(labels), ptr(lse), ptr(eb), ptr(g_lp), ptr(g_ent),
                                    T, H, V, float(temperature), int(label_offset), int(chunk_v), None, ptr(d_weight),
                                    ptr(ws), nbytes, int(accum_multicast_ptr), stream_ptr(hidden.device)))
    return d_weight


def cast_f32_bf16(src):
    _require_cuda(src)
    assert src.dtype == _F32 and src.is_contiguous()
    out = torch.empty(src.shape, dtype=torch.bfloat16, device=src.device)
    with on_device(src.device):
        check(_lib_().vpg_cast_f32_bf16(ptr(src), ptr(out), src.numel(), stream_ptr(src.device)))
    return out


def lce_finalize(partials, temperature: float):
    """partials [4, P, T] -> (logprobs, entropy, lse, eb), each fp32 [T]."""
    _require_cuda(partials)
    _, P, T = partials.shape
    out = torch.empty((4, T), dtype=_F32, device=partials.device)
    with on_device(partials.device):
        check(_lib_().vpg_lce_finalize(ptr(partials), P, T, float(temperature), ptr(out[0]), ptr(out[1]), ptr(out[2]),
                                       ptr(out[3]), stream_ptr(partials.device)))
    return out[0], out[1], out[2], out[3]


def lce_backward(hidden, weight, labels, lse, eb, g_lp, g_ent, temperature: float, label_offset: int = 0,
                 chunk_v: int = 0, d_hidden_ready_event: int = 0, reserve_cta_pairs: int = 0, allreduce=None,
                 chunk_t: int = 0):
    """-> (d_hidden [T,H], d_weight [V,H]) in the operands' dtype (fp32 gradients for fp32 operands).  One call of
    vpg_lce_bwd_ex: every option is an explicit field of its argument struct.

    d_hidden_ready_event (raw cudaEvent_t, 0 = none): recorded as soon as d_hidden is complete; the d_weight GEMM then
    runs last and spares SMs (at least reserve_cta_pairs CTA pairs) for the caller's all-reduce.
    allreduce: dict(out=[T,H] bf16 view of a symmetric buffer, mc=its multicast address, flags=uint32 tensor,
    peer_flags=[addresses], world, rank, epoch) — d_hidden is written into ``out`` and summed over the ranks by the
    d_weight GEMM kernel itself (in-kernel NVLS all-reduce); ``out`` is returned as d_hidden."""
    import ctypes

    _require_cuda(hidden, weight, labels, lse, eb)
    T, H = hidden.shape
    V = weight.shape[0]
    lib = _lib_()
    if chunk_t and not chunk_v:  # explicit row chunks (multiple of 256)
        nbytes = backward_plan(T, H, V, 0, chunk_t)["workspace_bytes"]
        ws = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=hidden.device)
    else:
        chunk_t = 0
        chunk_v, nbytes, ws = _bwd_workspace(T, H, V, hidden.device, chunk_v)
    dt = operand_dtype(hidden, weight)
    d_hidden = allreduce["out"] if allreduce else torch.empty_like(hidden)
    d_weight = torch.empty_like(weight)
    a = _lib.LceBwdArgs()
    a.struct_size = ctypes.sizeof(_lib.LceBwdArgs)
    a.dtype = dt
    keep = []
    if dt == 2:  # fp32: TF32 recompute on the fp32 operands, gradient GEMMs on bf16 copies, fp32 gradients
        h16, w16 = hidden.to(torch.bfloat16), weight.to(torch.bfloat16)
        keep += [h16, w16]
        a.hidden16, a.weight16 = ptr(h16), ptr(w16)
    a.hidden, a.weight, a.labels, a.lse, a.eb = ptr(hidden), ptr(weight), ptr(labels), ptr(lse), ptr(eb)
    a.g_lp, a.g_ent = ptr(g_lp), ptr(g_ent)
    a.T, a.H, a.V = T, H, V
    a.temperature, a.label_offset, a.chunk_v, a.chunk_t = float(temperature), int(label_offset), int(chunk_v), int(chunk_t)
    a.d_hidden, a.d_weight = ptr(d_hidden), ptr(d_weight)
    a.workspace, a.workspace_bytes = ptr(ws), nbytes
    a.stream = stream_ptr(hidden.device)
    if d_hidden_ready_event and dt != 2:
        a.d_hidden_ready_event, a.reserve_cta_pairs = int(d_hidden_ready_event), int(reserve_cta_pairs)
    if allreduce:
        assert dt == 0 and d_hidden.shape == (T, H) and d_hidden.dtype == torch.bfloat16 and d_hidden.is_contiguous()
        peers = [int(x) for x in allreduce["peer_flags"]]
        arr = (ctypes.c_void_p * len(peers))(*peers)
        keep.append(arr)
        a.ar_world, a.ar_rank, a.ar_epoch = int(allreduce["world"]), int(allreduce["rank"]), int(allreduce["epoch"])
        a.ar_d_hidden_mc, a.ar_flags = int(allreduce["mc"]), ptr(allreduce["flags"])
        a.ar_peer_flags = ctypes.cast(arr, ctypes.POINTER(ctypes.c_void_p))
    with on_device(hidden.device):
        check(lib.vpg_lce_bwd_ex(ctypes.byref(a)))
    return d_hidden, d_weight


def gemm_bf16(A, B, a_kmajor=True, b_kmajor=True, out_dtype=torch.bfloat16, split_k: bool = False):
    """C = op(A) @ op(B)^T on the tcgen05 mainloop.  A: [M,K] (k-major) or [K,M]; B: [N,K] (k-major) or [K,N]."""
    _require_cuda(A, B)
    M, K = (A.shape if a_kmajor else A.shape[::-1])
    N, K2 = (B.shape if b_kmajor else B.shape[::-1])
    assert K == K2
    C = torch.empty((M, N), dtype=out_dtype, device=A.device)
    mode = 0 if out_dtype == torch.bfloat16 else 1
    lib = _lib_()
    with on_device(A.device):
        if split_k:
            nbytes = int(lib.vpg_gemm_splitk_workspace_bytes())
            ws = torch.empty(nbytes, dtype=torch.uint8, device=A.device)
            check(lib.vpg_gemm_bf16_ws(ptr(A), ptr(B), M, N, K, A.stride(0), B.stride(0), int(a_kmajor), int(b_kmajor),
                                       ptr(C), C.stride(0), mode, None, 0, ptr(ws), nbytes, stream_ptr(A.device)))
        else:
            check(lib.vpg_gemm_bf16(ptr(A), ptr(B), M, N, K, A.stride(0), B.stride(0), int(a_kmajor), int(b_kmajor),
                                    ptr(C), C.stride(0), mode, None, 0, stream_ptr(A.device)))
    return C


# --------------------------------------------------------------------------------------------------- logits
def logits_stats(logits2d, labels=None, want_logprobs=True, want_entropy=True):
    """logits [T,V] (fp32 or bf16, last dim contiguous) -> (logprobs|None, entropy|None, lse), fp32 [T]."""
    _require_cuda(logits2d, labels)
    if logits2d.dtype not in (torch.float32, torch.bfloat16):
        logits2d = logits2d.float()
    if logits2d.stride(-1) != 1:
        logits2d = logits2d.contiguous()
    T, V = logits2d.shape
    dev = logits2d.device
    lp = torch.empty(T, dtype=_F32, device=dev) if (want_logprobs and labels is not None) else None
    ent = torch.empty(T, dtype=_F32, device=dev) if want_entropy else None
    lse = torch.empty(T, dtype=_F32, device=dev)
    ld = logits2d.stride(0) if T > 1 else V
    with on_device(dev):
        check(_lib_().vpg_logits_stats(logits2d.data_ptr(), int(logits2d.dtype == torch.bfloat16), T, V, ld,
                                       ptr(labels) if lp is not None else None, ptr(lp), ptr(ent), ptr(lse),
                                       stream_ptr(dev)))
    return lp, ent, lse


def logits_partials(logits2d, labels=None, label_offset: int = 0):
    """this rank's vocab shard of materialised logits [T,V_local] -> partials f32 [4,1,T] (the lm-head wire format)."""
    _require_cuda(logits2d, labels)
    if logits2d.dtype not in (torch.float32, torch.bfloat16):
        logits2d = logits2d.float()
    if logits2d.stride(-1) != 1:
        logits2d = logits2d.contiguous()
    T, V = logits2d.shape
    part = torch.empty((4, 1, T), dtype=_F32, device=logits2d.device)
    ld = logits2d.stride(0) if T > 1 else V
    with on_device(logits2d.device):
        check(_lib_().vpg_logits_partials(logits2d.data_ptr(), int(logits2d.dtype == torch.bfloat16), T, V, ld,
                                          ptr(labels), int(label_offset), ptr(part), stream_ptr(logits2d.device)))
    return part


def logits_stats_backward(logits2d, labels, lse, entropy, g_lp, g_ent):
    _require_cuda(logits2d)
    if logits2d.dtype not in (torch.float32, torch.bfloat16):
        raise TypeError(f"logits_stats_backward reads fp32 or bf16 logits, got {logits2d.dtype} (upcast first)")
    T, V = logits2d.shape
    assert logits2d.is_contiguous()
    d = torch.empty_like(logits2d)
    with on_device(logits2d.device):
        check(_lib_().vpg_logits_stats_bwd(ptr(logits2d), int(logits2d.dtype == torch.bfloat16), T, V, V, ptr(labels),
                                           ptr(lse), ptr(entropy), ptr(g_lp), ptr(g_ent), ptr(d),
                                           stream_ptr(logits2d.device)))
    return d


# --------------------------------------------------------------------------------------------------- rmpad
def unpad_indices_launch(mask2d):
    """Enqueue the two unpad kernels and return WITHOUT waiting: (indices_full i64 [B*S] — the first nnz entries are the
    result —, seqlens i32 [B], cu_seqlens i32 [B+1], slot).  ``slot.pair`` = {nnz, max_seqlen} in pinned host memory, valid
    once the stream has passed the launch (``unpad_indices`` waits for it; a caller with other work to enqueue first can
    wait later)."""
    _require_cuda(mask2d)
    if mask2d.dtype.is_floating_point:
        raise TypeError("attention masks must be integer or bool")
    if mask2d.dtype not in (torch.int64, torch.int32, torch.bool, torch.uint8):
        mask2d = mask2d.to(torch.int64)
    mask2d = mask2d.contiguous()
    m, mdt = mask_arg(mask2d)
    B, S = mask2d.shape
    dev = mask2d.device
    lib = _lib_()
    n = B * S
    indices_full = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
    seqlens = torch.empty(B, dtype=torch.int32, device=dev)
    cu = torch.empty(B + 1, dtype=torch.int32, device=dev)
    slot = _StatusSlot.get(dev)  # the kernels write {nnz, max_seqlen} straight into pinned host memory (UVA)
    nbytes = int(lib.vpg_unpad_workspace_bytes(n))
    ws = _ws(nbytes, dev)
    with on_device(dev):
        check(lib.vpg_unpad_indices(ptr(m), mdt, B, S, ptr(indices_full), ptr(seqlens), ptr(cu), slot.pair_ptr, ptr(ws),
                                    ws.numel() * 8, stream_ptr(dev)))
    return indices_full, seqlens, cu, slot


def unpad_indices(mask2d):
    """mask [B,S] -> (indices i64 [nnz], seqlens i32 [B], cu_seqlens i32 [B+1], max_seqlen int).

    One device->host read of {nnz, max_seqlen} (the reference syncs here too: nonzero() and .item()).
    """
    _require_cuda(mask2d)
    B, S = mask2d.shape
    dev = mask2d.device
    if B * S == 0:  # nothing to scan: no launch
        if mask2d.dtype.is_floating_point:
            raise TypeError("attention masks must be integer or bool")
        return (torch.empty(0, dtype=torch.int64, device=dev), torch.zeros(B, dtype=torch.int32, device=dev),
                torch.zeros(B + 1, dtype=torch.int32, device=dev), 0)
    indices_full, seqlens, cu, slot = unpad_indices_launch(mask2d)
    check(_lib_().vpg_stream_synchronize(stream_ptr(dev)))  # ctypes drops the GIL; no Stream object on the way
    nnz, max_seqlen = slot.read_pair()
    indices = indices_full[:nnz]
    mark_ascending(indices)  # provenance mark: pad_input may then fill gaps in one pass (see scatter_rows)
    return indices, seqlens, cu, max_seqlen


def mark_ascending(indices) -> None:
    """Record that `indices` is strictly ascending and in range AS OF NOW (its autograd version counter is stored with
    the mark; views share the counter with their base, so an in-place write through either one invalidates it)."""
    indices._vpg_ascending = ("version", indices._version)


def is_marked_ascending(indices) -> bool:
    return getattr(indices, "_vpg_ascending", None) == ("version", indices._version)


def response_rows(indices, ids_rmpad, S: int, L: int):
    """Response-window selection on the rmpad stream (vpg_response_rows + compaction).

    -> (sel i64 [T_sel] ascending stream positions inside the window, dst i64 [T_sel] rows of the [B*L] response
        matrix, labels i64 [T] = roll(ids_rmpad, -1)).  One device->host read (T_sel)."""
    _require_cuda(indices, ids_rmpad)
    T = indices.numel()
    dev = indices.device
    keep = torch.empty(max(T, 1), dtype=torch.uint8, device=dev)
    dst = torch.empty(max(T, 1), dtype=torch.int64, device=dev)
    labels = torch.empty(max(T, 1), dtype=torch.int64, device=dev)
    with on_device(dev):
        check(_lib_().vpg_response_rows(ptr(indices.contiguous()), ptr(ids_rmpad.contiguous()), T, int(S), int(L),
                                        ptr(keep), ptr(dst), ptr(labels), stream_ptr(dev)))
    sel, _, _, _ = unpad_indices(keep[:T].view(1, T)) if T else (indices[:0], None, None, 0)
    return sel, gather_rows(dst[:T], sel), labels[:T]


def gather_rows(src, indices):
    """src [N, ...] contiguous, indices i64 [n] -> [n, ...]  (index_first_axis)."""
    _require_cuda(src, indices)
    src = src.contiguous()
    row_shape = src.shape[1:]
    row_bytes = src.element_size() * int(math.prod(row_shape)) if len(row_shape) else src.element_size()
    out = torch.empty((indices.numel(), *row_shape), dtype=src.dtype, device=src.device)
    if indices.numel() and row_bytes:
        with on_device(src.device):
            check(_lib_().vpg_gather_rows(ptr(src), ptr(indices.contiguous()), indices.numel(), row_bytes, ptr(out),
                                          stream_ptr(src.device)))
    return out


def scatter_rows(src, indices, n_out: int):
    """zeros [n_out, ...] with out[indices[i]] = src[i]  (pad_input / index_put_first_axis).

    Indices that come straight out of ``unpad_indices`` (strictly ascending by construction; the tensor object carries
    the mark ``_vpg_ascending``) take the ONE-pass kernel: every output row — scattered row or zero gap — is written
    exactly once.  Any other index tensor (the API does not promise an order) takes memset + scatter — and so does a
    marked tensor that was edited in place afterwards: the mark records the tensor's version counter and lapses when
    the counter has moved (``indices.add_()``, ``indices[...] = ``, ``copy_``, an in-place sort)."""
    _require_cuda(src, indices)
    fill = 2 if is_marked_ascending(indices) else 1
    src = src.contiguous()
    row_shape = src.shape[1:]
    row_bytes = src.element_size() * int(math.prod(row_shape)) if len(row_shape) else src.element_size()
    out = torch.empty((n_out, *row_shape), dtype=src.dtype, device=src.device)
    if n_out and row_bytes:
        with on_device(src.device):
            check(_lib_().vpg_scatter_rows(ptr(src), ptr(indices.contiguous()), indices.numel(), row_bytes, ptr(out),
                                           n_out, fill, stream_ptr(src.device)))
    return out


def row_lengths(mask2d):
    """mask [B,S] -> int64 [B] exact integer row sums (attention_mask.sum(dim=1))."""
    _require_cuda(mask2d)
    if mask2d.dtype.is_floating_point:
        raise TypeError("attention masks must be integer or bool")
    if mask2d.dtype not in (torch.int64, torch.int32, torch.bool, torch.uint8):
        mask2d = mask2d.to(torch.int64)
    mask2d = mask2d.contiguous()
    m, mdt = mask_arg(mask2d)
    B, S = mask2d.shape
    out = torch.empty(B, dtype=torch.int64, device=mask2d.device)
    if B:
        with on_device(mask2d.device):
            check(_lib_().vpg_row_lengths(ptr(m), mdt, B, S, ptr(out), stream_ptr(mask2d.device)))
    return out


def seqlen_balanced_partitions(seqlen_list, k_partitions: int, equal_size: bool):
    """HOST scheduler (vpg_seqlen_balanced_partitions): Karmarkar-Karp partitions of `seqlen_list` into k index lists,
    bit-identical with verl/utils/seqlen_balancing.py:150-190 (each list sorted ascending)."""
    import ctypes as C

    n = len(seqlen_list)
    k = int(k_partitions)
    # the reference's assertions (seqlen_balancing.py:175,102)
    assert n >= k, f"number of items:[{n}] < k_partitions:[{k}]"
    if equal_size:
        assert n % k == 0, f"{n} % {k} != 0"
    lens = (C.c_int64 * max(n, 1))(*[int(x) for x in seqlen_list])
    idx = (C.c_int64 * max(n, 1))()
    off = (C.c_int64 * (k + 1))()
    rc = _lib_().vpg_seqlen_balanced_partitions(lens, n, k, 1 if equal_size else 0, idx, off)
    if rc == -1:
        raise AssertionError(_lib_().vpg_last_error().decode())
    check(rc)
    return [list(idx[off[p]:off[p + 1]]) for p in range(k)]


def packed_seq_layout(seqlens_i32, align: int):
    """seqlens i32 [B] -> (seqlens_padded i32 [B], cu_seqlens_padded i32 [B+1], sum_padded int, max_padded int).
    One device->host read (the reference reads the same two numbers with .item(), mcore/util.py:47,50)."""
    _require_cuda(seqlens_i32)
    B = seqlens_i32.numel()
    dev = seqlens_i32.device
    padded = torch.empty(B, dtype=torch.int32, device=dev)
    cu = torch.empty(B + 1, dtype=torch.int32, device=dev)
    totals = torch.empty(2, dtype=torch.int32, device=dev)
    with on_device(dev):
        check(_lib_().vpg_packed_seq_layout(ptr(seqlens_i32.contiguous()), B, int(align), ptr(padded), ptr(cu),
                                            ptr(totals), stream_ptr(dev)))
    total, mx = totals.tolist()
    return padded, cu, int(total), int(mx)


def packed_seq_map(cu_seqlens, cu_seqlens_padded, T: int, cp_size: int, cp_rank: int):
    """-> (gathered_pos i64 [T], local_dst i64 [T]) for the T valid tokens of the rmpad stream (see vpg_b200.h)."""
    _require_cuda(cu_seqlens, cu_seqlens_padded)
    dev = cu_seqlens.device
    gpos = torch.empty(T, dtype=torch.int64, device=dev)
    ldst = torch.empty(T, dtype=torch.int64, device=dev)
    if T:
        with on_device(dev):
            check(_lib_().vpg_packed_seq_map(ptr(cu_seqlens), ptr(cu_seqlens_padded), cu_seqlens.numel() - 1, T,
                                             int(cp_size), int(cp_rank), ptr(gpos), ptr(ldst), stream_ptr(dev)))
    return gpos, ldst


def permute_rows(src, src_idx, dst_idx, n_out: int, zero_fill: bool = True):
    """zeros [n_out, ...] with out[dst_idx[n]] = src[src_idx[n]] (None index = identity, dst < 0 skipped)."""
    _require_cuda(src, src_idx, dst_idx)
    src = src.contiguous()
    row_shape = src.shape[1:]
    row_bytes = src.element_size() * int(math.prod(row_shape)) if len(row_shape) else src.element_size()
    n = (src_idx if src_idx is not None else dst_idx if dst_idx is not None else src).shape[0]
    out = torch.empty((n_out, *row_shape), dtype=src.dtype, device=src.device)
    if n_out and row_bytes:
        with on_device(src.device):
            check(_lib_().vpg_permute_rows(ptr(src), ptr(src_idx), ptr(dst_idx), n, row_bytes, ptr(out), n_out,
                                           int(zero_fill), stream_ptr(src.device)))
    return out


# --------------------------------------------------------------------------------------------------- token ops
def masked_sums(values, mask):
    """-> float64 [2] = {sum(where(m,v,0)*m), sum(m)} over all elements."""
    _require_cuda(values, mask)
    v = _f32c(values)
    m, mdt = mask_arg(mask.contiguous())
    out = torch.empty(2, dtype=torch.float64, device=v.device)
    with on_device(v.device):
        check(_lib_().vpg_masked_sum(ptr(v), ptr(m), mdt, v.numel(), ptr(out), stream_ptr(v.device)))
    return out


def masked_row_sums(values, mask):
    """last-axis sums: -> (sum fp32 [rows], count fp32 [rows])."""
    _require_cuda(values, mask)
    v = _f32c(values)
    m, mdt = mask_arg(mask.contiguous())
    L = v.shape[-1]
    rows = v.numel() // max(L, 1)
    s = torch.empty(rows, dtype=_F32, device=v.device)
    c = torch.empty(rows, dtype=_F32, device=v.device)
    with on_device(v.device):
        check(_lib_().vpg_masked_row_sum(ptr(v), ptr(m), mdt, rows, L, ptr(s), ptr(c), stream_ptr(v.device)))
    return s, c


def kl_penalty_fwd(logprob, ref_logprob, kind: str, want_grad: bool):
    if kind == "full" or kind not in KL_KINDS:
        raise NotImplementedError  # verl/trainer/ppo/core_algos.py:1065-1069
    _require_cuda(logprob, ref_logprob)
    a, b = _f32c(logprob), _f32c(ref_logprob)
    out = torch.empty_like(a)
    g = torch.empty_like(a) if want_grad else None
    with on_device(a.device):
        check(_lib_().vpg_kl_penalty(ptr(a), ptr(b), KL_KINDS[kind], a.numel(), ptr(out), ptr(g), stream_ptr(a.device)))
    return out, g


def _as_2d(t):
    return t.reshape(-1, t.shape[-1]) if t.dim() != 2 else t


def ppo_loss(old_log_prob, log_prob, advantages, response_mask, *, cliprange_low, cliprange_high, clip_ratio_c=3.0,
             loss_agg_mode="token-mean", entropy=None, entropy_coeff=0.0, ref_log_prob=None, kl_loss_coef=0.0,
             kl_loss_type="low_var_kl", loss_scale=1.0, want_grad=True, pg_kind="clip"):
    """Fused actor loss forward + closed-form gradient (vpg_ppo_loss).

    -> (metrics fp32 [8], d_log_prob | None, d_entropy | None)
    metrics = [loss*scale, pg_loss, pg_clipfrac, ppo_kl, pg_clipfrac_lower, entropy_agg, kl_agg, mask_sum]
    """
    _require_cuda(old_log_prob, log_prob, advantages, response_mask)
    if loss_agg_mode not in AGG_MODES:
        raise ValueError(f"Invalid loss_agg_mode: {loss_agg_mode}")
    assert clip_ratio_c > 1.0, (
        "The lower bound of the clip_ratio_c for dual-clip PPO should be greater than 1.0,"
        + f" but get the value: {clip_ratio_c}."
    )
    if ref_log_prob is not None and (kl_loss_type == "full" or kl_loss_type not in KL_KINDS):
        raise NotImplementedError
    old, lp, adv = _f32c(_as_2d(old_log_prob)), _f32c(_as_2d(log_prob)), _f32c(_as_2d(advantages))
    m, mdt = mask_arg(_as_2d(response_mask).contiguous())
    ent = _f32c(_as_2d(entropy)) if entropy is not None else None
    ref = _f32c(_as_2d(ref_log_prob)) if ref_log_prob is not None else None
    B, L = lp.shape
    dev = lp.device
    lib = _lib_()
    metrics = torch.empty(8, dtype=_F32, device=dev)
    d_lp = torch.empty_like(lp) if want_grad else None
    d_ent = torch.empty_like(lp) if (want_grad and ent is not None) else None
    ws = _ws(int(lib.vpg_ppo_loss_workspace_bytes(B, L)), dev)
    with on_device(dev):
        check(lib.vpg_ppo_loss(ptr(old), ptr(lp), ptr(adv), ptr(m), mdt, ptr(ent), ptr(ref), B, L,
                               float(1 - cliprange_low), float(1 + cliprange_high), float(clip_ratio_c),
                               PG_KINDS[pg_kind], AGG_MODES[loss_agg_mode], float(entropy_coeff), float(kl_loss_coef),
                               KL_KINDS.get(kl_loss_type, -1) if ref is not None else -1, float(loss_scale),
                               ptr(metrics), ptr(d_lp), ptr(d_ent), ptr(ws), ws.numel() * 8, stream_ptr(dev)))
    return metrics, d_lp, d_ent


def _cov_ws(dev):
    return _ws(int(_lib_().vpg_cov_workspace_bytes()), dev)


def clip_cov_flags(old_log_prob, log_prob, advantages, response_mask, cliprange_low, cliprange_high, cov_lb, cov_ub):
    """-> (eligible u8 [n], workspace) — the workspace carries the masked sums into clip_cov_loss; its first double is
    sum(mask)."""
    old, lp, adv = _f32c(old_log_prob), _f32c(log_prob), _f32c(advantages)
    _require_cuda(old, lp, adv, response_mask)
    m, mdt = mask_arg(response_mask.contiguous())
    n = lp.numel()
    dev = lp.device
    ws = _cov_ws(dev)
    flags = torch.empty(n, dtype=torch.uint8, device=dev)
    with on_device(dev):
        check(_lib_().vpg_clip_cov_flags(ptr(old), ptr(lp), ptr(adv), ptr(m), mdt, n, float(cliprange_low),
                                         float(cliprange_high), float(cov_lb), float(cov_ub), ptr(flags), ptr(ws),
                                         ws.numel() * 8, stream_ptr(dev)))
    return flags, ws


def clip_cov_loss(old_log_prob, log_prob, advantages, response_mask, cliprange_low, cliprange_high, zeroed, ws):
    """-> (pg_losses f32 [n], d pg_losses / d log_prob f32 [n], out3 = {ppo_kl, pg_clipfrac, sum mask})"""
    old, lp, adv = _f32c(old_log_prob), _f32c(log_prob), _f32c(advantages)
    m, mdt = mask_arg(response_mask.contiguous())
    n = lp.numel()
    dev = lp.device
    pg = torch.empty(n, dtype=_F32, device=dev)
    dpg = torch.empty(n, dtype=_F32, device=dev)
    out3 = torch.empty(3, dtype=_F32, device=dev)
    with on_device(dev):
        check(_lib_().vpg_clip_cov_loss(ptr(old), ptr(lp), ptr(adv), ptr(m), mdt, n, float(cliprange_low),
                                        float(cliprange_high), ptr(zeroed), ptr(pg), ptr(dpg), ptr(out3), ptr(ws),
                                        ws.numel() * 8, stream_ptr(dev)))
    return pg, dpg, out3


def kl_cov_loss(old_log_prob, log_prob, advantages, response_mask, kl_cov_ratio: float, ppo_kl_coef: float):
    """-> (pg_losses f32 [n], d pg_losses / d log_prob f32 [n], out3 = {ppo_kl_abs, #selected, sum mask})"""
    old, lp, adv = _f32c(old_log_prob), _f32c(log_prob), _f32c(advantages)
    _require_cuda(old, lp, adv, response_mask)
    m, mdt = mask_arg(response_mask.contiguous())
    n = lp.numel()
    dev = lp.device
    ws = _cov_ws(dev)
    pg = torch.empty(n, dtype=_F32, device=dev)
    dpg = torch.empty(n, dtype=_F32, device=dev)
    out3 = torch.empty(3, dtype=_F32, device=dev)
    with on_device(dev):
        check(_lib_().vpg_kl_cov_loss(ptr(old), ptr(lp), ptr(adv), ptr(m), mdt, n, float(kl_cov_ratio),
                                      float(ppo_kl_coef), ptr(pg), ptr(dpg), ptr(out3), ptr(ws), ws.numel() * 8,
                                      stream_ptr(dev)))
    return pg, dpg, out3


def agg_loss_fwd(loss_mat, loss_mask, loss_agg_mode: str, want_grad: bool):
    if loss_agg_mode not in AGG_MODES:
        raise ValueError(f"Invalid loss_agg_mode: {loss_agg_mode}")
    _require_cuda(loss_mat, loss_mask)
    v = _f32c(_as_2d(loss_mat))
    m, mdt = mask_arg(_as_2d(loss_mask).contiguous())
    B, L = v.shape
    lib = _lib_()
    out = torch.empty(1, dtype=_F32, device=v.device)
    d = torch.empty_like(v) if want_grad else None
    ws = _ws(int(lib.vpg_ppo_loss_workspace_bytes(B, L)), v.device)
    with on_device(v.device):
        check(lib.vpg_agg_loss(ptr(v), ptr(m), mdt, B, L, AGG_MODES[loss_agg_mode], ptr(out), ptr(d), ptr(ws),
                               ws.numel() * 8, stream_ptr(v.device)))
    return out[0], d


def value_loss(vpreds, returns, values, response_mask, cliprange_value: float, loss_agg_mode: str, want_grad: bool):
    """compute_value_loss forward + closed-form gradient -> (out fp32 [2] = {vf_loss, vf_clipfrac}, d_vpreds | None)."""
    if loss_agg_mode not in AGG_MODES:
        raise ValueError(f"Invalid loss_agg_mode: {loss_agg_mode}")
    _require_cuda(vpreds, returns, values, response_mask)
    vp, rt, vo = _f32c(_as_2d(vpreds)), _f32c(_as_2d(returns)), _f32c(_as_2d(values))
    m, mdt = mask_arg(_as_2d(response_mask).contiguous())
    B, L = vp.shape
    lib = _lib_()
    out = torch.empty(2, dtype=_F32, device=vp.device)
    d = torch.empty_like(vp) if want_grad else None
    ws = _ws(int(lib.vpg_ppo_loss_workspace_bytes(B, L)), vp.device)
    with on_device(vp.device):
        check(lib.vpg_value_loss(ptr(vp), ptr(rt), ptr(vo), ptr(m), mdt, B, L, float(cliprange_value),
                                 AGG_MODES[loss_agg_mode], ptr(out), ptr(d), ptr(ws), ws.numel() * 8,
                                 stream_ptr(vp.device)))
    return out, d


def outcome_advantage(kind: str, token_level_rewards, response_mask, group_ids, num_groups: int, epsilon: float = 1e-6,
                      norm_by_std: bool = True, f_norm: float = 1.0):
    """RLOO / OPO / GPG / REINFORCE++-baseline / GRPO-pass@k (and GRPO) outcome advantages -> fp32 [B, L]."""
    _require_cuda(token_level_rewards, response_mask, group_ids)
    r = _f32c(token_level_rewards)
    m, mdt = mask_arg(response_mask.contiguous())
    B, L = r.shape
    lib = _lib_()
    adv = torch.empty_like(r)
    status = torch.empty(1, dtype=torch.int32, device=r.device)
    ws = _ws(int(lib.vpg_outcome_advantage_workspace_bytes(B, num_groups)), r.device)
    with on_device(r.device):
        check(lib.vpg_outcome_advantage(ADV_KINDS[kind], ptr(r), ptr(m), mdt, ptr(group_ids), B, L, int(num_groups),
                                        float(epsilon), int(bool(norm_by_std)), float(f_norm), ptr(adv), ptr(status),
                                        ptr(ws), ws.numel() * 8, stream_ptr(r.device)))
    if kind in ("reinforce_plus_plus_baseline", "grpo_passk"):  # the two kinds whose reference can raise
        code = int(status.item())
        if code == 3:
            raise ValueError("Pass@k requires at least 2 samples per group.")
        _raise_whiten_code(code)
    return adv


def reinforce_plus_plus(token_level_rewards, response_mask, gamma: float):
    _require_cuda(token_level_rewards, response_mask)
    r = _f32c(token_level_rewards)
    m, mdt = mask_arg(response_mask.contiguous())
    B, L = r.shape
    lib = _lib_()
    adv, ret = torch.empty_like(r), torch.empty_like(r)
    slot = _StatusSlot.get(r.device)
    ws = _ws(int(lib.vpg_reverse_scan_workspace_bytes(B, L)), r.device)
    with on_device(r.device):
        check(lib.vpg_reinforce_plus_plus(ptr(r), ptr(m), mdt, B, L, float(gamma), ptr(adv), ptr(ret), slot.word_ptr,
                                          ptr(ws), ws.numel() * 8, stream_ptr(r.device), slot.event_ptr))
    slot.wait_and_raise()
    return adv, ret


def remax(token_level_rewards, reward_baselines, response_mask):
    _require_cuda(token_level_rewards, reward_baselines, response_mask)
    r, bsl = _f32c(token_level_rewards), _f32c(reward_baselines)
    m, mdt = mask_arg(response_mask.contiguous())
    B, L = r.shape
    assert bsl.shape == (B,)
    adv, ret = torch.empty_like(r), torch.empty_like(r)
    with on_device(r.device):
        check(_lib_().vpg_remax(ptr(r), ptr(bsl), ptr(m), mdt, B, L, ptr(adv), ptr(ret), stream_ptr(r.device)))
    return adv, ret


def kl_reward(token_level_scores, old_log_probs, ref_log_prob, response_mask, kind: str, beta: float):
    """apply_kl_penalty's math in one pass -> (token_level_rewards fp32 [B,L], current_kl fp32 [1] on the device)."""
    if kind == "full" or kind not in KL_KINDS:
        raise NotImplementedError
    _require_cuda(token_level_scores, old_log_probs, ref_log_prob, response_mask)
    sc, lo, lr = _f32c(token_level_scores), _f32c(old_log_probs), _f32c(ref_log_prob)
    m, mdt = mask_arg(response_mask.contiguous())
    B, L = sc.shape
    lib = _lib_()
    out = torch.empty_like(sc)
    cur = torch.empty(1, dtype=_F32, device=sc.device)
    ws = _ws(int(lib.vpg_kl_reward_workspace_bytes(B)), sc.device)
    with on_device(sc.device):
        check(lib.vpg_kl_reward(ptr(sc), ptr(lo), ptr(lr), ptr(m), mdt, KL_KINDS[kind], float(beta), B, L, ptr(out),
                                ptr(cur), ptr(ws), ws.numel() * 8, stream_ptr(sc.device)))
    return out, cur


DATA_METRIC_SLOTS = {
    "critic/score": 0, "critic/rewards": 3, "critic/advantages": 6, "critic/returns": 9, "critic/values": 12,
    "critic/vf_explained_var": 15, "response_length": 16, "prompt_length": 20, "n_valid": 24,
}


def data_metrics(token_level_scores, token_level_rewards, advantages, returns, values, response_mask, attention_mask):
    """compute_data_metrics' statistics in one sweep -> fp32 [26] device tensor (layout: vpg_data_metrics)."""
    _require_cuda(token_level_scores, token_level_rewards, advantages, returns, response_mask, attention_mask)
    sc, rw, adv, ret = (_f32c(t) for t in (token_level_scores, token_level_rewards, advantages, returns))
    val = _f32c(values) if values is not None else None
    rm, rdt = mask_arg(response_mask.contiguous())
    am, adt = mask_arg(attention_mask.contiguous())
    B, L = sc.shape
    S = am.shape[-1]
    lib = _lib_()
    out = torch.empty(26, dtype=_F32, device=sc.device)
    ws = _ws(int(lib.vpg_data_metrics_workspace_bytes(B)), sc.device)
    with on_device(sc.device):
        check(lib.vpg_data_metrics(ptr(sc), ptr(rw), ptr(adv), ptr(ret), ptr(val), ptr(rm), rdt, ptr(am), adt, B, L, S,
                                   ptr(out), ptr(ws), ws.numel() * 8, stream_ptr(sc.device)))
    return out


def grpo_advantage(token_level_rewards, response_mask, group_ids, num_groups: int, epsilon: float, norm_by_std: bool):
    _require_cuda(token_level_rewards, response_mask, group_ids)
    r = _f32c(token_level_rewards)
    m, mdt = mask_arg(response_mask.contiguous())
    B, L = r.shape
    lib = _lib_()
    adv = torch.empty_like(r)
    ws = _ws(int(lib.vpg_grpo_workspace_bytes(B, num_groups)), r.device)
    with on_device(r.device):
        check(lib.vpg_grpo_advantage(ptr(r), ptr(m), mdt, ptr(group_ids), B, L, int(num_groups), float(epsilon),
                                     int(bool(norm_by_std)), ptr(adv), ptr(ws), ws.numel() * 8, stream_ptr(r.device)))
    return adv


class _StatusSlot:
    """Pinned host status word + CUDA event, one per (thread, device).  The whitening entry points write the status
    word from the last block of the kernel that produces the statistics, BEFORE their normalisation pass, and record the
    event right behind that kernel, so waiting for the event costs the statistics pass only: the ValueError of the
    reference (torch_functional.py:195-200) is still raised synchronously, but the normalisation pass overlaps the
    host's wake-up."""

    _tls = threading.local()

    def __init__(self, device):
        self.word = torch.zeros(1, dtype=torch.int32).pin_memory()  # UVA: the same pointer is valid on the device
        self.word_ptr = self.word.data_ptr()
        self.pair = torch.zeros(2, dtype=torch.int32).pin_memory()  # {nnz, max_seqlen} of unpad_indices
        self.pair_ptr = self.pair.data_ptr()
        # plain ctypes views of the pinned words: reading `int(tensor[0])` costs 3-4 us per value on a host-bound path
        self._word_c = ctypes.c_int32.from_address(self.word_ptr)
        self._pair_c = (ctypes.c_int32 * 2).from_address(self.pair_ptr)
        self.event = torch.cuda.Event()
        with on_device(device):
            self.event.record()  # torch creates the cudaEvent_t lazily, at the first record
        self.event.synchronize()
        self.event_ptr = int(self.event.cuda_event)
        assert self.event_ptr, "torch.cuda.Event did not expose its handle"

    @classmethod
    def get(cls, device) -> "_StatusSlot":
        slots = cls._tls.__dict__.setdefault("slots", {})
        idx = torch.device(device).index
        idx = torch.cuda.current_device() if idx is None else idx
        if idx not in slots:
            slots[idx] = cls(idx)
        return slots[idx]

    def read_pair(self):
        """{nnz, max_seqlen} as Python ints (valid once the stream has passed the unpad launch)."""
        return int(self._pair_c[0]), int(self._pair_c[1])

    def wait_and_raise(self):
        self.event.synchronize()
        _raise_whiten_code(int(self._word_c.value))


def _raise_whiten_code(code: int):
    if code == 1:
        raise ValueError("At least one element in the mask has to be 1.")
    if code == 2:
        raise ValueError("The sum of the mask is one, which can cause a division by zero.")


def gae(token_level_rewards, values, response_mask, gamma: float, lam: float):
    _require_cuda(token_level_rewards, values, response_mask)
    r, v = _f32c(token_level_rewards), _f32c(values)
    m, mdt = mask_arg(response_mask.contiguous())
    B, L = r.shape
    lib = _lib_()
    adv, ret = torch.empty_like(r), torch.empty_like(r)
    slot = _StatusSlot.get(r.device)
    ws = _ws(int(lib.vpg_gae_workspace_bytes(B, L)), r.device)
    with on_device(r.device):
        check(lib.vpg_gae(ptr(r), ptr(v), ptr(m), mdt, B, L, float(gamma), float(lam), ptr(adv), ptr(ret),
                          slot.word_ptr, ptr(ws), ws.numel() * 8, stream_ptr(r.device), slot.event_ptr))
    slot.wait_and_raise()
    return adv, ret


def masked_whiten(values, mask, shift_mean=True):
    _require_cuda(values, mask)
    v = _f32c(values)
    m, mdt = mask_arg(mask.contiguous())
    out = torch.empty_like(v)
    slot = _StatusSlot.get(v.device)
    ws = _ws(32, v.device)
    with on_device(v.device):
        check(_lib_().vpg_masked_whiten(ptr(v), ptr(m), mdt, v.numel(), int(bool(shift_mean)), ptr(out),
                                        slot.word_ptr, ptr(ws), ws.numel() * 8, stream_ptr(v.device), slot.event_ptr))
    slot.wait_and_raise()
    return out
