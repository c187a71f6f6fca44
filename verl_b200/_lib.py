"""ctypes binding of libvpg_b200.so (see include/vpg_b200.h).

There is no CPU fallback: importing this module without the built library, or calling an op on a non-CUDA tensor,
raises.  Build the library with ``python -m verl_b200.csrc.build`` (or ``__graft_entry__.build()``).
"""

from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import torch

_LIB_PATH = Path(__file__).resolve().parent / "lib" / "libvpg_b200.so"

c_void_p, c_int, c_int64, c_float, c_size_t = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t

# name -> (restype, argtypes); must list every symbol declared in include/vpg_b200.h
SIGNATURES: dict[str, tuple] = {
    "vpg_version": (C.c_char_p, []),
    "vpg_last_error": (C.c_char_p, []),
    "vpg_launch_count": (c_int64, []),
    "vpg_stream_synchronize": (c_int, [c_void_p]),
    "vpg_lce_num_splits": (c_int, [c_int64, c_int64, c_int64, c_int]),
    "vpg_lce_fwd_partials": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_float, c_int64, c_int, c_void_p, c_void_p],
    ),
    "vpg_lce_fwd_partials_dt": (
        c_int,
        [c_int, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_float, c_int64, c_int, c_void_p, c_void_p],
    ),
    "vpg_lce_bwd_dt": (
        c_int,
        [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64,
         c_float, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "vpg_lce_bwd_f32": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64,
         c_int64, c_float, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "vpg_lce_fwd_partials_push": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_float, c_int64, c_int, c_void_p, c_int, c_int,
         c_int, c_void_p, c_int, c_void_p],
    ),
    "vpg_lce_finalize": (
        c_int,
        [c_void_p, c_int, c_int64, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "vpg_lce_bwd_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int64, c_int64]),
    "vpg_lce_bwd_default_chunk": (c_int64, [c_int64, c_int64, c_int64]),
    "vpg_lce_bwd_chunk_for_budget": (c_int64, [c_int64, c_int64, c_int64, c_int64]),
    "vpg_lce_bwd": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_float,
         c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "vpg_lce_bwd_mcred": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_float,
         c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p, c_void_p],
    ),
    "vpg_cast_f32_bf16": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "vpg_gemm_bf16": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int64, c_int, c_int, c_void_p, c_int64, c_int,
         c_void_p, c_int64, c_void_p],
    ),
    "vpg_logits_stats": (
        c_int,
        [c_void_p, c_int, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "vpg_logits_partials": (
        c_int, [c_void_p, c_int, c_int64, c_int64, c_int64, c_void_p, c_int64, c_void_p, c_void_p]),
    "vpg_logits_stats_bwd": (
        c_int,
        [c_void_p, c_int, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
         c_void_p],
    ),
    "vpg_unpad_workspace_bytes": (c_size_t, [c_int64]),
    "vpg_unpad_indices": (
        c_int,
        [c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "vpg_response_rows": (
        c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "vpg_gather_rows": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "vpg_scatter_rows": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int, c_void_p]),
    "vpg_masked_sum": (c_int, [c_void_p, c_void_p, c_int, c_int64, c_void_p, c_void_p]),
    "vpg_masked_row_sum": (c_int, [c_void_p, c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p]),
    "vpg_kl_penalty": (c_int, [c_void_p, c_void_p, c_int, c_int64, c_void_p, c_void_p, c_void_p]),
    "vpg_ppo_loss_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "vpg_ppo_loss": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int64, c_int64, c_float, c_float,
         c_float, c_int, c_int, c_float, c_float, c_int, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
         c_void_p],
    ),
    "vpg_value_loss": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int64, c_float, c_int, c_void_p, c_void_p,
         c_void_p, c_size_t, c_void_p],
    ),
    "vpg_outcome_advantage_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "vpg_outcome_advantage": (
        c_int,
        [c_int, c_void_p, c_void_p, c_int, c_void_p, c_int64, c_int64, c_int64, c_float, c_int, c_float, c_void_p,
         c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "vpg_reverse_scan_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "vpg_reinforce_plus_plus": (
        c_int,
        [c_void_p, c_void_p, c_int, c_int64, c_int64, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
         c_void_p, c_void_p],
    ),
    "vpg_remax": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p]),
    "vpg_agg_loss": (
        c_int,
        [c_void_p, c_void_p, c_int, c_int64, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "vpg_grpo_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "vpg_grpo_advantage": (
        c_int,
        [c_void_p, c_void_p, c_int, c_void_p, c_int64, c_int64, c_int64, c_float, c_int, c_void_p, c_void_p,
         c_size_t, c_void_p],
    ),
    "vpg_grpo_advantage_static": (
        c_int,
        [c_void_p, c_void_p, c_int, c_void_p, c_int64, c_int64, c_int64, c_float, c_int, c_void_p, c_void_p],
    ),
    "vpg_kl_reward_workspace_bytes": (c_size_t, [c_int64]),
    "vpg_kl_reward": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_int64, c_int64, c_void_p, c_void_p,
         c_void_p, c_size_t, c_void_p],
    ),
    "vpg_data_metrics_workspace_bytes": (c_size_t, [c_int64]),
    "vpg_data_metrics": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_int64, c_int64,
         c_int64, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "vpg_gae_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "vpg_gae": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int64, c_float, c_float, c_void_p, c_void_p, c_void_p,
         c_void_p, c_size_t, c_void_p, c_void_p],
    ),
    "vpg_row_lengths": (c_int, [c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p]),
    "vpg_seqlen_balanced_partitions": (c_int, [c_void_p, c_int64, c_int64, c_int, c_void_p, c_void_p]),
    "vpg_packed_seq_layout": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "vpg_packed_seq_map": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "vpg_permute_rows": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int, c_void_p]),
    "vpg_cov_workspace_bytes": (c_size_t, []),
    "vpg_clip_cov_flags": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_float, c_float, c_float, c_float, c_void_p,
         c_void_p, c_size_t, c_void_p],
    ),
    "vpg_clip_cov_loss": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_float, c_float, c_void_p, c_void_p, c_void_p,
         c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "vpg_kl_cov_loss": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, C.c_double, c_float, c_void_p, c_void_p, c_void_p,
         c_void_p, c_size_t, c_void_p],
    ),
    "vpg_lce_bwd_ex": (c_int, [c_void_p]),  # const vpg_lce_bwd_args* (LceBwdArgs below)
    "vpg_lce_fwd_partials_ex": (
        c_int,
        [c_int, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_float, c_int64, c_int, c_void_p, c_int, c_int,
         c_int, c_void_p, c_int, c_void_p],
    ),
    "vpg_lce_bwd_plan": (c_int, [c_int64, c_int64, c_int64, c_int64, c_int64, c_void_p]),
    "vpg_timing_enable": (c_int, [c_int]),
    "vpg_timing_collect": (c_int, [c_void_p, c_void_p, c_int]),
    "vpg_gemm_splitk_workspace_bytes": (c_size_t, []),
    "vpg_gemm_bf16_ws": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int64, c_int, c_int, c_void_p, c_int64, c_int,
         c_void_p, c_int64, c_void_p, c_size_t, c_void_p],
    ),
    "vpg_masked_whiten": (
        c_int,
        [c_void_p, c_void_p, c_int, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p, c_void_p],
    ),
}



class LceBwdArgs(C.Structure):
    """vpg_lce_bwd_args (include/vpg_b200.h) — every option of the lm-head backward as an explicit field."""

    _fields_ = [
        ("struct_size", c_size_t), ("dtype", c_int), ("hidden", c_void_p), ("weight", c_void_p),
        ("hidden16", c_void_p), ("weight16", c_void_p), ("labels", c_void_p), ("lse", c_void_p), ("eb", c_void_p),
        ("g_lp", c_void_p), ("g_ent", c_void_p), ("T", c_int64), ("H", c_int64), ("V", c_int64),
        ("temperature", c_float), ("label_offset", c_int64), ("chunk_v", c_int64), ("chunk_t", c_int64), ("d_hidden", c_void_p),
        ("d_weight", c_void_p), ("workspace", c_void_p), ("workspace_bytes", c_size_t), ("stream", c_void_p),
        ("d_hidden_mc_f32", c_void_p), ("d_hidden_ready_event", c_void_p), ("reserve_cta_pairs", c_int),
        ("ar_world", c_int), ("ar_rank", c_int), ("ar_epoch", C.c_uint32), ("ar_d_hidden_mc", c_void_p),
        ("ar_flags", c_void_p), ("ar_peer_flags", C.POINTER(c_void_p)),
    ]


TIMING_KINDS = {0: "lm_head_fwd", 1: "dlogits_recompute", 2: "d_weight_gemm", 3: "d_hidden_gemm", 4: "gemm",
                5: "splitk_reduce"}

MASK_I64, MASK_I32, MASK_U8, MASK_F32 = 0, 1, 2, 3
AGG_MODES = {
    "token-mean": 0,
    "seq-mean-token-sum": 1,
    "seq-mean-token-mean": 2,
    "seq-mean-token-sum-norm": 3,
}
PG_KINDS = {"clip": 0, "gpg": 1}
ADV_KINDS = {"grpo": 0, "rloo": 1, "opo": 2, "gpg": 3, "reinforce_plus_plus_baseline": 4, "grpo_passk": 5}
KL_KINDS = {"kl": 0, "k1": 0, "abs": 1, "mse": 2, "k2": 2, "low_var_kl": 3, "k3": 3}

_lib = None


def lib_path() -> Path:
    return Path(os.environ.get("VPG_B200_LIB", str(_LIB_PATH)))


def load() -> C.CDLL:
    """Load the shared library and bind every declared symbol (raises if the library or a symbol is missing)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not path.exists():
        raise RuntimeError(
            f"{path} is missing: build it with `python -m verl_b200.csrc.build`. "
            "verl_b200 has no CPU / PyTorch fallback for its CUDA ops."
        )
    lib = C.CDLL(str(path))
    for name, (restype, argtypes) in SIGNATURES.items():
        if os.environ.get("VPG_B200_PARTIAL") == "1" and not hasattr(lib, name):
            continue  # developer probe of a partially built library only
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().vpg_last_error().decode()
        kind = {-1: ValueError, -3: RuntimeError, -4: NotImplementedError}.get(rc, RuntimeError)
        raise kind(f"libvpg_b200 error {rc}: {msg}")


def ptr(t: torch.Tensor | None):
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("verl_b200 ops require CUDA tensors (there is no CPU fallback)")
    if not t.is_contiguous():
        raise ValueError("verl_b200 ops require contiguous tensors")
    return t.data_ptr()


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr(device=None) -> int:
    """cudaStream_t of torch's current stream on `device`.  At the BASELINE micro-batch sizes the token ops are host-bound
    (12-45 us per call, `scripts/prof_host_overhead.py`), and `torch.cuda.current_stream(...)` alone costs 6 us of that
    (device-index normalisation + a Stream object); the raw-stream query the compiled-graph runtime uses is ~0.3 us."""
    if _raw_stream is not None:
        idx = getattr(device, "index", device)
        if idx is None:
            idx = torch.cuda.current_device()
        return _raw_stream(idx)
    return torch.cuda.current_stream(device).cuda_stream


class on_device:
    """``with on_device(dev):`` — `torch.cuda.device(dev)` that costs nothing when `dev` already is the current device
    (the normal case: one process per GPU).  The library launches on the current device, so the guard is only needed when
    a tensor lives elsewhere; entering torch's context manager unconditionally cost ~6 us per op call."""

    __slots__ = ("_ctx",)

    def __init__(self, device):
        idx = getattr(device, "index", device)
        self._ctx = None if (idx is None or idx == torch.cuda.current_device()) else torch.cuda.device(idx)

    def __enter__(self):
        if self._ctx is not None:
            self._ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if self._ctx is not None:
            return self._ctx.__exit__(*exc)
        return False


def mask_arg(mask: torch.Tensor):
    """(tensor, dtype enum) for a mask; bool is reinterpreted as uint8 without a copy."""
    if mask.dtype == torch.int64:
        return mask, MASK_I64
    if mask.dtype == torch.int32:
        return mask, MASK_I32
    if mask.dtype == torch.bool:
        return mask.view(torch.uint8), MASK_U8
    if mask.dtype == torch.uint8:
        return mask, MASK_U8
    if mask.dtype == torch.float32:
        return mask, MASK_F32
    return mask.to(torch.float32), MASK_F32


def launch_count() -> int:
    return int(load().vpg_launch_count())


def timing_enable(on: bool) -> None:
    check(load().vpg_timing_enable(int(bool(on))))


def timing_collect(max_n: int = 4096) -> list[tuple[str, float]]:
    """[(kernel kind, milliseconds)] of the tcgen05 launches recorded since the last call (synchronises their events)."""
    kinds = (C.c_int * max_n)()
    ms = (C.c_float * max_n)()
    n = int(load().vpg_timing_collect(kinds, ms, max_n))
    return [(TIMING_KINDS.get(int(kinds[i]), str(int(kinds[i]))), float(ms[i])) for i in range(n)]
