"""Host-tensor staging for the driver-side functions of the path.

The reference runs advantage estimation, the KL-in-reward and the data metrics on the DRIVER with CPU tensors
(verl/trainer/ppo/ray_trainer.py:1283-1305; compute_grpo_outcome_advantage builds CPU tensors itself,
core_algos.py:245-308), and its own acceptance tests call these functions on CPU tensors
(tests/trainer/ppo/test_core_algos_on_cpu.py, tests/utils/test_torch_functional.py:55-66).  A drop-in must accept the
same calls.  This module is INPUT STAGING, not a CPU fallback: no arithmetic of the path runs on the host.  A call whose
tensor arguments live on the host is executed as

    host tensor --memcpy--> pinned arena --cudaMemcpyAsync--> device --CUDA kernels--> device --cudaMemcpyAsync-->
    pinned arena --memcpy--> host tensor

on the current CUDA device and stream; results come back on the inputs' device.  Without a CUDA device the call raises
(there is nothing to fall back to).  Tensors that require grad are staged through autograd-aware copies, so the whole
graph lives on the device and only leaf gradients travel back.
"""

from __future__ import annotations

import functools
import threading

import torch

__all__ = ["staged", "call", "to_device", "to_host"]

_tls = threading.local()
_ALIGN = 256


class _Arena:
    """One growing pinned buffer per (thread, direction); bump-allocated within a call.  The arena of a direction is
    reused by the next call only after the copies of the previous one have completed (event)."""

    def __init__(self):
        self.buf = None
        self.off = 0
        self.event = None
        self.spill = []  # older, smaller buffers still referenced by in-flight copies

    def begin(self):
        if self.event is not None:
            self.event.synchronize()
            self.event = None
        self.off = 0
        self.spill.clear()

    def take(self, nbytes: int) -> torch.Tensor:
        start = (self.off + _ALIGN - 1) // _ALIGN * _ALIGN
        if self.buf is None or start + nbytes > self.buf.numel():
            if self.buf is not None:
                self.spill.append(self.buf)
            cap = max(1 << 20, 2 * (start + nbytes))
            self.buf = torch.empty(cap, dtype=torch.uint8).pin_memory()
            start = 0
        self.off = start + nbytes
        return self.buf[start:start + nbytes]

    def end(self, device):
        self.event = torch.cuda.Event()
        self.event.record(torch.cuda.current_stream(device))


def _arena(kind: str) -> _Arena:
    a = getattr(_tls, kind, None)
    if a is None:
        a = _Arena()
        setattr(_tls, kind, a)
    return a


def _device() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("verl_b200 runs its ops on a CUDA device (host tensors are staged to it); there is no CPU fallback "
                           "and no CUDA device is available")
    return torch.device("cuda", torch.cuda.current_device())


def _h2d(t: torch.Tensor, device) -> torch.Tensor:
    """one host memcpy into the pinned arena + one cudaMemcpyAsync"""
    src = t.detach()
    if not src.is_contiguous():
        src = src.contiguous()
    n = src.numel() * src.element_size()
    out = torch.empty(src.shape, dtype=src.dtype, device=device)
    if n == 0:
        return out
    pin = _arena("h2d").take(n).view(src.dtype).view(src.shape)
    pin.copy_(src)
    out.copy_(pin, non_blocking=True)
    return out


def _d2h(t: torch.Tensor) -> torch.Tensor:
    """cudaMemcpyAsync into the pinned arena; the caller synchronises once and clones into pageable memory"""
    src = t.detach()
    if not src.is_contiguous():
        src = src.contiguous()
    n = src.numel() * src.element_size()
    if n == 0:
        return torch.empty(src.shape, dtype=src.dtype)
    pin = _arena("d2h").take(n).view(src.dtype).view(src.shape)
    pin.copy_(src, non_blocking=True)
    return pin


class _StageIn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, t, device):
        return _h2d(t, device)

    @staticmethod
    def backward(ctx, g):
        return to_host(g), None


class _StageOut(torch.autograd.Function):
    @staticmethod
    def forward(ctx, t):
        ctx.device = t.device
        pin = _d2h(t)
        torch.cuda.current_stream(t.device).synchronize()
        return pin.clone()

    @staticmethod
    def backward(ctx, g):
        return to_device(g, ctx.device)


def to_device(t: torch.Tensor, device=None) -> torch.Tensor:
    device = device or _device()
    a = _arena("h2d")
    a.begin()
    out = _h2d(t, device)
    a.end(device)
    return out


def to_host(t: torch.Tensor) -> torch.Tensor:
    a = _arena("d2h")
    a.begin()
    pin = _d2h(t)
    torch.cuda.current_stream(t.device).synchronize()
    return pin.clone()


def _map(obj, fn, memo):
    if isinstance(obj, torch.Tensor):
        key = id(obj)
        if key not in memo:
            memo[key] = fn(obj)
        return memo[key]
    if isinstance(obj, tuple) and hasattr(obj, "_fields"):  # namedtuple
        return type(obj)(*(_map(o, fn, memo) for o in obj))
    if isinstance(obj, (tuple, list)):
        return type(obj)(_map(o, fn, memo) for o in obj)
    if isinstance(obj, dict):
        return {k: _map(v, fn, memo) for k, v in obj.items()}
    return obj


def _tensors(obj, out):
    if isinstance(obj, torch.Tensor):
        out.append(obj)
    elif isinstance(obj, (tuple, list)):
        for o in obj:
            _tensors(o, out)
    elif isinstance(obj, dict):
        for o in obj.values():
            _tensors(o, out)
    return out


def call(fn, *args, **kwargs):
    """fn(*args, **kwargs) with every host tensor argument staged to the current CUDA device; when ALL tensor arguments
    were host tensors the tensor results are returned on the host (same object identity structure: a tensor returned
    twice comes back as one host tensor returned twice)."""
    tens = _tensors((args, kwargs), [])
    if not tens or all(t.is_cuda for t in tens):
        return fn(*args, **kwargs)
    device = next((t.device for t in tens if t.is_cuda), None) or _device()
    all_host = not any(t.is_cuda for t in tens)
    grad_on = torch.is_grad_enabled()
    a = _arena("h2d")
    a.begin()

    def stage_in(t):
        if t.is_cuda:
            return t
        if grad_on and t.requires_grad:
            return _StageIn.apply(t, device)
        return _h2d(t, device)

    with torch.cuda.device(device):
        memo: dict = {}
        d_args, d_kwargs = _map(args, stage_in, memo), _map(kwargs, stage_in, memo)
        a.end(device)
        result = fn(*d_args, **d_kwargs)
        if not all_host:
            return result
        outs = _tensors(result, [])
        if not outs:
            return result
        b = _arena("d2h")
        b.begin()
        pins: dict = {}

        def stage_out(t):
            if not t.is_cuda:
                return t
            if grad_on and t.requires_grad:
                return _StageOut.apply(t)
            pins[id(t)] = _d2h(t)
            return t  # placeholder, replaced after the single synchronize below

        memo2: dict = {}
        staged_result = _map(result, stage_out, memo2)
        if pins:
            torch.cuda.current_stream(device).synchronize()
            memo3: dict = {}
            staged_result = _map(staged_result, lambda t: pins[id(t)].clone() if id(t) in pins else t, memo3)
        return staged_result


def staged(fn):
    """Decorator form of `call` for the reference-named public functions."""

    @functools.wraps(fn)
    def wrapper(*args, **kwargs):
        return call(fn, *args, **kwargs)

    wrapper.__wrapped__ = fn
    return wrapper
