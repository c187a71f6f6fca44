"""NVLink peer-memory workspace of one vocab-parallel group (one process per GPU, one 8xB200 NVSwitch box).

The fused lm-head kernels can do their collectives themselves instead of calling NCCL afterwards:

* forward  — the GEMM epilogue stores each token tile's partials {m, s, b, t} straight into the gathered
  ``[4, world*S, T]`` buffer of EVERY rank (``multimem.st`` on an NVSwitch multicast address, or plain stores to the
  peer-mapped buffers when multicast is unavailable): the all-gather happens tile by tile under the GEMM;
* backward — d_hidden is written into a symmetric bf16 ``[T, H]`` slot and summed over the ranks INSIDE the d_weight GEMM
  kernel: one otherwise idle warp per CTA runs a two-shot NVLS all-reduce (``multimem.ld_reduce`` of this rank's 1/world
  slice, ``multimem.st`` of the sum into every rank's slot) while the tensor cores work, so the all-reduce the reference
  leaves to its callers (tests/utils/test_linear_cross_entropy_tp.py:313-314) costs no SM and no extra launch.  (The
  older variant — ``multimem.red.add.f32`` from the d_hidden GEMM's epilogue into an fp32 buffer — is kept, opt-in.)

This module owns the buffers those kernels write: symmetric allocations made with torch's
``torch.distributed._symmetric_memory`` (CUDA VMM + fabric handles exchanged through the group's store — plumbing,
like the caching allocator), their peer / multicast addresses, and the signal-pad barrier between the phases.
Replaces the exchange of verl/utils/kernel/kernels.py:609-661.
"""

from __future__ import annotations

import os

import torch
import torch.distributed as dist

_workspaces: dict = {}
_AR_ERROR_WORD = 17  # kArErrorWord in csrc/tc_gemm.cuh
_AR_CHECK = os.environ.get("VPG_AR_CHECK", "0") == "1"  # one tiny all-reduce per backward: ranks agree on T, H, epoch


class PeerWorkspace:
    """Symmetric buffers of one (process group, device).  Every method that allocates is COLLECTIVE over the group and
    must be called with the same sizes in the same order on all ranks (true for the lm-head: T, H, S are replicated)."""

    def __init__(self, group, device: torch.device):
        import torch.distributed._symmetric_memory as symm_mem

        self._symm = symm_mem
        self.group = group
        self.device = device
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self._bufs: dict[str, tuple] = {}  # name -> (tensor, handle, numel)
        self._parity = 0
        try:  # older torch versions need the group to be enabled explicitly; newer ones treat this as a no-op
            import warnings

            with warnings.catch_warnings():  # deprecated no-op on newer torch
                warnings.simplefilter("ignore")
                symm_mem.enable_symm_mem_for_group(group.group_name)
        except Exception:
            pass

    # ------------------------------------------------------------------------------------------ allocation
    def _buffer(self, name: str, numel: int, zero: bool = False):
        cur = self._bufs.get(name)
        if cur is not None and cur[2] >= numel:
            return cur
        # (Re)allocation is COLLECTIVE.  Ranks that disagree on the size (different T) would deadlock in the rendezvous:
        # agree on it first and raise on every rank instead.
        if self.world > 1 and dist.is_initialized():
            want = torch.tensor([numel, -numel], dtype=torch.int64, device=self.device)
            dist.all_reduce(want, op=dist.ReduceOp.MAX, group=self.group)
            lo, hi = -int(want[1]), int(want[0])
            if lo != hi:
                raise RuntimeError(f"symmetric buffer '{name}': ranks ask for different sizes ({lo} .. {hi} elements); the "
                                   "vocab-parallel lm-head needs the same token count on every rank (hidden is replicated)")
        t = self._symm.empty(numel, dtype=torch.float32, device=self.device)
        hdl = self._symm.rendezvous(t, group=self.group.group_name)
        if zero:
            t.zero_()
            torch.cuda.current_stream(self.device).synchronize()
            hdl.barrier(channel=0)
        self._bufs[name] = (t, hdl, numel)
        return self._bufs[name]

    def reserve(self, max_tokens: int, hidden_size: int, num_splits: int = 16):
        """Explicit collective initialisation: allocate every symmetric buffer the lm-head needs for up to
        ``max_tokens`` tokens NOW, so that no forward / backward call allocates (or rendezvouses) later."""
        for parity in (0, 1):
            self._buffer(f"partials{parity}", 4 * self.world * num_splits * max_tokens)
            self._buffer(f"dh{parity}", (max_tokens * hidden_size + 1) // 2)
        self._buffer("ar_flags", 64, zero=True)
        self._buffer("probe", 1024)
        return self

    @staticmethod
    def _mc_ptr(hdl) -> int:
        try:
            return int(hdl.multicast_ptr or 0)
        except Exception:
            return 0

    @property
    def has_multicast(self) -> bool:
        t, hdl, _ = self._buffer("probe", 1024)
        return self._mc_ptr(hdl) != 0

    # ------------------------------------------------------------------------------------------ phases
    def barrier(self):
        """All ranks' prior work on the current stream is complete and visible before any rank continues."""
        _, hdl, _ = self._buffer("probe", 1024)
        hdl.barrier(channel=0)

    def partials(self, S: int, T: int):
        """-> (local view [4, world*S, T], multicast address | 0, [peer addresses of the same view]).
        Two slots alternate between calls, so rank A's next forward never overwrites a slot rank B is still reading."""
        numel = 4 * self.world * S * T
        self._parity ^= 1
        t, hdl, _ = self._buffer(f"partials{self._parity}", numel)
        view = t[:numel].view(4, self.world * S, T)
        return view, self._mc_ptr(hdl), [int(p) for p in hdl.buffer_ptrs]

    def d_hidden_slot(self, T: int, H: int) -> dict:
        """Arguments of the in-kernel all-reduce (ops.lce_backward(allreduce=...)): a bf16 [T,H] view of one of two
        alternating symmetric slots (the d_hidden of step i stays intact while step i+1 runs), its multicast address, the
        flag words and the call's epoch.  Same call order on all ranks (it is a collective's argument).

        Nothing is consumed yet: slot parity and epoch advance in ``commit_d_hidden_slot`` AFTER the launch went through,
        so a rank whose launch raises stays in step with its own state (its peers then run into the kernel's wait budget
        and fail too, see ``check_allreduce``)."""
        self.check_allreduce(wait=False)  # a failed exchange of an earlier call surfaces here at the latest
        parity = getattr(self, "_dh_parity", 0) ^ 1
        epoch = ((getattr(self, "_ar_epoch", 0) + 1) & 0x7FFFFFFF) or 1
        numel = (T * H + 1) // 2  # fp32 elements holding T*H bf16
        t, hdl, _ = self._buffer(f"dh{parity}", numel)
        ft, fh, _ = self._buffer("ar_flags", 64, zero=True)
        mc = self._mc_ptr(hdl)
        if not mc:
            raise RuntimeError("the in-kernel all-reduce needs NVSwitch multicast (no multicast address for the slot)")
        if _AR_CHECK:  # debug: agree on the call before anyone can start waiting inside a kernel
            sig = torch.tensor([T, -T, H, -H, epoch, -epoch], dtype=torch.int64, device=self.device)
            dist.all_reduce(sig, op=dist.ReduceOp.MAX, group=self.group)
            a = sig.tolist()
            if a[0] != -a[1] or a[2] != -a[3] or a[4] != -a[5]:
                raise RuntimeError(f"in-kernel all-reduce: ranks disagree on the call (T {-a[1]}..{a[0]}, H {-a[3]}..{a[2]}, "
                                   f"epoch {-a[5]}..{a[4]}); this rank has T={T} H={H} epoch={epoch}")
        return dict(out=t.view(torch.bfloat16)[:T * H].view(T, H), mc=mc, flags=ft.view(torch.int32),
                    peer_flags=[int(p) for p in fh.buffer_ptrs], world=self.world, rank=self.rank,
                    epoch=epoch, parity=parity)

    def commit_d_hidden_slot(self, ar: dict):
        """The backward that took `ar` has been enqueued: advance slot parity and epoch, and queue a 4-byte copy of this
        rank's error word (flags[17], written by the kernel when a wait for a peer ran out or met a poisoned flag) into
        pinned memory behind it.  No synchronisation here; ``check_allreduce`` looks at the word once the copy is done."""
        self._dh_parity, self._ar_epoch = ar["parity"], ar["epoch"]
        if getattr(self, "_err_host", None) is None:
            self._err_host = torch.zeros(1, dtype=torch.int32).pin_memory()
            self._err_event = torch.cuda.Event()
        self._err_host.copy_(ar["flags"][_AR_ERROR_WORD:_AR_ERROR_WORD + 1], non_blocking=True)
        self._err_event.record(torch.cuda.current_stream(self.device))
        self._err_pending = True

    def check_allreduce(self, wait: bool = True):
        """Raise if an in-kernel all-reduce on this rank gave up on a peer (it never launched the matching backward:
        raised, ran out of memory, or diverged in call order; budget: $VPG_AR_TIMEOUT_MS, default 60 s per wait).
        The returned d_hidden of that call is not the sum.  ``wait=False`` only looks if the kernel has finished;
        ``wait=True`` blocks on it (call it where you would synchronise anyway, e.g. before the optimizer step).  After a
        failure the group's epochs are out of step: ``peer.reset()`` on every rank before using the workspace again."""
        if not getattr(self, "_err_pending", False):
            return
        if wait:
            self._err_event.synchronize()
        elif not self._err_event.query():
            return
        self._err_pending = False
        bad = int(self._err_host[0])
        if bad:
            self._err_pending = True  # sticky: every later check raises as well
            raise RuntimeError(
                f"in-kernel d_hidden all-reduce failed on rank {self.rank} (call epoch {bad & 0x7FFFFFFF}): a peer did not "
                "reach the matching backward within the wait budget, or gave up itself; the gradient of that call is "
                "not the sum over the ranks")

    def d_hidden_accum(self, T: int, H: int):
        """-> (local fp32 view [T, H], multicast address | 0)"""
        numel = T * H
        t, hdl, _ = self._buffer("dh_accum", numel)
        return t[:numel].view(T, H), self._mc_ptr(hdl)


def workspace(group, device) -> PeerWorkspace:
    key = (group.group_name, torch.device(device).index)
    ws = _workspaces.get(key)
    if ws is None:
        ws = _workspaces[key] = PeerWorkspace(group, torch.device(device))
    return ws


def reset():
    _workspaces.clear()
