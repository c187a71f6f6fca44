"""A peer that never launches the matching backward must not hang the in-kernel d_hidden all-reduce.

torchrun, 2 ranks, VPG_AR_TIMEOUT_MS small (the wrapper test sets 1500):
  1. both ranks run fwd+bwd in "fused" mode: check_allreduce passes, d_hidden equals the NCCL-reduced one, and the
     returned d_hidden is a copy (a later backward does not overwrite it);
  2. rank 1 skips a backward (as if it had raised before the launch).  Rank 0's kernel gives up after the budget, its
     error word is set and check_allreduce raises; the kernel's wall time is bounded;
  3. rank 1 then launches that backward late: it finds rank 0's ready flag, but rank 0's done flag is poisoned, so rank 1
     fails too (at once, not after the budget) instead of returning a sum with rank 0's slice missing.
Prints one line per check; exit code 0 = every check held on every rank."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from verl_b200.utils.kernel import kernels, peer
from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    assert world == 2
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    group = dist.group.WORLD
    budget_s = int(os.environ.get("VPG_AR_TIMEOUT_MS", "60000")) / 1e3
    T, H, V = 2048, 1024, 16384
    g = torch.Generator(device=dev).manual_seed(7)
    hidden = (torch.rand(T, H, device=dev, generator=g) - 0.5).bfloat16()
    weight = (torch.rand(V, H, device=dev, generator=g) - 0.5).bfloat16()[rank * (V // 2):(rank + 1) * (V // 2)].contiguous()
    labels = torch.randint(0, V, (T,), device=dev, generator=g)
    g_lp = torch.rand(T, device=dev, generator=g) * 2 - 1
    g_ent = torch.rand(T, device=dev, generator=g) - 0.5
    kernels.set_vocab_parallel_exchange("nvlink")

    def step(mode, backward=True):
        kernels.set_reduce_d_hidden(mode is not None, mode=mode or "nccl")
        h = hidden.clone().requires_grad_(True)
        w = weight.clone().requires_grad_(True)
        lp, ent = linear_cross_entropy(h, w, labels, 1.0, "none", group)
        if not backward:
            return h, lp, ent
        torch.autograd.backward([lp, ent], [g_lp, g_ent])
        return h.grad

    ok = True
    px = peer.workspace(group, dev)
    # 1 ---------------------------------------------------------------------------------------------------------
    want = step(None)
    dist.all_reduce(want, group=group)
    got1 = step("fused")
    px.check_allreduce(wait=True)
    keep = got1.clone()
    for _ in range(3):  # two slots alternate: three more backwards overwrite both
        step("fused")
    px.check_allreduce(wait=True)
    torch.cuda.synchronize()
    c1 = torch.equal(got1, keep) and bool(((got1.float() - want.float()).abs() <= 2e-2 + 4e-2 * want.float().abs()).all())
    print(f"[rank {rank}] fused all-reduce ok, returned d_hidden is a stable copy: {c1}", flush=True)
    ok &= c1
    dist.barrier()
    torch.cuda.synchronize()
    # 2 ---------------------------------------------------------------------------------------------------------
    t0 = time.time()
    if rank == 0:
        step("fused")
        try:
            px.check_allreduce(wait=True)
            c2 = False
        except RuntimeError as e:
            c2 = "all-reduce failed" in str(e)
        dt = time.time() - t0
        c2 = c2 and dt < 3 * budget_s + 5.0  # ready wait + done wait, generous host slack
        print(f"[rank 0] peer skipped its backward: kernel returned after {dt:.2f} s and the failure was reported: {c2}",
              flush=True)
        ok &= c2
    else:
        held = step("fused", backward=False)  # the forward's exchange is collective; then "raise" before the backward
    torch.cuda.synchronize()
    dist.barrier()
    # 3 ---------------------------------------------------------------------------------------------------------
    if rank == 1:
        t0 = time.time()
        torch.autograd.backward([held[1], held[2]], [g_lp, g_ent])  # same epoch as rank 0's abandoned call
        try:
            px.check_allreduce(wait=True)
            c3 = False
        except RuntimeError as e:
            c3 = "all-reduce failed" in str(e)
        dt = time.time() - t0
        c3 = c3 and dt < 0.6 * budget_s + 0.5  # the poisoned flag is seen at once: no wait budget is spent
        print(f"[rank 1] late backward met the poisoned flag after {dt:.2f} s and failed as well: {c3}", flush=True)
        ok &= c3
    torch.cuda.synchronize()
    flag = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok = bool(flag.item() == 1.0)
    if rank == 0:
        print("RESULT", "OK" if ok else "MISMATCH", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
