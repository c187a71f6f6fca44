"""Vocab-parallel lm-head on N ranks (torchrun): every exchange / reduction mode (NCCL all-gather + caller all-reduce,
NVLink-fused partial push, multimem.red, NCCL overlapped, IN-KERNEL NVLS all-reduce) against an fp32 PyTorch reference
computed from the full (unsharded) weight — the protocol of the reference's own TP test
(tests/utils/test_linear_cross_entropy_tp.py:364-427: sharded kernel vs all-gather torch implementation, d_hidden
all-reduced outside) — and against each other and the full-vocab single-GPU kernel.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        scripts/tp_check.py [--tokens T] [--hidden H] [--vocab V] [--time]
Exit code 0 = all ranks agree within the parity tolerances."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from verl_b200.utils.kernel import kernels
from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, default=4099)
    ap.add_argument("--hidden", type=int, default=1024)
    ap.add_argument("--vocab", type=int, default=32768)
    ap.add_argument("--time", action="store_true")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    group = dist.group.WORLD
    T, H, V = args.tokens, args.hidden, args.vocab
    assert V % world == 0
    g = torch.Generator(device=dev).manual_seed(1234)
    hidden = (torch.rand(T, H, device=dev, generator=g) - 0.5).bfloat16()
    weight_full = (torch.rand(V, H, device=dev, generator=g) - 0.5).bfloat16()
    labels = torch.randint(0, V, (T,), device=dev, generator=g)
    g_lp = torch.rand(T, device=dev, generator=g) * 2 - 1
    g_ent = torch.rand(T, device=dev, generator=g) - 0.5
    vl = V // world
    w_local = weight_full[rank * vl:(rank + 1) * vl].contiguous()

    def run(mode, reduce_inside):
        kernels.set_vocab_parallel_exchange(mode)
        kernels.set_reduce_d_hidden(bool(reduce_inside), mode=reduce_inside or "nccl")
        h = hidden.clone().requires_grad_(True)
        w = w_local.clone().requires_grad_(True)
        lp, ent = linear_cross_entropy(h, w, labels, 1.5, "none", group)
        torch.autograd.backward([lp, ent], [g_lp, g_ent])
        dh = h.grad
        if not reduce_inside:
            dist.all_reduce(dh, group=group)
        return lp.detach(), ent.detach(), dh, w.grad

    # ---- fp32 PyTorch reference from the FULL weight (run_torch_entropy recipe, tests/utils/test_linear_cross_entropy.py:48-61)
    h32 = hidden.float().requires_grad_(True)
    w32 = weight_full.float().requires_grad_(True)
    logits = (h32 @ w32.t()) / 1.5
    lsm = torch.log_softmax(logits, dim=-1)
    lp_ref = lsm.gather(-1, labels[:, None]).squeeze(-1)
    ent_ref = -(lsm.exp() * lsm).sum(-1)
    torch.autograd.backward([lp_ref, ent_ref], [g_lp, g_ent])
    torch_ref = (lp_ref.detach(), ent_ref.detach(), h32.grad, w32.grad[rank * vl:(rank + 1) * vl])
    del logits, lsm

    def against_torch(got):
        e_lp = float((got[0] - torch_ref[0]).abs().max())
        e_ent = float((got[1] - torch_ref[1]).abs().max())
        bad_dh = float(((got[2].float() - torch_ref[2]).abs() - (2e-2 + 4e-2 * torch_ref[2].abs())).max())
        bad_dw = float(((got[3].float() - torch_ref[3]).abs() - (2e-2 + 4e-2 * torch_ref[3].abs())).max())
        return e_lp, e_ent, bad_dh, bad_dw, (e_lp < 2e-3 and e_ent < 2e-3 and bad_dh <= 0 and bad_dw <= 0)

    ref = run("nccl", False)
    ok = True
    e = against_torch(ref)
    ok &= e[4]
    if rank == 0:
        print(f"[nccl exchange + caller all-reduce vs fp32 torch, full weight] lp {e[0]:.2e} ent {e[1]:.2e} "
              f"d_hidden excess {e[2]:.2e} d_weight excess {e[3]:.2e} {'OK' if e[4] else 'MISMATCH'}", flush=True)
    for tag, mode, inside in (("nvlink", "nvlink", False), ("nvlink+mcred", "nvlink", "multimem"),
                              ("nvlink+nccl reduce", "nvlink", "nccl"), ("nccl+reduce", "nccl", "nccl"),
                              ("nvlink+overlapped reduce", "nvlink", "overlap"),
                              ("nvlink+in-kernel NVLS all-reduce", "nvlink", "fused")):
        for rep in range(3):  # repeated: exercises the double-buffered slots and the re-zeroed accumulator
            got = run(mode, inside)
            got = tuple(t.clone() for t in got)
        errs = [float((a.float() - b.float()).abs().max()) for a, b in zip(got, ref)]
        bitwise = [bool(torch.equal(a, b)) for a, b in zip(got[:2], ref[:2])]
        e = against_torch(got)
        good = errs[0] < 1e-5 and errs[1] < 1e-5 and errs[2] < 2e-2 and errs[3] < 2e-2 and e[4]  # d_weight: split-K plan follows the worker count
        ok &= good
        if rank == 0:
            print(f"[{tag}] vs nccl mode: lp {errs[0]:.2e} ent {errs[1]:.2e} d_hidden {errs[2]:.2e} d_weight {errs[3]:.2e} "
                  f"bitwise(lp,ent)={bitwise}; vs fp32 torch: lp {e[0]:.2e} ent {e[1]:.2e} excess dh {e[2]:.2e} dw {e[3]:.2e} "
                  f"{'OK' if good else 'MISMATCH'}", flush=True)
    # full-vocab single-GPU result on rank 0's device (same kernels, world = 1)
    h = hidden.clone().requires_grad_(True)
    w = weight_full.clone().requires_grad_(True)
    lp, ent = linear_cross_entropy(h, w, labels, 1.5, "none")
    torch.autograd.backward([lp, ent], [g_lp, g_ent])
    e = [float((lp - ref[0]).abs().max()), float((ent - ref[1]).abs().max()),
         float((h.grad.float() - ref[2].float()).abs().max())]
    good = e[0] < 2e-5 and e[1] < 2e-5 and e[2] < 2e-2
    ok &= good
    if rank == 0:
        from verl_b200.utils.kernel import peer
        print(f"[full-vocab vs vocab-parallel] lp {e[0]:.2e} ent {e[1]:.2e} d_hidden {e[2]:.2e} {'OK' if good else 'MISMATCH'}")
        print("multicast available:", peer.workspace(group, dev).has_multicast, flush=True)

    if args.time:
        def bench(mode, inside, n=10):
            for _ in range(3):
                run(mode, inside)
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                run(mode, inside)
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t)
        for tag, mode, inside in (("nccl (all-gather + caller all-reduce)", "nccl", False),
                                  ("nvlink push + NCCL all-reduce", "nvlink", "nccl"),
                                  ("nvlink push + in-kernel NVLS all-reduce", "nvlink", "fused"),
                                  ("nvlink push + multimem.red", "nvlink", "multimem")):
            ms = bench(mode, inside)
            if rank == 0:
                print(f"time {tag}: {ms:.3f} ms / fwd+bwd (T={T} H={H} V={V} world={world})", flush=True)
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    dist.destroy_process_group()
    sys.exit(0 if int(flag) == 0 else 1)


if __name__ == "__main__":
    main()
