"""torchrun check (>= 2 GPUs): vocab_parallel_entropy / vocab_parallel_log_probs_from_logits on real NCCL ranks and
postprocess_packed_seqs with a real context-parallel all-gather, against the single-GPU full computation.
Prints MISMATCH on failure.  usage: torchrun --nproc-per-node N scripts/tp_megatron_check.py"""

import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from verl_b200.models.mcore import util as mutil  # noqa: E402
from verl_b200.utils.megatron import tensor_parallel as tpmod  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl")
    ok = True
    g = torch.Generator(device=dev).manual_seed(1234)  # same data on every rank
    T, V = 4097, 1024 * world
    logits = torch.randn(T, V, device=dev, generator=g) * 3
    labels = torch.randint(0, V, (T,), device=dev, generator=g)
    g_e = torch.randn(T, device=dev, generator=g)
    g_l = torch.randn(T, device=dev, generator=g)
    full = logits.clone().requires_grad_(True)
    ent_f = tpmod.vocab_parallel_entropy(full, group=None)
    tpmod.set_tensor_model_parallel_group(None)
    lp_f = tpmod.vocab_parallel_log_probs_from_logits(full, labels)
    (ent_f * g_e + lp_f * g_l).sum().backward()
    v_local = V // world
    shard = logits[:, rank * v_local:(rank + 1) * v_local].contiguous().requires_grad_(True)
    tpmod.set_tensor_model_parallel_group(dist.group.WORLD)
    ent = tpmod.vocab_parallel_entropy(shard)
    lp = tpmod.vocab_parallel_log_probs_from_logits(shard, labels)
    (ent * g_e + lp * g_l).sum().backward()
    for name, a, b, tol in [("entropy", ent, ent_f, 2e-5), ("logprobs", lp, lp_f, 2e-5),
                            ("dlogits", shard.grad, full.grad[:, rank * v_local:(rank + 1) * v_local], 1e-5)]:
        err = (a.detach() - b.detach()).abs().max().item()
        if not err <= tol:
            ok = False
            print(f"MISMATCH rank {rank} {name}: {err}")
    # ---- context-parallel packed sequences: cp = world
    B, S, D = 8, 512, 64
    lens = torch.randint(4 * world, S + 1, (B,), device=dev, generator=g)
    attn = torch.arange(S, device=dev)[None] < lens[:, None]
    ids = torch.randint(0, 1000, (B, S), device=dev, generator=g)
    table = torch.randn(1000, D, device=dev, generator=g)
    packed, params = mutil.preprocess_packed_seqs(ids, attn, tp_size=1, cp_size=world, cp_rank=rank,
                                                  cp_group=dist.group.WORLD)
    out = table[packed[0]].unsqueeze(0).requires_grad_(True)
    padded = mutil.postprocess_packed_seqs(out, params, attn, B, S)
    want = table[ids] * attn[..., None]
    if not torch.equal(padded.detach(), want):
        ok = False
        print(f"MISMATCH rank {rank} packed round trip: {(padded.detach() - want).abs().max().item()}")
    padded.sum().backward()
    owned = params._vpg["local_dst"] >= 0
    g_want = torch.zeros(packed.shape[1], D, device=dev)
    g_want[params._vpg["local_dst"][owned]] = 1.0
    if not torch.equal(out.grad[0], g_want):
        ok = False
        print(f"MISMATCH rank {rank} packed backward")
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    if rank == 0:
        print("tp_megatron_check", "OK" if flag.item() == 0 else "FAILED", f"world={world}")
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 0 else 1)


if __name__ == "__main__":
    main()
