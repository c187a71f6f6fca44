"""Where one c3 micro-batch of fused_actor_step spends its time: every stage timed on the host with a device sync on both
sides (coarse, for finding stalls — not a benchmark)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from verl_b200 import ops  # noqa: E402,F401  (loads the library)
from verl_b200.trainer.ppo import core_algos  # noqa: E402
from verl_b200.workers.actor import dp_actor  # noqa: E402

dev = torch.device("cuda:0")
gen = torch.Generator(device=dev).manual_seed(1234)
H, V, mb, P, L = 3584, 152064, 16, 1024, 1024
S = P + L
weight = ((torch.rand(V, H, device=dev, generator=gen) - 0.5)).bfloat16().requires_grad_(True)
p_len = torch.randint(P // 2, P + 1, (mb,), device=dev, generator=gen)
r_len = torch.randint(L // 4, L + 1, (mb,), device=dev, generator=gen)
pos = torch.arange(S, device=dev)[None]
attn = ((pos >= (P - p_len)[:, None]) & (pos < (P + r_len)[:, None])).long()
Tm = int(attn.sum())
hidden = (torch.rand(Tm, H, device=dev, generator=gen) - 0.5).bfloat16()
ids = torch.randint(0, V, (mb, S), device=dev, generator=gen)
rmask = attn[:, P:].contiguous()
old = -torch.rand(mb, L, device=dev, generator=gen) * 3
adv = torch.randn(mb, L, device=dev, generator=gen)
ref = old + 0.05 * torch.randn(mb, L, device=dev, generator=gen)


def stage(name, fn, acc):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = fn()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    acc.setdefault(name, []).append(((t1 - t0) * 1e3, (t2 - t0) * 1e3))
    return out


acc = {}
for it in range(8):
    h = hidden.detach().requires_grad_(True)
    weight.grad = None
    win = stage("response_window", lambda: dp_actor.response_window(ids, attn, L), acc)
    ent_lp = stage("lm_head_fwd(+gather/scatter)", lambda: dp_actor.forward_micro_batch_from_hidden(h, weight, win, 1.0, calculate_entropy=False), acc)
    lm = stage("fused_actor_loss", lambda: core_algos.fused_actor_loss(old, ent_lp[1], adv, rmask, cliprange=0.2, ref_log_prob=ref, kl_loss_coef=0.01, kl_loss_type="low_var_kl", loss_scale=0.125), acc)
    stage("backward", lambda: lm[0].backward(), acc)
    stage("whole fused_actor_step", lambda: dp_actor.fused_actor_step(hidden.detach().requires_grad_(True), weight, ids, attn, old, adv, rmask, temperature=1.0, ref_log_prob=ref, clip_ratio=0.2, use_kl_loss=True, kl_loss_coef=0.01, loss_scale=0.125), acc)
for k, v in acc.items():
    v = v[2:]
    print(f"{k:32s} host-enqueue ms {sum(a for a, _ in v) / len(v):8.3f}   until-done ms {sum(b for _, b in v) / len(v):8.3f}   max {max(b for _, b in v):8.3f}")
