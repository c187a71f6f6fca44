"""BASELINE configs[2] micro-batch (Qwen2.5-VL-7B GRPO actor update: 16 sequences x (1k prompt + 1k response), KL loss,
token-mean): the row-selected fused actor step against the reference's dataflow on the same kernels (lm-head over
every rmpad token -> pad_input -> slice -> loss).  Prints one JSON line.  CUDA events, 3 warm-up + 10 timed."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from verl_b200 import bert_padding
from verl_b200.trainer.ppo import core_algos as ca
from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy
from verl_b200.workers.actor.dp_actor import fused_actor_step

dev = torch.device("cuda")
B, P, L, H, V = 16, 1024, 1024, 3584, 152064
g = torch.Generator(device=dev).manual_seed(0)
S = P + L
p_len = torch.randint(P // 2, P + 1, (B,), device=dev, generator=g)
r_len = torch.randint(L // 4, L + 1, (B,), device=dev, generator=g)
pos = torch.arange(S, device=dev)[None]
attn = ((pos >= (P - p_len)[:, None]) & (pos < (P + r_len)[:, None])).long()
rmask = attn[:, P:].contiguous()
ids = torch.randint(0, V, (B, S), device=dev, generator=g)
T = int(attn.sum())
hidden = (torch.rand(T, H, device=dev, generator=g) - 0.5).bfloat16()
weight = (torch.rand(V, H, device=dev, generator=g) - 0.5).bfloat16().requires_grad_(True)
old = -torch.rand(B, L, device=dev, generator=g) * 3
ref = old + 0.05 * torch.randn(B, L, device=dev, generator=g)
adv = torch.randn(B, L, device=dev, generator=g)
cfg = dict(clip_ratio=0.2, entropy_coeff=0.0, use_kl_loss=True, kl_loss_coef=0.01, loss_agg_mode="token-mean")


def fused():
    h = hidden.detach().requires_grad_(True)
    weight.grad = None
    return fused_actor_step(h, weight, ids, attn, old, adv, rmask, temperature=1.0, ref_log_prob=ref, **cfg)[1]


def all_rows():
    h = hidden.detach().requires_grad_(True)
    weight.grad = None
    ids_rmpad, indices, *_ = bert_padding.unpad_input(ids.unsqueeze(-1), attn)
    labels = torch.roll(ids_rmpad.squeeze(-1), -1, 0)
    lp_all, _ = linear_cross_entropy(h, weight, labels, 1.0, "none")
    lp = bert_padding.pad_input(lp_all.unsqueeze(-1), indices, B, S).squeeze(-1)[:, -L - 1:-1]
    loss, met = ca.fused_actor_loss(old, lp, adv, rmask, cliprange=0.2, ref_log_prob=ref, kl_loss_coef=0.01)
    loss.backward()
    return met


def time(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


t_sel = int(((pos[:, :] >= S - L - 1) & (pos < S - 1) & attn.bool()).sum())
ms_f, ms_a = time(fused), time(all_rows)
print(json.dumps({"workload": f"configs[2] micro-batch: {B} seq x ({P}+{L}), H={H}, V={V}", "stream_tokens": T,
                  "response_window_tokens": t_sel, "fused_actor_step_ms": ms_f, "all_rows_path_ms": ms_a,
                  "speedup": ms_a / ms_f, "sequences_per_s_fused": B / (ms_f * 1e-3),
                  "stream_tokens_per_s_fused": T / (ms_f * 1e-3)}))
