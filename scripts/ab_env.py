"""A/B of environment-controlled kernel knobs inside ONE process (same box, same thermal state): for every setting,
3 warm-up + N timed fwd+bwd steps of the lm-head at configs[1]; settings are interleaved over two rounds.
    python scripts/ab_env.py "VPG_DS_STCS=1" "VPG_HINT_A_DW=1,VPG_HINT_B_DW=2" ...   ("" = defaults, always included)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from verl_b200 import ops

T, H, V = 65536, 3584, 152064
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
hidden = (torch.rand(T, H, device=dev, generator=g) - 0.5).bfloat16()
weight = (torch.rand(V, H, device=dev, generator=g) - 0.5).bfloat16()
labels = torch.randint(0, V, (T,), device=dev, generator=g)
g_lp = torch.rand(T, device=dev, generator=g) * 2 - 1
g_ent = torch.rand(T, device=dev, generator=g) - 0.5


def step():
    part = ops.lce_forward_partials(hidden, weight, labels, 1.0)
    lp, ent, lse, eb = ops.lce_finalize(part, 1.0)
    return ops.lce_backward(hidden, weight, labels, lse, eb, g_lp, g_ent, 1.0)


def run(setting, n=6):
    keys = []
    for kv in filter(None, setting.split(",")):
        k, v = kv.split("=")
        os.environ[k] = v
        keys.append(k)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    fw = bw = 0.0
    for _ in range(n):
        e0.record()
        part = ops.lce_forward_partials(hidden, weight, labels, 1.0)
        lp, ent, lse, eb = ops.lce_finalize(part, 1.0)
        e1.record()
        dh, dw = ops.lce_backward(hidden, weight, labels, lse, eb, g_lp, g_ent, 1.0)
        e2.record()
        torch.cuda.synchronize()
        fw += e0.elapsed_time(e1)
        bw += e1.elapsed_time(e2)
    for k in keys:
        del os.environ[k]
    chk = (float(dh.float().abs().mean()), float(dw.float().abs().mean()))
    return fw / n, bw / n, chk


settings = [""] + sys.argv[1:]
for rnd in range(2):
    for s in settings:
        fw, bw, chk = run(s)
        print(f"round {rnd} [{s or 'defaults'}] fwd {fw:.2f} ms  bwd {bw:.2f} ms  step {fw + bw:.2f} ms  chk {chk[0]:.6f} {chk[1]:.6f}", flush=True)
