"""One forward + one backward of the fused lm-head at BASELINE configs[1] (for ncu captures)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from verl_b200 import ops

T, H, V = int(os.environ.get("PROF_T", 65536)), 3584, 152064
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
hidden = (torch.rand(T, H, device=dev, generator=g) - 0.5).bfloat16()
weight = (torch.rand(V, H, device=dev, generator=g) - 0.5).bfloat16()
labels = torch.randint(0, V, (T,), device=dev, generator=g)
g_lp = torch.rand(T, device=dev, generator=g) * 2 - 1
g_ent = torch.rand(T, device=dev, generator=g) - 0.5
part = ops.lce_forward_partials(hidden, weight, labels, 1.0)
lp, ent, lse, eb = ops.lce_finalize(part, 1.0)
dh, dw = ops.lce_backward(hidden, weight, labels, lse, eb, g_lp, g_ent, 1.0)
torch.cuda.synchronize()
print("ok", float(lp.mean()), float(ent.mean()), float(dh.float().abs().mean()), float(dw.float().abs().mean()))
