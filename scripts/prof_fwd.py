"""Forward-only launch of the fused lm-head at BASELINE configs[1] (for ncu metric sweeps)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from verl_b200 import ops
T, H, V = int(os.environ.get("PROF_T", 65536)), 3584, 152064
S = int(os.environ.get("PROF_S", 0))
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
hidden = (torch.rand(T, H, device=dev, generator=g) - 0.5).bfloat16()
weight = (torch.rand(V, H, device=dev, generator=g) - 0.5).bfloat16()
labels = torch.randint(0, V, (T,), device=dev, generator=g)
for _ in range(int(os.environ.get("PROF_N", 2))):
    part = ops.lce_forward_partials(hidden, weight, labels, 1.0, num_splits=S)
lp, ent, lse, eb = ops.lce_finalize(part, 1.0)
torch.cuda.synchronize()
print("ok", part.shape, float(lp.mean()), float(ent.mean()))
