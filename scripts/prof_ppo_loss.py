"""One fused-loss call with gradients at 2^25 tokens (the program ncu captures for the loss kernels):
    ncu --metrics ... -k regex:ppo_loss python scripts/prof_ppo_loss.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))

import torch  # noqa: E402

import bench_token_ops as b  # noqa: E402
from verl_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
L = 16384
B = (1 << 25) // L
d = b.make(B, L, dev)
for _ in range(3):
    ops.ppo_loss(d["old"], d["lp"], d["adv"], d["mask"], cliprange_low=0.2, cliprange_high=0.28, entropy=d["ent"],
                 entropy_coeff=1e-3, ref_log_prob=d["ref"], kl_loss_coef=0.01)
    ops.value_loss(d["values"], d["lp"], d["old"], d["mask"], 0.2, "token-mean", True)
torch.cuda.synchronize()
print("done")
