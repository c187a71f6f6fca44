"""Host-side cost of the token-op wrappers at a BASELINE micro-batch size (C5 16x16384): cProfile over 2000 calls each.
At these sizes the ops are host-bound, so this is where their latency is."""
import cProfile
import io
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))

import torch  # noqa: E402

import bench_token_ops as b  # noqa: E402
from verl_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
B, L = 16, 16384
d = b.make(B, L, dev)
cases = {
    "ppo_loss": lambda: ops.ppo_loss(d["old"], d["lp"], d["adv"], d["mask"], cliprange_low=0.2, cliprange_high=0.28,
                                     entropy=d["ent"], entropy_coeff=1e-3, ref_log_prob=d["ref"], kl_loss_coef=0.01),
    "masked_sums": lambda: ops.masked_sums(d["ent"], d["mask"]),
    "kl_penalty": lambda: ops.kl_penalty_fwd(d["lp"], d["ref"], "low_var_kl", False),
    "gae": lambda: ops.gae(d["rewards"], d["values"], d["mask"], 0.99, 0.95),
    "unpad_indices": lambda: ops.unpad_indices(d["mask"]),
}
for name, fn in cases.items():
    for _ in range(50):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(2000):
        fn()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"== {name}: host {1e6 * (t1 - t0) / 2000:.1f} us per call, with drain {1e6 * (t2 - t0) / 2000:.1f} us")
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(2000):
        fn()
    pr.disable()
    torch.cuda.synchronize()
    s = io.StringIO()
    pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(14)
    print("\n".join(l for l in s.getvalue().splitlines() if l.strip())[:2600])
