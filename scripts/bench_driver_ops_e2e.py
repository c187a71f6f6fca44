"""Driver-side token ops END TO END on host tensors — the way the reference's driver calls them
(verl/trainer/ppo/ray_trainer.py:1283-1305): host tensors in, host tensors out.

For each BASELINE batch shape (C1 40x512, C3 128x1024, C5 16x16384) and each op, times
  ours       verl_b200's reference-named function on CPU tensors: pinned-arena staging -> H2D -> CUDA kernels -> D2H
  reference  the reference's own function (snapshot oracle/_ref, unmodified) on the same CPU tensors, all host threads
and checks that the two results agree (1e-5 relative).  One JSON line per (shape, op).

    python scripts/bench_driver_ops_e2e.py [--iters 20]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402


def timeit(fn, iters, warmup=2):
    for _ in range(warmup):
        out = fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(iters):
        out = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / iters * 1e6, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    from oracle import ref_import

    ref = ref_import.load().core_algos  # the reference's module (imports verl.* from the snapshot)
    from verl_b200.trainer.ppo import core_algos as ours

    torch.set_num_threads(os.cpu_count() or 1)
    shapes = {"C1 40x512 (8 uid x 5)": (40, 512, 5), "C3 128x1024 (16 uid x 8)": (128, 1024, 8),
              "C5 16x16384 (4 uid x 4)": (16, 16384, 4)}
    for name, (B, L, n) in shapes.items():
        g = torch.Generator().manual_seed(B + L)
        lens = torch.randint(L // 4, L + 1, (B,), generator=g)
        mask = (torch.arange(L)[None] < lens[:, None]).long()
        rewards = torch.zeros(B, L)
        rewards[torch.arange(B), lens - 1] = torch.rand(B, generator=g)
        values = torch.randn(B, L, generator=g)
        old = -torch.rand(B, L, generator=g) * 3
        lp = old + 0.05 * torch.randn(B, L, generator=g)
        adv = torch.randn(B, L, generator=g)
        uid = np.repeat(np.array([f"u{i}" for i in range(B // n)], dtype=object), n)
        cases = {
            "compute_gae_advantage_return": lambda m: m.compute_gae_advantage_return(rewards, values, mask, 0.99, 0.95),
            "compute_grpo_outcome_advantage": lambda m: m.compute_grpo_outcome_advantage(rewards, mask, uid),
            "compute_rloo_outcome_advantage": lambda m: m.compute_rloo_outcome_advantage(rewards, mask, uid),
            "kl_penalty(low_var_kl)": lambda m: m.kl_penalty(lp, old, "low_var_kl"),
            "compute_policy_loss (forward)": lambda m: m.compute_policy_loss(old, lp, adv, mask, cliprange=0.2,
                                                                             cliprange_low=0.2, cliprange_high=0.28),
        }
        for op, call in cases.items():
            with torch.no_grad():
                us_ours, o = timeit(lambda: call(ours), args.iters)
                ref_iters = 1 if (op.startswith("compute_gae") and L > 4096) else max(2, args.iters // 4)
                t0 = time.perf_counter()
                for _ in range(ref_iters):
                    r = call(ref)
                us_ref = (time.perf_counter() - t0) / ref_iters * 1e6
            o0 = o[0] if isinstance(o, tuple) else o
            r0 = r[0] if isinstance(r, tuple) else r
            assert not o0.is_cuda, "host tensors in -> host tensors out"
            err = float((o0.float() - r0.float()).abs().max() / r0.float().abs().max().clamp_min(1e-12))
            print(json.dumps({"shape": name, "op": op, "ours_host_to_host_us": round(us_ours, 1),
                              "reference_cpu_us": round(us_ref, 1), "speedup": round(us_ref / us_ours, 2),
                              "max_rel_err": err, "ok": err < 1e-4, "host_threads": torch.get_num_threads()}), flush=True)


if __name__ == "__main__":
    main()
