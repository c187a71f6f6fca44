"""Per-kernel device times of ONE rank's share of the vocab-parallel lm-head, run on a single GPU (no collectives):
the forward partials + finalize and the backward of a [V_local, H] weight shard over T tokens, timed per launch with
the library's CUDA-event timing API.  Used to A/B kernel-level changes (split-K tail: VPG_SPLITK=0/1, read once per
process) at the 8-way shard shapes of BASELINE configs[1]/[3]/[4] without an 8-GPU box.

    python scripts/shard_kernel_times.py --tokens 65536 --hidden 3584 --vocab-local 19008 [--iters 5]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from verl_b200 import _lib, ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, default=65536)
    ap.add_argument("--hidden", type=int, default=3584)
    ap.add_argument("--vocab-local", type=int, default=19008)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--chunk-v", type=int, default=0)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    T, H, V = args.tokens, args.hidden, args.vocab_local
    g = torch.Generator(device=dev).manual_seed(1)
    hidden = (torch.rand(T, H, device=dev, generator=g) - 0.5).bfloat16()
    weight = (torch.rand(V, H, device=dev, generator=g) - 0.5).bfloat16()
    labels = torch.randint(0, V, (T,), device=dev, generator=g)
    g_lp = torch.rand(T, device=dev, generator=g) * 2 - 1
    g_ent = torch.rand(T, device=dev, generator=g) - 0.5

    def step():
        part = ops.lce_forward_partials(hidden, weight, labels, 1.0)
        lp, ent, lse, eb = ops.lce_finalize(part, 1.0)
        return ops.lce_backward(hidden, weight, labels, lse, eb, g_lp, g_ent, 1.0, chunk_v=args.chunk_v)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    _lib.timing_enable(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.iters):
        step()
    e1.record()
    torch.cuda.synchronize()
    rows = _lib.timing_collect()
    _lib.timing_enable(False)
    agg = {}
    for k, ms in rows:
        agg.setdefault(k, []).append(ms)
    flops = 2.0 * T * H * V
    out = {"shape": {"T": T, "H": H, "V_local": V}, "splitk_env": os.environ.get("VPG_SPLITK", "1"),
           "step_ms": e0.elapsed_time(e1) / args.iters, "ideal_flops_per_gemm": flops,
           "kernels": {k: {"launches_per_step": len(v) / args.iters, "ms_per_step": sum(v) / args.iters,
                           "tflops": (flops / (sum(v) / args.iters * 1e-3) / 1e12) if k != "splitk_reduce" else None}
                       for k, v in agg.items()},
           "workspace": ops.lce_backward_workspace_info(T, H, V, dev, args.chunk_v)}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
