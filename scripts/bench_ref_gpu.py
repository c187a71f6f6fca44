"""The reference's GPU path timed on the same B200, beside verl_b200 (SURVEY §2b / BASELINE.md §4).

Comparators (all from the UNMODIFIED reference snapshot oracle/_ref, loaded by oracle/ref_import.py):
  triton      verl.utils.kernel.linear_cross_entropy  (Triton kernels, kernels.py:474-663, 1345-1553; Triton JITs for sm_100)
  torch_fused verl.utils.experimental.torch_functional.FusedLinearForPPO  (the DEFAULT `torch` backend, chunk 512)
  ours        verl_b200.utils.kernel.linear_cross_entropy (tcgen05 / TMEM / TMA)
Protocol of the reference's own test (tests/utils/test_linear_cross_entropy.py:145-162, 289-348): inputs U(-0.5, 0.5)
bf16, g_logprobs U(-1, 1), g_entropy U(-0.5, 0.5); forward and backward timed separately with CUDA events; peak memory via
torch.cuda.max_memory_allocated after reset_peak_memory_stats.  Also cross-checks the three implementations' outputs.

    python scripts/bench_ref_gpu.py [--tokens 65536] [--hidden 3584] [--vocab 152064] [--iters 3] [--out FILE]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402


def measure(name, fwd, hidden, weight, labels, g_lp, g_ent, iters, warmup):
    out = {"impl": name}
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    f_ms, b_ms = [], []
    res = None
    for i in range(warmup + iters):
        h = hidden.detach().clone().requires_grad_(True)
        w = weight.detach().clone().requires_grad_(True)
        torch.cuda.synchronize()
        torch.cuda.reset_peak_memory_stats()
        base = torch.cuda.memory_allocated()
        e0, e1, e2 = ev(), ev(), ev()
        e0.record()
        lp, ent = fwd(h, w, labels)
        e1.record()
        torch.cuda.synchronize()
        peak_f = torch.cuda.max_memory_allocated() - base
        torch.cuda.reset_peak_memory_stats()
        e1b = ev()
        e1b.record()
        torch.autograd.backward([lp, ent], [g_lp.to(lp.dtype), g_ent.to(ent.dtype)])
        e2.record()
        torch.cuda.synchronize()
        peak_b = torch.cuda.max_memory_allocated() - base
        if i >= warmup:
            f_ms.append(e0.elapsed_time(e1))
            b_ms.append(e1b.elapsed_time(e2))
        res = (lp.detach().float().reshape(-1), ent.detach().float().reshape(-1), h.grad.float(), w.grad.float())
        out.update(peak_fwd_MB=peak_f / 2**20, peak_bwd_MB=peak_b / 2**20)
        del h, w, lp, ent
    out.update(fwd_ms=sum(f_ms) / len(f_ms), bwd_ms=sum(b_ms) / len(b_ms))
    out["fwd_bwd_ms"] = out["fwd_ms"] + out["bwd_ms"]
    return out, res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, default=65536)
    ap.add_argument("--hidden", type=int, default=3584)
    ap.add_argument("--vocab", type=int, default=152064)
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--temperature", type=float, default=1.0)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    T, H, V, temp = args.tokens, args.hidden, args.vocab, args.temperature
    torch.manual_seed(233376)
    hidden = torch.empty(T, H, dtype=torch.bfloat16, device=dev).uniform_(-0.5, 0.5)
    weight = torch.empty(V, H, dtype=torch.bfloat16, device=dev).uniform_(-0.5, 0.5)
    labels = torch.randint(0, V, (T,), device=dev)
    g_ent = torch.empty(T, dtype=torch.float32, device=dev).uniform_(-0.5, 0.5)
    g_lp = torch.empty(T, dtype=torch.float32, device=dev).uniform_(-1, 1)

    report = {"shape": {"T": T, "H": H, "V": V, "temperature": temp, "dtype": "bf16"},
              "gpu": torch.cuda.get_device_name(0), "impls": {}, "errors": {}}
    results = {}

    # ---- ours first (no dependency on the snapshot)
    from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy as ours

    m, results["ours"] = measure("verl_b200 tcgen05", lambda h, w, l: ours(h, w, l, temp, "none"), hidden, weight,
                                 labels, g_lp, g_ent, args.iters, args.warmup)
    report["impls"]["ours"] = m
    print(json.dumps(m), flush=True)

    from oracle import ref_import

    try:
        ns = ref_import.load()
        fused = ns.fused_torch.FusedLinearForPPO()
        m, results["torch_fused"] = measure(
            "reference FusedLinearForPPO (torch backend, chunk 512, bf16 inputs as dp_actor passes them)",
            lambda h, w, l: fused(h, w, l, temperature=temp), hidden, weight, labels, g_lp, g_ent, args.iters, 1)
        report["impls"]["torch_fused"] = m
        print(json.dumps(m), flush=True)
    except Exception as e:  # noqa: BLE001
        report["errors"]["torch_fused"] = f"{type(e).__name__}: {e}"
        print("torch_fused failed:", report["errors"]["torch_fused"], flush=True)
    try:
        _k, lce_mod = ref_import.load_kernel()
        ref_op = lce_mod.linear_cross_entropy
        m, results["triton"] = measure("reference Triton linear_cross_entropy (kernels.py, default _Split_Dlogits_N backward)",
                                       lambda h, w, l: ref_op(h, w, l, temp, "none"), hidden, weight, labels, g_lp,
                                       g_ent, args.iters, args.warmup)
        report["impls"]["triton"] = m
        print(json.dumps(m), flush=True)
    except Exception as e:  # noqa: BLE001
        import traceback

        report["errors"]["triton"] = f"{type(e).__name__}: {e}"
        traceback.print_exc()

    # ---- cross-check: ours vs each reference implementation on the same inputs
    names = ["lp", "ent", "d_hidden", "d_weight"]
    for ref_name in ("triton", "torch_fused"):
        if ref_name not in results:
            continue
        cmp = {}
        for n, a, b in zip(names, results["ours"], results[ref_name]):
            cmp[n + "_max_abs"] = float((a - b).abs().max())
        report[f"ours_vs_{ref_name}"] = cmp
    o = report["impls"].get("ours")
    for ref_name in ("triton", "torch_fused"):
        r = report["impls"].get(ref_name)
        if r and o:
            report[f"speedup_over_{ref_name}"] = {"fwd": r["fwd_ms"] / o["fwd_ms"], "bwd": r["bwd_ms"] / o["bwd_ms"],
                                                  "fwd_bwd": r["fwd_bwd_ms"] / o["fwd_bwd_ms"]}
    s = json.dumps(report, indent=1)
    print(s)
    if args.out:
        with open(args.out, "w") as f:
            f.write(s)


if __name__ == "__main__":
    main()
