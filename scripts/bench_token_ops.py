"""HBM-roofline evidence for the memory-bound token kernels: achieved GB/s at 2^25 tokens (inputs >> L2) and
latencies at the BASELINE config sizes.  Prints one JSON line per op.

    python scripts/bench_token_ops.py [--tokens-log2 25]
"""

from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from verl_b200 import bert_padding, ops  # noqa: E402
from verl_b200.trainer.ppo import core_algos as ca  # noqa: E402
from verl_b200.utils import torch_functional as F  # noqa: E402


def peak_hbm():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def timeit(fn, iters=10, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def timeit_graph(fn, iters=200):
    """Replay latency of `fn` captured into a CUDA graph (None if the op cannot be captured: it reads a value back)."""
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            keep = fn()  # noqa: F841  (outputs live in the graph's pool)
        for _ in range(5):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters
    except Exception:  # noqa: BLE001
        torch.cuda.synchronize()
        return None


# ops whose wrapper reads nothing back on the host (no ValueError semantics, no token counts): capturable
GRAPHABLE = ("ppo_loss", "agg_loss", "masked_sum", "kl_penalty", "grpo_advantage", "remax", "value_loss", "kl_reward",
             "gather_rows", "device time")


def make(B, L, dev, seed=0):
    g = torch.Generator(device=dev).manual_seed(seed)
    lens = torch.randint(L // 4, L + 1, (B,), device=dev, generator=g)
    mask = (torch.arange(L, device=dev)[None] < lens[:, None]).long()
    old = -torch.rand(B, L, device=dev, generator=g) * 3
    lp = old + 0.05 * torch.randn(B, L, device=dev, generator=g)
    ref = old + 0.05 * torch.randn(B, L, device=dev, generator=g)
    adv = torch.randn(B, L, device=dev, generator=g)
    ent = torch.rand(B, L, device=dev, generator=g)
    rewards = torch.zeros(B, L, device=dev)
    rewards[torch.arange(B, device=dev), lens - 1] = torch.rand(B, device=dev, generator=g)
    values = torch.randn(B, L, device=dev, generator=g)
    return dict(mask=mask, old=old, lp=lp, ref=ref, adv=adv, ent=ent, rewards=rewards, values=values, lens=lens)


def ops_table(d, B, L, n_uid):
    uid = np.repeat(np.arange((B + n_uid - 1) // n_uid), n_uid)[:B]
    gids = torch.from_numpy(uid.astype(np.int32)).to(d["mask"].device)
    ng = int(uid.max()) + 1
    attn = d["mask"]
    ids = torch.zeros_like(attn)
    idx, _, _, _ = ops.unpad_indices(attn)
    nnz = idx.numel()
    flat_ids = ids.reshape(-1, 1)
    vals = d["lp"].reshape(-1, 1)[:nnz].contiguous()
    n = B * L
    # (name, callable, algorithmic bytes)
    return [
        ("ppo_loss fwd+grad (clip + entropy + low_var_kl, token-mean)",
         lambda: ops.ppo_loss(d["old"], d["lp"], d["adv"], d["mask"], cliprange_low=0.2, cliprange_high=0.28,
                              entropy=d["ent"], entropy_coeff=1e-3, ref_log_prob=d["ref"], kl_loss_coef=0.01),
         n * (8 + 4 * 5 + 4 * 2)),  # SURVEY 8(d) floor: mask once, old/lp/adv/ent/ref, d_lp/d_ent (the kernels move 38)
        ("ppo_loss fwd only (vanilla)",
         lambda: ops.ppo_loss(d["old"], d["lp"], d["adv"], d["mask"], cliprange_low=0.2, cliprange_high=0.2,
                              want_grad=False),
         n * (8 + 12)),
        ("agg_loss token-mean (masked_mean)", lambda: ops.agg_loss_fwd(d["ent"], d["mask"], "token-mean", False),
         n * (8 + 4)),
        ("masked_sum", lambda: ops.masked_sums(d["ent"], d["mask"]), n * 12),
        ("kl_penalty low_var_kl", lambda: ops.kl_penalty_fwd(d["lp"], d["ref"], "low_var_kl", False), n * 12),
        ("grpo_advantage", lambda: ops.grpo_advantage(d["rewards"], d["mask"], gids, ng, 1e-6, True), n * 16),
        ("gae + whiten", lambda: ops.gae(d["rewards"], d["values"], d["mask"], 0.99, 0.95), n * (16 + 8 + 8)),
        ("masked_whiten", lambda: ops.masked_whiten(d["values"], d["mask"]), n * (12 + 8)),
        ("rloo_advantage (outcome_advantage skeleton)",
         lambda: ops.outcome_advantage("rloo", d["rewards"], d["mask"], gids, ng), n * (12 + 8 + 4)),
        ("reinforce++ (reverse scan + whiten*mask)", lambda: ops.reinforce_plus_plus(d["rewards"], d["mask"], 0.99),
         n * (12 + 4 + 12 + 4)),
        ("remax (reverse cumsum)", lambda: ops.remax(d["rewards"], d["lens"].float(), d["mask"]), n * (12 + 8)),
        ("value_loss fwd+grad", lambda: ops.value_loss(d["values"], d["lp"], d["old"], d["mask"], 0.2, "token-mean", True),
         n * (8 + 12 + 4)),
        ("kl_reward (apply_kl_penalty)", lambda: ops.kl_reward(d["rewards"], d["old"], d["ref"], d["mask"], "kl", 0.01),
         n * (8 + 12 + 4)),
        ("data_metrics (compute_data_metrics sweep)",
         lambda: ops.data_metrics(d["rewards"], d["rewards"], d["adv"], d["lp"], d["values"], d["mask"], attn),
         n * (8 + 20 + 8)),
        # mask read once + one int64 index per valid token (the 1-bit map adds n/4 bytes, not counted).  The
        # call ends in a host read of {nnz, max_seqlen} like the reference's nonzero() / .item(); the second row times the
        # two kernels alone (launches back to back, no host read).
        ("unpad_indices (int64 mask, incl. host read of nnz)", lambda: ops.unpad_indices(attn), n * 8 + nnz * 8),
        ("unpad_indices (int64 mask, device time)", lambda: ops.unpad_indices_launch(attn), n * 8 + nnz * 8),
        ("gather_rows int64 ids (index_first_axis)", lambda: ops.gather_rows(flat_ids, idx), nnz * 24),
        ("pad_input fp32 [nnz,1] -> [B*L,1]", lambda: ops.scatter_rows(vals, idx, n), nnz * 12 + n * 4),
    ]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens-log2", type=int, default=25)
    ap.add_argument("--only", default="", help="comma-separated substrings: run only the ops whose name contains one")
    ap.add_argument("--no-configs", action="store_true", help="skip the latencies at the BASELINE config sizes")
    ap.add_argument("--graphs", action="store_true",
                    help="config sizes: also the replay latency of every op that can be captured into a CUDA graph")
    args = ap.parse_args()
    only = [x for x in args.only.split(",") if x]
    dev = torch.device("cuda:0")
    peak, src = peak_hbm()
    out = []
    L = 16384
    B = (1 << args.tokens_log2) // L
    d = make(B, L, dev)
    for name, fn, nbytes in ops_table(d, B, L, 8):
        if only and not any(o in name for o in only):
            continue
        ms = timeit(fn)
        gbs = nbytes / (ms * 1e-3) / 1e9
        rec = {"op": name, "tokens": B * L, "ms": round(ms, 4), "algorithmic_bytes": nbytes, "achieved_gbs": round(gbs, 1),
               "peak_gbs": peak, "peak_source": src, "frac": round(gbs / peak, 3)}
        out.append(rec)
        print(json.dumps(rec), flush=True)
    del d
    if args.no_configs:
        return
    # latencies at config sizes: C1 40x512, C3 128x1024 (n=8), C5 16x16384
    for cname, (B, L, n_uid) in {"C1 40x512": (40, 512, 5), "C3 128x1024": (128, 1024, 8), "C5 16x16384": (16, 16384, 4)}.items():
        d = make(B, L, dev)
        for name, fn, nbytes in ops_table(d, B, L, n_uid):
            if only and not any(o in name for o in only):
                continue
            ms = timeit(fn, iters=50, warmup=5)
            rec = {"config": cname, "op": name, "tokens": B * L, "us": round(ms * 1e3, 2)}
            if args.graphs and any(g in name for g in GRAPHABLE):
                gms = timeit_graph(fn)
                rec["us_graph_replay"] = None if gms is None else round(gms * 1e3, 2)
            print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
