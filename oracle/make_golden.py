"""TEST INFRASTRUCTURE — generate tests/golden/*.npz by running the UNMODIFIED reference functions
(/root/reference, loaded through oracle.ref_import) on seeded inputs.  Run in the build container:

    python -m oracle.make_golden

The fixtures travel to the GPU box; the reference does not.  Every array is fp32 / int64 / int32 (bf16 inputs are
stored as their exact fp32 values).
"""

from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")


def _np(t):
    if isinstance(t, torch.Tensor):
        t = t.detach()
        if t.dtype == torch.bfloat16:
            t = t.float()
        return t.cpu().numpy()
    return np.asarray(t)


def T_(a):
    return torch.from_numpy(np.asarray(a))


def _bf16(shape, gen, lo=-0.5, hi=0.5):
    return (torch.rand(*shape, generator=gen) * (hi - lo) + lo).bfloat16()


def gen_lce(ref):
    """FusedLinearForPPO fed fp32 copies of bf16 data (the recipe of tests/utils/test_linear_cross_entropy.py:82-96)."""
    out = {}
    cases = {"a": (37, 128, 1000, 1.5), "b": (130, 256, 777, 0.7), "c": (64, 128, 1536, 1.0)}
    for name, (T, H, V, temp) in cases.items():
        g = torch.Generator().manual_seed(233376 + len(name) + T)
        hidden = _bf16((T, H), g)
        weight = _bf16((V, H), g)
        labels = torch.randint(0, V, (T,), generator=g)
        g_lp = torch.rand(T, generator=g) * 2 - 1
        g_ent = torch.rand(T, generator=g) - 0.5
        h32 = hidden.float().requires_grad_(True)
        w32 = weight.float().requires_grad_(True)
        lp, ent = ref.fused_torch.FusedLinearForPPO()(h32, w32, labels, temperature=temp)
        (lp * g_lp + ent * g_ent).sum().backward()
        out.update({
            f"{name}_shape": np.array([T, H, V]), f"{name}_temperature": np.array(temp),
            f"{name}_hidden": _np(hidden), f"{name}_weight": _np(weight), f"{name}_labels": _np(labels),
            f"{name}_g_lp": _np(g_lp), f"{name}_g_ent": _np(g_ent),
            f"{name}_logprobs": _np(lp), f"{name}_entropy": _np(ent),
            f"{name}_d_hidden": _np(h32.grad), f"{name}_d_weight": _np(w32.grad),
        })
    np.savez_compressed(os.path.join(OUT, "lce.npz"), **out)


def gen_logits(ref):
    g = torch.Generator().manual_seed(7)
    F_ = ref.torch_functional
    out = {}
    logits = torch.randn(5, 7, 333, generator=g) * 3
    labels = torch.randint(0, 333, (5, 7), generator=g)
    out["f32_logits"], out["labels"] = _np(logits), _np(labels)
    out["f32_logprobs_v2"] = _np(F_.logprobs_from_logits_v2(logits, labels))
    out["f32_logprobs_naive"] = _np(F_.logprobs_from_logits_naive(logits, labels))
    out["f32_entropy"] = _np(F_.entropy_from_logits(logits))
    out["f32_entropy_chunked"] = _np(F_.entropy_from_logits_with_chunking(logits.reshape(-1, 333), chunk_size=8))
    lb = logits.bfloat16()
    out["bf16_logits"] = _np(lb)
    out["bf16_logprobs_v2"] = _np(F_.logprobs_from_logits_v2(lb.reshape(-1, 333), labels.reshape(-1)))
    out["bf16_entropy"] = _np(F_.entropy_from_logits(lb))
    np.savez_compressed(os.path.join(OUT, "logits.npz"), **out)


def gen_masked(ref):
    g = torch.Generator().manual_seed(11)
    F_ = ref.torch_functional
    out = {}
    v = torch.randn(6, 40, generator=g)
    lens = torch.randint(3, 41, (6,), generator=g)
    m = (torch.arange(40)[None, :] < lens[:, None]).long()
    out["values"], out["mask"] = _np(v), _np(m)
    out["masked_sum"] = _np(F_.masked_sum(v, m))
    out["masked_mean"] = _np(F_.masked_mean(v, m))
    out["masked_mean_axis"] = _np(F_.masked_mean(v, m, axis=-1))
    out["masked_sum_axis"] = _np(F_.masked_sum(v, m, axis=-1))
    out["masked_var"] = _np(F_.masked_var(v, m))
    out["masked_var_biased"] = _np(F_.masked_var(v, m, unbiased=False))
    out["masked_whiten"] = _np(F_.masked_whiten(v, m))
    out["masked_whiten_noshift"] = _np(F_.masked_whiten(v, m, shift_mean=False))
    # known answers of tests/utils/test_torch_functional.py:55-66
    out["ka_values"] = np.array([1.0, 2.0, 3.0, 4.0], dtype=np.float32)
    out["ka_mask"] = np.array([1, 0, 0, 1], dtype=np.int64)
    out["ka_mean"] = _np(F_.masked_mean(torch.tensor([1.0, 2.0, 3.0, 4.0]), torch.tensor([1, 0, 0, 1])))
    out["ka_nan_outside"] = _np(F_.masked_mean(torch.tensor([1.0, float("nan"), 3.0, 4.0]),
                                               torch.tensor([True, False, False, True])))
    np.savez_compressed(os.path.join(OUT, "masked.npz"), **out)


def _response_mask(B, L, g, min_frac=0.25):
    lens = torch.randint(int(L * min_frac), L + 1, (B,), generator=g)
    return (torch.arange(L)[None, :] < lens[:, None]).long(), lens


def gen_grpo(ref):
    g = torch.Generator().manual_seed(13)
    ca = ref.core_algos
    out = {}
    B, L = 14, 24
    mask, lens = _response_mask(B, L, g)
    rewards = torch.zeros(B, L)
    rewards[torch.arange(B), lens - 1] = torch.rand(B, generator=g)  # one scalar at the last valid token
    uids = np.array([f"uid-{i // 3:02d}-é" for i in range(B)], dtype=object)  # 4 groups of 3 + a group of 2
    uids[-1] = "solo"                                                          # + one singleton
    perm = torch.randperm(B, generator=g).numpy()
    uids = uids[perm]
    # two rows of one group get identical scores (std 0 path needs eps)
    out["rewards"], out["mask"], out["uids"] = _np(rewards), _np(mask), np.array([str(u) for u in uids])
    for norm in (True, False):
        adv, ret = ca.compute_grpo_outcome_advantage(rewards.clone(), mask, uids, norm_adv_by_std_in_grpo=norm)
        out[f"adv_norm{int(norm)}"] = _np(adv)
    dense = torch.randn(B, L, generator=g) * mask
    out["dense_rewards"] = _np(dense)
    out["dense_adv"] = _np(ca.compute_grpo_outcome_advantage(dense.clone(), mask, uids)[0])
    eq = rewards.clone()
    eq[:] = 0
    eq[:, 0] = 0.5  # every score equal -> advantages exactly 0 for groups, (0.5-0)/(1+eps) for the singleton
    out["equal_adv"] = _np(ca.compute_grpo_outcome_advantage(eq.clone(), mask, uids)[0])
    np.savez_compressed(os.path.join(OUT, "grpo.npz"), **out)


def gen_gae(ref):
    g = torch.Generator().manual_seed(17)
    ca = ref.core_algos
    out = {}
    B, L = 5, 67
    # multi-turn style mask: response tokens interleaved with observation (0) spans, trailing padding
    mask = (torch.rand(B, L, generator=g) > 0.3).long()
    lens = torch.randint(20, L + 1, (B,), generator=g)
    mask = mask * (torch.arange(L)[None, :] < lens[:, None]).long()
    rewards = torch.randn(B, L, generator=g) * 0.1
    values = torch.randn(B, L, generator=g)
    out["rewards"], out["values"], out["mask"] = _np(rewards), _np(values), _np(mask)
    for i, (gamma, lam) in enumerate([(1.0, 1.0), (0.97, 0.9)]):
        adv, ret = ca.compute_gae_advantage_return(rewards, values, mask, gamma, lam)
        out[f"gamma{i}"], out[f"lam{i}"] = np.array(gamma), np.array(lam)
        out[f"adv{i}"], out[f"ret{i}"] = _np(adv), _np(ret)
    np.savez_compressed(os.path.join(OUT, "gae.npz"), **out)


def gen_loss(ref):
    g = torch.Generator().manual_seed(19)
    ca = ref.core_algos
    out = {}
    B, L = 8, 64
    mask, _ = _response_mask(B, L, g)
    old = -torch.rand(B, L, generator=g) * 3
    lp = old + 0.3 * torch.randn(B, L, generator=g)  # wide enough to hit clip and dual-clip regions
    lp[0, 0] = old[0, 0] + 25.0                      # beyond the +-20 clamp
    lp[0, 1] = old[0, 1] - 25.0
    ref_lp = old + 0.2 * torch.randn(B, L, generator=g)
    ref_lp[1, 0] = lp[1, 0] + 3.0                    # low_var_kl above its clamp of 10
    adv = torch.randn(B, L, generator=g)
    ent = torch.rand(B, L, generator=g) * 2
    out.update(mask=_np(mask), old=_np(old), lp=_np(lp), ref=_np(ref_lp), adv=_np(adv), ent=_np(ent))
    cfgs = {
        "vanilla": dict(lo=0.2, hi=0.2, c=3.0, ent_coeff=0.0, kl_coef=0.0, kl="low_var_kl", scale=1.0),
        "dapo": dict(lo=0.2, hi=0.28, c=10.0, ent_coeff=1e-3, kl_coef=0.0, kl="low_var_kl", scale=0.25),
        "grpo_kl": dict(lo=0.2, hi=0.2, c=3.0, ent_coeff=0.0, kl_coef=0.01, kl="low_var_kl", scale=0.5),
        "all": dict(lo=0.1, hi=0.3, c=2.0, ent_coeff=0.01, kl_coef=0.05, kl="mse", scale=1.0),
    }
    modes = ["token-mean", "seq-mean-token-sum", "seq-mean-token-mean", "seq-mean-token-sum-norm"]
    for cname, c in cfgs.items():
        for mode in modes:
            lpg = lp.clone().requires_grad_(True)
            eg = ent.clone().requires_grad_(True)
            pg_loss, clipfrac, ppo_kl, clipfrac_lower = ca.compute_policy_loss(
                old_log_prob=old, log_prob=lpg, advantages=adv, response_mask=mask, cliprange=c["lo"],
                cliprange_low=c["lo"], cliprange_high=c["hi"], clip_ratio_c=c["c"], loss_agg_mode=mode)
            # composition of verl/workers/actor/dp_actor.py:445-469
            policy_loss = pg_loss
            ent_loss = torch.zeros(())
            kl_loss = torch.zeros(())
            if c["ent_coeff"] != 0:
                ent_loss = ca.agg_loss(loss_mat=eg, loss_mask=mask, loss_agg_mode=mode)
                policy_loss = pg_loss - ent_loss * c["ent_coeff"]
            if c["kl_coef"] != 0:
                kld = ca.kl_penalty(logprob=lpg, ref_logprob=ref_lp, kl_penalty=c["kl"])
                kl_loss = ca.agg_loss(loss_mat=kld, loss_mask=mask, loss_agg_mode=mode)
                policy_loss = policy_loss + kl_loss * c["kl_coef"]
            loss = policy_loss * c["scale"]
            loss.backward()
            key = f"{cname}|{mode}"
            out[key + "|scalars"] = np.array([_np(loss), _np(pg_loss), _np(clipfrac), _np(ppo_kl), _np(clipfrac_lower),
                                              _np(ent_loss), _np(kl_loss)], dtype=np.float32)
            out[key + "|d_lp"] = _np(lpg.grad)
            out[key + "|d_ent"] = _np(eg.grad) if eg.grad is not None else np.zeros((B, L), np.float32)
        out[cname + "|cfg"] = np.array([c["lo"], c["hi"], c["c"], c["ent_coeff"], c["kl_coef"], c["scale"]])
        out[cname + "|kl"] = np.array(c["kl"])
    for kind in ["kl", "abs", "mse", "low_var_kl"]:
        out["kl_penalty|" + kind] = _np(ca.kl_penalty(lp, ref_lp, kind))
    for mode in modes:
        out["agg|" + mode] = _np(ca.agg_loss(ent, mask, mode))
    np.savez_compressed(os.path.join(OUT, "loss.npz"), **out)


def gen_cov(ref):
    """compute_policy_loss_clip_cov / compute_policy_loss_kl_cov (core_algos.py:818-972): the unmodified reference
    functions; clip_cov's random subset is made reproducible by seeding the host generator right before the call."""
    import types

    g = torch.Generator().manual_seed(41)
    ca = ref.core_algos
    out = {}
    B, L = 12, 40
    mask, _ = _response_mask(B, L, g)
    old = -torch.rand(B, L, generator=g) * 3
    lp = old + 0.3 * torch.randn(B, L, generator=g)
    adv = torch.randn(B, L, generator=g) * 2.5  # wide enough that some cov values land in (clip_cov_lb, clip_cov_ub)
    out.update(mask=_np(mask), old=_np(old), lp=_np(lp), adv=_np(adv))
    modes = ["token-mean", "seq-mean-token-sum", "seq-mean-token-mean", "seq-mean-token-sum-norm"]
    cases = {
        "default": dict(clip_ratio=0.2, clip_ratio_low=None, clip_ratio_high=None, clip_cov_ratio=None, clip_cov_lb=None,
                        clip_cov_ub=None, kl_cov_ratio=None, ppo_kl_coef=None),
        "wide": dict(clip_ratio=0.2, clip_ratio_low=0.2, clip_ratio_high=0.28, clip_cov_ratio=0.05, clip_cov_lb=0.5,
                     clip_cov_ub=6.0, kl_cov_ratio=0.04, ppo_kl_coef=0.7),
    }
    for cname, c in cases.items():
        cfg = types.SimpleNamespace(clip_ratio=c["clip_ratio"], clip_ratio_low=c["clip_ratio_low"],
                                    clip_ratio_high=c["clip_ratio_high"],
                                    policy_loss=types.SimpleNamespace(**{k: c[k] for k in (
                                        "clip_cov_ratio", "clip_cov_lb", "clip_cov_ub", "kl_cov_ratio", "ppo_kl_coef")}))
        for mode in modes:
            lpg = lp.clone().requires_grad_(True)
            torch.manual_seed(1234)
            loss, clipfrac, ppo_kl, _ = ca.compute_policy_loss_clip_cov(old, lpg, adv, mask, mode, cfg)
            loss.backward()
            out[f"clip_cov|{cname}|{mode}|scalars"] = np.array([_np(loss), _np(clipfrac), _np(ppo_kl)], dtype=np.float32)
            out[f"clip_cov|{cname}|{mode}|d_lp"] = _np(lpg.grad)
            lpg = lp.clone().requires_grad_(True)
            loss, _, ppo_kl_abs, _ = ca.compute_policy_loss_kl_cov(old, lpg, adv, mask, mode, cfg)
            loss.backward()
            out[f"kl_cov|{cname}|{mode}|scalars"] = np.array([_np(loss), _np(ppo_kl_abs)], dtype=np.float32)
            out[f"kl_cov|{cname}|{mode}|d_lp"] = _np(lpg.grad)
        out[cname + "|cfg"] = np.array([np.nan if v is None else v for v in c.values()], dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "cov.npz"), **out)


def gen_pfppo(ref):
    """compute_pf_ppo_reweight_data (core_algos.py:1072-1148), the unmodified reference function fed a DataProto
    stand-in (a dict subclass carrying batch_size, numpy object arrays, meta_info); the multinomial draw is made
    reproducible by seeding the host generator right before the call."""

    class TD(dict):
        batch_size = None

    class Data:
        pass

    g = torch.Generator().manual_seed(77)
    B, L = 24, 33
    out = {}
    scores = torch.zeros(B, L)
    lens = torch.randint(1, L + 1, (B,), generator=g)
    scores[torch.arange(B), lens - 1] = torch.randn(B, generator=g).round(decimals=1)  # ties in max / min on purpose
    other = torch.randn(B, L, 3, generator=g)
    ids = torch.randint(0, 1000, (B, L), generator=g)
    uid = np.array([f"u{i // 4}" for i in range(B)], dtype=object)
    out.update(token_level_scores=_np(scores), other=_np(other), ids=_np(ids))
    for method, pw in (("pow", 2.0), ("pow", 0.5), ("max_min", 2.0), ("max_random", 2.0)):
        d = Data()
        d.batch = TD(token_level_scores=scores.clone(), other=other.clone(), ids=ids.clone())
        d.batch.batch_size = [B]
        d.non_tensor_batch = {"uid": uid.copy(), "as_list": list(range(B))}
        d.meta_info = {"per_sample": list(range(100, 100 + B)), "scalar": 3, "short_list": [1, 2]}
        torch.manual_seed(4321)
        r = ref.core_algos.compute_pf_ppo_reweight_data(d, method, pw)
        key = f"{method}|{pw}"
        out[key + "|scores"] = _np(r.batch["token_level_scores"])
        out[key + "|other"] = _np(r.batch["other"])
        out[key + "|ids"] = _np(r.batch["ids"])
        out[key + "|uid"] = np.array([str(u) for u in r.non_tensor_batch["uid"]])
        out[key + "|as_list"] = np.array(r.non_tensor_batch["as_list"])
        out[key + "|per_sample"] = np.array(r.meta_info["per_sample"])
        assert r.meta_info["scalar"] == 3 and r.meta_info["short_list"] == [1, 2]
    np.savez_compressed(os.path.join(OUT, "pfppo.npz"), **out)


def gen_estimators(ref):
    """The other registered advantage estimators + value loss + GPG loss."""
    g = torch.Generator().manual_seed(29)
    ca = ref.core_algos
    out = {}
    B, L = 14, 24
    mask, lens = _response_mask(B, L, g)
    rewards = torch.zeros(B, L)
    rewards[torch.arange(B), lens - 1] = torch.rand(B, generator=g)
    rewards[3] = 0  # a zero score (GPG's alpha counts non-zero scores)
    uids = np.array([f"uid-{i // 3:02d}" for i in range(B)], dtype=object)  # groups of 3 + one of 2
    uids_solo = uids.copy()
    uids_solo[-1] = "solo"
    perm = torch.randperm(B, generator=g).numpy()
    uids, uids_solo = uids[perm], uids_solo[perm]
    dense = torch.randn(B, L, generator=g) * mask
    baselines = torch.rand(B, generator=g)
    out.update(rewards=_np(rewards), mask=_np(mask), uids=np.array([str(u) for u in uids]),
               uids_solo=np.array([str(u) for u in uids_solo]), dense=_np(dense), baselines=_np(baselines))
    cfg = {"norm_adv_by_std_in_grpo": True}
    for tag, rw in (("outcome", rewards), ("dense", dense)):
        for utag, u in (("", uids), ("_solo", uids_solo)):
            out[f"rloo_{tag}{utag}"] = _np(ca.compute_rloo_outcome_advantage(rw.clone(), mask, u)[0])
            out[f"opo_{tag}{utag}"] = _np(ca.compute_opo_outcome_advantage(rw.clone(), mask, u)[0])
            out[f"gpg_{tag}{utag}"] = _np(ca.compute_gpg_outcome_advantage(rw.clone(), mask, u, f_norm=1.5)[0])
            out[f"rfppb_{tag}{utag}"] = _np(
                ca.compute_reinforce_plus_plus_baseline_outcome_advantage(rw.clone(), mask, u)[0])
        out[f"passk_{tag}_norm1"] = _np(ca.compute_grpo_passk_outcome_advantage(rw.clone(), mask, uids, config=cfg)[0])
        out[f"passk_{tag}_norm0"] = _np(ca.compute_grpo_passk_outcome_advantage(
            rw.clone(), mask, uids, config={"norm_adv_by_std_in_grpo": False})[0])
    # reverse scans on a multi-turn style mask (holes inside the response)
    holes = (torch.rand(B, L, generator=g) > 0.25).long() * mask
    out["holes"] = _np(holes)
    for i, gamma in enumerate((1.0, 0.95)):
        cfgobj = type("Cfg", (), {"gamma": gamma})()
        adv, ret = ca.compute_reinforce_plus_plus_outcome_advantage(dense.clone(), holes, config=cfgobj)
        out[f"rfpp_gamma{i}"], out[f"rfpp_adv{i}"], out[f"rfpp_ret{i}"] = np.array(gamma), _np(adv), _np(ret)
    adv, ret = ca.compute_remax_outcome_advantage(dense.clone(), baselines, holes)
    out["remax_adv"], out["remax_ret"] = _np(adv), _np(ret)
    # value loss and GPG policy loss, every aggregation mode, with gradients
    vpreds = torch.randn(B, L, generator=g)
    values = vpreds + 0.3 * torch.randn(B, L, generator=g)
    rets = values + 0.5 * torch.randn(B, L, generator=g)
    lp = -torch.rand(B, L, generator=g) * 3
    adv = torch.randn(B, L, generator=g)
    out.update(vpreds=_np(vpreds), values=_np(values), rets=_np(rets), lp=_np(lp), adv=_np(adv))
    for mode in ["token-mean", "seq-mean-token-sum", "seq-mean-token-mean", "seq-mean-token-sum-norm"]:
        vp = vpreds.clone().requires_grad_(True)
        vf, frac = ca.compute_value_loss(vp, rets, values, mask, cliprange_value=0.2, loss_agg_mode=mode)
        vf.backward()
        out[f"vf|{mode}"] = np.array([_np(vf), _np(frac)], dtype=np.float32)
        out[f"vf_grad|{mode}"] = _np(vp.grad)
        lpg = lp.clone().requires_grad_(True)
        loss = ca.compute_policy_loss_gpg(None, lpg, adv, mask, loss_agg_mode=mode)[0]
        loss.backward()
        out[f"gpg_loss|{mode}"] = _np(loss)
        out[f"gpg_loss_grad|{mode}"] = _np(lpg.grad)
    np.savez_compressed(os.path.join(OUT, "estimators.npz"), **out)


def gen_driver(ref):
    """apply_kl_penalty / compute_data_metrics: the reference's function bodies, compiled out of files whose modules
    cannot be imported here (ray_trainer.py imports ray; metric_utils.py imports verl.DataProto)."""
    import json

    from oracle import ref_import

    g = torch.Generator().manual_seed(31)
    B, P, L = 10, 12, 20
    S = P + L
    p_len = torch.randint(1, P + 1, (B,), generator=g)
    r_len = torch.randint(1, L + 1, (B,), generator=g)
    r_len[0], p_len[1] = L, P  # one response / one prompt at the maximum length (clip_ratio > 0)
    pos = torch.arange(S)[None]
    attn = ((pos >= (P - p_len)[:, None]) & (pos < (P + r_len)[:, None])).long()
    rmask = attn[:, -L:].clone()
    scores = torch.zeros(B, L)
    scores[torch.arange(B), r_len - 1] = torch.rand(B, generator=g)
    old = -torch.rand(B, L, generator=g) * 3
    refp = old + 0.2 * torch.randn(B, L, generator=g)
    adv = torch.randn(B, L, generator=g) * rmask
    values = torch.randn(B, L, generator=g)
    returns = values + 0.5 * torch.randn(B, L, generator=g)

    class Batch(dict):
        batch_size = (B,)

    class Data:
        def __init__(self, **kw):
            self.batch, self.non_tensor_batch, self.meta_info = Batch(**kw), {}, {}

    fns = ref_import.load_functions("verl/trainer/ppo/ray_trainer.py", ["apply_kl_penalty"],
                                    dict(core_algos=ref.core_algos, masked_mean=ref.torch_functional.masked_mean, torch=torch))
    out = dict(attn=_np(attn), rmask=_np(rmask), scores=_np(scores), old=_np(old), ref=_np(refp), adv=_np(adv),
               values=_np(values), returns=_np(returns))
    for kind in ("kl", "abs", "mse", "low_var_kl"):
        ctrl = ref.core_algos.AdaptiveKLController(init_kl_coef=0.05, target_kl=0.1, horizon=1000)
        data = Data(response_mask=rmask, token_level_scores=scores, old_log_probs=old, ref_log_prob=refp)
        data, metrics = fns["apply_kl_penalty"](data, ctrl, kl_penalty=kind)
        out[f"klr|{kind}|rewards"] = _np(data.batch["token_level_rewards"])
        out[f"klr|{kind}|scalars"] = np.array([metrics["actor/reward_kl_penalty"], metrics["actor/reward_kl_penalty_coeff"],
                                               ctrl.value], dtype=np.float64)
    rewards = T_(out["klr|kl|rewards"])
    mfn = ref_import.load_functions("verl/trainer/ppo/metric_utils.py", ["_compute_response_info", "compute_data_metrics"],
                                    dict(torch=torch, Any=object, DataProto=object))
    data = Data(token_level_scores=scores, token_level_rewards=rewards, advantages=adv, returns=returns, values=values,
                responses=torch.zeros(B, L, dtype=torch.long), attention_mask=attn, response_mask=rmask)
    for critic in (True, False):
        m = mfn["compute_data_metrics"](data, use_critic=critic)
        out[f"dm|critic{int(critic)}"] = np.array(json.dumps({k: float(v) for k, v in m.items()}))
    np.savez_compressed(os.path.join(OUT, "driver.npz"), **out)


def gen_seqlen(ref):
    """karmarkar_karp / get_seqlen_balanced_partitions / rearrange_micro_batches (scheduling half) / get_reverse_idx:
    the reference's own function bodies compiled out of verl/utils/seqlen_balancing.py (the module imports
    verl.protocol -> ray / tensordict).  rearrange_micro_batches slices a TensorDict row by row; the stand-in batch
    below answers `batch[i:i+1]` with the row NUMBER, so the returned "micro-batches" are the scheduled row order."""
    import copy
    import heapq
    import json
    import random
    from itertools import chain

    from oracle import ref_import

    class _Dist:
        ReduceOp = None

        @staticmethod
        def is_initialized():
            return False

    fns = ref_import.load_functions(
        "verl/utils/seqlen_balancing.py",
        ["karmarkar_karp", "get_seqlen_balanced_partitions", "ceildiv", "roundup_divisible", "rearrange_micro_batches",
         "get_reverse_idx"],
        dict(heapq=heapq, copy=copy, chain=chain, torch=torch, dist=_Dist, get_device_name=lambda: "cpu"))

    class RowNumbers(dict):
        def __getitem__(self, key):
            if isinstance(key, slice):
                return torch.arange(key.start, key.stop)
            return dict.__getitem__(self, key)

    rnd = random.Random(20240607)
    kk = []
    for trial in range(60):
        n = rnd.randint(1, 48)
        k = rnd.randint(1, min(n, 8))
        hi = rnd.choice([2, 9, 100, 4096])  # small ranges force ties in sums, counts and item lists
        lens = [rnd.randint(0, hi) for _ in range(n)]
        for eq in (False, True):
            if eq and n % k:
                continue
            try:
                merged = fns["karmarkar_karp"](lens, k, eq)
                parts = fns["get_seqlen_balanced_partitions"](lens, k, eq)
            except AssertionError as e:  # an empty partition (many zero lengths)
                merged, parts = None, "AssertionError"
            kk.append(dict(lens=lens, k=k, equal_size=eq, merge_order=merged, partitions=parts))
    mb = []
    g = torch.Generator().manual_seed(11)
    for B, S, max_tok, min_mb, div, bal in [(20, 100, 300, None, None, True), (20, 100, 300, 4, None, True),
                                            (33, 64, 128, None, 2, True), (16, 512, 2048, None, None, False),
                                            (7, 30, 1000, None, None, True), (64, 256, 1024, 3, 4, True)]:
        lens = torch.randint(S // 4, S + 1, (B,), generator=g)
        left = torch.randint(0, S // 10 + 1, (B,), generator=g)
        pos = torch.arange(S)[None]
        attn = ((pos >= left[:, None]) & (pos < (left + lens)[:, None])).long()
        rows, idx = fns["rearrange_micro_batches"](RowNumbers(attention_mask=attn), max_tok, num_batches_divided_by=div,
                                                   min_num_micro_batch=min_mb, use_dynamic_bsz_balance=bal)
        assert [r.tolist() for r in rows] == idx
        flat = list(chain.from_iterable(idx))
        mb.append(dict(B=B, S=S, lens=lens.tolist(), left=left.tolist(), max_token_len=max_tok, min_num_micro_batch=min_mb,
                       num_batches_divided_by=div, use_dynamic_bsz_balance=bal, micro_bsz_idx=idx,
                       reverse_idx=fns["get_reverse_idx"](flat)))
    np.savez_compressed(os.path.join(OUT, "seqlen.npz"), karmarkar_karp=np.array(json.dumps(kk)),
                        micro_batches=np.array(json.dumps(mb)))


class _TorchProxy:
    """the real torch module with a few attributes replaced (used to stub collectives / torch.compile when a reference
    function body is compiled stand-alone)"""

    def __init__(self, **over):
        self._over = over

    def __getattr__(self, name):
        return self._over[name] if name in self._over else getattr(torch, name)


def gen_megatron(ref):
    """Megatron-side equivalents.
    * _VocabParallelEntropy (verl/utils/megatron/tensor_parallel.py:109-139): the reference class body, compiled
      stand-alone with a single-rank `dist` (all_reduce = no-op) and torch.compile -> identity, run on the FULL-vocab
      logits: forward + backward.  The vocab-sharded result must equal it (the merge is exact algebra).
    * vocab_parallel_log_probs_from_logits calls megatron.core.tensor_parallel.vocab_parallel_cross_entropy
      (third-party megatron-core, not in this image; reference pins megatron-core 0.12/0.13 in its Dockerfiles):
      anchored on the reference's own full-vocab logprobs_from_logits_v2 (same quantity by definition).
    * preprocess_packed_seqs / postprocess_packed_seqs (verl/models/mcore/util.py:23-142): the reference function
      bodies with a stub `mpu` (tp/cp sizes and rank) and a fake all_gather fed with the other ranks' outputs."""
    import types

    from oracle import ref_import

    out = {}
    g = torch.Generator().manual_seed(77)
    one_rank = types.SimpleNamespace(all_reduce=lambda *a, **k: None, ReduceOp=types.SimpleNamespace(MAX="max"))
    mpu1 = types.SimpleNamespace(get_tensor_model_parallel_group=lambda: None)
    cls = ref_import.load_functions("verl/utils/megatron/tensor_parallel.py", ["_VocabParallelEntropy"],
                                    dict(torch=_TorchProxy(compile=lambda **kw: (lambda f: f)), dist=one_rank, mpu=mpu1))
    T, V = 23, 1000
    logits = (torch.randn(T, V, generator=g) * 3)
    labels = torch.randint(0, V, (T,), generator=g)
    g_ent = torch.randn(T, generator=g)
    g_lp = torch.randn(T, generator=g)
    x = logits.clone().requires_grad_(True)
    ent = cls["_VocabParallelEntropy"].apply(x.clone())  # the class works in place on its input
    ent.backward(g_ent)
    out.update(vp_logits=_np(logits), vp_labels=_np(labels), vp_g_ent=_np(g_ent), vp_g_lp=_np(g_lp),
               vp_entropy=_np(ent), vp_entropy_dlogits=_np(x.grad))
    x2 = logits.clone().requires_grad_(True)
    lp = ref.torch_functional.logprobs_from_logits_v2(x2, labels)
    lp.backward(g_lp)
    out.update(vp_logprobs=_np(lp), vp_logprobs_dlogits=_np(x2.grad))

    # ---- packed sequences
    B, S, D = 6, 40, 4
    lens = torch.tensor([40, 8, 17, 9, 33, 26])
    left = torch.tensor([0, 5, 3, 0, 7, 2])
    pos = torch.arange(S)[None]
    attn = (pos >= left[:, None]) & (pos < (left + lens)[:, None])
    attn[2, 10] = False  # a hole inside a sequence: packing keeps the valid tokens in order
    ids = torch.randint(1, 500, (B, S), generator=g)
    table = torch.randn(500, D, generator=g)
    out.update(pk_ids=_np(ids), pk_attn=_np(attn.long()), pk_table=_np(table))

    class PackedSeqParams:
        def __init__(self, **kw):
            self.__dict__.update(kw)

    for tp, cp in [(1, 1), (2, 1), (4, 1), (1, 2), (2, 2), (1, 4)]:
        per_rank_out = []
        for rank in range(cp):
            mpu = types.SimpleNamespace(get_tensor_model_parallel_world_size=lambda tp=tp: tp,
                                        get_context_parallel_world_size=lambda cp=cp: cp,
                                        get_context_parallel_rank=lambda rank=rank: rank,
                                        get_context_parallel_group=lambda: None)
            fns = ref_import.load_functions("verl/models/mcore/util.py", ["preprocess_packed_seqs"],
                                            dict(torch=torch, mpu=mpu, PackedSeqParams=PackedSeqParams))
            packed, params = fns["preprocess_packed_seqs"](ids, attn, pre_process=True)
            out[f"pk|tp{tp}cp{cp}|r{rank}|packed"] = _np(packed)
            per_rank_out.append(table[packed[0]].unsqueeze(0))  # stand-in for the model: a row lookup
        out[f"pk|tp{tp}cp{cp}|cu_padded"] = _np(params.cu_seqlens_q_padded)
        out[f"pk|tp{tp}cp{cp}|max_seqlen"] = np.array(params.max_seqlen_q)
        for rank in range(cp):
            def fake_all_gather(lst, t, group=None):
                for j in range(cp):
                    lst[j] = per_rank_out[j].clone()

            mpu = types.SimpleNamespace(get_context_parallel_world_size=lambda cp=cp: cp,
                                        get_context_parallel_rank=lambda rank=rank: rank,
                                        get_context_parallel_group=lambda: None)
            tproxy = _TorchProxy(distributed=types.SimpleNamespace(all_gather=fake_all_gather))
            fns = ref_import.load_functions("verl/models/mcore/util.py", ["postprocess_packed_seqs"],
                                            dict(torch=tproxy, mpu=mpu, PackedSeqParams=PackedSeqParams))
            o = per_rank_out[rank].clone().requires_grad_(True)
            padded = fns["postprocess_packed_seqs"](o, params, attn, B, S, post_process=True)
            w = torch.arange(B * S * D, dtype=torch.float32).view(B, S, D) / 100
            (padded * w).sum().backward()
            out[f"pk|tp{tp}cp{cp}|r{rank}|padded"] = _np(padded)
            out[f"pk|tp{tp}cp{cp}|r{rank}|d_output"] = _np(o.grad)
    np.savez_compressed(os.path.join(OUT, "megatron.npz"), **out)


def gen_rmpad(ref):
    import flash_attn.bert_padding as bp  # third-party pure-torch file, v2.8.3 in this image

    out = {}
    torch.manual_seed(23)
    np.random.seed(23)  # create_random_mask (verl/utils/model.py:176-216) draws from numpy's global RNG
    B, S = 8, 64
    if ref.model is not None:
        mask = ref.model.create_random_mask(torch.zeros(B, S, dtype=torch.long), max_ratio_of_valid_token=0.8,
                                            max_ratio_of_left_padding=0.1, min_ratio_of_valid_token=0.5)
    else:
        mask = (torch.rand(B, S) > 0.3).long()
    ids = torch.randint(0, 50000, (B, S))
    x_rmpad, indices, cu, max_s, seqused = bp.unpad_input(ids.unsqueeze(-1), mask)
    idx2, cu2, max2 = ref.torch_functional.get_unpad_data(mask)
    assert torch.equal(indices, idx2) and torch.equal(cu, cu2) and max_s == max2
    vals = torch.randn(indices.numel(), 1)
    padded = bp.pad_input(vals, indices, B, S)
    hs = torch.randn(B, S, 6)
    hs_rmpad = bp.index_first_axis(hs.reshape(B * S, 6), indices)
    out.update(mask=_np(mask), ids=_np(ids), ids_rmpad=_np(x_rmpad.squeeze(-1)), indices=_np(indices),
               cu_seqlens=_np(cu), max_seqlen=np.array(max_s), seqused=_np(seqused), vals=_np(vals),
               padded=_np(padded), hs=_np(hs), hs_rmpad=_np(hs_rmpad))
    # edge cases: an empty row, a full row, all-empty
    edge = torch.zeros(4, 9, dtype=torch.long)
    edge[1] = 1
    edge[2, 3:7] = 1
    _, e_idx, e_cu, e_max, e_used = bp.unpad_input(torch.zeros(4, 9, 1), edge)
    out.update(edge_mask=_np(edge), edge_indices=_np(e_idx), edge_cu=_np(e_cu), edge_max=np.array(e_max))
    np.savez_compressed(os.path.join(OUT, "rmpad.npz"), **out)


def main():
    from oracle import ref_import

    if not ref_import.available():
        sys.exit("reference tree not available; golden vectors can only be regenerated in the build container")
    ref = ref_import.load()
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(min(8, os.cpu_count() or 1))
    only = set(sys.argv[1:])  # e.g. `python -m oracle.make_golden estimators` regenerates one fixture
    for fn in (gen_lce, gen_logits, gen_masked, gen_grpo, gen_gae, gen_loss, gen_cov, gen_pfppo, gen_estimators, gen_driver, gen_seqlen, gen_megatron, gen_rmpad):
        if only and fn.__name__[4:] not in only:
            continue
        fn(ref)
        print("wrote", fn.__name__)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
