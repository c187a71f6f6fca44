"""CPU: oracle restatements of the covariance-gated policy losses (core_algos.py:818-972) against golden vectors of
the unmodified reference functions (oracle/make_golden.py::gen_cov)."""

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port

MODES = ["token-mean", "seq-mean-token-sum", "seq-mean-token-mean", "seq-mean-token-sum-norm"]
KEYS = ["clip_ratio", "clip_ratio_low", "clip_ratio_high", "clip_cov_ratio", "clip_cov_lb", "clip_cov_ub", "kl_cov_ratio",
        "ppo_kl_coef"]


def T(a):
    return torch.from_numpy(np.asarray(a))


def cfg_of(g, cname):
    return {k: (None if np.isnan(v) else float(v)) for k, v in zip(KEYS, g[cname + "|cfg"])}


@pytest.mark.parametrize("cname", ["default", "wide"])
@pytest.mark.parametrize("mode", MODES)
def test_cov_losses_golden(cname, mode):
    g = load_golden("cov")
    c = cfg_of(g, cname)
    old, adv, mask = T(g["old"]), T(g["adv"]), T(g["mask"])
    lp = T(g["lp"]).clone().requires_grad_(True)
    kw = {k: c[k] for k in ("clip_cov_ratio", "clip_cov_lb", "clip_cov_ub") if c[k] is not None}
    torch.manual_seed(1234)
    loss, clipfrac, ppo_kl = port.policy_loss_clip_cov(old, lp, adv, mask, mode, clip_ratio=c["clip_ratio"],
                                                       clip_ratio_low=c["clip_ratio_low"],
                                                       clip_ratio_high=c["clip_ratio_high"], **kw)
    loss.backward()
    want = g[f"clip_cov|{cname}|{mode}|scalars"]
    np.testing.assert_allclose([loss.item(), clipfrac.item(), ppo_kl.item()], want, rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(lp.grad, T(g[f"clip_cov|{cname}|{mode}|d_lp"]), rtol=1e-6, atol=1e-9)
    lp = T(g["lp"]).clone().requires_grad_(True)
    kw = {k: c[k] for k in ("kl_cov_ratio", "ppo_kl_coef") if c[k] is not None}
    loss, ppo_kl_abs = port.policy_loss_kl_cov(old, lp, adv, mask, mode, **kw)
    loss.backward()
    want = g[f"kl_cov|{cname}|{mode}|scalars"]
    np.testing.assert_allclose([loss.item(), ppo_kl_abs.item()], want, rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(lp.grad, T(g[f"kl_cov|{cname}|{mode}|d_lp"]), rtol=1e-6, atol=1e-9)


@pytest.mark.parametrize("method,pw", [("pow", 2.0), ("pow", 0.5), ("max_min", 2.0), ("max_random", 2.0)])
def test_pf_ppo_draw_golden(method, pw):
    """compute_pf_ppo_reweight_data (core_algos.py:1072-1148): with the host generator seeded as the golden run was,
    the oracle draws the same rows as the unmodified reference function."""
    g = load_golden("pfppo")
    scores = T(g["token_level_scores"])
    torch.manual_seed(4321)
    idx = port.pf_ppo_sample_indices(scores, method, pw)
    key = f"{method}|{pw}"
    assert torch.equal(scores[idx], T(g[key + "|scores"]))
    assert torch.equal(T(g["other"])[idx], T(g[key + "|other"]))
    assert np.array_equal(np.arange(100, 100 + scores.shape[0])[idx.numpy()], g[key + "|per_sample"])
