"""GPU: compute_policy_loss_clip_cov / compute_policy_loss_kl_cov through the reference-named registry against the
reference goldens (loss, metrics, gradient w.r.t. log_prob) and against the oracle on larger seeded inputs."""

import types

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port
from verl_b200.trainer.ppo import core_algos as ca

pytestmark = pytest.mark.gpu

MODES = ["token-mean", "seq-mean-token-sum", "seq-mean-token-mean", "seq-mean-token-sum-norm"]
KEYS = ["clip_ratio", "clip_ratio_low", "clip_ratio_high", "clip_cov_ratio", "clip_cov_lb", "clip_cov_ub", "kl_cov_ratio",
        "ppo_kl_coef"]


def T(a):
    return torch.from_numpy(np.asarray(a))


def make_cfg(c):
    return types.SimpleNamespace(clip_ratio=c["clip_ratio"], clip_ratio_low=c["clip_ratio_low"],
                                 clip_ratio_high=c["clip_ratio_high"],
                                 policy_loss=types.SimpleNamespace(**{k: c[k] for k in KEYS[3:]}))


def close(a, b, rel=1e-5, abs_=1e-7):
    a, b = torch.as_tensor(a).detach().float().cpu(), torch.as_tensor(b).detach().float().cpu()
    torch.testing.assert_close(a, b, rtol=rel, atol=abs_)


@pytest.mark.parametrize("cname", ["default", "wide"])
@pytest.mark.parametrize("mode", MODES)
def test_cov_losses_golden(cname, mode, dev):
    g = load_golden("cov")
    c = {k: (None if np.isnan(v) else float(v)) for k, v in zip(KEYS, g[cname + "|cfg"])}
    cfg = make_cfg(c)
    old, adv, mask = T(g["old"]).to(dev), T(g["adv"]).to(dev), T(g["mask"]).to(dev)
    lp = T(g["lp"]).to(dev).requires_grad_(True)
    torch.manual_seed(1234)
    loss, clipfrac, ppo_kl, zero = ca.get_policy_loss_fn("clip_cov")(old, lp, adv, mask, mode, cfg)
    loss.backward()
    want = g[f"clip_cov|{cname}|{mode}|scalars"]
    close(torch.stack([loss.detach(), clipfrac, ppo_kl]), T(want), rel=1e-5)
    close(lp.grad, T(g[f"clip_cov|{cname}|{mode}|d_lp"]), rel=1e-5, abs_=1e-9)
    assert float(zero) == 0.0
    lp = T(g["lp"]).to(dev).requires_grad_(True)
    loss, z1, ppo_kl_abs, z2 = ca.get_policy_loss_fn("kl_cov")(old, lp, adv, mask, mode, cfg)
    loss.backward()
    want = g[f"kl_cov|{cname}|{mode}|scalars"]
    close(torch.stack([loss.detach(), ppo_kl_abs]), T(want), rel=1e-5)
    close(lp.grad, T(g[f"kl_cov|{cname}|{mode}|d_lp"]), rel=1e-5, abs_=1e-9)
    assert float(z1) == 0.0 and float(z2) == 0.0


@pytest.mark.parametrize("mask_dtype", [torch.int64, torch.bool, torch.float32])
@pytest.mark.parametrize("B,L", [(3, 7), (16, 1024), (5, 4099)])
def test_cov_losses_vs_oracle(B, L, mask_dtype, dev):
    torch.manual_seed(B * 1000 + L)
    lens = torch.randint(1, L + 1, (B,))
    mask = (torch.arange(L)[None] < lens[:, None])
    old = -torch.rand(B, L) * 3
    lp0 = old + 0.3 * torch.randn(B, L)
    adv = torch.randn(B, L) * 2.5
    c = dict(clip_ratio=0.2, clip_ratio_low=0.2, clip_ratio_high=0.28, clip_cov_ratio=0.01, clip_cov_lb=1.0,
             clip_cov_ub=5.0, kl_cov_ratio=0.002, ppo_kl_coef=1.0)
    cfg = make_cfg(c)
    m_cpu = mask.long()
    m_dev = mask.to(mask_dtype).to(dev)
    for mode in ("token-mean", "seq-mean-token-mean"):
        lp = lp0.clone().to(dev).requires_grad_(True)
        lpo = lp0.clone().requires_grad_(True)
        torch.manual_seed(77)
        loss, clipfrac, ppo_kl, _ = ca.compute_policy_loss_clip_cov(old.to(dev), lp, adv.to(dev), m_dev, mode, cfg)
        loss.backward()
        torch.manual_seed(77)
        lo, cf, kl = port.policy_loss_clip_cov(old, lpo, adv, m_cpu, mode, clip_ratio=0.2, clip_ratio_low=0.2,
                                               clip_ratio_high=0.28, clip_cov_ratio=0.01)
        lo.backward()
        close(loss, lo), close(clipfrac, cf), close(ppo_kl, kl, abs_=1e-6)
        close(lp.grad, lpo.grad, rel=1e-5, abs_=1e-9)
        lp = lp0.clone().to(dev).requires_grad_(True)
        lpo = lp0.clone().requires_grad_(True)
        loss, _, kl_abs, _ = ca.compute_policy_loss_kl_cov(old.to(dev), lp, adv.to(dev), m_dev, mode, cfg)
        loss.backward()
        lo, ka = port.policy_loss_kl_cov(old, lpo, adv, m_cpu, mode, kl_cov_ratio=0.002, ppo_kl_coef=1.0)
        lo.backward()
        close(loss, lo), close(kl_abs, ka, abs_=1e-6)
        close(lp.grad, lpo.grad, rel=1e-5, abs_=1e-9)


def test_kl_cov_selection_count_and_ties(dev):
    """k = max(1, int(N_valid * ratio)) tokens are selected exactly, also when every covariance ties (constant
    advantages): ties go to the lowest flat indices."""
    from verl_b200 import ops

    B, L = 8, 512
    mask = torch.ones(B, L, dtype=torch.long, device=dev)
    mask[:, 400:] = 0
    old = torch.zeros(B, L, device=dev)
    lp = torch.full((B, L), -0.5, device=dev)
    adv = torch.ones(B, L, device=dev)  # cov == 0 everywhere
    pg, dpg, out3 = ops.kl_cov_loss(old, lp, adv, mask, 0.01, 1.0)
    k = max(1, int(8 * 400 * 0.01))
    assert int(out3[1]) == k
    base = -torch.exp(torch.tensor(-0.5))
    picked = (pg.view(B, L) != base.to(dev)).nonzero()
    assert picked.shape[0] == k and picked[:, 0].eq(0).all() and picked[:, 1].tolist() == list(range(k))
    # a full-size case (C5: 16 x 16384): count only
    torch.manual_seed(0)
    B, L = 16, 16384
    mask = (torch.rand(B, L, device=dev) > 0.3).long()
    old = -torch.rand(B, L, device=dev)
    lp = old + 0.1 * torch.randn(B, L, device=dev)
    adv = torch.randn(B, L, device=dev)
    _, _, out3 = ops.kl_cov_loss(old, lp, adv, mask, 0.0002, 1.0)
    assert int(out3[1]) == max(1, int(int(mask.sum()) * 0.0002))
    # no valid token: nothing selected, loss matrix is the plain one
    _, _, out3 = ops.kl_cov_loss(old, lp, adv, torch.zeros_like(mask), 0.0002, 1.0)
    assert int(out3[1]) == 0


@pytest.mark.parametrize("method,pw", [("pow", 2.0), ("pow", 0.5), ("max_min", 2.0), ("max_random", 2.0)])
def test_pf_ppo_reweight_data_golden(method, pw):
    """compute_pf_ppo_reweight_data on a DataProto stand-in with CUDA tensors: same resampled rows (every tensor, the
    numpy / list side data, the per-sample meta lists) as the unmodified reference with the same host seed."""

    class TD(dict):
        batch_size = None

    class Data:
        pass

    g = load_golden("pfppo")
    dev = torch.device("cuda:0")
    B = g["token_level_scores"].shape[0]
    d = Data()
    d.batch = TD(token_level_scores=T(g["token_level_scores"]).to(dev), other=T(g["other"]).to(dev),
                 ids=T(g["ids"]).to(dev))
    d.batch.batch_size = [B]
    d.non_tensor_batch = {"uid": np.array([f"u{i // 4}" for i in range(B)], dtype=object), "as_list": list(range(B))}
    d.meta_info = {"per_sample": list(range(100, 100 + B)), "scalar": 3, "short_list": [1, 2]}
    torch.manual_seed(4321)
    r = ca.compute_pf_ppo_reweight_data(d, method, pw)
    key = f"{method}|{pw}"
    assert isinstance(r.batch, TD) and r.batch.batch_size == [B] and r is not d
    assert torch.equal(r.batch["token_level_scores"].cpu(), T(g[key + "|scores"]))
    assert torch.equal(r.batch["other"].cpu(), T(g[key + "|other"]))
    assert torch.equal(r.batch["ids"].cpu(), T(g[key + "|ids"]))
    assert [str(u) for u in r.non_tensor_batch["uid"]] == list(g[key + "|uid"])
    assert list(r.non_tensor_batch["as_list"]) == list(g[key + "|as_list"])
    assert r.meta_info["per_sample"] == list(g[key + "|per_sample"])
    assert r.meta_info["scalar"] == 3 and r.meta_info["short_list"] == [1, 2]
    assert d.batch["ids"].shape == r.batch["ids"].shape  # the input is left untouched
    with pytest.raises(ValueError):
        ca.compute_pf_ppo_reweight_data(d, "nope", 1.0)
