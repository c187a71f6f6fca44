"""GPU parity of the §8(f).2 rows — RLOO / OPO / GPG / REINFORCE++(-baseline) / ReMax / GRPO-pass@k advantages,
compute_value_loss, compute_policy_loss_gpg — through the reference-named Python API -> C ABI, against golden vectors
from the unmodified reference and against the CPU oracle at BASELINE config shapes.  Tolerance 1e-5 relative."""

import uuid

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port

pytestmark = pytest.mark.gpu

MODES = ["token-mean", "seq-mean-token-sum", "seq-mean-token-mean", "seq-mean-token-sum-norm"]


def T(a, dev):
    return torch.from_numpy(np.asarray(a)).to(dev)


def uid(a):
    return np.array([str(u) for u in a], dtype=object)


def close(a, b, rel=1e-5, abs_=1e-6):
    torch.testing.assert_close(a.detach().float().cpu(), b.detach().float().cpu(), rtol=rel, atol=abs_)


@pytest.mark.parametrize("tag", ["outcome", "dense"])
@pytest.mark.parametrize("utag", ["", "_solo"])
def test_group_estimators_golden(tag, utag, dev):
    from verl_b200.trainer.ppo import core_algos as ca

    g = load_golden("estimators")
    mask, u = T(g["mask"], dev), uid(g["uids" + utag])
    rw = T(g["rewards"] if tag == "outcome" else g["dense"], dev)
    adv, ret = ca.compute_rloo_outcome_advantage(rw, mask, u)
    assert adv is ret
    close(adv, T(g[f"rloo_{tag}{utag}"], dev))
    close(ca.compute_opo_outcome_advantage(rw, mask, u)[0], T(g[f"opo_{tag}{utag}"], dev))
    close(ca.compute_gpg_outcome_advantage(rw, mask, u, f_norm=1.5)[0], T(g[f"gpg_{tag}{utag}"], dev))
    close(ca.compute_reinforce_plus_plus_baseline_outcome_advantage(rw, mask, u)[0], T(g[f"rfppb_{tag}{utag}"], dev))
    # registry dispatch, as ray_trainer.compute_advantage does (ray_trainer.py:272-289)
    fn = ca.get_adv_estimator_fn(ca.AdvantageEstimator.RLOO)
    close(fn(token_level_rewards=rw, response_mask=mask, index=u, config=None)[0], T(g[f"rloo_{tag}{utag}"], dev))


@pytest.mark.parametrize("tag", ["outcome", "dense"])
def test_passk_golden(tag, dev):
    from verl_b200.trainer.ppo import core_algos as ca

    g = load_golden("estimators")
    mask, u = T(g["mask"], dev), uid(g["uids"])
    rw = T(g["rewards"] if tag == "outcome" else g["dense"], dev)
    close(ca.compute_grpo_passk_outcome_advantage(rw, mask, u, config={})[0], T(g[f"passk_{tag}_norm1"], dev))
    close(ca.compute_grpo_passk_outcome_advantage(rw, mask, u, config={"norm_adv_by_std_in_grpo": False})[0],
          T(g[f"passk_{tag}_norm0"], dev))
    with pytest.raises(ValueError, match="at least 2 samples"):
        ca.compute_grpo_passk_outcome_advantage(rw, mask, uid(g["uids_solo"]), config={})
    with pytest.raises(AssertionError):
        ca.compute_grpo_passk_outcome_advantage(rw, mask, u)  # config is required (core_algos.py:336)


def test_reverse_scans_golden(dev):
    from verl_b200.trainer.ppo import core_algos as ca

    g = load_golden("estimators")
    dense, holes = T(g["dense"], dev), T(g["holes"], dev)
    for i in (0, 1):
        cfg = type("Cfg", (), {"gamma": float(g[f"rfpp_gamma{i}"])})()
        adv, ret = ca.compute_reinforce_plus_plus_outcome_advantage(dense, holes, config=cfg)
        close(ret, T(g[f"rfpp_ret{i}"], dev))
        close(adv, T(g[f"rfpp_adv{i}"], dev))
    adv, ret = ca.compute_remax_outcome_advantage(dense, T(g["baselines"], dev), holes)
    close(ret, T(g["remax_ret"], dev))
    close(adv, T(g["remax_adv"], dev))


@pytest.mark.parametrize("B,L,n", [(40, 512, 5), (128, 1024, 8), (16, 16384, 4), (7, 33, 1), (9, 37, 3)])
def test_group_estimators_vs_oracle_config_shapes(B, L, n, dev):
    """BASELINE configs: C1 40x512 (8 prompts x 5), C3 128x1024 (n=8), C5 16x16384; all-singletons; unaligned rows."""
    from verl_b200.trainer.ppo import core_algos as ca

    torch.manual_seed(B * 7 + L)
    lens = torch.randint(max(L // 4, 1), L + 1, (B,))
    mask = (torch.arange(L)[None] < lens[:, None]).long()
    rewards = torch.zeros(B, L)
    rewards[torch.arange(B), lens - 1] = torch.rand(B)
    uids = np.repeat(np.array([str(uuid.uuid4()) for _ in range((B + n - 1) // n)], dtype=object), n)[:B]
    uids = uids[torch.randperm(B).numpy()]
    rd, md = rewards.to(dev), mask.to(dev)
    close(ca.compute_rloo_outcome_advantage(rd, md, uids)[0], port.rloo_outcome_advantage(rewards, mask, uids))
    close(ca.compute_opo_outcome_advantage(rd, md, uids)[0], port.opo_outcome_advantage(rewards, mask, uids))
    close(ca.compute_gpg_outcome_advantage(rd, md, uids)[0], port.gpg_outcome_advantage(rewards, mask, uids))
    close(ca.compute_reinforce_plus_plus_baseline_outcome_advantage(rd, md, uids)[0],
          port.reinforce_plus_plus_baseline_outcome_advantage(rewards, mask, uids), abs_=5e-6)
    if n >= 2 and B % n == 0:
        close(ca.compute_grpo_passk_outcome_advantage(rd, md, uids, config={})[0],
              port.grpo_passk_outcome_advantage(rewards, mask, uids), abs_=5e-6)


@pytest.mark.parametrize("B,L", [(16, 16384), (3, 9000), (4, 1), (2, 8193), (5, 37)])
@pytest.mark.parametrize("mask_dtype", [torch.int64, torch.bool])
def test_reverse_scans_vs_oracle_long_rows(B, L, mask_dtype, dev):
    from verl_b200.trainer.ppo import core_algos as ca

    torch.manual_seed(L)
    mask = (torch.rand(B, L) > 0.2).long()
    mask[:, 0] = 1
    if L > 1:
        mask[0, L // 2:] = 0
    r = torch.randn(B, L) * 0.1
    bsl = torch.rand(B)
    md = mask.to(dev).to(mask_dtype)
    adv, ret = ca.compute_remax_outcome_advantage(r.to(dev), bsl.to(dev), md)
    want = port.remax_outcome_advantage(r, bsl, mask)
    # fp32 suffix sums of up to 16k terms: the scan associates differently from the CPU cumsum, so the bar is 1e-5
    # relative to the magnitude of the running sums (|ret| reaches ~10 here)
    tol = 1e-5 * max(float(want[1].abs().max()), 1.0)
    close(ret, want[1], abs_=tol)
    close(adv, want[0], abs_=tol)
    if int(mask.sum()) > 1:
        cfg = type("Cfg", (), {"gamma": 0.99})()
        adv, ret = ca.compute_reinforce_plus_plus_outcome_advantage(r.to(dev), md, config=cfg)
        want = port.reinforce_plus_plus_outcome_advantage(r, mask, 0.99)
        tol = 1e-5 * max(float(want[1].abs().max()), 1.0)
        close(ret, want[1], abs_=tol)
        close(adv, want[0], abs_=2e-5)


def test_reverse_scan_errors(dev):
    from verl_b200.trainer.ppo import core_algos as ca

    cfg = type("Cfg", (), {"gamma": 1.0})()
    r = torch.randn(2, 8, device=dev)
    with pytest.raises(ValueError, match="At least one element"):
        ca.compute_reinforce_plus_plus_outcome_advantage(r, torch.zeros(2, 8, dtype=torch.long, device=dev), config=cfg)
    one = torch.zeros(2, 8, dtype=torch.long, device=dev)
    one[1, 3] = 1
    with pytest.raises(ValueError, match="sum of the mask is one"):
        ca.compute_reinforce_plus_plus_outcome_advantage(r, one, config=cfg)


@pytest.mark.parametrize("mode", MODES)
def test_value_and_gpg_loss_golden(mode, dev):
    from verl_b200.trainer.ppo import core_algos as ca

    g = load_golden("estimators")
    mask = T(g["mask"], dev)
    vp = T(g["vpreds"], dev).clone().requires_grad_(True)
    vf, frac = ca.compute_value_loss(vp, T(g["rets"], dev), T(g["values"], dev), mask, cliprange_value=0.2,
                                     loss_agg_mode=mode)
    vf.backward()
    close(torch.stack([vf.detach(), frac]), T(g[f"vf|{mode}"], dev))
    close(vp.grad, T(g[f"vf_grad|{mode}"], dev), abs_=1e-8)
    lp = T(g["lp"], dev).clone().requires_grad_(True)
    loss, z1, z2, z3 = ca.get_policy_loss_fn("gpg")(None, lp, T(g["adv"], dev), mask, loss_agg_mode=mode)
    (loss * 3.0).backward()
    assert float(z1) == float(z2) == float(z3) == 0.0
    close(loss, T(g[f"gpg_loss|{mode}"], dev))
    close(lp.grad / 3.0, T(g[f"gpg_loss_grad|{mode}"], dev), abs_=1e-8)


@pytest.mark.parametrize("B,L", [(128, 1024), (3, 37)])
def test_value_loss_vs_oracle(B, L, dev):
    from verl_b200.trainer.ppo import core_algos as ca

    torch.manual_seed(L)
    lens = torch.randint(1, L + 1, (B,))
    mask = (torch.arange(L)[None] < lens[:, None]).long()
    vp = torch.randn(B, L)
    values = vp + 0.3 * torch.randn(B, L)
    values[0, 0] = vp[0, 0] - 0.5  # exactly on the clip boundary vp == values + cliprange
    rets = values + torch.randn(B, L)
    a = vp.clone().requires_grad_(True)
    want, wfrac = port.value_loss(a, rets, values, mask, 0.5, "token-mean")
    want.backward()
    b = vp.to(dev).requires_grad_(True)
    got, gfrac = ca.compute_value_loss(b, rets.to(dev), values.to(dev), mask.to(dev), 0.5)
    got.backward()
    close(got, want), close(gfrac, wfrac)
    close(b.grad, a.grad, abs_=1e-9)
