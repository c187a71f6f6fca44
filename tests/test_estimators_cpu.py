"""CPU: oracle restatements of the §8(f).2 rows — the other advantage estimators, compute_value_loss and the GPG
policy loss — against golden vectors produced by the unmodified reference (oracle/make_golden.py::gen_estimators)
and, when /root/reference is present, against the live reference on fresh random inputs."""

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port, ref_import

MODES = ["token-mean", "seq-mean-token-sum", "seq-mean-token-mean", "seq-mean-token-sum-norm"]


def T(a):
    return torch.from_numpy(np.asarray(a))


def uid(a):
    return np.array([str(u) for u in a], dtype=object)


def close(a, b, rel=1e-5, abs_=1e-6):
    torch.testing.assert_close(a.float(), b.float(), rtol=rel, atol=abs_)


@pytest.mark.parametrize("tag", ["outcome", "dense"])
@pytest.mark.parametrize("utag", ["", "_solo"])
def test_group_estimators_golden(tag, utag):
    g = load_golden("estimators")
    mask, u = T(g["mask"]), uid(g["uids" + utag])
    rw = T(g["rewards"] if tag == "outcome" else g["dense"])
    close(port.rloo_outcome_advantage(rw, mask, u), T(g[f"rloo_{tag}{utag}"]))
    close(port.opo_outcome_advantage(rw, mask, u), T(g[f"opo_{tag}{utag}"]))
    close(port.gpg_outcome_advantage(rw, mask, u, f_norm=1.5), T(g[f"gpg_{tag}{utag}"]))
    close(port.reinforce_plus_plus_baseline_outcome_advantage(rw, mask, u), T(g[f"rfppb_{tag}{utag}"]))


@pytest.mark.parametrize("tag", ["outcome", "dense"])
def test_passk_golden(tag):
    g = load_golden("estimators")
    mask, u = T(g["mask"]), uid(g["uids"])
    rw = T(g["rewards"] if tag == "outcome" else g["dense"])
    close(port.grpo_passk_outcome_advantage(rw, mask, u), T(g[f"passk_{tag}_norm1"]))
    close(port.grpo_passk_outcome_advantage(rw, mask, u, norm_adv_by_std_in_grpo=False), T(g[f"passk_{tag}_norm0"]))
    with pytest.raises(ValueError):
        port.grpo_passk_outcome_advantage(rw, mask, uid(g["uids_solo"]))


def test_reverse_scans_golden():
    g = load_golden("estimators")
    dense, holes = T(g["dense"]), T(g["holes"])
    for i in (0, 1):
        adv, ret = port.reinforce_plus_plus_outcome_advantage(dense, holes, float(g[f"rfpp_gamma{i}"]))
        close(ret, T(g[f"rfpp_ret{i}"]))
        close(adv, T(g[f"rfpp_adv{i}"]))
    adv, ret = port.remax_outcome_advantage(dense, T(g["baselines"]), holes)
    close(ret, T(g["remax_ret"]))
    close(adv, T(g["remax_adv"]))


@pytest.mark.parametrize("mode", MODES)
def test_value_and_gpg_loss_golden(mode):
    g = load_golden("estimators")
    mask = T(g["mask"])
    vp = T(g["vpreds"]).clone().requires_grad_(True)
    vf, frac = port.value_loss(vp, T(g["rets"]), T(g["values"]), mask, 0.2, mode)
    vf.backward()
    close(torch.stack([vf.detach(), frac]), T(g[f"vf|{mode}"]))
    close(vp.grad, T(g[f"vf_grad|{mode}"]), abs_=1e-8)
    lp = T(g["lp"]).clone().requires_grad_(True)
    loss = port.policy_loss_gpg(lp, T(g["adv"]), mask, mode)
    loss.backward()
    close(loss.detach(), T(g[f"gpg_loss|{mode}"]))
    close(lp.grad, T(g[f"gpg_loss_grad|{mode}"]), abs_=1e-8)


@pytest.mark.skipif(not ref_import.available(), reason="reference tree only exists in the build container")
def test_estimators_against_live_reference_random():
    ca = ref_import.load().core_algos
    g = torch.Generator().manual_seed(101)
    B, L = 24, 40
    lens = torch.randint(5, L + 1, (B,), generator=g)
    mask = (torch.arange(L)[None] < lens[:, None]).long()
    rw = torch.randn(B, L, generator=g) * mask
    u = np.array([f"g{i % 7}" for i in range(B)], dtype=object)
    close(port.rloo_outcome_advantage(rw, mask, u), ca.compute_rloo_outcome_advantage(rw.clone(), mask, u)[0])
    close(port.opo_outcome_advantage(rw, mask, u), ca.compute_opo_outcome_advantage(rw.clone(), mask, u)[0])
    close(port.gpg_outcome_advantage(rw, mask, u), ca.compute_gpg_outcome_advantage(rw.clone(), mask, u)[0])
    close(port.reinforce_plus_plus_baseline_outcome_advantage(rw, mask, u),
          ca.compute_reinforce_plus_plus_baseline_outcome_advantage(rw.clone(), mask, u)[0])
    close(port.grpo_passk_outcome_advantage(rw, mask, u),
          ca.compute_grpo_passk_outcome_advantage(rw.clone(), mask, u, config={})[0])
    cfg = type("Cfg", (), {"gamma": 0.9})()
    a, r = ca.compute_reinforce_plus_plus_outcome_advantage(rw.clone(), mask, config=cfg)
    a2, r2 = port.reinforce_plus_plus_outcome_advantage(rw, mask, 0.9)
    close(a2, a), close(r2, r)
