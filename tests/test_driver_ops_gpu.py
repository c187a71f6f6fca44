"""GPU parity of the driver-side token ops: apply_kl_penalty, compute_response_mask,
compute_advantage dispatch, compute_data_metrics — against golden vectors from the reference's function bodies and the
CPU oracle at BASELINE config shapes.  Tolerance 1e-5 relative."""

import json

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port

pytestmark = pytest.mark.gpu


def T(a, dev):
    return torch.from_numpy(np.asarray(a)).to(dev)


class Data:
    def __init__(self, **kw):
        self.batch, self.non_tensor_batch, self.meta_info = dict(**kw), {}, {}


@pytest.mark.parametrize("kind", ["kl", "abs", "mse", "low_var_kl"])
def test_apply_kl_penalty_golden(kind, dev):
    from verl_b200.trainer.ppo import core_algos as ca
    from verl_b200.trainer.ppo.ray_trainer import apply_kl_penalty

    g = load_golden("driver")
    ctrl = ca.AdaptiveKLController(init_kl_coef=0.05, target_kl=0.1, horizon=1000)
    data = Data(response_mask=T(g["rmask"], dev), token_level_scores=T(g["scores"], dev), old_log_probs=T(g["old"], dev),
                ref_log_prob=T(g["ref"], dev))
    data, metrics = apply_kl_penalty(data, ctrl, kl_penalty=kind)
    want = g[f"klr|{kind}|scalars"]
    torch.testing.assert_close(data.batch["token_level_rewards"].cpu(), torch.from_numpy(g[f"klr|{kind}|rewards"]),
                               rtol=1e-5, atol=1e-7)
    assert abs(metrics["actor/reward_kl_penalty"] - want[0]) <= 1e-5 * max(1.0, abs(want[0]))
    assert metrics["actor/reward_kl_penalty_coeff"] == want[1]
    assert abs(ctrl.value - want[2]) <= 1e-7  # the adaptive controller saw the same KL
    with pytest.raises(NotImplementedError):
        apply_kl_penalty(data, ctrl, kl_penalty="full")


@pytest.mark.parametrize("critic", [1, 0])
def test_compute_data_metrics_golden(critic, dev):
    from verl_b200.trainer.ppo.metric_utils import compute_data_metrics

    g = load_golden("driver")
    want = json.loads(str(g[f"dm|critic{critic}"]))
    data = Data(token_level_scores=T(g["scores"], dev), token_level_rewards=T(g["klr|kl|rewards"], dev),
                advantages=T(g["adv"], dev), returns=T(g["returns"], dev), values=T(g["values"], dev),
                attention_mask=T(g["attn"], dev), response_mask=T(g["rmask"], dev))
    got = compute_data_metrics(data, use_critic=bool(critic))
    assert set(got) == set(want)
    for k, v in want.items():
        assert abs(got[k] - v) <= 1e-5 * max(1.0, abs(v)), (k, got[k], v)


@pytest.mark.parametrize("B,P,L", [(40, 512, 512), (128, 1024, 1024), (16, 2048, 16384), (3, 5, 37)])
def test_driver_ops_vs_oracle_config_shapes(B, P, L, dev):
    from verl_b200.trainer.ppo import core_algos as ca
    from verl_b200.trainer.ppo.metric_utils import compute_data_metrics
    from verl_b200.trainer.ppo.ray_trainer import apply_kl_penalty, compute_advantage, compute_response_mask

    torch.manual_seed(B + L)
    S = P + L
    p_len, r_len = torch.randint(1, P + 1, (B,)), torch.randint(1, L + 1, (B,))
    pos = torch.arange(S)[None]
    attn = ((pos >= (P - p_len)[:, None]) & (pos < (P + r_len)[:, None])).long()
    scores = torch.zeros(B, L)
    scores[torch.arange(B), r_len - 1] = torch.rand(B)
    old = -torch.rand(B, L) * 3
    ref = old + 0.1 * torch.randn(B, L)
    values = torch.randn(B, L)
    uid = np.array([f"u{i // 4}" for i in range(B)], dtype=object)
    data = Data(responses=torch.zeros(B, L, dtype=torch.long, device=dev), attention_mask=attn.to(dev),
                token_level_scores=scores.to(dev), old_log_probs=old.to(dev), ref_log_prob=ref.to(dev),
                values=values.to(dev))
    data.non_tensor_batch["uid"] = uid
    data.batch["response_mask"] = compute_response_mask(data)
    rmask = attn[:, -L:]
    assert torch.equal(data.batch["response_mask"].cpu(), rmask)
    ctrl = ca.FixedKLController(0.01)
    data, m = apply_kl_penalty(data, ctrl, "low_var_kl")
    want_r, want_kl = port.kl_reward(scores, old, ref, rmask, "low_var_kl", 0.01)
    torch.testing.assert_close(data.batch["token_level_rewards"].cpu(), want_r, rtol=1e-5, atol=1e-7)
    assert abs(m["actor/reward_kl_penalty"] - want_kl) <= 1e-5 * max(1e-3, abs(want_kl))
    # advantage dispatch (GAE needs values; RLOO goes through the registry)
    data = compute_advantage(data, ca.AdvantageEstimator.GAE, gamma=1.0, lam=0.95)
    want_adv, want_ret = port.gae_advantage_return(want_r, values, rmask, 1.0, 0.95)
    torch.testing.assert_close(data.batch["advantages"].cpu(), want_adv, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(data.batch["returns"].cpu(), want_ret, rtol=1e-5, atol=1e-5)
    got = compute_data_metrics(data, use_critic=True)
    want = port.data_metrics(scores, want_r, want_adv, want_ret, values, rmask, attn)
    for k, v in want.items():
        assert abs(got[k] - v) <= 2e-5 * max(1.0, abs(v)), (k, got[k], v)
    data = compute_advantage(data, "rloo")
    torch.testing.assert_close(data.batch["advantages"].cpu(), port.rloo_outcome_advantage(want_r, rmask, uid),
                               rtol=1e-5, atol=1e-6)


def test_driver_ops_accept_host_tensors_like_the_reference_driver(dev):
    """The reference runs these on the DRIVER with CPU tensors (verl/trainer/ppo/ray_trainer.py:1283-1305).  Host tensors
    in -> staged through pinned buffers -> kernels on the device -> host tensors out (verl_b200/staging.py); results
    identical to the device-tensor call."""
    from verl_b200.trainer.ppo import core_algos as ca
    from verl_b200.trainer.ppo.metric_utils import compute_data_metrics
    from verl_b200.trainer.ppo.ray_trainer import apply_kl_penalty, compute_advantage

    g = load_golden("driver")
    host = dict(response_mask=torch.from_numpy(g["rmask"]), token_level_scores=torch.from_numpy(g["scores"]),
                old_log_probs=torch.from_numpy(g["old"]), ref_log_prob=torch.from_numpy(g["ref"]),
                values=torch.from_numpy(g["values"]), attention_mask=torch.from_numpy(g["attn"]))
    B = host["token_level_scores"].shape[0]
    uid = np.array([f"u{i // 2}" for i in range(B)], dtype=object)
    outs = []
    for where in ("cpu", dev):
        data = Data(**{k: v.clone().to(where) for k, v in host.items()})
        data.non_tensor_batch["uid"] = uid
        data, m = apply_kl_penalty(data, ca.FixedKLController(0.05), "low_var_kl")
        data = compute_advantage(data, ca.AdvantageEstimator.GAE, gamma=0.99, lam=0.95)
        gae = (data.batch["advantages"], data.batch["returns"])
        data = compute_advantage(data, ca.AdvantageEstimator.GRPO)
        assert data.batch["advantages"] is data.batch["returns"]  # the reference returns the SAME tensor twice
        dm = compute_data_metrics(data, use_critic=True)
        for t in (data.batch["token_level_rewards"], *gae, data.batch["advantages"]):
            assert t.device.type == torch.device(where).type, (where, t.device)
        outs.append((m, data.batch["token_level_rewards"].cpu(), gae[0].cpu(), gae[1].cpu(),
                     data.batch["advantages"].cpu(), dm))
    (m0, r0, a0, t0, g0, d0), (m1, r1, a1, t1, g1, d1) = outs
    assert m0 == m1 and d0 == d1
    for x, y in ((r0, r1), (a0, a1), (t0, t1), (g0, g1)):
        assert torch.equal(x, y)
    torch.testing.assert_close(r0, torch.from_numpy(g["klr|low_var_kl|rewards"]) if "klr|low_var_kl|rewards" in g.files
                               else r1, rtol=1e-5, atol=1e-7)


def test_host_tensor_losses_keep_autograd(dev):
    """compute_policy_loss / agg_loss / masked_mean on HOST tensors that require grad: the graph runs on the device and the
    leaf gradient comes back on the host, equal to the device-tensor call."""
    from verl_b200.trainer.ppo import core_algos as ca
    from verl_b200.utils import torch_functional as F

    torch.manual_seed(5)
    B, L = 7, 33
    old = -torch.rand(B, L) * 3
    adv = torch.randn(B, L)
    mask = (torch.rand(B, L) < 0.8).long()
    grads = []
    for where in ("cpu", dev):
        lp = (old + 0.05 * torch.ones(B, L)).to(where).requires_grad_(True)
        loss, clipfrac, kl, clip_low = ca.compute_policy_loss(old.to(where), lp, adv.to(where), mask.to(where),
                                                              cliprange=0.2, clip_ratio_c=3.0)
        extra = F.masked_mean(lp * lp, mask.to(where))
        (loss + 0.1 * extra).backward()
        assert loss.device.type == torch.device(where).type and lp.grad.device.type == torch.device(where).type
        grads.append((float(loss), float(extra), lp.grad.cpu()))
    assert grads[0][0] == grads[1][0] and grads[0][1] == grads[1][1]
    assert torch.equal(grads[0][2], grads[1][2])
    with pytest.raises(NotImplementedError):
        F.masked_sum(torch.ones(2, 2, requires_grad=True), torch.ones(2, 2))
