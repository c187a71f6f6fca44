"""BASELINE configs[0] at FULL size on the GPU through the reference-named API: Qwen2.5-0.5B shape (H=896,
V=151936), GRPO batch 8 prompts x 5 rollouts x 512 response tokens (T = 20480 lm-head rows):
lm-head log-prob + entropy -> GRPO outcome advantage -> PPO clip loss, then the backward into hidden / weight.

The CPU oracle needs ~46 s for the full fp32 lm-head at this shape, so the lm-head is checked against it on a seeded
sample of 384 token rows (full H and V) within the north-star tolerance (2e-3 abs on fp32 log-probs / entropy);
advantages (bit-exact group assignment, 1e-5 relative values) and the loss (1e-5 relative) are checked on the whole
[40, 512] batch, the loss oracle being fed the GPU log-probs so that the loss tolerance is not diluted by the
lm-head's.  Gradients: d_hidden rows are independent of each other, so the sampled rows are compared with the
oracle's closed form; d_weight sums over all tokens, so it is covered by a size-independent property instead —
linearity of (d_hidden, d_weight) in the upstream gradient — plus "masked tokens carry no gradient"."""

import numpy as np
import pytest
import torch

from oracle import port

pytestmark = pytest.mark.gpu

B, N_ROLL, L, H, V = 8, 5, 512, 896, 151936


def _batch(seed=1234):
    g = torch.Generator().manual_seed(seed)
    rows = B * N_ROLL
    lens = torch.randint(L // 4, L + 1, (rows,), generator=g)
    lens[0] = L
    mask = (torch.arange(L)[None] < lens[:, None]).long()
    hidden = (torch.rand(rows * L, H, generator=g) - 0.5).bfloat16()
    weight = (torch.rand(V, H, generator=g) - 0.5).bfloat16() * 0.25
    labels = torch.randint(0, V, (rows * L,), generator=g)
    rewards = torch.zeros(rows, L)
    rewards[torch.arange(rows), lens - 1] = torch.rand(rows, generator=g)
    uid = np.repeat(np.array([f"prompt-{i}" for i in range(B)], dtype=object), N_ROLL)
    perm = torch.randperm(rows, generator=g).numpy()
    return dict(mask=mask, hidden=hidden, weight=weight, labels=labels, rewards=rewards, uid=uid[perm], lens=lens)


def test_config1_full_size_end_to_end(dev):
    from verl_b200.trainer.ppo import core_algos
    from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy

    d = _batch()
    rows, T = B * N_ROLL, B * N_ROLL * L
    h = d["hidden"].to(dev).requires_grad_(True)
    w = d["weight"].to(dev).requires_grad_(True)
    lab = d["labels"].to(dev)
    mask_d = d["mask"].to(dev)

    # ---- lm-head at full size, oracle on a row sample
    lp, ent = linear_cross_entropy(h, w, lab, 1.0, "none")
    assert lp.shape == (T,) and lp.dtype == torch.float32 and ent.dtype == torch.float32
    g = torch.Generator().manual_seed(7)
    sample = torch.randperm(T, generator=g)[:384].sort().values
    lp_o, ent_o = port.lce_forward(d["hidden"][sample], d["weight"], d["labels"][sample], 1.0)
    assert (lp.detach().cpu()[sample] - lp_o).abs().max() < 2e-3  # north-star tolerance, fp32 log-probs
    assert (ent.detach().cpu()[sample] - ent_o).abs().max() < 2e-3
    assert torch.isfinite(lp).all() and torch.isfinite(ent).all() and (lp <= 0).all() and (ent >= 0).all()

    # ---- GRPO advantage on the whole batch: group assignment bit-exact, values 1e-5 relative
    adv, ret = core_algos.compute_grpo_outcome_advantage(d["rewards"].to(dev), mask_d, d["uid"])
    adv_o = port.grpo_outcome_advantage(d["rewards"], d["mask"], d["uid"])
    torch.testing.assert_close(adv.cpu(), adv_o, rtol=1e-5, atol=1e-6)
    assert torch.equal((adv != 0).cpu() | (d["mask"] == 0), (adv_o != 0) | (d["mask"] == 0))
    assert ret.data_ptr() == adv.data_ptr() or torch.equal(ret, adv)  # the reference returns the same tensor twice

    # ---- PPO clip loss (vanilla, token-mean) on the whole batch; oracle fed the GPU log-probs
    lp2 = lp.view(rows, L)
    old = (lp2.detach() + 0.1 * torch.randn(rows, L, device=dev, generator=torch.Generator(device=dev).manual_seed(3)))
    pg_loss, clipfrac, ppo_kl, clipfrac_lower = core_algos.compute_policy_loss(
        old, lp2, adv, mask_d, cliprange=0.2, cliprange_low=0.2, cliprange_high=0.2, clip_ratio_c=3.0,
        loss_agg_mode="token-mean")
    lp_c = lp2.detach().cpu().requires_grad_(True)
    want = port.policy_loss(old.cpu(), lp_c, adv.cpu(), d["mask"], cliprange=0.2, cliprange_low=0.2, cliprange_high=0.2,
                            clip_ratio_c=3.0, loss_agg_mode="token-mean")
    for got, exp in zip((pg_loss, clipfrac, ppo_kl, clipfrac_lower), want):
        torch.testing.assert_close(got.detach().cpu().float(), exp.detach().float(), rtol=1e-5, atol=1e-7)

    # ---- backward: d loss / d log_prob against autograd of the oracle, then through the lm-head
    pg_loss.backward()
    want[0].backward()
    dh1, dw1 = h.grad.clone(), w.grad.clone()
    assert torch.isfinite(dh1.float()).all() and torch.isfinite(dw1.float()).all()
    # rows outside the response mask carry no gradient at all
    dead = (d["mask"].view(-1) == 0).to(dev)
    assert float(dh1[dead].float().abs().max()) == 0.0
    # closed-form d_hidden of the sampled rows from the oracle (d_hidden rows are independent of each other)
    g_lp = lp_c.grad.view(-1)
    dh_o, _ = port.lce_backward(d["hidden"][sample], d["weight"], d["labels"][sample], g_lp[sample],
                                torch.zeros(sample.numel()), 1.0)
    scale = float(dh_o.abs().max())
    assert (dh1.float().cpu()[sample] - dh_o).abs().max() <= 2e-2 * scale + 1e-9

    # linearity of the lm-head backward in the upstream gradient (size-independent property): grad(2 g) == 2 grad(g)
    h.grad = None
    w.grad = None
    lp_b, _ = linear_cross_entropy(h, w, lab, 1.0, "none")
    lp_b.backward(2.0 * g_lp.to(dev))
    torch.testing.assert_close(h.grad.float(), 2.0 * dh1.float(), rtol=2e-2, atol=2e-2 * float(dh1.float().abs().max()))
    torch.testing.assert_close(w.grad.float(), 2.0 * dw1.float(), rtol=2e-2, atol=2e-2 * float(dw1.float().abs().max()))
