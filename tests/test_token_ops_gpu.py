"""GPU parity of the memory-bound token kernels (through the reference-named Python API -> C ABI) against the
reference golden vectors and the CPU oracle.  Tolerance: 1e-5 relative on losses and advantages (north_star);
gradients 1e-5 relative with a small absolute floor."""

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port

pytestmark = pytest.mark.gpu

MODES = ["token-mean", "seq-mean-token-sum", "seq-mean-token-mean", "seq-mean-token-sum-norm"]
REL = 1e-5


def T(a, dev):
    return torch.from_numpy(np.asarray(a)).to(dev)


def close(a, b, rel=REL, abs_=1e-7):
    torch.testing.assert_close(a.detach().float().cpu(), b.detach().float().cpu(), rtol=rel, atol=abs_)


# ------------------------------------------------------------------------------------------------ actor loss
@pytest.mark.parametrize("cname", ["vanilla", "dapo", "grpo_kl", "all"])
@pytest.mark.parametrize("mode", MODES)
def test_fused_actor_loss_vs_reference_golden(cname, mode, dev):
    from verl_b200.trainer.ppo import core_algos as ca

    g = load_golden("loss")
    lo, hi, c, ent_coeff, kl_coef, scale = [float(x) for x in g[cname + "|cfg"]]
    kl = str(g[cname + "|kl"])
    lp = T(g["lp"], dev).requires_grad_(True)
    ent = T(g["ent"], dev).requires_grad_(True)
    loss, metrics = ca.fused_actor_loss(
        T(g["old"], dev), lp, T(g["adv"], dev), T(g["mask"], dev), cliprange_low=lo, cliprange_high=hi,
        clip_ratio_c=c, loss_agg_mode=mode, entropy=ent, entropy_coeff=ent_coeff,
        ref_log_prob=T(g["ref"], dev) if kl_coef != 0 else None, kl_loss_coef=kl_coef, kl_loss_type=kl,
        loss_scale=scale)
    loss.backward()
    want = g[f"{cname}|{mode}|scalars"]
    got = metrics.cpu().numpy()
    np.testing.assert_allclose(float(loss), want[0], rtol=REL, atol=1e-8)
    np.testing.assert_allclose(got[1:5], want[1:5], rtol=REL, atol=1e-8)
    if ent_coeff != 0:
        np.testing.assert_allclose(got[5], want[5], rtol=REL)
    if kl_coef != 0:
        np.testing.assert_allclose(got[6], want[6], rtol=REL)
    assert got[7] == g["mask"].sum()
    close(lp.grad, T(g[f"{cname}|{mode}|d_lp"], dev), rel=REL, abs_=1e-9)
    if ent_coeff != 0:
        close(ent.grad, T(g[f"{cname}|{mode}|d_ent"], dev), rel=REL, abs_=1e-10)


@pytest.mark.parametrize("mode", MODES)
def test_compute_policy_loss_api(mode, dev):
    from verl_b200.trainer.ppo import core_algos as ca

    g = load_golden("loss")
    lp = T(g["lp"], dev).requires_grad_(True)
    out = ca.compute_policy_loss(T(g["old"], dev), lp, T(g["adv"], dev), T(g["mask"], dev), cliprange=0.2,
                                 clip_ratio_c=3.0, loss_agg_mode=mode)
    want = g[f"vanilla|{mode}|scalars"]
    for a, b in zip(out, want[1:5]):
        assert a.dim() == 0
        np.testing.assert_allclose(float(a), b, rtol=REL, atol=1e-8)
    out[0].backward()
    close(lp.grad, T(g[f"vanilla|{mode}|d_lp"], dev), rel=REL, abs_=1e-9)
    with pytest.raises(AssertionError, match="clip_ratio_c"):
        ca.compute_policy_loss(T(g["old"], dev), lp, T(g["adv"], dev), T(g["mask"], dev), cliprange=0.2, clip_ratio_c=1.0)


@pytest.mark.parametrize("mask_dtype", [torch.int64, torch.int32, torch.bool, torch.float32])
@pytest.mark.parametrize("L", [64, 37])  # 37: rows not 16-byte aligned -> scalar path
def test_policy_loss_mask_dtypes_and_unaligned(mask_dtype, L, dev):
    from verl_b200.trainer.ppo import core_algos as ca

    torch.manual_seed(1)
    B = 9
    lens = torch.randint(1, L + 1, (B,))
    mask = (torch.arange(L)[None] < lens[:, None])
    old = -torch.rand(B, L) * 3
    lp = old + 0.3 * torch.randn(B, L)
    adv = torch.randn(B, L)
    lp_c = lp.clone().requires_grad_(True)
    want = port.policy_loss(old, lp_c, adv, mask.long(), cliprange=0.2, cliprange_high=0.28, clip_ratio_c=3.0)
    want[0].backward()
    lp_g = lp.to(dev).requires_grad_(True)
    got = ca.compute_policy_loss(old.to(dev), lp_g, adv.to(dev), mask.to(mask_dtype).to(dev), cliprange=0.2,
                                 cliprange_high=0.28, clip_ratio_c=3.0)
    got[0].backward()
    for a, b in zip(got, want):
        np.testing.assert_allclose(float(a), float(b), rtol=REL, atol=1e-8)
    close(lp_g.grad, lp_c.grad, rel=REL, abs_=1e-9)


@pytest.mark.parametrize("mask_kind", ["binary_i64", "float_weights", "int_two", "int_wide"])
@pytest.mark.parametrize("kl", [False, True])
@pytest.mark.parametrize("shape", [(9, 64), (3, 4096), (700, 1024)])
def test_fused_actor_loss_weighted_masks(mask_kind, kl, shape, dev):
    """Token-mean loss WITH gradients over masks that are weights rather than 0/1 flags (the reference's masked_mean takes
    any mask values): float weights, an integer 2, a value that does not fit a byte.  The vector path of this call is the
    cooperative single-pass kernel (mask bits kept in shared memory for 0/1 masks, the mask re-read otherwise); the three
    shapes give one slice unit per block, several units per block, and more blocks than one wave of SMs."""
    from verl_b200.trainer.ppo import core_algos as ca

    torch.manual_seed(11)
    B, L = shape
    lens = torch.randint(1, L + 1, (B,))
    on = (torch.arange(L)[None] < lens[:, None])
    if mask_kind == "binary_i64":
        mask = on.long()
    elif mask_kind == "float_weights":
        mask = on.float() * (0.25 + torch.rand(B, L))
    elif mask_kind == "int_two":
        mask = on.long() * torch.randint(1, 3, (B, L))
    else:
        mask = on.long() * torch.where(torch.rand(B, L) < 0.1, 300, 1)
    old = -torch.rand(B, L) * 3
    lp = old + 0.3 * torch.randn(B, L)
    ref = old + 0.2 * torch.randn(B, L)
    adv = torch.randn(B, L)
    ent = torch.rand(B, L)
    kw = dict(cliprange_low=0.2, cliprange_high=0.28, clip_ratio_c=3.0, entropy_coeff=1e-3, kl_loss_coef=0.01 if kl else 0.0,
              loss_scale=0.5)
    lp_c, ent_c = lp.clone().requires_grad_(True), ent.clone().requires_grad_(True)
    want, sc = port.actor_loss(old, lp_c, adv, mask, entropy=ent_c, ref_log_prob=ref if kl else None, **kw)
    want.backward()
    lp_g, ent_g = lp.to(dev).requires_grad_(True), ent.to(dev).requires_grad_(True)
    loss, metrics = ca.fused_actor_loss(old.to(dev), lp_g, adv.to(dev), mask.to(dev), entropy=ent_g,
                                        ref_log_prob=ref.to(dev) if kl else None, **kw)
    loss.backward()
    np.testing.assert_allclose(float(loss), float(want), rtol=REL, atol=1e-8)
    got = metrics.cpu().numpy()
    np.testing.assert_allclose(got[1:5], [float(sc[k]) for k in ("pg_loss", "pg_clipfrac", "ppo_kl", "pg_clipfrac_lower")],
                               rtol=REL, atol=1e-8)
    np.testing.assert_allclose(got[7], float(mask.sum()), rtol=1e-6)
    close(lp_g.grad, lp_c.grad, rel=REL, abs_=1e-9)
    close(ent_g.grad, ent_c.grad, rel=REL, abs_=1e-10)


def test_agg_loss_and_kl_penalty(dev):
    from verl_b200.trainer.ppo import core_algos as ca

    g = load_golden("loss")
    ent, mask = T(g["ent"], dev), T(g["mask"], dev)
    for mode in MODES:
        e = ent.clone().requires_grad_(True)
        out = ca.agg_loss(e, mask, mode)
        np.testing.assert_allclose(float(out), float(g["agg|" + mode]), rtol=REL)
        out.backward()
        e_c = T(g["ent"], "cpu").requires_grad_(True)
        port.agg_loss(e_c, T(g["mask"], "cpu"), mode).backward()
        close(e.grad, e_c.grad, rel=REL, abs_=1e-10)
    with pytest.raises(ValueError, match="Invalid loss_agg_mode"):
        ca.agg_loss(ent, mask, "nope")
    lp, ref = T(g["lp"], dev), T(g["ref"], dev)
    for kind in ["kl", "abs", "mse", "low_var_kl", "k1", "k2", "k3"]:
        base = {"k1": "kl", "k2": "mse", "k3": "low_var_kl"}.get(kind, kind)
        lpg = lp.clone().requires_grad_(True)
        out = ca.kl_penalty(lpg, ref, kind)
        # Which bar governs per-token kl_penalty: the 1e-5 RELATIVE bar of north_star is stated
        # for "loss and advantages", i.e. for the aggregated scalars and O(1) per-token values.  A per-token
        # low_var_kl is exp(k) - k - 1 with k near 0: both implementations cancel ~1.0 - 1.0 in fp32, so one element
        # is only good to ~1 ulp of 1.0 (6e-8) in ABSOLUTE terms and its relative error is unbounded as kl -> 0 (the
        # 1.37e-5 max_rel_err of profiles/r2_final2_driver_ops_e2e.jsonl is one such element of size ~4e-3).  The
        # per-token bar is therefore absolute (5e-7 = a few ulp of 1.0) with a tight relative term for large values;
        # the 1e-5 relative bar is applied where the reference consumes the values: the aggregated kl loss below.
        close(out, T(g["kl_penalty|" + base], dev), rel=1e-6, abs_=5e-7)
        want_agg = port.agg_loss(T(g["kl_penalty|" + base], "cpu"), T(g["mask"], "cpu"), "token-mean")
        np.testing.assert_allclose(float(ca.agg_loss(out.detach(), mask, "token-mean")), float(want_agg), rtol=REL)
        out.sum().backward()
        lpc = T(g["lp"], "cpu").requires_grad_(True)
        port.kl_penalty(lpc, T(g["ref"], "cpu"), base).sum().backward()
        close(lpg.grad, lpc.grad, rel=1e-5, abs_=1e-6)
    with pytest.raises(NotImplementedError):
        ca.kl_penalty(lp, ref, "full")


# ------------------------------------------------------------------------------------------------ masked ops
def test_masked_ops_golden(dev):
    from verl_b200.utils import torch_functional as F

    g = load_golden("masked")
    v, m = T(g["values"], dev), T(g["mask"], dev)
    close(F.masked_sum(v, m), T(g["masked_sum"], dev))
    close(F.masked_mean(v, m), T(g["masked_mean"], dev))
    close(F.masked_mean(v, m, axis=-1), T(g["masked_mean_axis"], dev))
    close(F.masked_sum(v, m, axis=-1), T(g["masked_sum_axis"], dev))
    close(F.masked_var(v, m), T(g["masked_var"], dev))
    close(F.masked_var(v, m, unbiased=False), T(g["masked_var_biased"], dev))
    close(F.masked_whiten(v, m), T(g["masked_whiten"], dev), abs_=1e-6)
    close(F.masked_whiten(v, m, shift_mean=False), T(g["masked_whiten_noshift"], dev), abs_=1e-6)
    # masked_mean gradient
    vg = v.clone().requires_grad_(True)
    F.masked_mean(vg, m).backward()
    vc = T(g["values"], "cpu").requires_grad_(True)
    port.masked_mean(vc, T(g["mask"], "cpu")).backward()
    close(vg.grad, vc.grad)


def test_masked_mean_known_answers(dev):
    """the reference's tests/utils/test_torch_functional.py:55-66"""
    from verl_b200.utils import torch_functional as F

    v = torch.tensor([1.0, 2.0, 3.0, 4.0], device=dev)
    assert F.masked_mean(v, torch.tensor([1, 0, 0, 1], device=dev)).item() == pytest.approx(2.5)
    vn = torch.tensor([1.0, float("nan"), 3.0, 4.0], device=dev)
    assert F.masked_mean(vn, torch.tensor([True, False, False, True], device=dev)).item() == pytest.approx(2.5)
    assert torch.isnan(F.masked_mean(vn, torch.tensor([True, True, False, False], device=dev)))
    with pytest.raises(ValueError, match="At least one element"):
        F.masked_whiten(v, torch.zeros(4, dtype=torch.long, device=dev))
    with pytest.raises(ValueError, match="sum of the mask is one"):
        F.masked_whiten(v, torch.tensor([0, 1, 0, 0], device=dev))


# ------------------------------------------------------------------------------------------------ advantages
def test_grpo_golden(dev):
    from verl_b200.trainer.ppo import core_algos as ca

    g = load_golden("grpo")
    uids = np.array([str(u) for u in g["uids"]], dtype=object)
    r, m = T(g["rewards"], dev), T(g["mask"], dev)
    adv, ret = ca.compute_grpo_outcome_advantage(r, m, uids)
    assert adv is ret  # the reference returns the same tensor twice (core_algos.py:308)
    close(adv, T(g["adv_norm1"], dev), abs_=1e-6)
    close(ca.compute_grpo_outcome_advantage(r, m, uids, norm_adv_by_std_in_grpo=False)[0], T(g["adv_norm0"], dev), abs_=1e-6)
    close(ca.compute_grpo_outcome_advantage(T(g["dense_rewards"], dev), m, uids)[0], T(g["dense_adv"], dev), abs_=1e-6)
    eq = torch.zeros_like(r)
    eq[:, 0] = 0.5
    got = ca.compute_grpo_outcome_advantage(eq, m, uids)[0]
    close(got, T(g["equal_adv"], dev), abs_=1e-7)


@pytest.mark.parametrize("B,L,n", [(40, 512, 5), (128, 1024, 8), (16, 16384, 4), (7, 33, 1), (20000, 8, 8), (9001, 4, 3)])
def test_grpo_vs_oracle_config_shapes(B, L, n, dev):
    """BASELINE configs: C1 40x512 (8 prompts x 5), C3 128x1024 (n=8), C5 16x16384; plus all-singletons; plus batches
    beyond one block of the group kernel (cooperative multi-block sweep, shuffled uids: groups span warps and blocks)."""
    import uuid

    from verl_b200.trainer.ppo import core_algos as ca

    torch.manual_seed(B)
    lens = torch.randint(L // 4, L + 1, (B,))
    mask = (torch.arange(L)[None] < lens[:, None]).long()
    rewards = torch.zeros(B, L)
    # the big batches draw graded scores (0, 1/4 .. 1) like a rubric reward: thousands of random continuous groups contain
    # ill-conditioned ones (std << mean) where the fp32 reference itself is only good to a few 1e-5
    rewards[torch.arange(B), lens - 1] = torch.rand(B) if B <= 4096 else torch.randint(0, 5, (B,)).float() / 4
    uids = np.repeat(np.array([str(uuid.uuid4()) for _ in range((B + n - 1) // n)], dtype=object), n)[:B]
    uids = uids[torch.randperm(B).numpy()]
    want = port.grpo_outcome_advantage(rewards, mask, uids)
    got = ca.compute_grpo_outcome_advantage(rewards.to(dev), mask.to(dev), uids)[0]
    close(got, want, abs_=1e-6)


@pytest.mark.parametrize("i", [0, 1])
def test_gae_golden(i, dev):
    from verl_b200.trainer.ppo import core_algos as ca

    g = load_golden("gae")
    adv, ret = ca.compute_gae_advantage_return(T(g["rewards"], dev), T(g["values"], dev), T(g["mask"], dev),
                                               float(g[f"gamma{i}"]), float(g[f"lam{i}"]))
    close(adv, T(g[f"adv{i}"], dev), abs_=2e-6)
    close(ret, T(g[f"ret{i}"], dev), abs_=2e-6)


@pytest.mark.parametrize("B,L", [(16, 16384), (3, 9000), (4, 1), (2, 8193)])
def test_gae_vs_oracle_long_rows(B, L, dev):
    from verl_b200.trainer.ppo import core_algos as ca

    torch.manual_seed(L)
    mask = (torch.rand(B, L) > 0.2).long()
    mask[:, 0] = 1
    if L > 1:
        mask[0, L // 2:] = 0
    r, v = torch.randn(B, L) * 0.1, torch.randn(B, L)
    if B * L < 100000:
        want = port.gae_advantage_return(r, v, mask, 0.99, 0.95) if mask.sum() > 1 else None
    else:  # python loop over 16384 steps is the reference's slow path; still seconds at this size
        want = port.gae_advantage_return(r, v, mask, 0.99, 0.95)
    if want is None:
        return
    adv, ret = ca.compute_gae_advantage_return(r.to(dev), v.to(dev), mask.to(dev), 0.99, 0.95)
    close(ret, want[1], rel=1e-5, abs_=5e-6)
    close(adv, want[0], rel=1e-5, abs_=5e-6)


def test_gae_multi_turn_invariance(dev):
    """the reference's tests/trainer/ppo/test_core_algos_on_cpu.py:134-188 (torch.equal on response tokens)"""
    from verl_b200.trainer.ppo import core_algos as ca

    torch.manual_seed(0)
    B, L = 3, 40
    mask = (torch.rand(B, L) > 0.4).long().to(dev)
    mask[:, 0] = 1
    r, v = torch.randn(B, L, device=dev), torch.randn(B, L, device=dev)
    v2 = torch.where(mask.bool(), v, torch.randn(B, L, device=dev) * 100)
    a1, ret1 = ca.compute_gae_advantage_return(r, v, mask, 0.95, 0.9)
    a2, ret2 = ca.compute_gae_advantage_return(r, v2, mask, 0.95, 0.9)
    assert torch.equal(a1 * mask, a2 * mask)
    assert torch.equal(ret1 * mask, ret2 * mask)


def test_large_batch_properties(dev):
    """2^25 tokens (the size the HBM roofline is evidenced at): idempotence / consistency properties —
    loss of a batch duplicated along B equals the loss of the batch (token-mean), gradients halve."""
    from verl_b200.trainer.ppo import core_algos as ca

    torch.manual_seed(3)
    B, L = 1024, 16384
    g = torch.Generator(device=dev).manual_seed(3)
    lens = torch.randint(L // 4, L + 1, (B,), device=dev, generator=g)
    mask = (torch.arange(L, device=dev)[None] < lens[:, None]).long()
    old = -torch.rand(B, L, device=dev, generator=g) * 3
    lp = (old + 0.05 * torch.randn(B, L, device=dev, generator=g)).requires_grad_(True)
    adv = torch.randn(B, L, device=dev, generator=g)
    loss, m = ca.fused_actor_loss(old, lp, adv, mask, cliprange=0.2)
    loss.backward()
    half = B // 2
    lp2 = lp.detach()[:half].clone().requires_grad_(True)
    loss2, m2 = ca.fused_actor_loss(old[:half], lp2, adv[:half], mask[:half], cliprange=0.2)
    loss2.backward()
    n1, n2 = float(m[7]), float(m2[7])
    # token-mean: per-token gradient scales with 1/N
    close(lp.grad[:half] * (n1 / n2), lp2.grad, rel=1e-5, abs_=1e-12)
    assert n1 == float(mask.sum()) and torch.isfinite(loss)


def test_masked_ops_any_axis(dev):
    """masked_sum / masked_mean over every axis choice of a 3-D tensor (the reference accepts any `axis`)."""
    from verl_b200.utils import torch_functional as F

    torch.manual_seed(5)
    v = torch.randn(3, 5, 7)
    m = (torch.rand(3, 5, 7) > 0.4).long()
    for axis in (0, 1, 2, -1, -2, (0, 1), (1, 2), (0, 2), [0]):
        ax = tuple(axis) if isinstance(axis, list) else axis
        close(F.masked_sum(v.to(dev), m.to(dev), axis=axis), port.masked_sum(v, m, axis=ax), abs_=1e-6)
        close(F.masked_mean(v.to(dev), m.to(dev), axis=axis), port.masked_mean(v, m, axis=ax), abs_=1e-6)


def test_nan_propagates_like_torch(dev):
    """torch.clamp / maximum / min propagate NaN: a NaN log-prob on a response token must give a NaN loss (that is how
    a diverged run is noticed); a NaN on a masked-out token must not (masked_mean zeroes it first)."""
    from verl_b200.trainer.ppo import core_algos as ca

    B, L = 2, 8
    old = -torch.rand(B, L, device=dev)
    adv = torch.randn(B, L, device=dev)
    mask = torch.ones(B, L, dtype=torch.long, device=dev)
    mask[1, 5:] = 0
    lp = old.clone()
    lp[1, 6] = float("nan")  # masked out
    loss, _ = ca.fused_actor_loss(old, lp, adv, mask, cliprange=0.2)
    want, _ = port.actor_loss(old.cpu(), lp.cpu(), adv.cpu(), mask.cpu(), cliprange_low=0.2, cliprange_high=0.2)
    assert torch.isfinite(loss) and torch.isfinite(want)
    close(loss, want)
    lp[0, 2] = float("nan")  # a response token
    loss, _ = ca.fused_actor_loss(old, lp, adv, mask, cliprange=0.2)
    want, _ = port.actor_loss(old.cpu(), lp.cpu(), adv.cpu(), mask.cpu(), cliprange_low=0.2, cliprange_high=0.2)
    assert torch.isnan(want) and torch.isnan(loss)
    kl = ca.kl_penalty(lp, old, "low_var_kl")
    assert torch.isnan(kl[0, 2]) and torch.isfinite(kl[0, 3])


def test_token_ops_are_cuda_graph_capturable(dev):
    """At the BASELINE micro-batch sizes the token ops are host-bound (8-25 us of wrapper + launch API per call), so the
    ones without a host read — fused loss with gradients (the cooperative single-pass kernel), value loss, masked sums,
    KL penalty / KL reward, GRPO advantage, index_first_axis — must be capturable into ONE CUDA graph: capture on one
    batch, replay on another, compare with the eager calls on that batch."""
    from verl_b200 import ops

    B, L, n = 16, 2048, 4

    def data(seed):
        g = torch.Generator(device=dev).manual_seed(seed)
        lens = torch.randint(L // 4, L + 1, (B,), device=dev, generator=g)
        d = dict(mask=(torch.arange(L, device=dev)[None] < lens[:, None]).long(),
                 old=-torch.rand(B, L, device=dev, generator=g) * 3, adv=torch.randn(B, L, device=dev, generator=g),
                 ent=torch.rand(B, L, device=dev, generator=g), values=torch.randn(B, L, device=dev, generator=g),
                 rewards=torch.rand(B, L, device=dev, generator=g))
        d["lp"] = d["old"] + 0.1 * torch.randn(B, L, device=dev, generator=g)
        d["ref"] = d["old"] + 0.1 * torch.randn(B, L, device=dev, generator=g)
        d["rows"] = torch.randperm(B * L, device=dev, generator=g)[:4096].sort().values
        return d

    gids = torch.arange(B, device=dev, dtype=torch.int32) // n

    def run(d):
        m, dlp, dent = ops.ppo_loss(d["old"], d["lp"], d["adv"], d["mask"], cliprange_low=0.2, cliprange_high=0.28,
                                    entropy=d["ent"], entropy_coeff=1e-3, ref_log_prob=d["ref"], kl_loss_coef=0.01)
        vl, dvp = ops.value_loss(d["values"], d["lp"], d["old"], d["mask"], 0.2, "token-mean", True)
        ms = ops.masked_sums(d["ent"], d["mask"])
        kl, _ = ops.kl_penalty_fwd(d["lp"], d["ref"], "low_var_kl", False)
        rw, cur = ops.kl_reward(d["rewards"], d["old"], d["ref"], d["mask"], "kl", 0.01)
        ga = ops.grpo_advantage(d["rewards"], d["mask"], gids, B // n, 1e-6, True)
        picked = ops.gather_rows(d["lp"].reshape(-1, 1), d["rows"])
        return [m, dlp, dent, vl, dvp, ms, kl, rw, cur, ga, picked]

    static = data(0)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):  # warm-up outside the capture (library initialisation, allocator pools)
        run(static)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        outs = run(static)
    fresh = data(1)
    for k, v in fresh.items():
        static[k].copy_(v)
    graph.replay()
    torch.cuda.synchronize()
    want = run(fresh)
    for got, ref in zip(outs, want):
        torch.testing.assert_close(got.double().cpu(), ref.double().cpu(), rtol=1e-6, atol=1e-9)
    graph.replay()  # a second replay gives the same result (every reduction slot is re-zeroed inside the graph)
    torch.cuda.synchronize()
    for got, ref in zip(outs, want):
        torch.testing.assert_close(got.double().cpu(), ref.double().cpu(), rtol=1e-6, atol=1e-9)
