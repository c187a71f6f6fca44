"""CPU: the oracle restatement (oracle/port.py) against golden vectors produced by the unmodified reference
(oracle/make_golden.py), against the reference tests' own known answers, and — when /root/reference is present —
against the live reference on fresh random inputs."""

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port, ref_import


def T(a):
    return torch.from_numpy(np.asarray(a))


# ---------------------------------------------------------------------------------------------- lm-head
@pytest.mark.parametrize("case", ["a", "b", "c"])
def test_lce_port_matches_reference_golden(case):
    g = load_golden("lce")
    hidden, weight = T(g[f"{case}_hidden"]).bfloat16(), T(g[f"{case}_weight"]).bfloat16()
    labels = T(g[f"{case}_labels"])
    temp = float(g[f"{case}_temperature"])
    lp, ent = port.lce_forward(hidden, weight, labels, temp)
    # torch-vs-torch tolerance of the reference test (tests/utils/test_linear_cross_entropy.py:210-216)
    torch.testing.assert_close(lp, T(g[f"{case}_logprobs"]), atol=1e-4, rtol=1e-4)
    torch.testing.assert_close(ent, T(g[f"{case}_entropy"]), atol=1e-4, rtol=1e-4)
    dh, dw = port.lce_backward(hidden, weight, labels, T(g[f"{case}_g_lp"]), T(g[f"{case}_g_ent"]), temp)
    torch.testing.assert_close(dh, T(g[f"{case}_d_hidden"]), atol=1e-4, rtol=1e-4)
    torch.testing.assert_close(dw, T(g[f"{case}_d_weight"]), atol=1e-4, rtol=1e-4)


@pytest.mark.parametrize("world", [2, 8])
def test_vocab_parallel_merge_equals_full(world):
    g = load_golden("lce")
    hidden, weight = T(g["c_hidden"]).bfloat16(), T(g["c_weight"]).bfloat16()  # V = 1536 divides by 2 and 8
    labels = T(g["c_labels"])
    shards = list(weight.chunk(world, dim=0))
    lp, ent = port.lce_forward_vocab_parallel(hidden, shards, labels, 1.0)
    torch.testing.assert_close(lp, T(g["c_logprobs"]), atol=2e-5, rtol=1e-5)
    torch.testing.assert_close(ent, T(g["c_entropy"]), atol=2e-5, rtol=1e-5)


@pytest.mark.parametrize("splits", [1, 3])
def test_partial_format_roundtrip(splits):
    g = load_golden("lce")
    hidden, weight = T(g["b_hidden"]).bfloat16(), T(g["b_weight"]).bfloat16()
    labels, temp = T(g["b_labels"]), float(g["b_temperature"])
    part = port.lce_partials(hidden, weight, labels, temp, num_splits=splits)
    lp, ent = port.lce_finalize(part, temp)
    torch.testing.assert_close(lp, T(g["b_logprobs"]), atol=2e-5, rtol=1e-5)
    torch.testing.assert_close(ent, T(g["b_entropy"]), atol=2e-5, rtol=1e-5)


def test_logits_functions():
    g = load_golden("logits")
    logits, labels = T(g["f32_logits"]), T(g["labels"])
    torch.testing.assert_close(port.logprobs_from_logits(logits, labels), T(g["f32_logprobs_v2"]), atol=1e-6, rtol=1e-6)
    torch.testing.assert_close(port.logprobs_from_logits(logits, labels), T(g["f32_logprobs_naive"]), atol=1e-5, rtol=1e-5)
    torch.testing.assert_close(port.entropy_from_logits(logits), T(g["f32_entropy"]), atol=1e-6, rtol=1e-6)
    torch.testing.assert_close(port.entropy_from_logits(logits).reshape(-1), T(g["f32_entropy_chunked"]), atol=1e-5, rtol=1e-5)
    lb = T(g["bf16_logits"]).bfloat16()
    torch.testing.assert_close(port.logprobs_from_logits(lb.reshape(-1, 333), labels.reshape(-1)).float(),
                               T(g["bf16_logprobs_v2"]), atol=0, rtol=0)


# ---------------------------------------------------------------------------------------------- masked ops
def test_masked_ops_golden():
    g = load_golden("masked")
    v, m = T(g["values"]), T(g["mask"])
    torch.testing.assert_close(port.masked_sum(v, m), T(g["masked_sum"]))
    torch.testing.assert_close(port.masked_mean(v, m), T(g["masked_mean"]))
    torch.testing.assert_close(port.masked_mean(v, m, axis=-1), T(g["masked_mean_axis"]))
    torch.testing.assert_close(port.masked_var(v, m), T(g["masked_var"]))
    torch.testing.assert_close(port.masked_var(v, m, unbiased=False), T(g["masked_var_biased"]))
    torch.testing.assert_close(port.masked_whiten(v, m), T(g["masked_whiten"]))
    torch.testing.assert_close(port.masked_whiten(v, m, shift_mean=False), T(g["masked_whiten_noshift"]))


def test_masked_mean_known_answers():
    """tests/utils/test_torch_functional.py:55-66 of the reference"""
    assert port.masked_mean(torch.tensor([1.0, 2.0, 3.0, 4.0]), torch.tensor([1, 0, 0, 1])).item() == pytest.approx(2.5)
    v = torch.tensor([1.0, float("nan"), 3.0, 4.0])
    assert port.masked_mean(v, torch.tensor([True, False, False, True])).item() == pytest.approx(2.5)
    assert torch.isnan(port.masked_mean(v, torch.tensor([True, True, False, False])))
    with pytest.raises(ValueError):
        port.masked_var(torch.ones(4), torch.zeros(4))
    with pytest.raises(ValueError):
        port.masked_var(torch.ones(4), torch.tensor([0, 1, 0, 0]))


# ---------------------------------------------------------------------------------------------- rmpad
def test_rmpad_golden_bit_exact():
    g = load_golden("rmpad")
    idx, seqlens, cu, mx = port.unpad_indices(T(g["mask"]))
    assert idx.dtype == torch.int64 and cu.dtype == torch.int32
    assert torch.equal(idx, T(g["indices"])) and torch.equal(cu, T(g["cu_seqlens"])) and mx == int(g["max_seqlen"])
    assert torch.equal(seqlens, T(g["seqused"]))
    ids = T(g["ids"]).reshape(-1)
    assert torch.equal(port.gather_rows(ids, idx), T(g["ids_rmpad"]))
    padded = port.pad_rows(T(g["vals"]), idx, ids.numel()).view(*g["mask"].shape, 1)
    assert torch.equal(padded, T(g["padded"]))
    e_idx, _, e_cu, e_mx = port.unpad_indices(T(g["edge_mask"]))
    assert torch.equal(e_idx, T(g["edge_indices"])) and torch.equal(e_cu, T(g["edge_cu"])) and e_mx == int(g["edge_max"])


# ---------------------------------------------------------------------------------------------- advantages
def test_grpo_golden():
    g = load_golden("grpo")
    uids = np.array([str(u) for u in g["uids"]], dtype=object)
    r, m = T(g["rewards"]), T(g["mask"])
    torch.testing.assert_close(port.grpo_outcome_advantage(r, m, uids), T(g["adv_norm1"]), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(port.grpo_outcome_advantage(r, m, uids, norm_adv_by_std_in_grpo=False), T(g["adv_norm0"]),
                               rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(port.grpo_outcome_advantage(T(g["dense_rewards"]), m, uids), T(g["dense_adv"]),
                               rtol=1e-5, atol=1e-6)
    eq = torch.zeros_like(r)
    eq[:, 0] = 0.5
    torch.testing.assert_close(port.grpo_outcome_advantage(eq, m, uids), T(g["equal_adv"]), rtol=1e-5, atol=1e-7)


def test_group_ids_first_seen_order():
    ids, n = port.group_ids(np.array(["b", "a", "b", "c", "a"], dtype=object))
    assert ids.tolist() == [0, 1, 0, 2, 1] and n == 3


@pytest.mark.parametrize("i", [0, 1])
def test_gae_golden(i):
    g = load_golden("gae")
    adv, ret = port.gae_advantage_return(T(g["rewards"]), T(g["values"]), T(g["mask"]), float(g[f"gamma{i}"]),
                                         float(g[f"lam{i}"]))
    torch.testing.assert_close(adv, T(g[f"adv{i}"]), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(ret, T(g[f"ret{i}"]), rtol=1e-5, atol=1e-6)


def test_gae_multi_turn_invariance():
    """tests/trainer/ppo/test_core_algos_on_cpu.py:134-188 of the reference: values on masked (observation) tokens
    must not change advantages / returns on response tokens."""
    torch.manual_seed(0)
    B, L = 3, 40
    mask = (torch.rand(B, L) > 0.4).long()
    mask[:, 0] = 1
    r, v = torch.randn(B, L), torch.randn(B, L)
    v2 = torch.where(mask.bool(), v, torch.randn(B, L) * 100)
    a1, ret1 = port.gae_advantage_return(r, v, mask, 0.95, 0.9)
    a2, ret2 = port.gae_advantage_return(r, v2, mask, 0.95, 0.9)
    assert torch.equal(a1 * mask, a2 * mask)
    assert torch.equal(ret1 * mask, ret2 * mask)


# ---------------------------------------------------------------------------------------------- losses
MODES = ["token-mean", "seq-mean-token-sum", "seq-mean-token-mean", "seq-mean-token-sum-norm"]


@pytest.mark.parametrize("cname", ["vanilla", "dapo", "grpo_kl", "all"])
@pytest.mark.parametrize("mode", MODES)
def test_actor_loss_golden(cname, mode):
    g = load_golden("loss")
    lo, hi, c, ent_coeff, kl_coef, scale = [float(x) for x in g[cname + "|cfg"]]
    kl = str(g[cname + "|kl"])
    lp = T(g["lp"]).clone().requires_grad_(True)
    ent = T(g["ent"]).clone().requires_grad_(True)
    loss, logged = port.actor_loss(
        T(g["old"]), lp, T(g["adv"]), T(g["mask"]), cliprange_low=lo, cliprange_high=hi, clip_ratio_c=c,
        loss_agg_mode=mode, entropy=ent, entropy_coeff=ent_coeff, ref_log_prob=T(g["ref"]) if kl_coef != 0 else None,
        kl_loss_coef=kl_coef, kl_loss_type=kl, loss_scale=scale)
    loss.backward()
    want = g[f"{cname}|{mode}|scalars"]
    got = [loss, logged["pg_loss"], logged["pg_clipfrac"], logged["ppo_kl"], logged["pg_clipfrac_lower"],
           logged["entropy_loss"], logged["kl_loss"]]
    for a, b in zip(got, want):
        assert float(a) == pytest.approx(float(b), rel=1e-6, abs=1e-8)
    torch.testing.assert_close(lp.grad, T(g[f"{cname}|{mode}|d_lp"]), rtol=1e-5, atol=1e-9)
    if ent.grad is not None:
        torch.testing.assert_close(ent.grad, T(g[f"{cname}|{mode}|d_ent"]), rtol=1e-5, atol=1e-9)


def test_kl_and_agg_golden():
    g = load_golden("loss")
    for kind in ["kl", "abs", "mse", "low_var_kl"]:
        torch.testing.assert_close(port.kl_penalty(T(g["lp"]), T(g["ref"]), kind), T(g["kl_penalty|" + kind]),
                                   rtol=1e-6, atol=1e-7)
    with pytest.raises(NotImplementedError):
        port.kl_penalty(T(g["lp"]), T(g["ref"]), "full")
    for mode in MODES:
        assert float(port.agg_loss(T(g["ent"]), T(g["mask"]), mode)) == pytest.approx(float(g["agg|" + mode]), rel=1e-6)
    with pytest.raises(ValueError):
        port.agg_loss(T(g["ent"]), T(g["mask"]), "nope")


# ---------------------------------------------------------------------------------------------- live reference
@pytest.mark.skipif(not ref_import.available(), reason="reference tree only exists in the build container")
def test_port_against_live_reference_random():
    ref = ref_import.load()
    torch.manual_seed(5)
    B, L = 6, 50
    lens = torch.randint(5, L + 1, (B,))
    mask = (torch.arange(L)[None] < lens[:, None]).long()
    old = -torch.rand(B, L) * 2
    lp = old + 0.2 * torch.randn(B, L)
    adv = torch.randn(B, L)
    for mode in MODES:
        a = ref.core_algos.compute_policy_loss(old, lp, adv, mask, cliprange=0.2, clip_ratio_c=3.0, loss_agg_mode=mode)
        b = port.policy_loss(old, lp, adv, mask, cliprange=0.2, clip_ratio_c=3.0, loss_agg_mode=mode)
        for x, y in zip(a, b):
            assert float(x) == pytest.approx(float(y), rel=1e-6, abs=1e-9)
    r, v = torch.randn(B, L), torch.randn(B, L)
    a = ref.core_algos.compute_gae_advantage_return(r, v, mask, 0.99, 0.95)
    b = port.gae_advantage_return(r, v, mask, 0.99, 0.95)
    torch.testing.assert_close(a[0], b[0], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(a[1], b[1], rtol=1e-5, atol=1e-6)
    uid = np.array(["x", "y", "x", "z", "y", "x"], dtype=object)
    a = ref.core_algos.compute_grpo_outcome_advantage(r.clone(), mask, uid)[0]
    torch.testing.assert_close(a, port.grpo_outcome_advantage(r, mask, uid), rtol=1e-5, atol=1e-6)
