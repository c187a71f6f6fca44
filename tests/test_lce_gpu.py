"""GPU parity of the fused lm-head op (tcgen05 kernels through the C ABI) against the oracle and the reference's
golden vectors.  Tolerances: north_star's 2e-3 abs on fp32 log-probs / entropy for bf16 inputs (the reference's own
kernel test allows logprobs atol 1e-3 rtol 2e-4, entropy atol 5e-3 rtol 5e-4, grads atol 2e-2 rtol 4e-2:
tests/utils/test_linear_cross_entropy.py:218-223,270-275)."""

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port

pytestmark = pytest.mark.gpu

LP_ATOL = 2e-3
GRAD_ATOL, GRAD_RTOL = 2e-2, 4e-2


def T(a, dev):
    return torch.from_numpy(np.asarray(a)).to(dev)


def _inputs(Tn, H, V, seed, dev):
    """distributions of the reference test (tests/utils/test_linear_cross_entropy.py:145-162)"""
    g = torch.Generator(device="cpu").manual_seed(seed)
    hidden = (torch.rand(Tn, H, generator=g) - 0.5).bfloat16().to(dev)
    weight = (torch.rand(V, H, generator=g) - 0.5).bfloat16().to(dev)
    labels = torch.randint(0, V, (Tn,), generator=g).to(dev)
    g_lp = (torch.rand(Tn, generator=g) * 2 - 1).to(dev)
    g_ent = (torch.rand(Tn, generator=g) - 0.5).to(dev)
    return hidden, weight, labels, g_lp, g_ent


def _gpu_fp32_reference(hidden, weight, labels, temperature, g_lp=None, g_ent=None):
    """plain PyTorch fp32 reference of the same op (run_torch_entropy recipe,
    tests/utils/test_linear_cross_entropy.py:48-61), chunked so that [T,V] fp32 stays small"""
    h = hidden.float().requires_grad_(g_lp is not None)
    w = weight.float().requires_grad_(g_lp is not None)
    lps, ents = [], []
    for s in range(0, h.shape[0], 1024):
        logits = (h[s:s + 1024] @ w.t()) / temperature
        lse = torch.logsumexp(logits, -1)
        p = torch.softmax(logits, -1)
        ents.append(lse - (p * logits).sum(-1))
        lps.append(logits.gather(-1, labels[s:s + 1024, None]).squeeze(-1) - lse)
    lp, ent = torch.cat(lps), torch.cat(ents)
    if g_lp is None:
        return lp, ent, None, None
    (lp * g_lp + ent * g_ent).sum().backward()
    return lp.detach(), ent.detach(), h.grad, w.grad


@pytest.mark.parametrize("case", ["a", "b", "c"])
def test_forward_backward_vs_reference_golden(case, dev):
    from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy

    g = load_golden("lce")
    hidden = T(g[f"{case}_hidden"], dev).bfloat16().requires_grad_(True)
    weight = T(g[f"{case}_weight"], dev).bfloat16().requires_grad_(True)
    labels = T(g[f"{case}_labels"], dev)
    temp = float(g[f"{case}_temperature"])
    lp, ent = linear_cross_entropy(hidden, weight, labels, temp, "none")
    assert lp.dtype == torch.float32 and lp.shape == (hidden.shape[0],)
    torch.testing.assert_close(lp, T(g[f"{case}_logprobs"], dev), atol=LP_ATOL, rtol=0)
    torch.testing.assert_close(ent, T(g[f"{case}_entropy"], dev), atol=LP_ATOL, rtol=0)
    (lp * T(g[f"{case}_g_lp"], dev) + ent * T(g[f"{case}_g_ent"], dev)).sum().backward()
    assert hidden.grad.dtype == torch.bfloat16 and weight.grad.dtype == torch.bfloat16
    torch.testing.assert_close(hidden.grad.float(), T(g[f"{case}_d_hidden"], dev), atol=GRAD_ATOL, rtol=GRAD_RTOL)
    torch.testing.assert_close(weight.grad.float(), T(g[f"{case}_d_weight"], dev), atol=GRAD_ATOL, rtol=GRAD_RTOL)


# the reference's five test shapes (tests/utils/test_linear_cross_entropy.py:116-140); V of the first two scaled by
# 1/8 and 1/4 to keep the fp32 torch reference's [T,V] buffers small, the tile-tail structure (V % 256 != 0) is kept
SHAPES = [
    (1937, 3584, 152064 // 8),
    (2169, 896, 151936 // 4),
    (1530, 2048, 32256),
    (1388, 4096, 102400 // 4),
    (300, 128, 1000),
    (128, 64, 256),
    (1, 128, 7),
]


@pytest.mark.parametrize("Tn,H,V", SHAPES)
def test_forward_backward_vs_oracle_shapes(Tn, H, V, dev):
    from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy

    temp = 1.5
    hidden, weight, labels, g_lp, g_ent = _inputs(Tn, H, V, 233376, dev)
    hidden.requires_grad_(True), weight.requires_grad_(True)
    lp, ent = linear_cross_entropy(hidden, weight, labels, temp, "none")
    (lp * g_lp + ent * g_ent).sum().backward()
    lp_r, ent_r, dh_r, dw_r = _gpu_fp32_reference(hidden.detach(), weight.detach(), labels, temp, g_lp, g_ent)
    torch.testing.assert_close(lp, lp_r, atol=LP_ATOL, rtol=0)
    torch.testing.assert_close(ent, ent_r, atol=LP_ATOL, rtol=0)
    torch.testing.assert_close(hidden.grad.float(), dh_r, atol=GRAD_ATOL, rtol=GRAD_RTOL)
    torch.testing.assert_close(weight.grad.float(), dw_r, atol=GRAD_ATOL, rtol=GRAD_RTOL)
    if Tn <= 300:  # the CPU oracle itself at sizes it finishes in seconds
        lp_o, ent_o = port.lce_forward(hidden.detach().cpu(), weight.detach().cpu(), labels.cpu(), temp)
        torch.testing.assert_close(lp.cpu(), lp_o, atol=LP_ATOL, rtol=0)
        torch.testing.assert_close(ent.cpu(), ent_o, atol=LP_ATOL, rtol=0)


def test_reductions_and_3d_shapes(dev):
    from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy

    hidden, weight, labels, g_lp, g_ent = _inputs(4 * 96, 256, 1500, 1, dev)
    lp, ent = linear_cross_entropy(hidden, weight, labels, 1.0, "none")
    lp_sum, _ = linear_cross_entropy(hidden, weight, labels, 1.0, "sum")
    lp_mean, _ = linear_cross_entropy(hidden, weight, labels, 1.0, "mean")
    assert lp_sum.dim() == 0 and lp_mean.dim() == 0
    torch.testing.assert_close(lp_sum, lp.sum(), rtol=1e-5, atol=1e-3)
    torch.testing.assert_close(lp_mean, lp.mean(), rtol=1e-5, atol=1e-5)
    # [B,S,H] with B>1 -> [B,S]; [1,T,H] -> flat [T] like the reference's Triton op
    lp3, ent3 = linear_cross_entropy(hidden.view(4, 96, 256), weight, labels.view(4, 96), 1.0, "none")
    assert lp3.shape == (4, 96) and torch.equal(lp3.reshape(-1), lp) and torch.equal(ent3.reshape(-1), ent)
    lp1, _ = linear_cross_entropy(hidden.view(1, -1, 256), weight, labels.view(1, -1), 1.0, "none")
    assert lp1.shape == (4 * 96,)
    # backward through the mean reduction and without an entropy gradient
    h = hidden.clone().requires_grad_(True)
    w = weight.clone().requires_grad_(True)
    lp_mean, _ = linear_cross_entropy(h, w, labels, 1.0, "mean")
    lp_mean.backward()
    ones = torch.full_like(g_lp, 1.0 / h.shape[0])
    _, _, dh_r, dw_r = _gpu_fp32_reference(hidden, weight, labels, 1.0, ones, torch.zeros_like(g_ent))
    torch.testing.assert_close(h.grad.float(), dh_r, atol=GRAD_ATOL, rtol=GRAD_RTOL)
    torch.testing.assert_close(w.grad.float(), dw_r, atol=GRAD_ATOL, rtol=GRAD_RTOL)


@pytest.mark.parametrize("chunk", [256, 512, 1024])
def test_backward_chunking_is_consistent(chunk, dev):
    """d_hidden accumulated in fp32 over vocab chunks must agree with the single-chunk result"""
    from verl_b200 import ops

    hidden, weight, labels, g_lp, g_ent = _inputs(500, 256, 2000, 3, dev)
    part = ops.lce_forward_partials(hidden, weight, labels, 1.3)
    _, _, lse, eb = ops.lce_finalize(part, 1.3)
    dh0, dw0 = ops.lce_backward(hidden, weight, labels, lse, eb, g_lp, g_ent, 1.3, chunk_v=2048)
    dh1, dw1 = ops.lce_backward(hidden, weight, labels, lse, eb, g_lp, g_ent, 1.3, chunk_v=chunk)
    torch.testing.assert_close(dw0.float(), dw1.float(), atol=0, rtol=0)  # d_weight rows do not depend on chunking
    torch.testing.assert_close(dh0.float(), dh1.float(), atol=4e-3, rtol=2e-2)


@pytest.mark.parametrize("splits", [1, 2, 5, 8])
def test_split_invariance(splits, dev):
    from verl_b200 import ops

    hidden, weight, labels, _, _ = _inputs(257, 128, 2100, 4, dev)
    base = ops.lce_finalize(ops.lce_forward_partials(hidden, weight, labels, 1.0, num_splits=1), 1.0)
    out = ops.lce_finalize(ops.lce_forward_partials(hidden, weight, labels, 1.0, num_splits=splits), 1.0)
    for a, b in zip(base, out):
        torch.testing.assert_close(a, b, atol=2e-5, rtol=1e-5)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_vocab_parallel_emulated_on_one_gpu(world, dev):
    """vocab-parallel path with the ranks emulated serially: shard r runs the real kernel with label_offset =
    r*V_local (verl/utils/kernel/kernels.py:232), partials are concatenated as exchange_partials would, finalize
    merges.  Must equal the single-shard result; d_hidden partials must sum to the full d_hidden
    (tests/utils/test_linear_cross_entropy_tp.py:313-314)."""
    from verl_b200 import ops

    Tn, H, V = 700, 512, 8 * 2004
    hidden, weight, labels, g_lp, g_ent = _inputs(Tn, H, V, 5, dev)
    full = ops.lce_finalize(ops.lce_forward_partials(hidden, weight, labels, 1.5), 1.5)
    v_local = V // world
    parts = [ops.lce_forward_partials(hidden, weight[r * v_local:(r + 1) * v_local].contiguous(), labels, 1.5,
                                      label_offset=r * v_local, num_splits=2) for r in range(world)]
    merged = torch.cat(parts, dim=1)
    lp, ent, lse, eb = ops.lce_finalize(merged, 1.5)
    torch.testing.assert_close(lp, full[0], atol=2e-5, rtol=1e-5)
    torch.testing.assert_close(ent, full[1], atol=2e-5, rtol=1e-5)
    dh_full, dw_full = ops.lce_backward(hidden, weight, labels, full[2], full[3], g_lp, g_ent, 1.5)
    dh_sum = torch.zeros(Tn, H, device=dev)
    for r in range(world):
        w_r = weight[r * v_local:(r + 1) * v_local].contiguous()
        dh_r, dw_r = ops.lce_backward(hidden, w_r, labels, lse, eb, g_lp, g_ent, 1.5, label_offset=r * v_local)
        dh_sum += dh_r.float()
        torch.testing.assert_close(dw_r.float(), dw_full[r * v_local:(r + 1) * v_local].float(), atol=1e-2, rtol=2e-2)
    torch.testing.assert_close(dh_sum, dh_full.float(), atol=GRAD_ATOL, rtol=GRAD_RTOL)


def test_full_size_properties(dev):
    """BASELINE configs[1] at full size (T=65536, H=3584, V=152064): size-independent properties —
    (1) probabilities: logprob <= 0, 0 <= entropy <= ln V;  (2) a random subset of rows equals the fp32 reference;
    (3) linearity of the backward in (g_lp, g_ent) checked on d_hidden rows."""
    from verl_b200 import ops

    Tn, H, V = 65536, 3584, 152064
    g = torch.Generator(device=dev).manual_seed(9)
    hidden = (torch.rand(Tn, H, device=dev, generator=g) - 0.5).bfloat16()
    weight = (torch.rand(V, H, device=dev, generator=g) - 0.5).bfloat16()
    labels = torch.randint(0, V, (Tn,), device=dev, generator=g)
    lp, ent, lse, eb = ops.lce_finalize(ops.lce_forward_partials(hidden, weight, labels, 1.0), 1.0)
    assert torch.isfinite(lp).all() and torch.isfinite(ent).all()
    assert (lp <= 1e-4).all() and (ent >= -1e-3).all() and (ent <= np.log(V) + 1e-3).all()
    rows = torch.randperm(Tn, device=dev, generator=g)[:512]
    lp_r, ent_r, _, _ = _gpu_fp32_reference(hidden[rows], weight, labels[rows], 1.0)
    torch.testing.assert_close(lp[rows], lp_r, atol=LP_ATOL, rtol=0)
    torch.testing.assert_close(ent[rows], ent_r, atol=LP_ATOL, rtol=0)
    del lp_r, ent_r
    # backward on a 4096-token slice at full H and V: linearity d(a*g1 + g2) == a*d(g1) + d(g2)
    sl = slice(0, 4096)
    h, lab = hidden[sl].contiguous(), labels[sl].contiguous()
    g1 = torch.rand(4096, device=dev, generator=g) * 2 - 1
    g2 = torch.rand(4096, device=dev, generator=g) - 0.5
    dh_a, _ = ops.lce_backward(h, weight, lab, lse[sl].contiguous(), eb[sl].contiguous(), g1, None, 1.0)
    dh_b, _ = ops.lce_backward(h, weight, lab, lse[sl].contiguous(), eb[sl].contiguous(), None, g2, 1.0)
    dh_c, _ = ops.lce_backward(h, weight, lab, lse[sl].contiguous(), eb[sl].contiguous(), 2 * g1, g2, 1.0)
    torch.testing.assert_close(dh_c.float(), 2 * dh_a.float() + dh_b.float(), atol=GRAD_ATOL, rtol=GRAD_RTOL)


@pytest.mark.parametrize("chunk_v", [0, 9472, -1])
def test_full_size_backward_sampled_rows_vs_fp32(chunk_v, dev):
    """BASELINE configs[1] at FULL size through the autograd op: d_hidden for 256 sampled tokens and d_weight for 256
    sampled vocab rows against an fp32 PyTorch reference (the reference's own gradient bar: atol 2e-2, rtol 4e-2,
    tests/utils/test_linear_cross_entropy.py:270-275), for the default workspace budget (chunk_v 0), the reference's
    default split width of 9 504 columns rounded to the tile (9 472, 17 chunks) and the whole vocabulary in one chunk
    (-1: 20 GB of dlogits).  Catches what linearity cannot: a dropped k-block in the 65 536-deep K loop, a mis-strided
    MN-major tensor map at ldds = 152 064, a tail bug in the last vocab chunk or in the split-K tail."""
    from lce_reference import excess, sampled_d_weight_reference, sampled_rows_reference
    from verl_b200.utils.kernel import kernels
    from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy

    Tn, H, V, temp = 65536, 3584, 152064, 1.0
    g = torch.Generator(device=dev).manual_seed(21)
    hidden = (torch.rand(Tn, H, device=dev, generator=g) - 0.5).bfloat16().requires_grad_(True)
    weight = (torch.rand(V, H, device=dev, generator=g) - 0.5).bfloat16().requires_grad_(True)
    labels = torch.randint(0, V, (Tn,), device=dev, generator=g)
    g_lp = torch.rand(Tn, device=dev, generator=g) * 2 - 1
    g_ent = torch.rand(Tn, device=dev, generator=g) - 0.5
    tok = torch.randperm(Tn, device=dev, generator=g)[:256]
    voc = torch.cat([torch.randperm(V, device=dev, generator=g)[:192], labels[tok[:32]],  # label rows carry the one-hot term
                     torch.arange(V - 32, V, device=dev)])                                 # the last rows of the last chunk
    kernels.set_backward_chunk(((V + 255) // 256) * 256 if chunk_v < 0 else chunk_v)
    try:
        lp, ent = linear_cross_entropy(hidden, weight, labels, temp, "none")
        torch.autograd.backward([lp, ent], [g_lp, g_ent])
    finally:
        kernels.set_backward_chunk(0)
    lp, ent = lp.detach(), ent.detach()
    lp_r, ent_r, lse_r, eb_r, dh_r = sampled_rows_reference(hidden.detach(), weight.detach(), labels, temp, g_lp, g_ent, tok)
    torch.testing.assert_close(lp[tok], lp_r, atol=LP_ATOL, rtol=0)
    torch.testing.assert_close(ent[tok], ent_r, atol=LP_ATOL, rtol=0)
    assert excess(hidden.grad[tok], dh_r, GRAD_ATOL, GRAD_RTOL) <= 0
    # per-token scalars of ALL tokens from the kernel's outputs (verified on the sample just above)
    tgt = (hidden.detach().float() * weight.detach()[labels].float()).sum(-1) / temp
    lse = tgt - lp
    eb = lse - ent
    torch.testing.assert_close(lse[tok], lse_r, atol=LP_ATOL, rtol=0)
    dw_r = sampled_d_weight_reference(hidden.detach(), weight.detach()[voc], voc, labels, lse, eb, temp, g_lp, g_ent)
    assert excess(weight.grad[voc], dw_r, GRAD_ATOL, GRAD_RTOL) <= 0


@pytest.mark.parametrize("name,Tn,H,V,world", [
    ("configs[3] Llama-3-70B, 8-way shard", 65536, 8192, 128256, 8),
    ("configs[4] Qwen2.5-32B DAPO micro-batch, 8-way shard", 262144, 5120, 152064, 8),
])
def test_full_size_vocab_shard_shapes(name, Tn, H, V, world, dev):
    """BASELINE configs[3] / configs[4] at full size, one rank's shard of the vocab-parallel path (V/8 = 16032 and
    19008 columns: neither is a multiple of the 256-column tile, so the masked vocab tail runs at full size).  The
    other ranks' partials are replaced by the exact fp32 statistics of a row sample, so the merged result must equal
    the full-vocab fp32 reference on those rows."""
    from verl_b200 import ops

    v_local = V // world
    rank = 5
    g = torch.Generator(device=dev).manual_seed(11)
    hidden = (torch.rand(Tn, H, device=dev, generator=g) - 0.5).bfloat16()
    w_local = (torch.rand(v_local, H, device=dev, generator=g) - 0.5).bfloat16()
    labels = torch.randint(0, V, (Tn,), device=dev, generator=g)
    part = ops.lce_forward_partials(hidden, w_local, labels, 1.0, label_offset=rank * v_local)  # [4, S, T]
    assert torch.isfinite(part).all()
    lp, ent, lse, eb = ops.lce_finalize(part, 1.0)  # as if this shard were the whole vocabulary
    in_shard = (labels >= rank * v_local) & (labels < (rank + 1) * v_local)
    assert (ent >= -1e-3).all() and (ent <= np.log(v_local) + 1e-3).all()
    assert (lp[in_shard] <= 1e-4).all()
    rows = torch.randperm(Tn, device=dev, generator=g)[:384]
    logits = hidden[rows].float() @ w_local.float().t()
    lse_r = torch.logsumexp(logits, -1)
    p = torch.softmax(logits, -1)
    torch.testing.assert_close(lse[rows], lse_r, atol=LP_ATOL, rtol=0)
    torch.testing.assert_close(ent[rows], lse_r - (p * logits).sum(-1), atol=LP_ATOL, rtol=0)
    hit = in_shard[rows]
    tgt = logits.gather(-1, (labels[rows] - rank * v_local).clamp(0, v_local - 1)[:, None]).squeeze(-1)
    torch.testing.assert_close(lp[rows][hit], (tgt - lse_r)[hit], atol=LP_ATOL, rtol=0)
    del logits, p
    # backward on a 2048-token slice at full H and shard width: linearity in (g_lp, g_ent), d_weight rows of labels
    sl = slice(0, 2048)
    h, lab = hidden[sl].contiguous(), labels[sl].contiguous()
    g1 = torch.rand(2048, device=dev, generator=g) * 2 - 1
    g2 = torch.rand(2048, device=dev, generator=g) - 0.5
    args = (h, w_local, lab, lse[sl].contiguous(), eb[sl].contiguous())
    dh_a, dw_a = ops.lce_backward(*args, g1, None, 1.0, label_offset=rank * v_local)
    dh_b, dw_b = ops.lce_backward(*args, None, g2, 1.0, label_offset=rank * v_local)
    dh_c, dw_c = ops.lce_backward(*args, 2 * g1, g2, 1.0, label_offset=rank * v_local)
    torch.testing.assert_close(dh_c.float(), 2 * dh_a.float() + dh_b.float(), atol=GRAD_ATOL, rtol=GRAD_RTOL)
    torch.testing.assert_close(dw_c.float(), 2 * dw_a.float() + dw_b.float(), atol=GRAD_ATOL, rtol=GRAD_RTOL)
    del dh_a, dh_b, dh_c, dw_a, dw_b, dw_c
    # FULL-size backward of this shard (all T tokens; per-token scalars of the shard-as-vocabulary statistics, which is a
    # consistent soft-max over the shard's columns) vs fp32 torch on 256 sampled tokens / 256 sampled shard rows
    from lce_reference import excess, sampled_d_weight_reference

    g_lp = torch.rand(Tn, device=dev, generator=g) * 2 - 1
    g_ent = torch.rand(Tn, device=dev, generator=g) - 0.5
    dh, dw = ops.lce_backward(hidden, w_local, labels, lse, eb, g_lp, g_ent, 1.0, label_offset=rank * v_local)
    tok = torch.randperm(Tn, device=dev, generator=g)[:256]
    voc = torch.cat([torch.randperm(v_local, device=dev, generator=g)[:224], torch.arange(v_local - 32, v_local, device=dev)])
    z = hidden[tok].float() @ w_local.float().t()
    p = torch.exp(z - lse[tok, None])
    onehot = ((labels[tok] - rank * v_local)[:, None] == torch.arange(v_local, device=dev)[None]).float()
    d = g_lp[tok, None] * (onehot - p) - g_ent[tok, None] * p * (z - eb[tok, None])
    assert excess(dh[tok], d @ w_local.float(), GRAD_ATOL, GRAD_RTOL) <= 0
    del z, p, onehot, d
    dw_r = sampled_d_weight_reference(hidden, w_local[voc], voc + rank * v_local, labels, lse, eb, 1.0, g_lp, g_ent)
    assert excess(dw[voc], dw_r, GRAD_ATOL, GRAD_RTOL) <= 0


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_backward_token_chunks_match_vocab_chunks(dtype, dev):
    """The three ways of cutting the dlogits workspace — one block, column chunks (fp32 d_hidden accumulator; the
    reference's _Split_Dlogits_N, kernels.py:1492-1548) and row chunks (fp32 d_weight accumulator; its _Split_Dlogits_M)
    — give the same gradients, and the row-chunked result equals the fp32 reference.  Ragged T and V (neither a
    multiple of the 256 tile), three chunks each way."""
    from verl_b200 import ops

    Tn, H, V, temp = 700, 256, 1900, 1.3
    hidden, weight, labels, g_lp, g_ent = _inputs(Tn, H, V, 17, dev)
    hidden, weight = hidden.to(dtype), weight.to(dtype)
    lp_r, ent_r, dh_r, dw_r = _gpu_fp32_reference(hidden, weight, labels, temp, g_lp, g_ent)
    part = ops.lce_forward_partials(hidden, weight, labels, temp)
    _, _, lse, eb = ops.lce_finalize(part, temp)
    one = ops.lce_backward(hidden, weight, labels, lse, eb, g_lp, g_ent, temp, chunk_v=2048)
    cols = ops.lce_backward(hidden, weight, labels, lse, eb, g_lp, g_ent, temp, chunk_v=768)
    rows = ops.lce_backward(hidden, weight, labels, lse, eb, g_lp, g_ent, temp, chunk_t=256)
    assert ops.backward_plan(Tn, H, V, 0, 256)["mode"].startswith("token") and ops.backward_plan(Tn, H, V, 0, 256)["chunks"] == 3
    for got in (one, cols, rows):
        torch.testing.assert_close(got[0].float(), dh_r, atol=GRAD_ATOL, rtol=GRAD_RTOL)
        torch.testing.assert_close(got[1].float(), dw_r, atol=GRAD_ATOL, rtol=GRAD_RTOL)
    # d_hidden rows are final per row chunk (no accumulation across chunks): equal to the one-block result up to the
    # summation order of the split-K tail tiles
    torch.testing.assert_close(rows[0].float(), one[0].float(), atol=4e-3, rtol=1e-2)


def test_split_k_tail_gemm(dev):
    """Plain tcgen05 GEMM with the split-K tail: the tiles of the last partial round are cut along K into fp32 partial
    tiles and summed in a fixed order — results equal the unsplit GEMM to bf16 rounding, fp32 output to 1e-4 rel, for
    every operand-major combination and a ragged shape (M, N, K not multiples of the tile)."""
    from verl_b200 import ops

    g = torch.Generator(device=dev).manual_seed(4)
    for (M, N, K) in [(19008, 3584, 8192), (1000, 704, 4160), (256 * 80, 256, 2048)]:
        A = (torch.rand(M, K, device=dev, generator=g) - 0.5).bfloat16()
        B = (torch.rand(N, K, device=dev, generator=g) - 0.5).bfloat16()
        ref = A.float() @ B.float().t()
        for a_k, b_k in ((True, True), (False, True), (True, False), (False, False)):
            Aop = A if a_k else A.t().contiguous()
            Bop = B if b_k else B.t().contiguous()
            if (not a_k and M % 8) or (not b_k and N % 8):
                continue  # MN-major operands need a 16-byte row stride
            c32 = ops.gemm_bf16(Aop, Bop, a_k, b_k, out_dtype=torch.float32, split_k=True)
            c16 = ops.gemm_bf16(Aop, Bop, a_k, b_k, out_dtype=torch.bfloat16, split_k=True)
            base = ops.gemm_bf16(Aop, Bop, a_k, b_k, out_dtype=torch.float32, split_k=False)
            torch.testing.assert_close(c32, ref, atol=2e-3, rtol=1e-4)
            torch.testing.assert_close(c32, base, atol=2e-3, rtol=1e-4)
            torch.testing.assert_close(c16.float(), ref, atol=2e-1, rtol=1e-2)


def test_fp16_operands(dev):
    """fp16 hidden / weight (the reference kernel takes any float dtype, verl/utils/kernel/kernels.py:485-490): same
    tcgen05 kind::f16 path with the fp16 operand format; gradients come back in fp16 (kernels.py:1393-1394).
    Multi-chunk backward so that the fp32 d_hidden accumulation and the fp16 conversions of the last chunk run."""
    from verl_b200.utils.kernel import kernels
    from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy

    Tn, H, V, temp = 777, 384, 3000, 1.5
    g = torch.Generator(device="cpu").manual_seed(3)
    hidden = ((torch.rand(Tn, H, generator=g) - 0.5)).half().to(dev).requires_grad_(True)
    weight = ((torch.rand(V, H, generator=g) - 0.5)).half().to(dev).requires_grad_(True)
    labels = torch.randint(0, V, (Tn,), generator=g).to(dev)
    g_lp = (torch.rand(Tn, generator=g) * 2 - 1).to(dev)
    g_ent = (torch.rand(Tn, generator=g) - 0.5).to(dev)
    lp_r, ent_r, dh_r, dw_r = _gpu_fp32_reference(hidden.detach(), weight.detach(), labels, temp, g_lp, g_ent)
    kernels.set_backward_chunk(1024)
    try:
        lp, ent = linear_cross_entropy(hidden, weight, labels, temp, "none")
        (lp * g_lp + ent * g_ent).sum().backward()
    finally:
        kernels.set_backward_chunk(0)
    assert lp.dtype == torch.float32 and hidden.grad.dtype == torch.float16 and weight.grad.dtype == torch.float16
    torch.testing.assert_close(lp, lp_r, atol=LP_ATOL, rtol=0)
    torch.testing.assert_close(ent, ent_r, atol=LP_ATOL, rtol=0)
    torch.testing.assert_close(hidden.grad.float(), dh_r, atol=GRAD_ATOL, rtol=GRAD_RTOL)
    torch.testing.assert_close(weight.grad.float(), dw_r, atol=GRAD_ATOL, rtol=GRAD_RTOL)
    with pytest.raises((AssertionError, TypeError)):
        linear_cross_entropy(hidden.detach().bfloat16(), weight, labels, temp, "none")
    with pytest.raises((AssertionError, TypeError)):
        linear_cross_entropy(hidden.detach().double(), weight.detach().double(), labels, temp, "none")


def test_fp32_operands_tf32(dev):
    """fp32 hidden / weight: the reference's tl.dot on fp32 inputs runs TF32 (verl/utils/kernel/kernels.py:189-200);
    here the forward and the backward's recompute are tcgen05 kind::tf32 GEMMs on the fp32 operands, the gradient
    GEMMs read bf16 copies, gradients come back in fp32.  Tolerances: north_star's 2e-3 is stated for bf16 inputs
    (exactly representable operands); with fp32 inputs the 10-bit TF32 mantissa is the error source.  The TMA load
    rounds to TF32 (no truncation bias): measured 8e-5 against the fp32 reference fed TF32-rounded operands and 8e-4
    against the exact fp32 reference, both inside the 2e-3 bar."""
    from verl_b200.utils.kernel import kernels
    from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy

    Tn, H, V, temp = 600, 320, 2500, 1.5
    g = torch.Generator(device="cpu").manual_seed(4)
    hidden = (torch.rand(Tn, H, generator=g) - 0.5).to(dev).requires_grad_(True)
    weight = (torch.rand(V, H, generator=g) - 0.5).to(dev).requires_grad_(True)
    labels = torch.randint(0, V, (Tn,), generator=g).to(dev)
    g_lp = (torch.rand(Tn, generator=g) * 2 - 1).to(dev)
    g_ent = (torch.rand(Tn, generator=g) - 0.5).to(dev)
    lp_r, ent_r, dh_r, dw_r = _gpu_fp32_reference(hidden.detach(), weight.detach(), labels, temp, g_lp, g_ent)
    kernels.set_backward_chunk(1024)
    try:
        lp, ent = linear_cross_entropy(hidden, weight, labels, temp, "none")
        (lp * g_lp + ent * g_ent).sum().backward()
    finally:
        kernels.set_backward_chunk(0)
    assert hidden.grad.dtype == torch.float32 and weight.grad.dtype == torch.float32

    def tf32(x):  # round to nearest (ties away) at 10 mantissa bits, as the TMA TFLOAT32 load does
        return ((x.detach().view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)

    lp_t, ent_t, _, _ = _gpu_fp32_reference(tf32(hidden), tf32(weight), labels, temp)
    print("tf32 max |d lp| vs tf32-rounded ref", float((lp - lp_t).abs().max()), "vs exact", float((lp - lp_r).abs().max()))
    torch.testing.assert_close(lp, lp_t, atol=LP_ATOL, rtol=0)
    torch.testing.assert_close(ent, ent_t, atol=LP_ATOL, rtol=0)
    torch.testing.assert_close(lp, lp_r, atol=LP_ATOL, rtol=0)
    torch.testing.assert_close(ent, ent_r, atol=LP_ATOL, rtol=0)
    torch.testing.assert_close(hidden.grad, dh_r, atol=GRAD_ATOL, rtol=GRAD_RTOL)
    torch.testing.assert_close(weight.grad, dw_r, atol=GRAD_ATOL, rtol=GRAD_RTOL)


def test_fused_linear_for_ppo_torch_backend_contract(dev):
    """FusedLinearForPPO (verl/utils/experimental/torch_functional.py:201-216, the reference's default fused backend):
    [B,S,H] in -> [B,S] out in the hidden dtype, gradients in the operand dtypes; values against the fp32 reference at
    bf16 resolution (the outputs are rounded to bf16 exactly as the reference's new_zeros buffers are)."""
    from verl_b200.utils.experimental.torch_functional import FusedLinearForPPO

    B, S, H, V, temp = 3, 50, 256, 1500, 1.5
    g = torch.Generator().manual_seed(21)
    hidden = (torch.rand(B, S, H, generator=g) - 0.5).bfloat16().to(dev).requires_grad_(True)
    weight = (torch.rand(V, H, generator=g) - 0.5).bfloat16().to(dev).requires_grad_(True)
    ids = torch.randint(0, V, (B, S), generator=g).to(dev)
    g_lp = (torch.rand(B, S, generator=g) * 2 - 1).to(dev)
    g_ent = (torch.rand(B, S, generator=g) - 0.5).to(dev)
    lp, ent = FusedLinearForPPO()(hidden, weight, ids.int(), temperature=temp)
    assert lp.shape == (B, S) and ent.shape == (B, S) and lp.dtype == torch.bfloat16 and ent.dtype == torch.bfloat16
    (lp.float() * g_lp + ent.float() * g_ent).sum().backward()
    lp_r, ent_r, dh_r, dw_r = _gpu_fp32_reference(hidden.detach().view(-1, H), weight.detach(), ids.view(-1), temp,
                                                  g_lp.view(-1), g_ent.view(-1))
    torch.testing.assert_close(lp.float().view(-1), lp_r, atol=6e-2, rtol=8e-3)   # one bf16 ulp of a value <= 16
    torch.testing.assert_close(ent.float().view(-1), ent_r, atol=6e-2, rtol=8e-3)
    assert hidden.grad.shape == (B, S, H) and hidden.grad.dtype == torch.bfloat16
    torch.testing.assert_close(hidden.grad.float().view(-1, H), dh_r, atol=GRAD_ATOL, rtol=GRAD_RTOL)
    torch.testing.assert_close(weight.grad.float(), dw_r, atol=GRAD_ATOL, rtol=GRAD_RTOL)
    lp2, ent2 = FusedLinearForPPO()(hidden.detach().view(-1, H), weight.detach(), ids.view(-1))  # 2-D in -> flat out
    assert lp2.shape == (B * S,) and ent2.shape == (B * S,)


def test_hf_forward_shims(dev):
    """forward_with_{torch,triton}_backend (verl/models/transformers/dense_common.py:73-193) on a stand-in causal LM:
    `.log_probs` / `.entropy` of the rolled labels, the rmpad ([1,T]) and padded ([B,S]) call shapes of
    verl/workers/actor/dp_actor.py:163-178,254-256, and the backend switch of monkey_patch.py:148-192."""
    import types

    from verl_b200.models.transformers import dense_common as dc

    H, V = 128, 700

    class Body(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.emb = torch.nn.Embedding(V, H)

        def forward(self, input_ids=None, inputs_embeds=None, **kw):
            return torch.tanh(self.emb(input_ids) if inputs_embeds is None else inputs_embeds)

    class LM(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.config = types.SimpleNamespace(output_attentions=False, output_hidden_states=False, model_type="x")
            self.body = Body()
            self.lm_head = torch.nn.Linear(H, V, bias=False)

        def model(self, **kw):
            h = self.body(**kw)

            class Out(tuple):
                past_key_values = None
                hidden_states = None
                attentions = None

            return Out((h,))

    torch.manual_seed(0)
    lm = LM().to(dev).bfloat16()
    ids = torch.randint(0, V, (2, 24), device=dev)
    h = torch.tanh(lm.body.emb(ids)).float()
    logits = (h @ lm.lm_head.weight.float().t()) / 1.3
    rolled = torch.roll(ids, -1, -1)
    want_lp = torch.log_softmax(logits, -1).gather(-1, rolled[..., None]).squeeze(-1)
    dc.patch_forward_with_backends(lm, use_fused_kernels=True, fused_kernels_backend="triton")
    out = lm(input_ids=ids, temperature=1.3, return_dict=True)
    assert isinstance(out, dc.CausalLMOutputForPPO) and out.log_probs.shape == (2, 24) and out.log_probs.dtype == torch.float32
    torch.testing.assert_close(out.log_probs, want_lp, atol=LP_ATOL, rtol=0)
    out1 = lm(input_ids=ids.view(1, -1), temperature=1.3, return_dict=True)  # rmpad call: [1,T] -> flat [T]
    assert out1.log_probs.shape == (48,) and out1.entropy.shape == (48,)
    dc.patch_forward_with_backends(lm, use_fused_kernels=True, fused_kernels_backend="torch")
    out2 = lm(input_ids=ids, temperature=1.3, return_dict=True)
    assert out2.log_probs.shape == (2, 24) and out2.log_probs.dtype == torch.bfloat16
    torch.testing.assert_close(out2.log_probs.float(), want_lp, atol=6e-2, rtol=8e-3)
    with pytest.raises(NotImplementedError):
        lm(input_ids=ids, temperature=1.3, return_dict=False)
    with pytest.raises(RuntimeError):
        lm(inputs_embeds=h.bfloat16(), temperature=1.3, return_dict=True)  # neither labels nor input_ids


def test_gemm_all_majors(dev):
    from verl_b200 import ops

    torch.manual_seed(0)
    # N = 904 / 1496 / 1024 take the 256 x 512 tiles of CTA pairs (ragged last tile across both column groups)
    for (M, N, K) in [(128, 256, 64), (304, 520, 192), (1000, 1024, 1024), (304, 904, 192), (520, 1496, 320)]:
        A = (torch.rand(M, K, device=dev) - 0.5).bfloat16()
        B = (torch.rand(N, K, device=dev) - 0.5).bfloat16()
        ref = A.float() @ B.float().t()
        for a_k in (True, False):
            for b_k in (True, False):
                Ain = A if a_k else A.t().contiguous()
                Bin = B if b_k else B.t().contiguous()
                C = ops.gemm_bf16(Ain, Bin, a_k, b_k, out_dtype=torch.float32)
                torch.testing.assert_close(C, ref, atol=1e-3, rtol=1e-4)
                Cb = ops.gemm_bf16(Ain, Bin, a_k, b_k, out_dtype=torch.bfloat16)
                torch.testing.assert_close(Cb.float(), ref, atol=5e-2, rtol=2e-2)
