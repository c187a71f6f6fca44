"""The c3 parity probe of bench.py (`actor_parity_probe`) is itself a checker: this test feeds it an INDEPENDENT plain-
PyTorch implementation of one actor micro-batch (reference flow: lm-head over the rmpad stream, labels = roll(ids, -1),
pad back to [B,S], slice [:, -L-1:-1], clip + low_var_kl token-mean loss, autograd) in place of `fused_actor_step` and
expects agreement — and expects it to FAIL when that implementation is perturbed (wrong label shift, gradient leaking
outside the response window), so a green probe on the GPU means something."""

import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import bench  # noqa: E402


def _micro_batch(mb=4, P=8, L=8, H=32, V=64):
    torch.manual_seed(0)
    S = P + L
    p_len = torch.randint(P // 2, P + 1, (mb,))
    r_len = torch.randint(L // 4, L + 1, (mb,))
    pos = torch.arange(S)[None]
    attn = ((pos >= (P - p_len)[:, None]) & (pos < (P + r_len)[:, None])).long()
    Tm = int(attn.sum())
    m = dict(hidden=(torch.rand(Tm, H) - 0.5).bfloat16(), ids=torch.randint(0, V, (mb, S)), attn=attn,
             rmask=attn[:, P:].contiguous(), old=-torch.rand(mb, L) * 3, adv=torch.randn(mb, L))
    m["ref"] = m["old"] + 0.05 * torch.randn(mb, L)
    weight = (torch.rand(V, H) - 0.5).bfloat16()
    return m, weight, dict(loss=bench.LOSS_C3, P=P, L=L)


def _plain_step(label_shift=-1, leak=0.0):
    def step(h, w, ids, attn, old, adv, rmask, temperature, ref_log_prob, loss_scale, clip_ratio, entropy_coeff,
             use_kl_loss, kl_loss_coef, kl_loss_type, loss_agg_mode):
        L = rmask.shape[-1]
        idx = attn.reshape(-1).nonzero().squeeze(-1)
        lab = torch.roll(ids.reshape(-1)[idx], label_shift)
        z = (h.float() @ w.float().t()) / temperature
        lp = z.gather(-1, lab[:, None]).squeeze(-1) - torch.logsumexp(z, -1)
        lpm = torch.zeros(attn.numel()).index_put((idx,), lp).view(attn.shape)[:, -L - 1:-1]
        mask = rmask.float()
        mm = lambda x: (x * mask).sum() / (mask.sum() + 1e-8)  # noqa: E731
        ratio = (lpm - old).clamp(-20, 20).exp()
        l1, l2 = -adv * ratio, -adv * ratio.clamp(1 - clip_ratio, 1 + clip_ratio)
        c1 = torch.maximum(l1, l2)
        pg = torch.where(adv < 0, torch.minimum(-adv * 3.0, c1), c1)
        k = (ref_log_prob - lpm).clamp(-20, 20)
        kl = (k.exp() - k - 1).clamp(-10, 10)
        loss = (mm(pg) + kl_loss_coef * mm(kl)) * loss_scale + leak * h.float().sum()
        loss.backward()
        met = torch.tensor([loss.item(), mm(pg).item(), 0, 0, 0, 0, mm(kl).item(), mask.sum().item()])
        return loss.detach(), met, lpm.detach(), None
    return step


def test_probe_accepts_an_independent_implementation():
    m, weight, cfg = _micro_batch()
    out = bench.actor_parity_probe(torch, _plain_step(), m, weight, cfg, 0.125, torch.device("cpu"), n_rows=40)
    assert out["ok"], out
    assert out["window_rows"] < out["stream_rows"] and out["dh_outside_window_max_abs"] == 0.0


@pytest.mark.parametrize("kw", [dict(label_shift=+1), dict(leak=1e-3)])
def test_probe_rejects_a_wrong_implementation(kw):
    m, weight, cfg = _micro_batch()
    out = bench.actor_parity_probe(torch, _plain_step(**kw), m, weight, cfg, 0.125, torch.device("cpu"), n_rows=40)
    assert not out["ok"], out


def _plain_lce(scale_bug=1.0, drop_entropy_grad=False):
    def lce(h, w, labels, temperature, reduction, group):
        z = (h.float() @ w.float().t()) / temperature * scale_bug
        lse = torch.logsumexp(z, -1)
        p = torch.softmax(z, -1)
        lp = z.gather(-1, labels[:, None]).squeeze(-1) - lse
        ent = lse - (p * z).sum(-1)
        return lp, (ent.detach() if drop_entropy_grad else ent)
    return lce


def _lce_inputs(T=384, H=64, V=512):
    torch.manual_seed(1)
    return ((torch.rand(T, H) - 0.5).bfloat16(), (torch.rand(V, H) - 0.5).bfloat16(), torch.randint(0, V, (T,)))


def test_lce_probe_accepts_an_independent_implementation():
    """the lm-head probe (the `parity` key of every bench line) against plain PyTorch autograd, world size 1"""
    h, w, lab = _lce_inputs()

    class _Lce:  # d_hidden / d_weight come back in the operand dtype, like the op
        def __call__(self, hh, ww, labels, t, red, group):
            return _plain_lce()(hh, ww, labels, t, red, group)

    out = bench.parity_probe(torch, None, _Lce(), h, w, lab, None, 0, 1, torch.device("cpu"))
    assert out["ok"], out


@pytest.mark.parametrize("kw", [dict(scale_bug=1.05), dict(drop_entropy_grad=True)])
def test_lce_probe_rejects_a_wrong_implementation(kw):
    h, w, lab = _lce_inputs()
    if kw.get("drop_entropy_grad"):
        # torch.autograd.backward([lp, ent], ...) needs both outputs to require grad: detach only the value path
        def lce(hh, ww, labels, t, red, group):
            lp, ent = _plain_lce()(hh, ww, labels, t, red, group)
            return lp, ent.detach() + 0.0 * lp
    else:
        lce = _plain_lce(**kw)
    out = bench.parity_probe(torch, None, lce, h, w, lab, None, 0, 1, torch.device("cpu"))
    assert not out["ok"], out
