"""fp32 PyTorch reference of the lm-head BACKWARD on a SAMPLE of rows, usable at full size (test helper).

d_hidden[t, :] for sampled tokens needs that token's whole logits row (autograd over ``hidden[rows] @ W^T``).
d_weight[v, :] for sampled vocab rows needs column v of dlogits for ALL tokens:
    dlogit[t, v] = ( g_lp[t] * (1[v == y_t] - p) - g_ent[t] * p * (z - eb[t]) ) / temp,   z = x[t,v]/temp, p = exp(z - lse[t])
(verl/utils/kernel/kernels.py:795-802, verl/utils/experimental/torch_functional.py:52-67) — the per-token scalars
``lse`` / ``eb`` come from the caller (the kernel's own, cross-checked on the sampled token rows by the caller), so the
cost is one [T, n_v] fp32 GEMM instead of the [T, V] logits."""

import torch


def sampled_rows_reference(hidden, weight, labels, temperature, g_lp, g_ent, tok_rows):
    """-> (lp, ent, lse, eb, d_hidden) fp32 for the sampled token rows, full vocabulary (weight = the FULL weight)."""
    h = hidden[tok_rows].float().requires_grad_(True)
    w = weight.float()
    z = (h @ w.t()) / temperature
    lse = torch.logsumexp(z, -1)
    p = torch.softmax(z, -1)
    eb = (p * z).sum(-1)
    lp = z.gather(-1, labels[tok_rows, None]).squeeze(-1) - lse
    ent = lse - eb
    gl = g_lp[tok_rows] if g_lp is not None else torch.zeros_like(lp)
    ge = g_ent[tok_rows] if g_ent is not None else torch.zeros_like(lp)
    (lp * gl + ent * ge).sum().backward()
    return lp.detach(), ent.detach(), lse.detach(), eb.detach(), h.grad


def sampled_d_weight_reference(hidden, w_rows, row_ids, labels, lse, eb, temperature, g_lp, g_ent, chunk=16384):
    """-> d_weight rows fp32 [n_v, H] for the vocab rows ``row_ids`` (GLOBAL vocabulary ids; w_rows = weight[row_ids])."""
    T, H = hidden.shape
    wv = w_rows.float()
    out = torch.zeros(wv.shape[0], H, dtype=torch.float32, device=hidden.device)
    zeros = torch.zeros(T, dtype=torch.float32, device=hidden.device)
    gl = g_lp if g_lp is not None else zeros
    ge = g_ent if g_ent is not None else zeros
    for s in range(0, T, chunk):
        e = min(T, s + chunk)
        h = hidden[s:e].float()
        z = (h @ wv.t()) / temperature
        p = torch.exp(z - lse[s:e, None])
        onehot = (labels[s:e, None] == row_ids[None, :]).float()
        d = (gl[s:e, None] * (onehot - p) - ge[s:e, None] * p * (z - eb[s:e, None])) / temperature
        out += d.t() @ h
    return out


def excess(got, ref, atol, rtol):
    """max over elements of |got - ref| - (atol + rtol*|ref|)  (<= 0 means torch.testing.assert_close would pass)"""
    return float(((got.float() - ref).abs() - (atol + rtol * ref.abs())).max())
