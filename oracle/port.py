"""TEST INFRASTRUCTURE ONLY — CPU restatement (PyTorch-CPU / numpy) of the reference algorithms on the hot path.

Each function cites the reference lines it restates (paths relative to /root/reference).  The restatement is
validated against the unmodified reference (oracle/make_golden.py -> tests/golden/*.npz, tests/test_oracle_cpu.py).
Parity status: pinned (see oracle/__init__.py).  Never imported by verl_b200.
"""

from __future__ import annotations

from collections import OrderedDict

import numpy as np
import torch
import torch.nn.functional as F


# ------------------------------------------------------------------------------------------------ lm-head
def lce_forward(hidden, weight, labels, temperature=1.0, chunk=512, compute_dtype=torch.float32):
    """Per-token log-prob and entropy of softmax(hidden @ weight.T / temperature).

    Restates verl/utils/experimental/torch_functional.py:20-37 (chunked, :109-117) and the fp32 recipe of
    tests/utils/test_linear_cross_entropy.py:48-61: logits in fp32, log_softmax gather, entropy = lse - sum p*logit.
    compute_dtype=float32 upcasts bf16 inputs first (the kernel tests' reference); bf16 keeps the matmul in bf16 like
    the reference's production torch backend.
    """
    T = hidden.shape[0]
    w = weight.to(compute_dtype)
    lp = torch.empty(T, dtype=torch.float32)
    ent = torch.empty(T, dtype=torch.float32)
    for s in range(0, T, chunk):
        e = min(s + chunk, T)
        logits = (hidden[s:e].to(compute_dtype) @ w.t()) / temperature
        logits = logits.float()
        lse = torch.logsumexp(logits, dim=-1)
        p = torch.softmax(logits, dim=-1)
        lp[s:e] = logits.gather(-1, labels[s:e, None]).squeeze(-1) - lse
        ent[s:e] = lse - (p * logits).sum(-1)
    return lp, ent


def lce_backward(hidden, weight, labels, g_lp, g_ent, temperature=1.0, chunk=512):
    """d_hidden, d_weight (fp32) of lce_forward for upstream grads g_lp, g_ent (either may be None).

    Restates verl/utils/experimental/torch_functional.py:40-72:
      dlogits = g_lp*(onehot - p) + p*(log p + entropy)*(-g_ent), / temperature; dH = dlogits@W; dW = dlogits.T@H
    (same expression as the Triton kernel's, verl/utils/kernel/kernels.py:1327-1334).
    """
    T = hidden.shape[0]
    h32, w32 = hidden.float(), weight.float()
    dh = torch.zeros_like(h32)
    dw = torch.zeros_like(w32)
    for s in range(0, T, chunk):
        e = min(s + chunk, T)
        logits = (h32[s:e] @ w32.t()) / temperature
        p = torch.softmax(logits, dim=-1)
        d = torch.zeros_like(logits)
        if g_lp is not None:
            onehot = torch.zeros_like(logits).scatter_(-1, labels[s:e, None], 1.0)
            d += g_lp[s:e, None].float() * (onehot - p)
        if g_ent is not None:
            logp = torch.log_softmax(logits, dim=-1)
            entropy = torch.logsumexp(logits, dim=-1) - (p * logits).sum(-1)
            d += p * (logp + entropy[:, None]) * (-g_ent[s:e, None].float())
        d = d / temperature
        dh[s:e] = d @ w32
        dw += d.t() @ h32[s:e]
    return dh, dw


def lce_forward_vocab_parallel(hidden, weight_shards, labels, temperature=1.0):
    """Vocab-parallel combine, as verl/utils/kernel/kernels.py:609-661 does with 3 all-reduces: every shard r gives
    {m_r, s_r = sum exp(l - m_r), b_r = sum l*exp(l - m_r), t_r = l[label] if the label is local}; merge with the
    log-sum-exp rule.  Equal shards, label offset rank*V_local (kernels.py:232)."""
    ms, ss, bs, ts = [], [], [], []
    for r, w in enumerate(weight_shards):
        logits = (hidden.float() @ w.float().t()) / temperature
        m = logits.max(-1).values
        ex = torch.exp(logits - m[:, None])
        ms.append(m), ss.append(ex.sum(-1)), bs.append((logits * ex).sum(-1))
        local = labels - r * w.shape[0]
        ok = (local >= 0) & (local < w.shape[0])
        t = torch.zeros_like(m)
        t[ok] = logits[ok, local[ok]]
        ts.append(t)
    M = torch.stack(ms).max(0).values
    s = sum(si * torch.exp(mi - M) for si, mi in zip(ss, ms))
    b = sum(bi * torch.exp(mi - M) for bi, mi in zip(bs, ms))
    lse = M + torch.log(s)
    return sum(ts) - lse, lse - b / s


def logprobs_from_logits(logits, labels):
    """verl/utils/torch_functional.py:116-133 (fp32 branch: gather - logsumexp; bf16 branch: log_softmax then gather).
    Output in logits' dtype."""
    if logits.dtype in (torch.float32, torch.float64):
        picked = torch.gather(logits, -1, labels.unsqueeze(-1)).squeeze(-1)
        return picked - torch.logsumexp(logits, dim=-1)
    return F.log_softmax(logits, dim=-1).gather(-1, labels.unsqueeze(-1)).squeeze(-1)


def entropy_from_logits(logits):
    """verl/utils/torch_functional.py:145-149"""
    p = torch.softmax(logits, dim=-1)
    return torch.logsumexp(logits, dim=-1) - (p * logits).sum(-1)


# ------------------------------------------------------------------------------------------------ masked ops
def masked_sum(values, mask, axis=None):
    """verl/utils/torch_functional.py:163-168"""
    kept = torch.where(mask.bool(), values, torch.zeros((), dtype=values.dtype))
    return (kept * mask).sum(axis)


def masked_mean(values, mask, axis=None):
    """verl/utils/torch_functional.py:171-185"""
    return masked_sum(values, mask, axis) / (mask.sum(axis) + 1e-8)


def masked_var(values, mask, unbiased=True):
    """verl/utils/torch_functional.py:188-203"""
    mu = masked_mean(values, mask)
    var = masked_mean((values - mu) ** 2, mask)
    if unbiased:
        n = mask.sum()
        if n == 0:
            raise ValueError("At least one element in the mask has to be 1.")
        if n == 1:
            raise ValueError("The sum of the mask is one, which can cause a division by zero.")
        var = var * (n / (n - 1))
    return var


def masked_whiten(values, mask, shift_mean=True):
    """verl/utils/torch_functional.py:206-223"""
    mu, var = masked_mean(values, mask), masked_var(values, mask)
    out = (values - mu) * torch.rsqrt(var + 1e-8)
    return out if shift_mean else out + mu


# ------------------------------------------------------------------------------------------------ rmpad
def unpad_indices(attention_mask):
    """flash_attn.bert_padding.unpad_input's index part (third-party, v2.8.3; call sites
    verl/workers/actor/dp_actor.py:110-125) == verl/utils/torch_functional.py:629-638 (get_unpad_data):
    -> indices int64 [nnz] ascending, seqlens int32 [B], cu_seqlens int32 [B+1], max_seqlen int."""
    m = np.asarray(attention_mask.cpu())
    seqlens = m.sum(axis=-1).astype(np.int32)
    indices = np.flatnonzero(m.reshape(-1)).astype(np.int64)
    cu = np.zeros(m.shape[0] + 1, dtype=np.int32)
    cu[1:] = np.cumsum(seqlens, dtype=np.int32)
    return (torch.from_numpy(indices), torch.from_numpy(seqlens), torch.from_numpy(cu),
            int(seqlens.max()) if seqlens.size else 0)


def gather_rows(x, indices):
    """index_first_axis (flash_attn.bert_padding)"""
    return x[indices]


def pad_rows(x, indices, n_rows):
    """pad_input / index_put_first_axis (flash_attn.bert_padding): zeros everywhere else"""
    out = torch.zeros((n_rows, *x.shape[1:]), dtype=x.dtype)
    out[indices] = x
    return out


# ------------------------------------------------------------------------------------------------ advantages
def group_ids(index):
    """First-seen integer id per uid — the grouping verl/trainer/ppo/core_algos.py:286-291 performs with a dict."""
    table = OrderedDict()
    out = np.empty(len(index), dtype=np.int32)
    for i, k in enumerate(index):
        k = k.item() if isinstance(k, np.generic) else k
        out[i] = table.setdefault(k, len(table))
    return out, len(table)


def grpo_outcome_advantage(token_level_rewards, response_mask, index, epsilon=1e-6, norm_adv_by_std_in_grpo=True):
    """verl/trainer/ppo/core_algos.py:245-308: score = row sum; group mean and UNBIASED std (torch.std); a group
    of one uses mean 0 / std 1; adv = (score-mean)/(std+eps) or score-mean; broadcast * mask."""
    scores = token_level_rewards.sum(dim=-1)
    gid, n = group_ids(index)
    out = scores.clone()
    for g in range(n):
        rows = np.flatnonzero(gid == g)
        vals = scores[rows]
        if len(rows) == 1:
            mean, std = torch.tensor(0.0), torch.tensor(1.0)
        else:
            mean, std = vals.mean(), vals.std()  # unbiased
        out[rows] = (vals - mean) / (std + epsilon) if norm_adv_by_std_in_grpo else vals - mean
    return out.unsqueeze(-1) * response_mask


def _group_rows(index):
    gid, n = group_ids(index)
    return [np.flatnonzero(gid == g) for g in range(n)]


def rloo_outcome_advantage(token_level_rewards, response_mask, index):
    """verl/trainer/ppo/core_algos.py:425-476: leave-one-out baseline, written as score*n/(n-1) - mean*n/(n-1)."""
    scores = token_level_rewards.sum(dim=-1)
    out = scores.clone()
    for rows in _group_rows(index):
        n = len(rows)
        if n > 1:
            vals = scores[rows]
            out[rows] = vals * n / (n - 1) - vals.mean() * n / (n - 1)
    return out.unsqueeze(-1) * response_mask


def opo_outcome_advantage(token_level_rewards, response_mask, index):
    """verl/trainer/ppo/core_algos.py:479-530: baseline = sum(len*score)/sum(len) of the uid group (0 for singletons)."""
    scores = token_level_rewards.sum(dim=-1)
    lens = response_mask.sum(dim=-1)
    out = scores.clone()
    for rows in _group_rows(index):
        if len(rows) > 1:
            ln = lens[rows]
            out[rows] = scores[rows] - (ln * scores[rows]).sum() / ln.sum()
    return out.unsqueeze(-1) * response_mask


def gpg_outcome_advantage(token_level_rewards, response_mask, index, f_norm=1.0):
    """verl/trainer/ppo/core_algos.py:608-667: alpha*(score - group mean)/f_norm, alpha = B / max(#nonzero scores, 1)."""
    scores = token_level_rewards.sum(dim=-1)
    alpha = scores.shape[0] / torch.count_nonzero(scores).clamp(min=1)
    out = scores.clone()
    for rows in _group_rows(index):
        mean = scores[rows].mean() if len(rows) > 1 else torch.tensor(0.0)
        out[rows] = alpha * (scores[rows] - mean) / f_norm
    return out.unsqueeze(-1) * response_mask


def reinforce_plus_plus_baseline_outcome_advantage(token_level_rewards, response_mask, index):
    """verl/trainer/ppo/core_algos.py:372-422: score - group mean, tiled * mask, masked_whiten, * mask."""
    L = token_level_rewards.shape[-1]
    scores = token_level_rewards.sum(dim=-1)
    out = scores.clone()
    for rows in _group_rows(index):
        if len(rows) > 1:
            out[rows] = scores[rows] - scores[rows].mean()
    tiled = out.unsqueeze(-1).tile([1, L]) * response_mask
    return masked_whiten(tiled, response_mask) * response_mask


def grpo_passk_outcome_advantage(token_level_rewards, response_mask, index, epsilon=1e-6, norm_adv_by_std_in_grpo=True):
    """verl/trainer/ppo/core_algos.py:309-369: the best response of each group gets r_max - r_second_max
    (/ (std + eps) when normalised), every other response 0; a group of fewer than 2 raises ValueError."""
    scores = token_level_rewards.sum(dim=-1)
    out = torch.zeros_like(scores)
    for rows in _group_rows(index):
        if len(rows) < 2:
            raise ValueError("Pass@k requires at least 2 samples per group.")
        vals = scores[rows]
        order = torch.argsort(vals, descending=True, stable=True)
        a = vals[order[0]] - vals[order[1]]
        if norm_adv_by_std_in_grpo:
            a = a / (vals.std() + epsilon)
        out[rows[order[0]]] = a
    return out.unsqueeze(-1) * response_mask


def reinforce_plus_plus_outcome_advantage(token_level_rewards, response_mask, gamma):
    """verl/trainer/ppo/core_algos.py:533-569: R_t = r_t + gamma * run; run = R_t * m_t (reset after EOS);
    advantages = masked_whiten(returns) * mask."""
    B, L = token_level_rewards.shape
    returns = torch.zeros_like(token_level_rewards)
    run = torch.zeros(B)
    for t in range(L - 1, -1, -1):
        run = token_level_rewards[:, t] + gamma * run
        returns[:, t] = run
        run = run * response_mask[:, t]
    return masked_whiten(returns, response_mask) * response_mask, returns


def remax_outcome_advantage(token_level_rewards, reward_baselines, response_mask):
    """verl/trainer/ppo/core_algos.py:572-605: suffix sums of the masked rewards minus the per-row baseline."""
    masked = (token_level_rewards * response_mask).numpy()
    returns = torch.from_numpy(np.cumsum(masked[:, ::-1], axis=-1, dtype=np.float32)[:, ::-1].copy())
    return returns - reward_baselines.unsqueeze(-1) * response_mask, returns


def gae_advantage_return(token_level_rewards, values, response_mask, gamma, lam):
    """verl/trainer/ppo/core_algos.py:193-241: reverse recurrence with masked carry-through, returns = adv + values,
    then masked_whiten(adv)."""
    B, L = token_level_rewards.shape
    next_v = torch.zeros(B)
    carry = torch.zeros(B)
    adv = torch.empty(B, L)
    for t in range(L - 1, -1, -1):
        m = response_mask[:, t]
        delta = token_level_rewards[:, t] + gamma * next_v - values[:, t]
        cand = delta + gamma * lam * carry
        next_v = values[:, t] * m + (1 - m) * next_v
        carry = cand * m + (1 - m) * carry
        adv[:, t] = carry
    returns = adv + values
    return masked_whiten(adv, response_mask), returns


# ------------------------------------------------------------------------------------------------ losses
def agg_loss(loss_mat, loss_mask, loss_agg_mode):
    """verl/trainer/ppo/core_algos.py:686-719"""
    if loss_agg_mode == "token-mean":
        return masked_mean(loss_mat, loss_mask)
    per_seq = (loss_mat * loss_mask).sum(-1)
    if loss_agg_mode == "seq-mean-token-sum":
        return per_seq.mean()
    if loss_agg_mode == "seq-mean-token-mean":
        return (per_seq / loss_mask.sum(-1)).mean()
    if loss_agg_mode == "seq-mean-token-sum-norm":
        return per_seq.sum() / loss_mask.shape[-1]
    raise ValueError(f"Invalid loss_agg_mode: {loss_agg_mode}")


def policy_loss(old_log_prob, log_prob, advantages, response_mask, cliprange=None, cliprange_low=None,
                cliprange_high=None, clip_ratio_c=3.0, loss_agg_mode="token-mean"):
    """verl/trainer/ppo/core_algos.py:722-794 -> (pg_loss, pg_clipfrac, ppo_kl, pg_clipfrac_lower)"""
    assert clip_ratio_c > 1.0
    lo = cliprange if cliprange_low is None else cliprange_low
    hi = cliprange if cliprange_high is None else cliprange_high
    log_ratio = torch.clamp(log_prob - old_log_prob, min=-20.0, max=20.0)
    ratio = torch.exp(log_ratio)
    ppo_kl = masked_mean(-log_ratio, response_mask)
    unclipped = -advantages * ratio
    clipped = -advantages * torch.clamp(ratio, 1 - lo, 1 + hi)
    worst = torch.maximum(unclipped, clipped)
    pg_clipfrac = masked_mean(torch.gt(clipped, unclipped).float(), response_mask)
    dual = -advantages * clip_ratio_c
    capped = torch.min(dual, worst)
    pg_clipfrac_lower = masked_mean(torch.gt(worst, dual) * (advantages < 0).float(), response_mask)
    per_token = torch.where(advantages < 0, capped, worst)
    return agg_loss(per_token, response_mask, loss_agg_mode), pg_clipfrac, ppo_kl, pg_clipfrac_lower


def policy_loss_clip_cov(old_log_prob, log_prob, advantages, response_mask, loss_agg_mode="token-mean", clip_ratio=0.2,
                         clip_ratio_low=None, clip_ratio_high=None, clip_cov_ratio=0.0002, clip_cov_lb=1.0,
                         clip_cov_ub=5.0):
    """verl/trainer/ppo/core_algos.py:818-905 -> (pg_loss, pg_clipfrac, ppo_kl).  The random subset is drawn with
    torch.randperm on the global host generator at the same point as the reference, so seeding reproduces it."""
    lo = clip_ratio if clip_ratio_low is None else clip_ratio_low
    hi = clip_ratio if clip_ratio_high is None else clip_ratio_high
    delta = log_prob - old_log_prob
    ratio = torch.exp(delta)
    ppo_kl = masked_mean(-delta, response_mask)
    unclipped = -advantages * ratio
    clipped = -advantages * torch.clamp(ratio, 1 - lo, 1 + hi)
    valid = response_mask > 0
    cut_by_ppo = (clipped > unclipped) & valid
    cov = (advantages - masked_mean(advantages, response_mask)) * (
        log_prob.detach() - masked_mean(log_prob.detach(), response_mask))
    cov = torch.where((response_mask == 0) | cut_by_ppo, torch.full_like(cov, -torch.inf), cov)
    n_drop = max(int(clip_cov_ratio * response_mask.sum().item()), 1)
    eligible = torch.nonzero(((cov < clip_cov_ub) & (cov > clip_cov_lb) & valid).reshape(-1)).squeeze(-1)
    keep = torch.ones(cov.numel())
    if eligible.numel() > 0:
        order = torch.randperm(eligible.numel())
        keep[eligible[order[: min(n_drop, eligible.numel())]]] = 0
    keep = keep.view_as(cov)
    pg_clipfrac = masked_mean((keep == 0).float(), response_mask)
    per_token = torch.maximum(unclipped, clipped) * keep
    return agg_loss(per_token, response_mask, loss_agg_mode), pg_clipfrac, ppo_kl


def policy_loss_kl_cov(old_log_prob, log_prob, advantages, response_mask, loss_agg_mode="token-mean",
                       kl_cov_ratio=0.0002, ppo_kl_coef=1.0):
    """verl/trainer/ppo/core_algos.py:908-972 -> (pg_loss, ppo_kl_abs)"""
    delta = log_prob - old_log_prob
    ppo_kl_abs = masked_mean(delta.abs(), response_mask)
    plain = -advantages * torch.exp(delta)
    with_kl = plain + ppo_kl_coef * delta.abs()
    valid = (response_mask > 0).reshape(-1)
    pos = torch.nonzero(valid).squeeze(-1)
    a = advantages.detach().reshape(-1)[pos]
    l = log_prob.detach().reshape(-1)[pos]
    pick = torch.zeros(valid.numel(), dtype=torch.bool)
    if pos.numel() > 0:
        cov = (a - a.mean()) * (l - l.mean())
        k = max(1, int(pos.numel() * kl_cov_ratio))
        pick[pos[torch.topk(cov, k, largest=True).indices]] = True
    per_token = torch.where(pick.view_as(plain), with_kl, plain)
    return agg_loss(per_token, response_mask, loss_agg_mode), ppo_kl_abs


def policy_loss_gpg(log_prob, advantages, response_mask, loss_agg_mode="token-mean"):
    """verl/trainer/ppo/core_algos.py:797-815"""
    return agg_loss(-log_prob * advantages, response_mask, loss_agg_mode)


def value_loss(vpreds, returns, values, response_mask, cliprange_value, loss_agg_mode="token-mean"):
    """verl/trainer/ppo/core_algos.py:992-1031 (clip_by_value: verl/utils/torch_functional.py:136-142)"""
    clipped = torch.max(torch.min(vpreds, values + cliprange_value), values - cliprange_value)
    plain_err = (vpreds - returns) ** 2
    clipped_err = (clipped - returns) ** 2
    vf_loss = 0.5 * agg_loss(torch.max(plain_err, clipped_err), response_mask, loss_agg_mode)
    return vf_loss, masked_mean(torch.gt(clipped_err, plain_err).float(), response_mask)


def kl_penalty(logprob, ref_logprob, kind):
    """verl/trainer/ppo/core_algos.py:1034-1069"""
    if kind in ("kl", "k1"):
        return logprob - ref_logprob
    if kind == "abs":
        return (logprob - ref_logprob).abs()
    if kind in ("mse", "k2"):
        return 0.5 * (logprob - ref_logprob).square()
    if kind in ("low_var_kl", "k3"):
        k = torch.clamp(ref_logprob - logprob, min=-20, max=20)
        return torch.clamp(torch.exp(k) - k - 1, min=-10, max=10)
    raise NotImplementedError


def actor_loss(old_log_prob, log_prob, advantages, response_mask, *, cliprange_low, cliprange_high, clip_ratio_c=3.0,
               loss_agg_mode="token-mean", entropy=None, entropy_coeff=0.0, ref_log_prob=None, kl_loss_coef=0.0,
               kl_loss_type="low_var_kl", loss_scale=1.0):
    """Loss composition of verl/workers/actor/dp_actor.py:422-469 -> (loss, dict of the logged scalars)."""
    pg, clipfrac, ppo_kl, clipfrac_lower = policy_loss(
        old_log_prob, log_prob, advantages, response_mask, cliprange_low=cliprange_low, cliprange_high=cliprange_high,
        clip_ratio_c=clip_ratio_c, loss_agg_mode=loss_agg_mode)
    total = pg
    ent_loss = torch.zeros(())
    kl_loss = torch.zeros(())
    if entropy is not None and entropy_coeff != 0:
        ent_loss = agg_loss(entropy, response_mask, loss_agg_mode)
        total = pg - ent_loss * entropy_coeff
    if ref_log_prob is not None:
        kl_loss = agg_loss(kl_penalty(log_prob, ref_log_prob, kl_loss_type), response_mask, loss_agg_mode)
        total = total + kl_loss * kl_loss_coef
    return total * loss_scale, dict(pg_loss=pg, pg_clipfrac=clipfrac, ppo_kl=ppo_kl, pg_clipfrac_lower=clipfrac_lower,
                                    entropy_loss=ent_loss, kl_loss=kl_loss)


def actor_micro_batch(hidden_rmpad, weight, input_ids, attention_mask, old_log_probs, advantages, response_mask, *,
                      temperature, ref_log_prob=None, cliprange_low=0.2, cliprange_high=0.2, clip_ratio_c=3.0,
                      entropy_coeff=0.0, kl_loss_coef=0.0, kl_loss_type="low_var_kl", loss_agg_mode="token-mean",
                      loss_scale=1.0):
    """One micro-batch of the reference actor update from the transformer's rmpad hidden states, in the reference's
    order (verl/workers/actor/dp_actor.py): lm-head over ALL stream tokens with labels = roll(ids_rmpad, -1) (:135,
    dense_common.py:169-185) -> pad_input to [B,S] (:221-232) -> slice [:, -L-1:-1] (:236-237) -> loss composition
    (:422-469) -> backward.  -> dict(loss, scalars, log_prob, entropy, d_hidden, d_weight)."""
    B, S = input_ids.shape
    L = response_mask.shape[-1]
    idx, _, _, _ = unpad_indices(attention_mask)
    ids_rmpad = input_ids.reshape(-1)[idx]
    labels = torch.roll(ids_rmpad, shifts=-1, dims=0)
    lp_all, ent_all = lce_forward(hidden_rmpad, weight, labels, temperature)
    lp_leaf = lp_all.clone().requires_grad_(True)
    ent_leaf = ent_all.clone().requires_grad_(True)

    def window(x):
        full = torch.zeros(B * S, dtype=x.dtype).index_put((idx,), x)
        return full.view(B, S)[:, -L - 1:-1]

    log_prob, entropy = window(lp_leaf), window(ent_leaf)
    loss, scalars = actor_loss(old_log_probs, log_prob, advantages, response_mask, cliprange_low=cliprange_low,
                               cliprange_high=cliprange_high, clip_ratio_c=clip_ratio_c, loss_agg_mode=loss_agg_mode,
                               entropy=entropy if entropy_coeff != 0 else None, entropy_coeff=entropy_coeff,
                               ref_log_prob=ref_log_prob, kl_loss_coef=kl_loss_coef, kl_loss_type=kl_loss_type,
                               loss_scale=loss_scale)
    loss.backward()
    g_ent = ent_leaf.grad if ent_leaf.grad is not None else None
    dh, dw = lce_backward(hidden_rmpad, weight, labels, lp_leaf.grad, g_ent, temperature)
    return dict(loss=loss.detach(), scalars=scalars, log_prob=log_prob.detach(), entropy=entropy.detach(),
                d_hidden=dh, d_weight=dw)


# ------------------------------------------------------------------------------------------------ driver-side ops
def kl_reward(token_level_scores, old_log_probs, ref_log_prob, response_mask, kind, beta):
    """verl/trainer/ppo/ray_trainer.py:153-193 (apply_kl_penalty without the DataProto / controller plumbing):
    -> (token_level_rewards, current_kl as a python float)"""
    kld = kl_penalty(old_log_probs, ref_log_prob, kind) * response_mask
    rewards = token_level_scores - beta * kld
    per_seq = masked_mean(kld, response_mask, axis=-1)
    return rewards, torch.mean(per_seq, dim=0).item()


def data_metrics(token_level_scores, token_level_rewards, advantages, returns, values, response_mask, attention_mask):
    """verl/trainer/ppo/metric_utils.py:80-180 -> dict with the reference's keys (values=None: use_critic=False)."""
    L = response_mask.shape[-1]
    seq_score, seq_reward = token_level_scores.sum(-1), token_level_rewards.sum(-1)
    keep = response_mask.bool()
    prompt_len = attention_mask[:, :-L].sum(-1).float()
    resp_len = attention_mask[:, -L:].sum(-1).float()
    out = {}

    def stats(name, x):
        out[f"{name}/mean"], out[f"{name}/max"], out[f"{name}/min"] = x.mean().item(), x.max().item(), x.min().item()

    stats("critic/score", seq_score)
    stats("critic/rewards", seq_reward)
    stats("critic/advantages", advantages[keep])
    stats("critic/returns", returns[keep])
    if values is not None:
        stats("critic/values", values[keep])
        out["critic/vf_explained_var"] = (1.0 - torch.var(returns[keep] - values[keep]) /
                                          (torch.var(returns[keep]) + 1e-5)).item()
    stats("response_length", resp_len)
    out["response_length/clip_ratio"] = torch.eq(resp_len, L).float().mean().item()
    stats("prompt_length", prompt_len)
    out["prompt_length/clip_ratio"] = torch.eq(prompt_len, attention_mask.shape[-1] - L).float().mean().item()
    return out


# ------------------------------------------------------------------------------------------------ partial format
def lce_partials(hidden, weight, labels, temperature, label_offset=0, num_splits=1):
    """CPU restatement of the PARTIAL format the CUDA forward emits (include/vpg_b200.h, vpg_lce_fwd_partials):
    [4][S][T] = {m (log2 domain), s, b (raw-dot weighted), t}.  Used to test the vocab-parallel exchange on CPU
    (gloo); the statistics themselves are the reference's (verl/utils/kernel/kernels.py:226-254)."""
    T, V = hidden.shape[0], weight.shape[0]
    x = hidden.float() @ weight.float().t()
    c = 1.4426950408889634 / temperature
    n_tiles = -(-V // 256)
    tps = -(-n_tiles // num_splits)
    out = torch.zeros(4, num_splits, T)
    out[0].fill_(float("-inf"))
    local = labels - label_offset
    for sp in range(num_splits):
        lo, hi = sp * tps * 256, min((sp + 1) * tps * 256, V)
        if lo >= hi:
            continue
        xs = x[:, lo:hi]
        m = (xs * c).max(-1).values
        e = torch.exp2(xs * c - m[:, None])
        out[0, sp], out[1, sp], out[2, sp] = m, e.sum(-1), (xs * e).sum(-1)
        ok = (local >= lo) & (local < hi)
        out[3, sp][ok] = x[ok, local[ok]]
    return out


def lce_finalize(partials, temperature):
    """CPU restatement of vpg_lce_finalize: merge [4][P][T] partials -> (logprobs, entropy)."""
    m, s, b, t = partials[0], partials[1], partials[2], partials[3]
    M = m.max(0).values
    f = torch.where(torch.isinf(m), torch.zeros_like(m), torch.exp2(m - M[None]))
    S, Bv = (s * f).sum(0), (b * f).sum(0)
    lse = (M + torch.log2(S)) * 0.6931471805599453
    return t.sum(0) / temperature - lse, lse - (Bv / S) / temperature


# ------------------------------------------------------------------------------------------------ dynamic batches
class _Entry:
    """heap entry ordered by a static key only (states never change while they sit in the queue)"""

    __slots__ = ("key", "sets")

    def __init__(self, sets):
        # sets: descending list of (sum, [(idx, len), ...]).  Pops first: the largest spread, then the largest first
        # set by (sum, count, items) — verl/utils/seqlen_balancing.py:77-83 with Set.__lt__ :42-47.  Equal-length
        # item lists compare in reverse when every integer is negated.
        top_sum, top_items = sets[0]
        self.sets = sets
        self.key = (-(top_sum - sets[-1][0]), -top_sum, -len(top_items), [(-i, -v) for i, v in top_items])

    def __lt__(self, other):
        return self.key < other.key


def _set_key(s):
    return (s[0], len(s[1]), s[1])


def karmarkar_karp(seqlen_list, k_partitions, equal_size):
    """Largest-differencing partition, verl/utils/seqlen_balancing.py:26-127: every state holds k sets in descending
    order; the two states with the largest spread are merged largest-with-smallest until one state is left."""
    import heapq

    order = sorted((int(v), i) for i, v in enumerate(seqlen_list))
    if equal_size:
        assert len(order) % k_partitions == 0, f"{len(order)} % {k_partitions} != 0"
        groups = [order[o:o + k_partitions] for o in range(0, len(order), k_partitions)]
    else:
        groups = [[item] for item in order]
    queue = []
    for grp in groups:
        sets = [(v, [(i, v)]) for v, i in grp] + [(0, [])] * (k_partitions - len(grp))
        heapq.heappush(queue, _Entry(sorted(sets, key=_set_key, reverse=True)))
    while len(queue) > 1:
        a, b = heapq.heappop(queue), heapq.heappop(queue)
        merged = [(sa + sb, ia + ib) for (sa, ia), (sb, ib) in zip(a.sets, reversed(b.sets))]
        heapq.heappush(queue, _Entry(sorted(merged, key=_set_key, reverse=True)))
    return [[i for i, _ in items] for _, items in queue[0].sets]


def seqlen_balanced_partitions(seqlen_list, k_partitions, equal_size):
    """verl/utils/seqlen_balancing.py:150-190"""
    assert len(seqlen_list) >= k_partitions
    parts = karmarkar_karp(seqlen_list, k_partitions, equal_size)
    assert all(len(p) > 0 for p in parts)
    assert sorted(i for p in parts for i in p) == list(range(len(seqlen_list)))
    return [sorted(p) for p in parts]


def micro_batch_indices(attention_mask, max_token_len, min_num_micro_batch=None, num_batches_divided_by=None,
                        use_dynamic_bsz_balance=True):
    """The scheduling half of rearrange_micro_batches, verl/utils/seqlen_balancing.py:250-306 (single process)."""
    lens = attention_mask.sum(dim=1).tolist()
    assert max_token_len >= attention_mask.shape[-1]
    n = min(len(lens), -(sum(lens) // -max_token_len))
    if min_num_micro_batch is not None:
        n = max(min_num_micro_batch, n)
    if num_batches_divided_by is not None:
        n = -(n // -num_batches_divided_by) * num_batches_divided_by
    assert n <= len(lens)
    parts = seqlen_balanced_partitions(lens, n, equal_size=False)
    if use_dynamic_bsz_balance:
        parts.sort(key=lambda p: (sum(lens[i] ** 2 for i in p), min(p) if p else 0), reverse=True)
    return parts


def reverse_idx(idx_map):
    """verl/utils/seqlen_balancing.py:322-337"""
    out = [0] * len(idx_map)
    for i, j in enumerate(idx_map):
        out[j] = i
    return out


def restore_dynamic_batch(data, batch_idx_list):
    """verl/utils/seqlen_balancing.py:362-375"""
    flat = [i for p in batch_idx_list for i in p]
    return data[torch.tensor(reverse_idx(flat), dtype=torch.long)]


# ------------------------------------------------------------------------------------------------ Megatron-side
def vocab_parallel_entropy(shards):
    """verl/utils/megatron/tensor_parallel.py:109-130 with the tensor-parallel all-reduces written out over the list of
    per-rank logit shards [T, V/tp]: MAX of the row maxima, SUM of sum-exp, SUM of sum softmax*logit."""
    shards = [s.float() for s in shards]
    row_max = torch.stack([s.max(dim=-1).values for s in shards]).max(dim=0).values[:, None]
    exps = [(s - row_max).exp() for s in shards]
    sum_exp = sum(e.sum(dim=-1, keepdim=True) for e in exps)
    sum_p_x = sum(((e / sum_exp) * s).sum(dim=-1, keepdim=True) for e, s in zip(exps, shards))
    return (row_max + sum_exp.log() - sum_p_x).squeeze(-1)


def vocab_parallel_log_probs(shards, labels):
    """-vocab_parallel_cross_entropy of megatron-core (third-party; the reference installs core_v0.12.1 /
    core_r0.13.0, docker/*/Dockerfile.app.*mcore*) as called at tensor_parallel.py:154-158.  Published algorithm:
    all-reduce MAX of the row maxima, subtract; each rank gathers the label logit if rank*V_local <= label <
    (rank+1)*V_local (else 0) and the gathered values are SUM-reduced; sum-exp SUM-reduced;
    loss = log(sum_exp) - predicted."""
    shards = [s.float() for s in shards]
    v_local = shards[0].shape[-1]
    row_max = torch.stack([s.max(dim=-1).values for s in shards]).max(dim=0).values
    predicted = torch.zeros_like(row_max)
    sum_exp = torch.zeros_like(row_max)
    for rank, s in enumerate(shards):
        z = s - row_max[:, None]
        local = labels - rank * v_local
        hit = (local >= 0) & (local < v_local)
        picked = z.gather(-1, local.clamp(0, v_local - 1)[:, None]).squeeze(-1)
        predicted = predicted + torch.where(hit, picked, torch.zeros_like(picked))
        sum_exp = sum_exp + z.exp().sum(dim=-1)
    return predicted - sum_exp.log()


def packed_layout(attention_mask, tp_size, cp_size):
    """verl/models/mcore/util.py:36-47: per-sequence lengths padded to a multiple of tp (tp*cp*2 with cp) + cumsum"""
    seqlens = attention_mask.bool().sum(dim=-1).to(torch.int32)
    align = tp_size * cp_size * 2 if cp_size > 1 else tp_size
    padded = seqlens + (align - seqlens % align) % align
    cu_padded = torch.zeros(len(seqlens) + 1, dtype=torch.int32)
    cu_padded[1:] = torch.cumsum(padded, dim=0)
    return seqlens, padded, cu_padded


def preprocess_packed_seqs(input_ids, attention_mask, tp_size=1, cp_size=1, cp_rank=0):
    """verl/models/mcore/util.py:23-87: this cp rank's packed stream [1, sum_padded/cp, ...].  Token j of a padded
    sequence of Lp tokens belongs to chunk j // (Lp / (2 cp)); rank r holds chunks r and 2cp-1-r back to back."""
    seqlens, padded, cu_padded = packed_layout(attention_mask, tp_size, cp_size)
    out = torch.zeros((int(padded.sum()) // cp_size, *input_ids.shape[2:]), dtype=input_ids.dtype)
    for i in range(input_ids.shape[0]):
        d = input_ids[i][attention_mask[i].bool()]
        if cp_size <= 1:
            out[cu_padded[i]:cu_padded[i] + len(d)] = d
            continue
        lp = int(padded[i])
        half = lp // (2 * cp_size)
        base = int(cu_padded[i]) // cp_size
        for j in range(len(d)):
            chunk = j // half
            owner = chunk if chunk < cp_size else 2 * cp_size - 1 - chunk
            if owner == cp_rank:
                out[base + (0 if chunk < cp_size else half) + j % half] = d[j]
    return out.unsqueeze(0), cu_padded, int(padded.max())


def postprocess_packed_seqs(outputs_per_rank, attention_mask, tp_size=1, cp_size=1):
    """verl/models/mcore/util.py:90-142 with the cp all-gather written out: outputs_per_rank[r] is rank r's
    [1, sum_padded/cp, ...] stream; -> [B, S, ...] with zeros at masked positions."""
    seqlens, padded, cu_padded = packed_layout(attention_mask, tp_size, cp_size)
    B, S = attention_mask.shape
    o0 = outputs_per_rank[0]
    out = torch.zeros((B, S, *o0.shape[2:]), dtype=o0.dtype)
    for i in range(B):
        lp = int(padded[i])
        half = lp // (2 * cp_size) if cp_size > 1 else 0
        base = int(cu_padded[i]) // cp_size
        seq = []
        for j in range(int(seqlens[i])):
            if cp_size <= 1:
                seq.append(o0[0][base + j])
                continue
            chunk = j // half
            owner = chunk if chunk < cp_size else 2 * cp_size - 1 - chunk
            seq.append(outputs_per_rank[owner][0][base + (0 if chunk < cp_size else half) + j % half])
        if seq:
            out[i, attention_mask[i].bool()] = torch.stack(seq)
    return out


def pf_ppo_sample_indices(token_level_scores, reweight_method="pow", weight_pow=2.0):
    """compute_pf_ppo_reweight_data's draw — verl/trainer/ppo/core_algos.py:1103-1121: importance weights from the row
    sums of token_level_scores, then B indices drawn with replacement from the host generator."""
    scores = token_level_scores.sum(dim=-1)
    if reweight_method == "pow":
        weights = torch.pow(torch.abs(scores), weight_pow)
    elif reweight_method == "max_min":
        weights = torch.where((scores == scores.max()) | (scores == scores.min()), 1.0, 0.0)
    elif reweight_method == "max_random":
        weights = torch.where(scores == scores.max(), 0.4, 0.1)
    else:
        raise ValueError(f"Unsupported reweight_method: {reweight_method}")
    weights = torch.clamp(weights + 1e-8, min=1e-8)
    return torch.multinomial(weights, scores.shape[0], replacement=True)
