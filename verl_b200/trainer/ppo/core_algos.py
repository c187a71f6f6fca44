"""Token-level RL math of verl/trainer/ppo/core_algos.py on libvpg_b200, same names and signatures.

Hot-path functions (CUDA kernels, no torch fallback):
  compute_gae_advantage_return   core_algos.py:193-241
  compute_grpo_outcome_advantage core_algos.py:245-308
  the other registered estimators: grpo_passk :309-369, reinforce_plus_plus_baseline :372-422,
  rloo :425-476, opo :479-530, reinforce_plus_plus :533-569, remax :572-605, gpg :608-667
  compute_policy_loss_gpg        core_algos.py:797-815
  compute_value_loss             core_algos.py:992-1031
  agg_loss                       core_algos.py:686-719
  compute_policy_loss            core_algos.py:722-794
  kl_penalty                     core_algos.py:1034-1069
  compute_rewards                core_algos.py:~670-684 (elementwise, built on kl_penalty's k1 kernel)
plus fused_actor_loss(): compute_policy_loss + entropy bonus + KL loss + agg_loss + loss scale in one launch pair,
the composition verl/workers/actor/dp_actor.py:422-469 performs with ~60 torch kernels.

Host-side API kept for drop-in use: the policy-loss / advantage-estimator registries (:33-108), the
AdvantageEstimator enum (:111-128) and the KL controllers (:131-190).
"""

from __future__ import annotations

from enum import Enum
from typing import Any, Optional

import numpy as np
import torch

from ... import ops
from ...staging import staged

__all__ = [
    "register_policy_loss",
    "get_policy_loss_fn",
    "register_adv_est",
    "get_adv_estimator_fn",
    "AdvantageEstimator",
    "AdaptiveKLController",
    "FixedKLController",
    "get_kl_controller",
    "compute_gae_advantage_return",
    "compute_grpo_outcome_advantage",
    "compute_grpo_passk_outcome_advantage",
    "compute_reinforce_plus_plus_baseline_outcome_advantage",
    "compute_rloo_outcome_advantage",
    "compute_opo_outcome_advantage",
    "compute_reinforce_plus_plus_outcome_advantage",
    "compute_remax_outcome_advantage",
    "compute_gpg_outcome_advantage",
    "compute_policy_loss_gpg",
    "compute_value_loss",
    "compute_rewards",
    "agg_loss",
    "compute_policy_loss",
    "kl_penalty",
    "fused_actor_loss",
    "compute_pf_ppo_reweight_data",
    "group_ids_from_index",
]

# ------------------------------------------------------------------------------------------------ registries
POLICY_LOSS_REGISTRY: dict[str, Any] = {}
ADV_ESTIMATOR_REGISTRY: dict[str, Any] = {}


def register_policy_loss(name):
    """Decorator: make ``fn`` retrievable as policy loss ``name`` (core_algos.py:36-50)."""

    def _wrap(fn):
        POLICY_LOSS_REGISTRY[name] = fn
        return fn

    return _wrap


def get_policy_loss_fn(name):
    """Look a registered policy loss up by name (core_algos.py:53-68)."""
    if name not in POLICY_LOSS_REGISTRY:
        raise ValueError(f"Unsupported loss mode: {name}. Supported modes are: {list(POLICY_LOSS_REGISTRY.keys())}")
    return POLICY_LOSS_REGISTRY[name]


def _est_key(name_or_enum):
    return name_or_enum.value if isinstance(name_or_enum, Enum) else name_or_enum


def register_adv_est(name_or_enum):
    """Decorator registering an advantage estimator; re-registering a different fn under a taken name is an error
    (core_algos.py:74-91)."""

    def _wrap(fn):
        key = _est_key(name_or_enum)
        prev = ADV_ESTIMATOR_REGISTRY.get(key)
        if prev is not None and prev != fn:
            raise ValueError(f"Adv estimator {key} has already been registered: {prev} vs {fn}")
        ADV_ESTIMATOR_REGISTRY[key] = fn
        return fn

    return _wrap


def get_adv_estimator_fn(name_or_enum):
    """core_algos.py:94-108"""
    key = _est_key(name_or_enum)
    if key not in ADV_ESTIMATOR_REGISTRY:
        raise ValueError(f"Unknown advantage estimator simply: {key}")
    return ADV_ESTIMATOR_REGISTRY[key]


class AdvantageEstimator(str, Enum):
    """Names of the reference's estimators (core_algos.py:111-128); every one of them has kernels in this build."""

    GAE = "gae"
    GRPO = "grpo"
    REINFORCE_PLUS_PLUS = "reinforce_plus_plus"
    REINFORCE_PLUS_PLUS_BASELINE = "reinforce_plus_plus_baseline"
    REMAX = "remax"
    RLOO = "rloo"
    OPO = "opo"
    GRPO_PASSK = "grpo_passk"
    GPG = "gpg"


# ------------------------------------------------------------------------------------------------ KL controllers
class AdaptiveKLController:
    """Adaptive KL coefficient, https://arxiv.org/pdf/1909.08593.pdf (core_algos.py:131-152)."""

    def __init__(self, init_kl_coef, target_kl, horizon):
        self.value = init_kl_coef
        self.target = target_kl
        self.horizon = horizon

    def update(self, current_kl, n_steps):
        err = np.clip(current_kl / self.target - 1, -0.2, 0.2)
        self.value *= 1 + err * n_steps / self.horizon


class FixedKLController:
    """Constant KL coefficient (core_algos.py:155-168)."""

    def __init__(self, kl_coef):
        self.value = kl_coef

    def update(self, current_kl, n_steps):
        pass


def get_kl_controller(kl_ctrl):
    """core_algos.py:171-190"""
    if kl_ctrl.type == "fixed":
        return FixedKLController(kl_coef=kl_ctrl.kl_coef)
    if kl_ctrl.type == "adaptive":
        assert kl_ctrl.horizon > 0, f"horizon must be larger than 0. Got {kl_ctrl.horizon}"
        return AdaptiveKLController(init_kl_coef=kl_ctrl.kl_coef, target_kl=kl_ctrl.target_kl, horizon=kl_ctrl.horizon)
    raise NotImplementedError


# ------------------------------------------------------------------------------------------------ advantages
def group_ids_from_index(index) -> tuple[np.ndarray, int]:
    """uid array (numpy object array of strings, list, or tensor) -> (int32 ids in first-seen order, num_groups).

    This is the host half of GRPO's "group assignment": the reference groups by dictionary key
    (``id2score[index[i]]``, core_algos.py:286-291); insertion order of that dict is first-seen order.
    """
    if isinstance(index, torch.Tensor):
        index = index.detach().cpu().numpy()
    table: dict[Any, int] = {}
    out = np.empty(len(index), dtype=np.int32)
    for i, key in enumerate(index):
        if isinstance(key, np.generic):
            key = key.item()
        gid = table.get(key)
        if gid is None:
            gid = len(table)
            table[key] = gid
        out[i] = gid
    return out, len(table)


@register_adv_est(AdvantageEstimator.GAE)
@staged
def compute_gae_advantage_return(
    token_level_rewards: torch.Tensor,
    values: torch.Tensor,
    response_mask: torch.Tensor,
    gamma: torch.Tensor,
    lam: torch.Tensor,
):
    """GAE advantages and returns, then masked whitening of the advantages (core_algos.py:193-241).

    One kernel evaluates the reverse recurrence of every row as a block-parallel scan of affine maps, a second one
    normalises.  Returns (advantages, returns), both (bs, response_length) fp32.
    """
    with torch.no_grad():
        advantages, returns = ops.gae(token_level_rewards, values, response_mask, float(gamma), float(lam))
    return advantages, returns


@register_adv_est(AdvantageEstimator.GRPO)
@staged
def compute_grpo_outcome_advantage(
    token_level_rewards: torch.Tensor,
    response_mask: torch.Tensor,
    index: np.ndarray,
    epsilon: float = 1e-6,
    norm_adv_by_std_in_grpo: bool = True,
    config: Optional[Any] = None,
) -> tuple[torch.Tensor, torch.Tensor]:
    """GRPO outcome advantage (core_algos.py:245-308): per-uid group mean / unbiased std of the sequence scores,
    singleton groups use (0, 1); the SAME tensor is returned twice, as the reference does (:308)."""
    gids, n_groups = group_ids_from_index(index)
    with torch.no_grad():
        gid_dev = torch.from_numpy(gids).to(token_level_rewards.device, non_blocking=True)
        scores = ops.grpo_advantage(
            token_level_rewards, response_mask, gid_dev, n_groups, float(epsilon), bool(norm_adv_by_std_in_grpo)
        )
    return scores, scores


def _outcome(kind, token_level_rewards, response_mask, index, **kw):
    gids, n_groups = group_ids_from_index(index)
    with torch.no_grad():
        gid_dev = torch.from_numpy(gids).to(token_level_rewards.device, non_blocking=True)
        adv = ops.outcome_advantage(kind, token_level_rewards, response_mask, gid_dev, n_groups, **kw)
    return adv, adv


def _cfg_get(config, key, default):
    """``config.get(key, default)`` for dict-like / dataclass-like algorithm configs (AlgoConfig.get, base_config.py:26-91)."""
    if config is None:
        return default
    if hasattr(config, "get"):
        return config.get(key, default)
    return getattr(config, key, default)


@register_adv_est(AdvantageEstimator.GRPO_PASSK)
@staged
def compute_grpo_passk_outcome_advantage(
    token_level_rewards: torch.Tensor,
    response_mask: torch.Tensor,
    index: np.ndarray,
    epsilon: float = 1e-6,
    norm_adv_by_std_in_grpo: bool = True,
    config: Optional[Any] = None,
    **kwargs,
) -> tuple[torch.Tensor, torch.Tensor]:
    """Pass@k advantage (core_algos.py:309-369): only the best response of each uid group gets r_max - r_second_max
    (optionally / (std + eps)); groups with fewer than two samples raise ValueError."""
    assert config is not None
    norm = _cfg_get(config, "norm_adv_by_std_in_grpo", True)
    return _outcome("grpo_passk", token_level_rewards, response_mask, index, epsilon=float(epsilon), norm_by_std=bool(norm))


@register_adv_est(AdvantageEstimator.REINFORCE_PLUS_PLUS_BASELINE)
@staged
def compute_reinforce_plus_plus_baseline_outcome_advantage(
    token_level_rewards: torch.Tensor,
    response_mask: torch.Tensor,
    index: torch.Tensor,
    epsilon: float = 1e-6,
    config: Optional[Any] = None,
    **kwargs,
) -> tuple[torch.Tensor, torch.Tensor]:
    """RF++-baseline (core_algos.py:372-422): score minus the uid-group mean, broadcast over the response, then
    masked_whiten over the batch and * mask.  The whitening of the broadcast matrix is evaluated in closed form from
    the row lengths (response masks are 0/1)."""
    return _outcome("reinforce_plus_plus_baseline", token_level_rewards, response_mask, index, epsilon=float(epsilon))


@register_adv_est(AdvantageEstimator.RLOO)
@staged
def compute_rloo_outcome_advantage(
    token_level_rewards: torch.Tensor,
    response_mask: torch.Tensor,
    index: np.ndarray,
    epsilon: float = 1e-6,
    config: Optional[Any] = None,
    **kwargs,
) -> tuple[torch.Tensor, torch.Tensor]:
    """RLOO (core_algos.py:425-476): score * n/(n-1) - group_mean * n/(n-1) for groups of n > 1."""
    return _outcome("rloo", token_level_rewards, response_mask, index, epsilon=float(epsilon))


@register_adv_est(AdvantageEstimator.OPO)
@staged
def compute_opo_outcome_advantage(
    token_level_rewards: torch.Tensor,
    response_mask: torch.Tensor,
    index: np.ndarray,
    epsilon: float = 1e-6,
    config: Optional[Any] = None,
    **kwargs,
) -> tuple[torch.Tensor, torch.Tensor]:
    """OPO (core_algos.py:479-530): baseline = length-weighted mean score of the uid group."""
    return _outcome("opo", token_level_rewards, response_mask, index, epsilon=float(epsilon))


@register_adv_est(AdvantageEstimator.REINFORCE_PLUS_PLUS)
@staged
def compute_reinforce_plus_plus_outcome_advantage(
    token_level_rewards: torch.Tensor, response_mask: torch.Tensor, config: Optional[Any] = None, **kwargs
) -> tuple[torch.Tensor, torch.Tensor]:
    """REINFORCE++ (core_algos.py:533-569): discounted reverse return that resets after masked-out tokens (one
    block-parallel scan per row instead of a Python loop over time steps), advantages = masked_whiten(returns) * mask."""
    assert config is not None
    gamma = config.gamma if hasattr(config, "gamma") else config["gamma"]
    with torch.no_grad():
        advantages, returns = ops.reinforce_plus_plus(token_level_rewards, response_mask, float(gamma))
    return advantages, returns


@register_adv_est(AdvantageEstimator.REMAX)
@staged
def compute_remax_outcome_advantage(
    token_level_rewards: torch.Tensor,
    reward_baselines: torch.Tensor,
    response_mask: torch.Tensor,
    config: Optional[Any] = None,
    **kwargs,
) -> tuple[torch.Tensor, torch.Tensor]:
    """ReMax (core_algos.py:572-605): returns = reverse cumulative sum of the masked rewards, advantages = returns -
    reward_baselines[:, None] * mask."""
    with torch.no_grad():
        advantages, returns = ops.remax(token_level_rewards, reward_baselines, response_mask)
    return advantages, returns


@register_adv_est(AdvantageEstimator.GPG)
@staged
def compute_gpg_outcome_advantage(
    token_level_rewards: torch.Tensor,
    response_mask: torch.Tensor,
    index: np.ndarray,
    epsilon: float = 1e-6,
    f_norm: float = 1.0,
    alpha: float = 1.0,
    config=None,
    **kwargs,
):
    """GPG (core_algos.py:608-667): alpha * (score - group_mean) / f_norm with alpha = bsz / count_nonzero(scores)
    (the ``alpha`` argument is overwritten by the reference too, :647)."""
    return _outcome("gpg", token_level_rewards, response_mask, index, epsilon=float(epsilon), f_norm=float(f_norm))


@staged
def compute_rewards(token_level_scores, old_log_prob, ref_log_prob, kl_ratio):
    """token_level_scores - (old_log_prob - ref_log_prob) * kl_ratio (core_algos.py:670-684)."""
    kl, _ = ops.kl_penalty_fwd(old_log_prob, ref_log_prob, "kl", want_grad=False)
    return token_level_scores - kl * kl_ratio


# ------------------------------------------------------------------------------------------------ losses
class _AggLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, loss_mat, loss_mask, mode):
        out, d = ops.agg_loss_fwd(loss_mat, loss_mask, mode, want_grad=loss_mat.requires_grad)
        ctx.shape, ctx.in_dtype = loss_mat.shape, loss_mat.dtype
        if d is not None:
            ctx.save_for_backward(d)
        return out

    @staticmethod
    def backward(ctx, g):
        (d,) = ctx.saved_tensors
        return (d * g).view(ctx.shape).to(ctx.in_dtype), None, None


@staged
def agg_loss(loss_mat: torch.Tensor, loss_mask: torch.Tensor, loss_agg_mode: str):
    """Aggregate a (bs, response_length) loss matrix into a scalar (core_algos.py:686-719).

    Modes: token-mean, seq-mean-token-sum, seq-mean-token-mean, seq-mean-token-sum-norm; anything else raises
    ValueError like the reference.
    """
    if loss_agg_mode not in ops.AGG_MODES:
        raise ValueError(f"Invalid loss_agg_mode: {loss_agg_mode}")
    return _AggLoss.apply(loss_mat, loss_mask, loss_agg_mode)


class _KlPenalty(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logprob, ref_logprob, kind):
        need = logprob.requires_grad or ref_logprob.requires_grad
        out, g = ops.kl_penalty_fwd(logprob, ref_logprob, kind, want_grad=need)
        ctx.dt = (logprob.dtype, ref_logprob.dtype)
        ctx.need = (logprob.requires_grad, ref_logprob.requires_grad)
        if need:
            ctx.save_for_backward(g)
        return out.view(logprob.shape)

    @staticmethod
    def backward(ctx, go):
        (g,) = ctx.saved_tensors
        d = go.reshape(g.shape) * g  # d out / d logprob; every kind is antisymmetric in (logprob, ref_logprob)
        d_lp = d.view(go.shape).to(ctx.dt[0]) if ctx.need[0] else None
        d_ref = (-d).view(go.shape).to(ctx.dt[1]) if ctx.need[1] else None
        return d_lp, d_ref, None


@staged
def kl_penalty(logprob: torch.FloatTensor, ref_logprob: torch.FloatTensor, kl_penalty) -> torch.FloatTensor:
    """kl / k1, abs, mse / k2, low_var_kl / k3 (core_algos.py:1034-1069); "full" and unknown kinds raise
    NotImplementedError like the reference."""
    if kl_penalty == "full" or kl_penalty not in ops.KL_KINDS:
        raise NotImplementedError
    return _KlPenalty.apply(logprob, ref_logprob, kl_penalty)


class _PolicyLoss(torch.autograd.Function):
    """Forward = fused kernel, backward = the closed-form d loss / d log_prob the same launch produced."""

    @staticmethod
    def forward(ctx, log_prob, old_log_prob, advantages, response_mask, entropy, ref_log_prob, cfg):
        need = log_prob.requires_grad or (entropy is not None and entropy.requires_grad)
        metrics, d_lp, d_ent = ops.ppo_loss(
            old_log_prob, log_prob, advantages, response_mask, entropy=entropy, ref_log_prob=ref_log_prob,
            want_grad=need, **cfg,
        )
        ctx.lp_meta = (log_prob.shape, log_prob.dtype, log_prob.requires_grad)
        ctx.ent_meta = None if entropy is None else (entropy.shape, entropy.dtype, entropy.requires_grad)
        saved = [t for t in (d_lp, d_ent) if t is not None]
        ctx.has = (d_lp is not None, d_ent is not None)
        ctx.save_for_backward(*saved)
        loss = metrics[0].clone()
        stats = metrics
        ctx.mark_non_differentiable(stats)
        return loss, stats

    @staticmethod
    def backward(ctx, g_loss, _g_stats):
        saved = list(ctx.saved_tensors)
        d_lp = saved.pop(0) if ctx.has[0] else None
        d_ent = saved.pop(0) if ctx.has[1] else None
        g_lp = None
        if d_lp is not None and ctx.lp_meta[2]:
            g_lp = (d_lp * g_loss).view(ctx.lp_meta[0]).to(ctx.lp_meta[1])
        g_ent = None
        if d_ent is not None and ctx.ent_meta is not None and ctx.ent_meta[2]:
            g_ent = (d_ent * g_loss).view(ctx.ent_meta[0]).to(ctx.ent_meta[1])
        return g_lp, None, None, None, g_ent, None, None


@staged
def compute_policy_loss(
    old_log_prob,
    log_prob,
    advantages,
    response_mask,
    cliprange=None,
    cliprange_low=None,
    cliprange_high=None,
    clip_ratio_c=3.0,
    loss_agg_mode: str = "token-mean",
):
    """Clipped PPO objective with dual-clip (core_algos.py:722-794).

    Returns (pg_loss, pg_clipfrac, ppo_kl, pg_clipfrac_lower) as 0-dim fp32 tensors; pg_loss carries the gradient
    wrt log_prob (the three statistics are detached — nothing in verl differentiates them).
    """
    assert clip_ratio_c > 1.0, (
        "The lower bound of the clip_ratio_c for dual-clip PPO should be greater than 1.0,"
        + f" but get the value: {clip_ratio_c}."
    )
    if cliprange_low is None:
        cliprange_low = cliprange
    if cliprange_high is None:
        cliprange_high = cliprange
    cfg = dict(cliprange_low=cliprange_low, cliprange_high=cliprange_high, clip_ratio_c=clip_ratio_c,
               loss_agg_mode=loss_agg_mode)
    loss, stats = _PolicyLoss.apply(log_prob, old_log_prob, advantages, response_mask, None, None, cfg)
    return loss, stats[2], stats[3], stats[4]


@register_policy_loss("gpg")
@staged
def compute_policy_loss_gpg(old_log_prob, log_prob, advantages, response_mask, loss_agg_mode="token-mean", config=None):
    """GPG policy loss (core_algos.py:797-815): agg_loss(-log_prob * advantages); the three statistics are 0."""
    cfg = dict(cliprange_low=0.2, cliprange_high=0.2, clip_ratio_c=3.0, loss_agg_mode=loss_agg_mode, pg_kind="gpg")
    if old_log_prob is None:  # unused by GPG (the reference never reads it); the kernel still wants a valid pointer
        old_log_prob = log_prob.detach()
    loss, _ = _PolicyLoss.apply(log_prob, old_log_prob, advantages, response_mask, None, None, cfg)
    zero = torch.tensor(0.0)
    return loss, zero, zero.clone(), zero.clone()


class _LossMatrix(torch.autograd.Function):
    """A per-token loss matrix whose elementwise derivative w.r.t. log_prob came out of the same kernel."""

    @staticmethod
    def forward(ctx, log_prob, pg_losses, d_pg_losses):
        ctx.save_for_backward(d_pg_losses)
        ctx.meta = (log_prob.shape, log_prob.dtype)
        return pg_losses.view(log_prob.shape)

    @staticmethod
    def backward(ctx, g):
        (d,) = ctx.saved_tensors
        return (g.reshape(-1) * d).view(ctx.meta[0]).to(ctx.meta[1]), None, None


def _policy_loss_cfg(config, name, default):
    """config.policy_loss.<name> (None -> default); attribute- or dict-style configs (core_algos.py:861-866,940-941)."""
    pl = config.get("policy_loss") if isinstance(config, dict) else getattr(config, "policy_loss", None)
    val = pl.get(name) if isinstance(pl, dict) else getattr(pl, name, None)
    return default if val is None else val


@register_policy_loss("clip_cov")
@staged
def compute_policy_loss_clip_cov(old_log_prob, log_prob, advantages, response_mask, loss_agg_mode="token-mean",
                                 config=None):
    """Clip-Cov policy loss (core_algos.py:818-905).  A random subset (size max(int(clip_cov_ratio * sum mask), 1)) of the
    tokens with clip_cov_lb < cov < clip_cov_ub that PPO's clip left alone is detached from the loss (corr = 0).

    The eligible set and the loss run on the device; the draw is `torch.randperm(n_eligible)` on the host generator,
    exactly the reference's call, so a seeded run selects the same tokens.  Host syncs: one (n_eligible and sum(mask),
    the reference has the same two).  Returns (pg_loss, pg_clipfrac, ppo_kl, 0)."""
    clip_cov_ratio = _policy_loss_cfg(config, "clip_cov_ratio", 0.0002)
    cliprange = _cfg_get(config, "clip_ratio", None)
    low, high = _cfg_get(config, "clip_ratio_low", None), _cfg_get(config, "clip_ratio_high", None)
    cliprange_low = low if low is not None else cliprange
    cliprange_high = high if high is not None else cliprange
    clip_cov_ub = _policy_loss_cfg(config, "clip_cov_ub", 5.0)
    clip_cov_lb = _policy_loss_cfg(config, "clip_cov_lb", 1.0)
    assert clip_cov_ratio > 0, "clip_ratio should be larger than 0."

    lp_d = log_prob.detach()
    flags, ws = ops.clip_cov_flags(old_log_prob, lp_d, advantages, response_mask, cliprange_low, cliprange_high,
                                   clip_cov_lb, clip_cov_ub)
    eligible, _, _, _ = ops.unpad_indices(flags.view(1, -1))  # ascending flat positions == torch.nonzero order
    clip_num = max(int(clip_cov_ratio * float(ws[0].item())), 1)
    n = flags.numel()
    if eligible.numel() > 0:
        perm = torch.randperm(eligible.numel())
        chosen = ops.gather_rows(eligible, perm[: min(clip_num, eligible.numel())].to(eligible.device))
        ones = torch.ones((chosen.numel(), 1), dtype=torch.uint8, device=flags.device)
        zeroed = ops.scatter_rows(ones, chosen, n).view(-1)
    else:
        zeroed = torch.zeros(n, dtype=torch.uint8, device=flags.device)
    pg, dpg, out3 = ops.clip_cov_loss(old_log_prob, lp_d, advantages, response_mask, cliprange_low, cliprange_high,
                                      zeroed, ws)
    pg_losses = _LossMatrix.apply(log_prob, pg, dpg)
    pg_loss = agg_loss(loss_mat=pg_losses, loss_mask=response_mask, loss_agg_mode=loss_agg_mode)
    return pg_loss, out3[1], out3[0], torch.tensor(0.0)


@register_policy_loss("kl_cov")
@staged
def compute_policy_loss_kl_cov(old_log_prob, log_prob, advantages, response_mask, loss_agg_mode="token-mean",
                               config=None):
    """KL-Cov policy loss (core_algos.py:908-972): the max(1, int(N_valid * kl_cov_ratio)) valid tokens with the largest
    covariance (A - mean A)(log_prob - mean log_prob) pay an extra ppo_kl_coef * |log_prob - old_log_prob|.

    The reference moves the valid tokens to the CPU for torch.topk; here the k-th largest is found on the device
    (radix select), no host sync at all.  Returns (pg_loss, 0, ppo_kl_abs, 0)."""
    kl_cov_ratio = _policy_loss_cfg(config, "kl_cov_ratio", 0.0002)
    ppo_kl_coef = _policy_loss_cfg(config, "ppo_kl_coef", 1.0)
    assert kl_cov_ratio > 0, "kl_cov_ratio should be larger than 0."
    pg, dpg, out3 = ops.kl_cov_loss(old_log_prob, log_prob.detach(), advantages, response_mask, kl_cov_ratio, ppo_kl_coef)
    pg_losses = _LossMatrix.apply(log_prob, pg, dpg)
    pg_loss = agg_loss(loss_mat=pg_losses, loss_mask=response_mask, loss_agg_mode=loss_agg_mode)
    return pg_loss, torch.tensor(0.0), out3[0], torch.tensor(0.0)


@staged
def compute_entropy_loss(logits, response_mask, loss_agg_mode: str = "token-mean"):
    """core_algos.py:975-989 (kept by the reference for backward compatibility): agg_loss of the per-token entropy of
    materialised logits [bs, response_length, vocab]."""
    from verl_b200.utils import torch_functional as verl_F

    token_entropy = verl_F.entropy_from_logits(logits)
    return agg_loss(loss_mat=token_entropy, loss_mask=response_mask, loss_agg_mode=loss_agg_mode)


class _ValueLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, vpreds, returns, values, response_mask, cliprange_value, mode):
        out, d = ops.value_loss(vpreds, returns, values, response_mask, cliprange_value, mode,
                                want_grad=vpreds.requires_grad)
        ctx.meta = (vpreds.shape, vpreds.dtype)
        if d is not None:
            ctx.save_for_backward(d)
        clipfrac = out[1]
        ctx.mark_non_differentiable(clipfrac)
        return out[0], clipfrac

    @staticmethod
    def backward(ctx, g, _g_clip):
        (d,) = ctx.saved_tensors
        return (d * g).view(ctx.meta[0]).to(ctx.meta[1]), None, None, None, None, None


@staged
def compute_value_loss(
    vpreds: torch.Tensor,
    returns: torch.Tensor,
    values: torch.Tensor,
    response_mask: torch.Tensor,
    cliprange_value: float,
    loss_agg_mode: str = "token-mean",
):
    """Clipped value loss (core_algos.py:992-1031) -> (vf_loss, vf_clipfrac); one fused forward + gradient pass."""
    if loss_agg_mode not in ops.AGG_MODES:
        raise ValueError(f"Invalid loss_agg_mode: {loss_agg_mode}")
    return _ValueLoss.apply(vpreds, returns, values, response_mask, float(cliprange_value), loss_agg_mode)


@staged
def fused_actor_loss(
    old_log_prob,
    log_prob,
    advantages,
    response_mask,
    *,
    cliprange=0.2,
    cliprange_low=None,
    cliprange_high=None,
    clip_ratio_c=3.0,
    loss_agg_mode: str = "token-mean",
    entropy=None,
    entropy_coeff: float = 0.0,
    ref_log_prob=None,
    kl_loss_coef: float = 0.0,
    kl_loss_type: str = "low_var_kl",
    loss_scale: float = 1.0,
):
    """Whole actor-loss composition of dp_actor.py:422-469 in one fused pass:

        loss = loss_scale * ( pg_loss - entropy_coeff * agg(entropy) + kl_loss_coef * agg(kl_penalty(lp, ref)) )

    Returns (loss, metrics) where ``loss`` is differentiable wrt log_prob and entropy and ``metrics`` is ONE detached
    fp32[8] device tensor [loss, pg_loss, pg_clipfrac, ppo_kl, pg_clipfrac_lower, entropy_loss, kl_loss, mask_sum]
    (a single .tolist() replaces the reference's four .item() syncs, dp_actor.py:472-479).
    """
    if cliprange_low is None:
        cliprange_low = cliprange
    if cliprange_high is None:
        cliprange_high = cliprange
    use_ent = entropy is not None and entropy_coeff != 0
    cfg = dict(cliprange_low=cliprange_low, cliprange_high=cliprange_high, clip_ratio_c=clip_ratio_c,
               loss_agg_mode=loss_agg_mode, entropy_coeff=entropy_coeff if use_ent else 0.0,
               kl_loss_coef=kl_loss_coef, kl_loss_type=kl_loss_type, loss_scale=loss_scale)
    return _PolicyLoss.apply(log_prob, old_log_prob, advantages, response_mask, entropy if use_ent else None,
                             ref_log_prob, cfg)


def compute_pf_ppo_reweight_data(data, reweight_method: str = "pow", weight_pow: float = 2.0):
    """verl/trainer/ppo/core_algos.py:1072-1148: resample the batch with replacement, rows weighted by their score.

    ``data`` is anything with DataProto's ``.batch`` (mapping of [B, ...] tensors incl. "token_level_scores"; host
    tensors, as on the reference's driver, are staged through verl_b200.staging and come back on the host),
    ``.non_tensor_batch`` (numpy arrays or lists) and ``.meta_info``.  The row sums run on the device, the B weights and
    the multinomial draw stay on the host generator exactly as in the reference (a seeded run resamples the same rows),
    and every tensor of the batch is resampled by ONE row-gather pass.  Returns a shallow copy of ``data`` with the three
    fields replaced (the reference deep-copies everything first and then overwrites the same three fields)."""
    import copy

    from ... import staging

    tls = data.batch["token_level_scores"]
    on_host = not tls.is_cuda
    if on_host:  # stage the whole batch once: every tensor is row-gathered on the device
        dev_batch = {k: staging.to_device(v) if not v.is_cuda else v for k, v in data.batch.items()}
        tls = dev_batch["token_level_scores"]
    else:
        dev_batch = data.batch
    ones = torch.ones(tls.shape, dtype=torch.uint8, device=tls.device)
    scores = ops.masked_row_sums(tls, ones)[0].view(tls.shape[:-1]).cpu()
    if reweight_method == "pow":
        weights = torch.pow(torch.abs(scores), weight_pow)
    elif reweight_method == "max_min":
        weights = torch.where((scores == torch.max(scores)) | (scores == torch.min(scores)), 1.0, 0.0)
    elif reweight_method == "max_random":
        weights = torch.where(scores == torch.max(scores), 0.4, 0.1)
    else:
        raise ValueError(f"Unsupported reweight_method: {reweight_method}")
    weights = torch.clamp(weights + 1e-8, min=1e-8)
    batch_size = scores.shape[0]
    sample_indices = torch.multinomial(weights, batch_size, replacement=True)
    idx_dev = sample_indices.to(tls.device)
    resampled_batch = {key: ops.gather_rows(tensor, idx_dev) for key, tensor in dev_batch.items()}
    if on_host:
        resampled_batch = {key: staging.to_host(tensor) for key, tensor in resampled_batch.items()}
    sample_indices_np = sample_indices.numpy()
    resampled_non_tensor_batch = {}
    for key, array in getattr(data, "non_tensor_batch", {}).items():
        if isinstance(array, np.ndarray):
            resampled_non_tensor_batch[key] = array[sample_indices_np]
        else:
            resampled_non_tensor_batch[key] = [array[i] for i in sample_indices_np]
    resampled_meta_info = {}
    for key, value in getattr(data, "meta_info", {}).items():
        if isinstance(value, list) and len(value) == batch_size:
            resampled_meta_info[key] = [value[i] for i in sample_indices_np]
        else:
            resampled_meta_info[key] = value
    resampled_data = copy.copy(data)
    try:
        new_batch = type(data.batch)(resampled_batch)
        if hasattr(data.batch, "batch_size"):
            new_batch.batch_size = data.batch.batch_size
    except Exception:  # a mapping type that cannot be rebuilt from a dict
        new_batch = resampled_batch
    resampled_data.batch = new_batch
    resampled_data.non_tensor_batch = resampled_non_tensor_batch
    resampled_data.meta_info = resampled_meta_info
    return resampled_data
