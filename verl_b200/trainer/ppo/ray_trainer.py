"""Driver-side token ops of verl/trainer/ppo/ray_trainer.py on the batch tensors:
apply_kl_penalty :153-193, compute_response_mask :196-211, compute_advantage :214-291.

The Ray trainer itself (RayPPOTrainer, resource pools, worker groups) is out of scope.  ``data`` is duck-typed: any
object with ``.batch`` (mapping name -> tensor; a TensorDict or a plain dict), ``.non_tensor_batch`` (mapping) and
``.meta_info`` works — verl's DataProto does.  The reference runs these on the driver's CPU tensors
(ray_trainer.py:1283-1305): host tensors are staged to the current CUDA device through pinned buffers, the kernels run
there, and the results are returned on the host (verl_b200/staging.py — input staging, no host arithmetic)."""

from __future__ import annotations

import torch

from ... import ops, staging
from . import core_algos
from .core_algos import AdvantageEstimator

__all__ = ["apply_kl_penalty", "compute_response_mask", "compute_advantage"]


def apply_kl_penalty(data, kl_ctrl, kl_penalty="kl"):
    """token_level_rewards = token_level_scores - beta * kl_penalty(old_log_probs, ref_log_prob) * response_mask, and
    the controller update with the sequence-averaged KL (ray_trainer.py:153-193).  One kernel + one scalar read."""
    batch = data.batch
    scores = batch["token_level_scores"]
    batch_size = scores.shape[0]
    beta = kl_ctrl.value
    with torch.no_grad():
        rewards, cur = staging.call(ops.kl_reward, scores, batch["old_log_probs"], batch["ref_log_prob"],
                                    batch["response_mask"], kl_penalty, float(beta))
    current_kl = cur.item()
    kl_ctrl.update(current_kl=current_kl, n_steps=batch_size)
    batch["token_level_rewards"] = rewards
    return data, {"actor/reward_kl_penalty": current_kl, "actor/reward_kl_penalty_coeff": beta}


def compute_response_mask(data):
    """attention_mask[:, -response_length:] (ray_trainer.py:196-211) — a view, no kernel."""
    response_length = data.batch["responses"].size(1)
    return data.batch["attention_mask"][:, -response_length:]


def compute_advantage(data, adv_estimator, gamma=1.0, lam=1.0, num_repeat=1, norm_adv_by_std_in_grpo=True, config=None):
    """Advantage dispatch of ray_trainer.py:214-291 (GAE, GRPO, and every registered estimator by name)."""
    batch = data.batch
    if "response_mask" not in batch.keys():
        batch["response_mask"] = compute_response_mask(data)
    if adv_estimator == AdvantageEstimator.GAE:
        advantages, returns = core_algos.compute_gae_advantage_return(
            token_level_rewards=batch["token_level_rewards"], values=batch["values"],
            response_mask=batch["response_mask"], gamma=gamma, lam=lam)
        batch["advantages"] = advantages
        batch["returns"] = returns
        if config is not None and core_algos._cfg_get(config, "use_pf_ppo", False):  # ray_trainer.py:256-261
            pf = core_algos._cfg_get(config, "pf_ppo", None)
            return core_algos.compute_pf_ppo_reweight_data(
                data, core_algos._cfg_get(pf, "reweight_method", "pow"), core_algos._cfg_get(pf, "weight_pow", 2.0))
        return data
    elif adv_estimator == AdvantageEstimator.GRPO:
        advantages, returns = core_algos.compute_grpo_outcome_advantage(
            token_level_rewards=batch["token_level_rewards"], response_mask=batch["response_mask"],
            index=data.non_tensor_batch["uid"], norm_adv_by_std_in_grpo=norm_adv_by_std_in_grpo)
    else:
        fn = core_algos.get_adv_estimator_fn(adv_estimator)
        kwargs = {"token_level_rewards": batch["token_level_rewards"], "response_mask": batch["response_mask"],
                  "config": config}
        if "uid" in data.non_tensor_batch:
            kwargs["index"] = data.non_tensor_batch["uid"]
        if "reward_baselines" in batch.keys():
            kwargs["reward_baselines"] = batch["reward_baselines"]
        advantages, returns = fn(**kwargs)
    batch["advantages"] = advantages
    batch["returns"] = returns
    return data
