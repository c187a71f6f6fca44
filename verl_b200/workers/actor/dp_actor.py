"""The lm-head half of DataParallelPPOActor fused end to end.

Reference flow (verl/workers/actor/dp_actor.py):
  _forward_micro_batch :80-270   unpad ids -> model -> lm-head over ALL rmpad tokens (dense_common.py:150-193) ->
                                 pad_input back to [B,S] (:221-232) -> slice [:, -L-1:-1] (:236-237)
  update_policy        :393-470  compute_policy_loss + entropy bonus + KL loss + agg_loss, scaled, .backward(),
                                 four .item() syncs for the metrics (:472-479)

The transformer body is out of scope: these functions start from its output, the rmpad hidden states
``[T, H]`` (what ``self.model(...)`` returns inside forward_with_*_backend), and end at the gradient wrt those hidden
states and the lm-head weight.  Differences from the reference flow, none of which changes a result:
  * the lm-head runs only on the stream rows whose output survives the ``[:, -L-1:-1]`` slice (the reference computes
    every prompt row and throws it away — about half of T at BASELINE configs[2]); dropped rows get an exactly-zero
    hidden-state gradient, as they do in the reference;
  * no [B,S] intermediates: selected rows scatter straight into the [B,L] response matrix;
  * loss, its gradient and all logged scalars come from one fused launch pair; metrics stay on the device as ONE
    fp32[8] tensor.
"""

from __future__ import annotations

from typing import Optional

import torch

from ... import bert_padding, ops
from ...trainer.ppo import core_algos
from ...utils.kernel.linear_cross_entropy import linear_cross_entropy

__all__ = ["response_window", "forward_micro_batch_from_hidden", "fused_actor_step"]


def response_window(input_ids: torch.Tensor, attention_mask: torch.Tensor, response_length: int):
    """Index plumbing of one micro-batch, all on the device (one host read of the two token counts).

    -> dict(indices [T] padded flat positions of the valid tokens, cu_seqlens, max_seqlen,
            sel [T_sel] stream rows inside the response window, dst [T_sel] rows of the [B*L] response matrix,
            labels [T] = roll(input_ids_rmpad, -1))
    """
    B, S = input_ids.shape
    ids_rmpad, indices, cu_seqlens, max_seqlen, _ = bert_padding.unpad_input(input_ids.unsqueeze(-1), attention_mask)
    sel, dst, labels = ops.response_rows(indices, ids_rmpad.squeeze(-1), S, int(response_length))
    return dict(indices=indices, cu_seqlens=cu_seqlens, max_seqlen=max_seqlen, sel=sel, dst=dst, labels=labels,
                batch=B, seqlen=S, response_length=int(response_length))


def forward_micro_batch_from_hidden(hidden_rmpad, lm_head_weight, window, temperature: float,
                                    calculate_entropy: bool = True, dist_process_group=None):
    """(entropy [B,L] | None, log_probs [B,L]) like DataParallelPPOActor._forward_micro_batch (dp_actor.py:80-270),
    from the rmpad hidden states ``[T,H]`` (or ``[1,T,H]``).  Differentiable wrt hidden_rmpad and lm_head_weight."""
    if hidden_rmpad.dim() == 3:
        hidden_rmpad = hidden_rmpad.squeeze(0)
    B, L = window["batch"], window["response_length"]
    h_sel = bert_padding.index_first_axis(hidden_rmpad, window["sel"])
    lab_sel = ops.gather_rows(window["labels"], window["sel"])
    lp_sel, ent_sel = linear_cross_entropy(h_sel, lm_head_weight, lab_sel, float(temperature), "none",
                                           dist_process_group)
    log_probs = bert_padding.index_put_first_axis(lp_sel.unsqueeze(-1), window["dst"], B * L).view(B, L)
    entropy = None
    if calculate_entropy:
        entropy = bert_padding.index_put_first_axis(ent_sel.unsqueeze(-1), window["dst"], B * L).view(B, L)
    return entropy, log_probs


def fused_actor_step(
    hidden_rmpad: torch.Tensor,
    lm_head_weight: torch.Tensor,
    input_ids: torch.Tensor,
    attention_mask: torch.Tensor,
    old_log_probs: torch.Tensor,
    advantages: torch.Tensor,
    response_mask: torch.Tensor,
    *,
    temperature: float,
    ref_log_prob: Optional[torch.Tensor] = None,
    clip_ratio: float = 0.2,
    clip_ratio_low: Optional[float] = None,
    clip_ratio_high: Optional[float] = None,
    clip_ratio_c: float = 3.0,
    entropy_coeff: float = 0.0,
    use_kl_loss: bool = False,
    kl_loss_coef: float = 0.001,
    kl_loss_type: str = "low_var_kl",
    loss_agg_mode: str = "token-mean",
    loss_scale: float = 1.0,
    dist_process_group=None,
    window: Optional[dict] = None,
    backward: bool = True,
):
    """One micro-batch of update_policy (dp_actor.py:393-470) from the transformer's rmpad hidden states.

    ``loss_scale`` is the caller's ``B_micro / ppo_mini_batch_size`` or ``1 / gradient_accumulation`` (:465-469).
    Returns ``(loss, metrics, log_prob, entropy)``; with ``backward=True`` the gradients are already accumulated into
    ``hidden_rmpad.grad`` / ``lm_head_weight.grad``.  ``metrics`` is one detached fp32[8] device tensor
    [loss, pg_loss, pg_clipfrac, ppo_kl, pg_clipfrac_lower, entropy_loss, kl_loss, mask_sum].
    """
    L = response_mask.shape[-1]
    if window is None:
        window = response_window(input_ids, attention_mask, L)
    entropy, log_prob = forward_micro_batch_from_hidden(
        hidden_rmpad, lm_head_weight, window, temperature, calculate_entropy=entropy_coeff != 0,
        dist_process_group=dist_process_group)
    loss, metrics = core_algos.fused_actor_loss(
        old_log_probs, log_prob, advantages, response_mask,
        cliprange=clip_ratio, cliprange_low=clip_ratio_low, cliprange_high=clip_ratio_high, clip_ratio_c=clip_ratio_c,
        loss_agg_mode=loss_agg_mode, entropy=entropy, entropy_coeff=entropy_coeff,
        ref_log_prob=ref_log_prob if use_kl_loss else None, kl_loss_coef=kl_loss_coef if use_kl_loss else 0.0,
        kl_loss_type=kl_loss_type, loss_scale=loss_scale)
    if backward:
        loss.backward()
    return loss.detach(), metrics, log_prob.detach(), None if entropy is None else entropy.detach()
