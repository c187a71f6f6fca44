"""compute_data_metrics of verl/trainer/ppo/metric_utils.py:80-180 as one sweep over the batch tensors (SURVEY.md §8
f.2): the reference issues ~30 tiny torch reductions and as many ``.item()`` syncs; here three launches fill one
fp32[26] device vector that is read back once."""

from __future__ import annotations

from typing import Any

from ... import ops, staging

__all__ = ["compute_data_metrics"]


def compute_data_metrics(batch, use_critic: bool = True) -> dict[str, Any]:
    b = batch.batch
    out = staging.call(ops.data_metrics, b["token_level_scores"], b["token_level_rewards"], b["advantages"], b["returns"],
                           b["values"] if use_critic else None, b["response_mask"], b["attention_mask"]).tolist()
    metrics: dict[str, Any] = {}
    for name in ("critic/score", "critic/rewards", "critic/advantages", "critic/returns"):
        o = ops.DATA_METRIC_SLOTS[name]
        metrics[f"{name}/mean"], metrics[f"{name}/max"], metrics[f"{name}/min"] = out[o], out[o + 1], out[o + 2]
    if use_critic:
        o = ops.DATA_METRIC_SLOTS["critic/values"]
        metrics["critic/values/mean"], metrics["critic/values/max"], metrics["critic/values/min"] = out[o:o + 3]
        metrics["critic/vf_explained_var"] = out[ops.DATA_METRIC_SLOTS["critic/vf_explained_var"]]
    for name in ("response_length", "prompt_length"):
        o = ops.DATA_METRIC_SLOTS[name]
        metrics[f"{name}/mean"], metrics[f"{name}/max"], metrics[f"{name}/min"] = out[o], out[o + 1], out[o + 2]
        metrics[f"{name}/clip_ratio"] = out[o + 3]
    if "__num_turns__" in getattr(batch, "non_tensor_batch", {}):
        num_turns = batch.non_tensor_batch["__num_turns__"]
        metrics["num_turns/min"] = num_turns.min()
        metrics["num_turns/max"] = num_turns.max()
        metrics["num_turns/mean"] = num_turns.mean()
    return metrics
