"""Drop-in for the HF forward shims of verl/models/transformers/dense_common.py (CausalLMOutputForPPO :22-26,
forward_with_torch_backend :73-130, forward_with_triton_backend :133-193) and for the backend switch
verl/models/transformers/monkey_patch.py:148-192 (patch_forward_with_backends).

The shims replace `Model.forward` of a HF causal LM: run the transformer body (`self.model`), take the last hidden
states, roll the labels by one and hand `lm_head.weight` to the fused lm-head op, so the [tokens x vocab] logits are
never built.  Contract kept from the reference (consumed at verl/workers/actor/dp_actor.py:163-178,254-256): called
with `temperature=` and `return_dict=True`, returns an object with `.log_probs` and `.entropy`.
Both backend names end in the same sm_100a kernels; they differ only in the output contract the reference gives them:
  "triton": fp32, flat [T] for [1,T,H] hidden states (rmpad path);  "torch": hidden dtype, shaped like the labels.
"b200" is accepted as an alias of "triton".
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch

try:  # same base class as the reference when transformers is present, a plain dataclass otherwise
    from transformers.modeling_outputs import CausalLMOutputWithPast as _Base
except Exception:  # pragma: no cover

    @dataclass
    class _Base:  # type: ignore[no-redef]
        loss: Optional[torch.Tensor] = None
        logits: Optional[torch.Tensor] = None
        past_key_values: Optional[object] = None
        hidden_states: Optional[tuple] = None
        attentions: Optional[tuple] = None


@dataclass
class CausalLMOutputForPPO(_Base):
    log_probs: Optional[torch.Tensor] = None
    entropy: Optional[torch.Tensor] = None


_BODY_KEYS = ("input_ids", "attention_mask", "position_ids", "past_key_values", "inputs_embeds", "use_cache",
              "output_attentions", "output_hidden_states", "cache_position")


def forward_base_model(self, **kw):
    """dense_common.py:29-70: the transformer body only (`self.model`), config defaults for the two output flags."""
    cfg = getattr(self, "config", None)
    if kw.get("output_attentions") is None:
        kw["output_attentions"] = getattr(cfg, "output_attentions", None)
    if kw.get("output_hidden_states") is None:
        kw["output_hidden_states"] = getattr(cfg, "output_hidden_states", None)
    return self.model(**{k: kw.get(k) for k in _BODY_KEYS if k in kw})


def _rolled_labels(labels, input_ids, who: str):
    if labels is not None:
        return torch.roll(labels, shifts=-1, dims=-1)
    if input_ids is not None:
        return torch.roll(input_ids, shifts=-1, dims=-1)
    raise RuntimeError(f"To use {who}, either labels or input_ids must be provided.")


def _forward_fused(self, who: str, backend: str, input_ids, attention_mask, position_ids, past_key_values,
                   inputs_embeds, labels, use_cache, output_attentions, output_hidden_states, return_dict,
                   cache_position, temperature):
    outputs = forward_base_model(
        self, input_ids=input_ids, attention_mask=attention_mask, position_ids=position_ids,
        past_key_values=past_key_values, inputs_embeds=inputs_embeds, use_cache=use_cache,
        output_attentions=output_attentions, output_hidden_states=output_hidden_states, cache_position=cache_position)
    hidden_states = outputs[0]
    if not return_dict:
        raise NotImplementedError(f"{who} has to return_dict")
    rolled = _rolled_labels(labels, input_ids, who)
    if backend == "torch":
        from ...utils.experimental.torch_functional import FusedLinearForPPO

        log_probs, entropy = FusedLinearForPPO().forward(hidden_states=hidden_states, vocab_weights=self.lm_head.weight,
                                                         input_ids=rolled, temperature=temperature)
    else:
        from ...utils.kernel.linear_cross_entropy import linear_cross_entropy

        log_probs, entropy = linear_cross_entropy(hidden_states, self.lm_head.weight, rolled, float(temperature), "none")
    return CausalLMOutputForPPO(log_probs=log_probs, entropy=entropy,
                                past_key_values=getattr(outputs, "past_key_values", None),
                                hidden_states=getattr(outputs, "hidden_states", None),
                                attentions=getattr(outputs, "attentions", None))


def forward_with_torch_backend(self, input_ids=None, attention_mask=None, position_ids=None, past_key_values=None,
                               inputs_embeds=None, labels=None, use_cache=None, output_attentions=None,
                               output_hidden_states=None, return_dict=None, cache_position=None, logits_to_keep=0,
                               temperature: float = 1.0, **loss_kwargs):
    return _forward_fused(self, "forward_with_torch_backend", "torch", input_ids, attention_mask, position_ids,
                          past_key_values, inputs_embeds, labels, use_cache, output_attentions, output_hidden_states,
                          return_dict, cache_position, temperature)


def forward_with_triton_backend(self, input_ids=None, attention_mask=None, position_ids=None, past_key_values=None,
                                inputs_embeds=None, labels=None, use_cache=None, output_attentions=None,
                                output_hidden_states=None, return_dict=None, cache_position=None, logits_to_keep=0,
                                temperature: float = 1.0, **loss_kwargs):
    return _forward_fused(self, "forward_with_triton_backend", "triton", input_ids, attention_mask, position_ids,
                          past_key_values, inputs_embeds, labels, use_cache, output_attentions, output_hidden_states,
                          return_dict, cache_position, temperature)


def patch_forward_with_backends(model, use_fused_kernels: bool = False, fused_kernels_backend: str = None):
    """monkey_patch.py:148-192 for text-only causal LMs: swap `model.__class__.forward` for the fused shim.  The
    Qwen2-VL / Qwen2.5-VL variants of the reference (qwen2_vl.py:416-559, qwen2_5_vl.py:141-288) differ only in how
    the vision tower feeds the body; they reach the same op through the shadowed import (compat.install)."""
    if not use_fused_kernels or fused_kernels_backend not in ("triton", "torch", "b200"):
        print(f"Skipping monkey patch for {model.__class__.__name__} as use_fused_kernels is {use_fused_kernels} or "
              f"fused_kernels_backend is {fused_kernels_backend}")
        return
    if fused_kernels_backend == "torch":
        model.__class__.forward = forward_with_torch_backend
    else:
        model.__class__.forward = forward_with_triton_backend
    print(f"Using the sm_100a lm-head kernels ({fused_kernels_backend} contract) in {model.__class__.__name__}")
