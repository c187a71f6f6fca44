"""Install verl_b200 behind the reference's own import paths.

``install()`` makes the hot-path names of an (importable or absent) ``verl`` resolve to this package:

  verl.utils.kernel.linear_cross_entropy.linear_cross_entropy   (used by verl/models/transformers/dense_common.py:
      179-185, qwen2_vl.py:544, qwen2_5_vl.py:273-279, verl/models/mcore/model_forward_fused.py:302-309)
  verl.utils.kernel.kernels.{set_backward_method, BackwardEnum, EntropyReductionEnum, ...}
  verl.utils.torch_functional.{logprobs_from_logits, entropy_from_logits, masked_mean, ...}
  verl.trainer.ppo.core_algos.{compute_policy_loss, agg_loss, kl_penalty, compute_value_loss, compute_policy_loss_gpg,
      compute_grpo_outcome_advantage, compute_gae_advantage_return and the seven other estimators}
      (+ every ADV_ESTIMATOR_REGISTRY entry and the "gpg" / "clip_cov" / "kl_cov" POLICY_LOSS_REGISTRY entries)
  verl.utils.experimental.torch_functional.{FusedLinearForPPO, FusedLinearForPPOFunction}   (the DEFAULT "torch"
      fused backend, verl/models/transformers/dense_common.py:88-121)
  verl.models.transformers.dense_common.{forward_with_torch_backend, forward_with_triton_backend, ...}
  verl.trainer.ppo.metric_utils.compute_data_metrics and verl.trainer.ppo.ray_trainer.{apply_kl_penalty,
      compute_advantage, compute_response_mask, compute_data_metrics}   (the driver-side call sites,
      verl/trainer/ppo/ray_trainer.py:1283-1305; ray_trainer binds compute_data_metrics by name at import, so an
      importable ray_trainer gets that name patched too)
  flash_attn.bert_padding.{unpad_input, pad_input, index_first_axis}   (only with patch_flash_attn=True)

If the real ``verl`` package is importable its modules are patched attribute by attribute (everything outside the
hot path keeps working); otherwise light namespace modules are registered so that
``from verl.utils.kernel.linear_cross_entropy import linear_cross_entropy`` works stand-alone.
"""

from __future__ import annotations

import importlib
import sys
import types

from . import bert_padding as _bp
from .trainer.ppo import core_algos as _ca
from .utils import torch_functional as _tf
from .utils.kernel import kernels as _k
from .utils.kernel import linear_cross_entropy as _lce
from .utils.experimental import torch_functional as _xtf
from .models.transformers import dense_common as _dc
from .trainer.ppo import metric_utils as _mu
from .trainer.ppo import ray_trainer as _rt

_TF_NAMES = [
    "logprobs_from_logits", "logprobs_from_logits_v2", "logprobs_from_logits_naive", "logprobs_from_logits_flash_attn",
    "entropy_from_logits", "entropy_from_logits_with_chunking", "masked_sum", "masked_mean", "masked_var",
    "masked_whiten", "get_unpad_data", "log_probs_from_logits_response_rmpad", "log_probs_from_logits_all_rmpad",
    "gather_from_labels", "log_probs_from_logits_response", "distributed_masked_mean", "distributed_mean_max_min_std",
]
_CA_NAMES = ["compute_policy_loss", "agg_loss", "kl_penalty", "compute_grpo_outcome_advantage",
             "compute_gae_advantage_return", "compute_rewards", "compute_grpo_passk_outcome_advantage",
             "compute_reinforce_plus_plus_baseline_outcome_advantage", "compute_rloo_outcome_advantage",
             "compute_opo_outcome_advantage", "compute_reinforce_plus_plus_outcome_advantage",
             "compute_remax_outcome_advantage", "compute_gpg_outcome_advantage", "compute_policy_loss_gpg",
             "compute_value_loss", "compute_policy_loss_clip_cov", "compute_policy_loss_kl_cov",
             "compute_entropy_loss", "compute_pf_ppo_reweight_data"]


def _try_import(name):
    try:
        return importlib.import_module(name)
    except Exception:
        return None


def _namespace(name):
    mod = sys.modules.get(name)
    if mod is None:
        mod = types.ModuleType(name)
        mod.__path__ = []  # namespace-like
        sys.modules[name] = mod
        parent, _, leaf = name.rpartition(".")
        if parent:
            setattr(_namespace(parent), leaf, mod)
    return mod


def _rebind_by_name_imports(replaced: dict) -> list:
    """`from verl.trainer.ppo.core_algos import agg_loss, kl_penalty` executed BEFORE install() leaves the caller holding
    the reference function (verl/workers/actor/dp_actor.py:29,36, megatron_actor.py:40, ray_trainer.py:47): patching the
    defining module's attribute does not reach it.  Walk the already-imported `verl.*` modules and swap every module
    global that still IS one of the replaced reference objects.  -> ["module.name", ...] of what was re-bound."""
    if not replaced:
        return []
    rebound = []
    for mod_name, mod in list(sys.modules.items()):
        if mod is None or not (mod_name == "verl" or mod_name.startswith("verl.")):
            continue
        for attr, val in list(vars(mod).items()):
            hit = replaced.get(id(val))
            if hit is not None and hit[0] is val:
                setattr(mod, attr, hit[1])
                rebound.append(f"{mod_name}.{attr}")
    return sorted(rebound)


_EXTRA_TARGETS = {
    # §8(f).3 / (f).4 callers' plumbing: module -> (verl_b200 module, names)
    "verl.utils.seqlen_balancing": ("verl_b200.utils.seqlen_balancing", [
        "karmarkar_karp", "get_seqlen_balanced_partitions", "rearrange_micro_batches", "prepare_dynamic_batch",
        "restore_dynamic_batch", "get_reverse_idx", "log_seqlen_unbalance", "ceildiv", "roundup_divisible"]),
    "verl.utils.ulysses": ("verl_b200.utils.ulysses", [
        "gather_outputs_and_unpad", "ulysses_pad", "ulysses_pad_and_slice_inputs", "slice_input_tensor", "Gather"]),
    "verl.utils.megatron.tensor_parallel": ("verl_b200.utils.megatron.tensor_parallel", [
        "vocab_parallel_entropy", "vocab_parallel_log_probs_from_logits",
        "vocab_parallel_log_probs_from_logits_response_rmpad"]),
    "verl.models.mcore.util": ("verl_b200.models.mcore.util", ["preprocess_packed_seqs", "postprocess_packed_seqs"]),
}


def install(patch_flash_attn: bool = False, patch_extras: bool = False) -> dict:
    """Route the reference's hot-path imports to verl_b200.  Returns {module name: 'patched' | 'shadowed'} (+ the key
    'rebound': module globals of already-imported `verl.*` callers that held a reference function by name and were
    swapped too, so calling install() after the workers were imported still takes effect).

    The §8(f).3/(f).4 plumbing modules (seqlen_balancing, ulysses, megatron.tensor_parallel, mcore.util) are shadowed
    when the reference module is not importable; an importable one is only patched with patch_extras=True (its
    callers then receive plain dicts of tensor views instead of TensorDicts from rearrange_micro_batches)."""
    report = {}
    for name, (src_name, names) in _EXTRA_TARGETS.items():
        src = importlib.import_module(src_name)
        real = _try_import(name)
        if real is not None and real is not src:
            if patch_extras:
                for n in names:
                    setattr(real, n, getattr(src, n))
                report[name] = "patched"
        else:
            parent, _, leaf = name.rpartition(".")
            sys.modules[name] = src
            setattr(_namespace(parent), leaf, src)
            report[name] = "shadowed"
    targets = {
        "verl.utils.kernel.linear_cross_entropy": (_lce, ["LinearCrossEntropy", "linear_cross_entropy"]),
        "verl.utils.kernel.kernels": (_k, [n for n in dir(_k) if not n.startswith("__")]),
        "verl.utils.torch_functional": (_tf, _TF_NAMES),
        "verl.trainer.ppo.core_algos": (_ca, _CA_NAMES),
        # the reference's DEFAULT fused backend ("torch", ppo_trainer.yaml:90-94) and the HF forward shims that pick one
        "verl.utils.experimental.torch_functional": (_xtf, ["FusedLinearForPPOFunction", "FusedLinearForPPO"]),
        "verl.models.transformers.dense_common": (_dc, ["CausalLMOutputForPPO", "forward_base_model",
                                                        "forward_with_torch_backend", "forward_with_triton_backend"]),
        # driver-side call sites (ray_trainer.py:1283-1305)
        "verl.trainer.ppo.metric_utils": (_mu, ["compute_data_metrics"]),
        "verl.trainer.ppo.ray_trainer": (_rt, ["apply_kl_penalty", "compute_response_mask", "compute_advantage"]),
    }
    replaced = {}  # id(reference object) -> (reference object, replacement): for callers that imported it BY NAME
    for name, (src, names) in targets.items():
        real = _try_import(name)
        if real is not None and real is not src:
            for n in names:
                old = getattr(real, n, None)
                if old is not None and old is not getattr(src, n) and callable(old):
                    replaced[id(old)] = (old, getattr(src, n))
                setattr(real, n, getattr(src, n))
            if name.endswith("ray_trainer"):  # `from verl.trainer.ppo.metric_utils import compute_data_metrics` at its top
                real.compute_data_metrics = _mu.compute_data_metrics
            if name.endswith("core_algos"):
                for key, fn in _ca.ADV_ESTIMATOR_REGISTRY.items():
                    real.ADV_ESTIMATOR_REGISTRY[key] = fn
                for key, fn in _ca.POLICY_LOSS_REGISTRY.items():  # "gpg", "clip_cov", "kl_cov"
                    real.POLICY_LOSS_REGISTRY[key] = fn
            report[name] = "patched"
        else:
            parent, _, leaf = name.rpartition(".")
            sys.modules[name] = src
            setattr(_namespace(parent), leaf, src)
            report[name] = "shadowed"
    rebound = _rebind_by_name_imports(replaced)
    if rebound:
        report["rebound"] = rebound
    if patch_flash_attn:
        fa = _try_import("flash_attn.bert_padding")
        if fa is not None:
            for n in ("unpad_input", "pad_input", "index_first_axis", "index_put_first_axis"):
                setattr(fa, n, getattr(_bp, n))
            report["flash_attn.bert_padding"] = "patched"
    return report
