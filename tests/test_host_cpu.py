"""CPU: the C-ABI library loads and exports every symbol include/vpg_b200.h declares; host-side logic (group ids,
registries, enums, argument validation that needs no device); no CPU fallback exists."""

import os
import re
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "vpg_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vpg_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from verl_b200 import _lib

    lib = _lib.load()  # binds all SIGNATURES, raises on a missing symbol
    declared = _header_symbols()
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/vpg_b200.h but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature"
    assert set(_lib.SIGNATURES) == set(declared)
    assert b"sm_100a" in lib.vpg_version()


def test_host_only_entry_points():
    from verl_b200 import _lib

    lib = _lib.load()
    # schedule / workspace sizing functions do not touch the device
    assert 1 <= lib.vpg_lce_num_splits(65536, 3584, 152064, 148) <= 64
    assert lib.vpg_lce_num_splits(128, 896, 151936, 148) >= 37  # one token tile must still fill the machine
    c = lib.vpg_lce_bwd_default_chunk(65536, 3584, 152064)
    assert c % 256 == 0 and 256 <= c <= 152064 + 255
    assert lib.vpg_lce_bwd_workspace_bytes(65536, 3584, 152064, c) >= 65536 * c * 2
    assert lib.vpg_ppo_loss_workspace_bytes(16, 1024) >= 16 * 8 * 4
    assert lib.vpg_unpad_workspace_bytes(1 << 20) > 0
    # argument validation happens before any CUDA call
    rc = lib.vpg_lce_fwd_partials(None, None, None, 128, 100, 1000, 1.0, 0, 1, None, None)
    assert rc == -1 and b"multiple of 64" in lib.vpg_last_error()
    rc = lib.vpg_ppo_loss(None, None, None, None, 0, None, None, 4, 4, 0.8, 1.2, 3.0, 0, 0, 0.0, 0.0, -1, 1.0, None, None,
                          None, None, 0, None)
    assert rc == -1


def test_no_cpu_fallback():
    from verl_b200 import ops
    from verl_b200.trainer.ppo import core_algos
    from verl_b200.utils import torch_functional as F

    x = torch.randn(4, 8)
    m = torch.ones(4, 8, dtype=torch.long)
    with pytest.raises(RuntimeError, match="CUDA"):
        F.masked_mean(x, m)
    with pytest.raises(RuntimeError, match="CUDA"):
        core_algos.compute_policy_loss(x, x, x, m, cliprange=0.2)
    with pytest.raises(RuntimeError, match="CUDA"):
        F.logprobs_from_logits(x, torch.zeros(4, dtype=torch.long))
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.unpad_indices(m)


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from verl_b200 import _lib

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setenv("VPG_B200_LIB", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU"):
        _lib.load()
    monkeypatch.delenv("VPG_B200_LIB")
    monkeypatch.setattr(_lib, "_lib", None)
    _lib.load()


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "verl_b200")):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{f} imports the oracle"


def test_group_ids_bit_exact_with_oracle():
    from oracle import port
    from verl_b200.trainer.ppo.core_algos import group_ids_from_index

    import uuid

    rng = np.random.default_rng(0)
    uids = np.repeat(np.array([str(uuid.uuid4()) for _ in range(16)], dtype=object), 8)
    rng.shuffle(uids)
    a, na = group_ids_from_index(uids)
    b, nb = port.group_ids(uids)
    assert a.dtype == np.int32 and na == nb == 16 and np.array_equal(a, b)
    # tensors and ints work as ids too
    a, na = group_ids_from_index(torch.tensor([5, 3, 5, 9]))
    assert a.tolist() == [0, 1, 0, 2] and na == 3


def test_registries_and_enums():
    """behaviour pinned by the reference's tests/trainer/ppo/test_core_algos_on_cpu.py:29-131"""
    from verl_b200.trainer.ppo import core_algos as ca

    assert ca.get_adv_estimator_fn("gae") is ca.compute_gae_advantage_return
    assert ca.get_adv_estimator_fn(ca.AdvantageEstimator.GRPO) is ca.compute_grpo_outcome_advantage
    with pytest.raises(ValueError, match="Unknown advantage estimator"):
        ca.get_adv_estimator_fn("nope")

    @ca.register_adv_est("test_custom")
    def est(*a, **k):
        return None

    assert ca.get_adv_estimator_fn("test_custom") is est
    ca.register_adv_est("test_custom")(est)  # same fn again is fine
    with pytest.raises(ValueError, match="already been registered"):
        ca.register_adv_est("test_custom")(lambda: None)
    ca.ADV_ESTIMATOR_REGISTRY.pop("test_custom")

    @ca.register_policy_loss("test_loss")
    def loss(*a, **k):
        return None

    assert ca.get_policy_loss_fn("test_loss") is loss
    with pytest.raises(ValueError, match="Unsupported loss mode"):
        ca.get_policy_loss_fn("nope")
    ca.POLICY_LOSS_REGISTRY.pop("test_loss")
    assert ca.AdvantageEstimator.GRPO == "grpo" and ca.AdvantageEstimator("gae") is ca.AdvantageEstimator.GAE

    class Ctrl:
        type, kl_coef, target_kl, horizon = "adaptive", 0.1, 0.01, 100

    k = ca.get_kl_controller(Ctrl)
    k.update(0.02, 10)
    assert k.value == pytest.approx(0.1 * (1 + 0.2 * 10 / 100))
    Ctrl.type = "fixed"
    assert ca.get_kl_controller(Ctrl).value == 0.1


def test_kernel_enums_and_validation():
    from verl_b200.utils.kernel import kernels
    from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy

    assert kernels.get_entropy_reduction_enum_number("none") == 0
    assert kernels.get_entropy_reduction_enum_number("sum") == 1
    assert kernels.get_entropy_reduction_enum_number("mean") == 2
    with pytest.raises(ValueError):
        kernels.get_entropy_reduction_enum_number("max")
    kernels.set_backward_method(kernels.BackwardEnum._Total_Separate)
    kernels.set_backward_method(kernels.BackwardEnum._Split_Dlogits_N)
    h = torch.zeros(4, 128, dtype=torch.bfloat16)
    w = torch.zeros(16, 128, dtype=torch.bfloat16)
    with pytest.raises(AssertionError, match="temperature must be a float"):
        linear_cross_entropy(h, w, torch.zeros(4, dtype=torch.long), 1, "none")
    with pytest.raises(AssertionError):  # CPU tensors are rejected (kernels.py:485 of the reference asserts is_cuda)
        linear_cross_entropy(h, w, torch.zeros(4, dtype=torch.long), 1.0, "none")


def test_compat_install_shadows_reference_paths():
    import subprocess

    code = (
        "import verl_b200.compat as c; r = c.install();"
        "from verl.utils.kernel.linear_cross_entropy import linear_cross_entropy as a;"
        "from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy as b;"
        "import verl.trainer.ppo.core_algos as ca, verl.utils.torch_functional as F;"
        "from verl.utils.seqlen_balancing import get_seqlen_balanced_partitions as kk;"
        "from verl.models.mcore.util import preprocess_packed_seqs; from verl.utils.ulysses import ulysses_pad;"
        "from verl.utils.megatron.tensor_parallel import vocab_parallel_entropy;"
        "from verl.utils.experimental.torch_functional import FusedLinearForPPO;"
        "from verl.models.transformers.dense_common import forward_with_torch_backend, forward_with_triton_backend;"
        "from verl.trainer.ppo.ray_trainer import compute_advantage, apply_kl_penalty;"
        "from verl.trainer.ppo.metric_utils import compute_data_metrics;"
        "assert a is b and hasattr(ca, 'compute_policy_loss') and hasattr(F, 'masked_mean'); print('ok', len(r))"
    )
    out = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "ok 12" in out.stdout, out.stderr


def test_d_hidden_allreduce_overlap_plan():
    """Cost model behind set_reduce_d_hidden(True, "auto"): hiding the d_hidden all-reduce under the last d_weight GEMM
    costs that GEMM extra scheduling rounds (16 SMs are kept free for NCCL); it must pick the overlap for the 8-way
    shards of the BASELINE shapes (measured win, profiles/r1_tp8_dhidden_allreduce_overlap_ab.json) and keep the serial
    all-reduce at 2 and 4 GPUs, where the extra rounds cost more than the all-reduce takes."""
    from verl_b200.utils.kernel.kernels import _overlap_plan

    w = 74  # CTA pairs of a 148-SM B200
    for world, want in ((2, False), (4, False), (8, True)):
        assert _overlap_plan(65536, 3584, 152064 // world, world, None, workers=w)[0] is want   # configs[1]
        assert _overlap_plan(65536, 8192, 128256 // world, world, None, workers=w)[0] is want   # configs[3]
    pays, pairs = _overlap_plan(262144, 5120, 19008, 8, None, workers=w)                        # configs[4]
    assert pays and pairs == 8
    assert _overlap_plan(1024, 3584, 19008, 8, None, workers=4) == (False, 0)  # fewer pairs than the reserve


def test_backward_plan_at_the_baseline_shapes():
    """The backward's workspace plan under the default 3 GiB dlogits budget (host-only function): an 8-way vocab shard of
    the BASELINE shapes stays ONE block (no fp32 accumulator pass), the single-GPU headline shape is cut along the
    vocabulary, a vocab shard of a long micro-batch along the tokens (the smaller accumulator), and no plan exceeds the
    budget by more than its one fp32 accumulator."""
    from verl_b200 import ops

    if os.environ.get("VPG_BWD_WS_MB"):
        pytest.skip("VPG_BWD_WS_MB overrides the default budget")
    budget = 3072 << 20
    plans = {
        "c2 n1": ((65536, 3584, 152064), "vocab", 7),
        "c2 n8": ((65536, 3584, 19008), "single", 1),
        "c4 n8": ((65536, 8192, 16032), "single", 1),
        "c2 n4": ((65536, 3584, 38016), "token", 2),
        "c5 n8": ((262144, 5120, 19008), "token", 4),
    }
    for name, ((T, H, V), mode, chunks) in plans.items():
        p = ops.backward_plan(T, H, V)
        assert p["mode"].startswith(mode) and p["chunks"] == chunks, (name, p)
        acc = 0 if mode == "single" else 4 * H * (T if mode == "vocab" else V)
        assert p["workspace_bytes"] <= budget + acc + (64 << 20), (name, p)


def test_ascending_mark_lapses_on_in_place_edits():
    """the one-pass pad_input path is only taken while the index tensor is untouched since unpad_indices made it."""
    from verl_b200 import ops

    base = torch.arange(10)
    idx = base[:6]
    assert not ops.is_marked_ascending(idx)
    ops.mark_ascending(idx)
    assert ops.is_marked_ascending(idx)
    assert not ops.is_marked_ascending(idx.clone())  # a different object carries no mark
    for edit in (lambda t: t.add_(1), lambda t: t.__setitem__(0, 5), lambda t: t.copy_(torch.arange(6).flip(0)),
                 lambda t: base.mul_(2)):  # the last one writes through the base the view shares its counter with
        ops.mark_ascending(idx)
        edit(idx)
        assert not ops.is_marked_ascending(idx)


def test_backward_chunk_under_memory_pressure(monkeypatch):
    """the fallback counts torch's cached-but-unused memory as available, narrows the vocabulary chunks when
    the planned workspace exceeds half of that, and raises a clear error when even 256-column chunks cannot fit because
    the fp32 [T,H] accumulator does not shrink with the chunk width."""
    from verl_b200 import ops

    T, H, V = 65536, 3584, 152064
    GiB = 2**30

    def pretend(free, reserved=0, allocated=0):
        monkeypatch.setattr(torch.cuda, "mem_get_info", lambda device=None: (free, 180 * GiB))
        monkeypatch.setattr(torch.cuda, "memory_reserved", lambda device=None: reserved)
        monkeypatch.setattr(torch.cuda, "memory_allocated", lambda device=None: allocated)

    pretend(100 * GiB)
    assert ops.backward_chunk(T, H, V, "cuda:0") == 0  # plenty: the library plans
    assert ops.backward_chunk(T, H, V, "cuda:0", chunk_v=4096) == 4096  # an explicit width always wins
    pretend(1 * GiB, reserved=120 * GiB, allocated=20 * GiB)
    assert ops.backward_chunk(T, H, V, "cuda:0") == 0  # the cache counts as available
    pretend(4 * GiB)
    narrow = ops.backward_chunk(T, H, V, "cuda:0")
    assert narrow >= 256 and narrow % 256 == 0
    assert ops.backward_plan(T, H, V, narrow)["workspace_bytes"] <= 4 * GiB
    pretend(GiB // 2)  # less than the 896 MiB accumulator alone
    with pytest.raises(RuntimeError, match="accumulator"):
        ops.backward_chunk(T, H, V, "cuda:0")


def test_compat_install_rebinds_callers_that_imported_by_name():
    """a caller that did `from verl.trainer.ppo.core_algos import agg_loss` before install() must not
    keep the reference function.  Fake reference modules stand in for verl (separate process: install() edits
    sys.modules)."""
    import subprocess

    code = '''
import sys, types
def mk(name, **attrs):
    m = types.ModuleType(name); m.__dict__.update(attrs); sys.modules[name] = m
    parent, _, leaf = name.rpartition(".")
    if parent: setattr(sys.modules[parent], leaf, m)
    return m
for pkg in ("verl", "verl.trainer", "verl.trainer.ppo", "verl.workers", "verl.workers.actor"):
    mk(pkg).__path__ = []
def agg_loss(*a, **k): return "reference"
def kl_penalty(*a, **k): return "reference"
ca = mk("verl.trainer.ppo.core_algos", agg_loss=agg_loss, kl_penalty=kl_penalty, ADV_ESTIMATOR_REGISTRY={}, POLICY_LOSS_REGISTRY={})
caller = mk("verl.workers.actor.dp_actor", agg_loss=agg_loss, kl_penalty=kl_penalty, unrelated=len)
import verl_b200.compat as c
from verl_b200.trainer.ppo import core_algos as ours
r = c.install()
assert r["verl.trainer.ppo.core_algos"] == "patched", r
assert ca.agg_loss is ours.agg_loss and caller.agg_loss is ours.agg_loss and caller.kl_penalty is ours.kl_penalty
assert caller.unrelated is len
assert "verl.workers.actor.dp_actor.agg_loss" in r["rebound"], r
print("ok")
'''
    out = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "ok" in out.stdout, out.stdout + out.stderr
