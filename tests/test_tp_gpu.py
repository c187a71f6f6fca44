"""Multi-GPU parity of the vocab-parallel lm-head with the NVLink-fused exchange (needs >= 2 GPUs; skipped on the
single-GPU box).  Runs scripts/tp_check.py under torchrun: fused push / multimem.red paths vs the NCCL exchange vs the
full-vocab single-GPU result."""

import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("shape", [(4099, 1024, 32768), (1000, 896, 8192)])
def test_nvlink_fused_exchange_matches_nccl(shape):
    n = 2 if torch.cuda.device_count() < 4 else 4
    T, H, V = shape
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "scripts", "tp_check.py"), "--tokens", str(T),
           "--hidden", str(H), "--vocab", str(V)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "MISMATCH" not in r.stdout


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_megatron_side_equivalents_on_real_ranks():
    """vocab_parallel_entropy / log-probs over NCCL and the context-parallel packed-seq round trip"""
    n = 2 if torch.cuda.device_count() < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", "29537", os.path.join(ROOT, "scripts", "tp_megatron_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "MISMATCH" not in r.stdout


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.skipif(not os.path.isfile(os.path.join(ROOT, "oracle", "_ref", "tests", "utils",
                                                    "test_linear_cross_entropy_tp.py")),
                    reason="reference snapshot (oracle/_ref) not built")
@pytest.mark.parametrize("exchange", ["nccl", "nvlink"])
def test_reference_own_tensor_parallel_test_passes(exchange):
    """The reference's OWN tensor-parallel kernel test (tests/utils/test_linear_cross_entropy_tp.py, all five shapes,
    unmodified, from the oracle/_ref snapshot) with `verl.utils.kernel.linear_cross_entropy` resolved to verl_b200: the
    sharded op against PyTorch on the all-gathered weight, d_hidden all-reduced by the script, on two real ranks; once
    with the NCCL exchange of the partials and once with the exchange fused into the GEMM epilogue over NVLink."""
    env = dict(os.environ, VPG_TP_EXCHANGE=exchange, VERL_REFERENCE_ROOT=os.path.join(ROOT, "oracle", "_ref"))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "scripts", "run_ref_lce_test.py"), "--mode", "ours",
           "--tp"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("[PASS]: Verified kernel forward & backward correctness.") == 5, r.stdout[-3000:]


@pytest.mark.skipif(not os.path.isfile(os.path.join(ROOT, "oracle", "_ref", "tests", "utils",
                                                    "test_linear_cross_entropy.py")),
                    reason="reference snapshot (oracle/_ref) not built")
def test_reference_own_kernel_test_passes_on_one_gpu():
    """The reference's OWN single-GPU kernel test (tests/utils/test_linear_cross_entropy.py, unmodified, from the
    oracle/_ref snapshot) with `verl.utils.kernel.linear_cross_entropy` resolved to verl_b200: its five shapes, its three
    PyTorch comparators (plain, FusedLinearForPPO, chunked entropy) and its tolerances (log-probs atol 1e-3, entropy 5e-3,
    gradients atol 2e-2 / rtol 4e-2) judge the tcgen05 kernels, forward and backward."""
    env = dict(os.environ, VERL_REFERENCE_ROOT=os.path.join(ROOT, "oracle", "_ref"))
    cmd = [sys.executable, os.path.join(ROOT, "scripts", "run_ref_lce_test.py"), "--mode", "ours"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "mode=ours: all cases passed" in r.stdout, r.stdout[-3000:]
    assert r.stdout.count("Verified forward & backward correctness.") == 5, r.stdout[-3000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_in_kernel_allreduce_gives_up_on_a_missing_peer():
    """The fused d_hidden all-reduce waits for its peers inside the d_weight GEMM kernel.  A peer that never launches the
    matching backward must cost the others a bounded wait and a RuntimeError, not a hung GPU; a peer that arrives late
    must fail too rather than return a partial sum (scripts/tp_timeout_check.py)."""
    env = dict(os.environ, VPG_AR_TIMEOUT_MS="1500")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29545", os.path.join(ROOT, "scripts", "tp_timeout_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "RESULT OK" in r.stdout and "MISMATCH" not in r.stdout
