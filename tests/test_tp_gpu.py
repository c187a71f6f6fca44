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
