"""GPU: materialised-logits log-prob / entropy (verl/utils/torch_functional.py:64-160) vs golden and oracle."""

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port

pytestmark = pytest.mark.gpu


def T(a, dev):
    return torch.from_numpy(np.asarray(a)).to(dev)


def test_logits_golden(dev):
    from verl_b200.utils import torch_functional as F

    g = load_golden("logits")
    logits, labels = T(g["f32_logits"], dev), T(g["labels"], dev)
    lp = F.logprobs_from_logits(logits, labels)
    assert lp.shape == labels.shape and lp.dtype == torch.float32
    torch.testing.assert_close(lp.cpu(), T(g["f32_logprobs_v2"], "cpu"), atol=2e-6, rtol=1e-5)
    torch.testing.assert_close(F.logprobs_from_logits_v2(logits, labels).cpu(), T(g["f32_logprobs_naive"], "cpu"),
                               atol=1e-5, rtol=1e-5)
    torch.testing.assert_close(F.entropy_from_logits(logits).cpu(), T(g["f32_entropy"], "cpu"), atol=2e-6, rtol=1e-5)
    torch.testing.assert_close(F.entropy_from_logits_with_chunking(logits.reshape(-1, 333)).cpu(),
                               T(g["f32_entropy_chunked"], "cpu"), atol=1e-5, rtol=1e-5)
    lb = T(g["bf16_logits"], dev).bfloat16()
    out = F.logprobs_from_logits(lb.reshape(-1, 333), labels.reshape(-1))
    assert out.dtype == torch.bfloat16
    # the reference rounds log_softmax to bf16 before the gather; one bf16 ulp at |x|<16 is 0.0625
    torch.testing.assert_close(out.float().cpu(), T(g["bf16_logprobs_v2"], "cpu"), atol=0.0625, rtol=0)
    torch.testing.assert_close(F.entropy_from_logits(lb).float().cpu(), T(g["bf16_entropy"], "cpu"), atol=0.05, rtol=2e-2)


@pytest.mark.parametrize("Tn,V", [(1, 1), (7, 5), (33, 1000), (4, 151936), (16, 152064)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_logits_vs_oracle_with_grad(Tn, V, dtype, dev):
    from verl_b200.utils import torch_functional as F

    torch.manual_seed(V)
    logits = (torch.randn(Tn, V) * 4).to(dtype)
    labels = torch.randint(0, V, (Tn,))
    g_lp, g_ent = torch.randn(Tn), torch.randn(Tn)
    # fp64 oracle: at V ~ 150k the fp32 CPU evaluation of entropy_from_logits is itself ~4e-4 off
    x_c = logits.double().requires_grad_(True)
    lp_c = port.logprobs_from_logits(x_c, labels)
    ent_c = port.entropy_from_logits(x_c)
    (lp_c * g_lp.double() + ent_c * g_ent.double()).sum().backward()
    lp_c, ent_c, x_c_grad = lp_c.float(), ent_c.float(), x_c.grad.float()
    x_g = logits.detach().clone().to(dev).requires_grad_(True)
    lp_g = F.logprobs_from_logits(x_g, labels.to(dev))
    ent_g = F.entropy_from_logits(x_g)
    (lp_g.float() * g_lp.to(dev) + ent_g.float() * g_ent.to(dev)).sum().backward()
    tol = dict(atol=2e-5, rtol=1e-5) if dtype == torch.float32 else dict(atol=0.07, rtol=2e-2)
    torch.testing.assert_close(lp_g.float().cpu(), lp_c.detach(), **tol)
    torch.testing.assert_close(ent_g.float().cpu(), ent_c.detach(), **tol)
    gtol = dict(atol=1e-5, rtol=1e-4) if dtype == torch.float32 else dict(atol=2e-2, rtol=4e-2)
    torch.testing.assert_close(x_g.grad.float().cpu(), x_c_grad, **gtol)


def test_extreme_logits(dev):
    from verl_b200.utils import torch_functional as F

    x = torch.tensor([[1e4, -1e4, 0.0, 5.0], [-3e4, -3e4, -3e4, -3e4], [0.0, 0.0, 0.0, 0.0]], device=dev)
    labels = torch.tensor([0, 2, 3], device=dev)
    lp, ent = F.logprobs_from_logits(x, labels), F.entropy_from_logits(x)
    torch.testing.assert_close(lp.cpu(), port.logprobs_from_logits(x.cpu(), labels.cpu()), atol=1e-5, rtol=1e-5)
    torch.testing.assert_close(ent.cpu(), port.entropy_from_logits(x.cpu()), atol=1e-5, rtol=1e-5)


def test_log_probs_from_logits_response(dev):
    """verl/utils/torch_functional.py:421-435 against the reference's own composition (slice + log_softmax gather)"""
    from verl_b200.utils import torch_functional as F

    g = torch.Generator().manual_seed(12)
    B, S, V, L = 3, 17, 301, 6
    logits = torch.randn(B, S, V, generator=g) * 3
    ids = torch.randint(0, V, (B, S), generator=g)
    want = torch.log_softmax(logits[:, -L - 1:-1], -1).gather(-1, ids[:, -L:, None]).squeeze(-1)
    got = F.log_probs_from_logits_response(ids.to(dev), logits.to(dev), L)
    assert got.shape == (B, L)
    torch.testing.assert_close(got.cpu(), want, atol=1e-5, rtol=1e-5)


def test_distributed_stats_single_rank_group():
    """distributed_masked_mean / distributed_mean_max_min_std (verl/utils/torch_functional.py:709-771) over a one-rank
    NCCL group in a child process (the multi-rank case runs in scripts/tp_megatron_check.py under torchrun)."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import torch, torch.distributed as dist\n"
        "from verl_b200.utils import torch_functional as F\n"
        "dist.init_process_group('nccl', init_method='tcp://127.0.0.1:29571', world_size=1, rank=0)\n"
        "torch.cuda.set_device(0)\n"
        "g = torch.Generator().manual_seed(1)\n"
        "x = torch.randn(7, 33, generator=g); m = (torch.rand(7, 33, generator=g) < 0.6).long()\n"
        "mm = F.distributed_masked_mean(x.cuda(), m.cuda())\n"
        "assert abs(float(mm) - float((x * m).sum() / m.sum())) < 1e-6, float(mm)\n"
        "mean, mx, mn, std = F.distributed_mean_max_min_std(x.cuda())\n"
        "assert abs(float(mean) - float(x.mean())) < 1e-6 and float(mx) == float(x.max()) and float(mn) == float(x.min())\n"
        "assert abs(float(std) - float(x.std())) < 1e-5\n"
        "assert F.distributed_mean_max_min_std(x.cuda(), False, False, False)[1:] == (None, None, None)\n"
        "dist.destroy_process_group(); print('DIST_OK')\n"
    )
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, cwd=root)
    assert r.returncode == 0 and "DIST_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
