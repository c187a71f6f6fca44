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
