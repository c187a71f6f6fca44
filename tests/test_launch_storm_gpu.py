"""Launch storm: the stand-in for compute-sanitizer's racecheck, which is closed on this pool.

The lm-head kernels hand-roll their synchronisation (mbarrier pipelines between the TMA, MMA and epilogue warps, TMEM
double buffering, split-K partials, persistent schedulers).  A missing arrive, a stale phase bit or a reused buffer shows
up as a result that depends on what else is in flight.  So: 2 000+ forward+backward calls over randomised shapes, enqueued
back to back on three streams with no host synchronisation in between, every output compared BITWISE on the device with
the result of the same call made alone on an idle GPU.  The kernels have a fixed reduction order (no float atomics on
this path), so any differing bit is a race, not rounding."""

import random

import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(rng, dev, gen):
    T = rng.choice([1, 7, 64, 129, 255, 256, 257, 300, 511, 640])
    H = rng.choice([64, 128, 192, 256, 512])
    V = rng.choice([256, 300, 1000, 2048, 2049, 4096, 5000])
    hidden = (torch.rand(T, H, device=dev, generator=gen) - 0.5).bfloat16()
    weight = (torch.rand(V, H, device=dev, generator=gen) - 0.5).bfloat16()
    labels = torch.randint(0, V, (T,), device=dev, generator=gen)
    g_lp = torch.rand(T, device=dev, generator=gen) * 2 - 1
    g_ent = torch.rand(T, device=dev, generator=gen) - 0.5
    return hidden, weight, labels, g_lp, g_ent, rng.choice([1.0, 0.7, 1.5])


def _run(case):
    from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy

    hidden, weight, labels, g_lp, g_ent, temp = case
    h = hidden.clone().requires_grad_(True)
    w = weight.clone().requires_grad_(True)
    lp, ent = linear_cross_entropy(h, w, labels, temp, "none")
    torch.autograd.backward([lp, ent], [g_lp, g_ent])
    return lp.detach(), ent.detach(), h.grad, w.grad


def test_launch_storm_is_bitwise_equal_to_serial_runs(dev):
    rng = random.Random(20261106)
    gen = torch.Generator(device=dev).manual_seed(99)
    cases = [_case(rng, dev, gen) for _ in range(32)]
    serial = []
    for c in cases:  # each call alone on an idle GPU
        torch.cuda.synchronize()
        serial.append(_run(c))
        torch.cuda.synchronize()
    again = _run(cases[0])
    torch.cuda.synchronize()
    assert all(torch.equal(a, b) for a, b in zip(again, serial[0])), "the op is not run-to-run deterministic when serial"

    streams = [torch.cuda.Stream(device=dev) for _ in range(3)]
    mismatches = [torch.zeros((), dtype=torch.int64, device=dev) for _ in streams]
    n_calls = 2048
    for i in range(n_calls):
        k = rng.randrange(len(cases))
        si = rng.randrange(len(streams))
        with torch.cuda.stream(streams[si]):
            got = _run(cases[k])
            for a, b in zip(got, serial[k]):
                # bit patterns, so that NaNs would compare too
                av = a.view(torch.int16) if a.dtype == torch.bfloat16 else a.view(torch.int32)
                bv = b.view(torch.int16) if b.dtype == torch.bfloat16 else b.view(torch.int32)
                mismatches[si] += (av != bv).sum()
    torch.cuda.synchronize()
    total = sum(int(m) for m in mismatches)
    assert total == 0, f"{total} elements differ from the serial results over {n_calls} overlapped calls"
