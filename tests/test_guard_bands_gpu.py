"""GPU: out-of-bounds WRITE check with guard bands (the pool's compute-sanitizer is closed, profiles/r2_sanitizer_status.txt).

Every device buffer `verl_b200.ops` allocates while a case runs (outputs, partials, dlogits / accumulator / split-K
workspaces, index arrays) is carved out of ONE sentinel-filled arena with a 64 KiB guard band on either side.  After the
case every byte outside the carved regions must still hold the sentinel, and the results must equal the ones computed
with ordinary allocations.  Shapes are ragged on purpose (T, V, row counts not multiples of any tile) and every way the
lm-head backward cuts its workspace is covered for both tile variants' epilogues (TMA tensor stores of dlogits, fp32
accumulators, split-K partial tiles).  Reads out of bounds are not visible to this check (they fault or they do not).
"""

import math

import pytest
import torch

pytestmark = pytest.mark.gpu

SENTINEL = 0xA5
GUARD = 64 * 1024


class _Arena:
    def __init__(self, nbytes, dev):
        self.buf = torch.full((nbytes,), SENTINEL, dtype=torch.uint8, device=dev)
        self.off = 0
        self.regions = []

    def alloc(self, shape, dtype):
        if isinstance(shape, int):
            shape = (shape,)
        shape = tuple(int(s) for s in shape)
        n = math.prod(shape) * torch.empty((), dtype=dtype).element_size()
        start = (self.off + GUARD + 511) // 512 * 512
        assert start + n + GUARD <= self.buf.numel(), "guard arena too small for this case"
        self.off = start + n
        self.regions.append((start, n))
        return self.buf[start:start + n].view(dtype).view(shape)

    def check(self):
        outside = torch.ones(self.buf.numel(), dtype=torch.bool, device=self.buf.device)
        for start, n in self.regions:
            outside[start:start + n] = False
        bad = (self.buf != SENTINEL) & outside
        if bool(bad.any()):
            first = int(bad.nonzero()[0])
            owner = max((r for r in self.regions if r[0] <= first), default=None, key=lambda r: r[0])
            raise AssertionError(f"{int(bad.sum())} guard bytes overwritten, first at arena offset {first} "
                                 f"(after region {owner}, {len(self.regions)} regions)")


class _TorchProxy:
    """`torch` as seen by verl_b200.ops, with CUDA allocations redirected into the arena"""

    def __init__(self, arena):
        self._arena = arena

    def __getattr__(self, name):
        return getattr(torch, name)

    def empty(self, *shape, dtype=None, device=None, **kw):
        if len(shape) == 1 and isinstance(shape[0], (tuple, list, torch.Size)):
            shape = tuple(shape[0])
        if device is None or torch.device(device).type != "cuda":
            return torch.empty(*shape, dtype=dtype, device=device, **kw)
        return self._arena.alloc(shape, dtype or torch.float32)

    def zeros(self, *shape, dtype=None, device=None, **kw):
        t = self.empty(*shape, dtype=dtype, device=device, **kw)
        return t.zero_() if t.is_cuda else torch.zeros(*shape, dtype=dtype, device=device, **kw)

    def empty_like(self, t, dtype=None, **kw):
        if not t.is_cuda:
            return torch.empty_like(t, dtype=dtype, **kw)
        return self._arena.alloc(tuple(t.shape), dtype or t.dtype)

    def zeros_like(self, t, dtype=None, **kw):
        return self.empty_like(t, dtype=dtype).zero_() if t.is_cuda else torch.zeros_like(t, dtype=dtype, **kw)


@pytest.fixture
def guarded(monkeypatch, dev):
    from verl_b200 import ops

    arena = _Arena(768 << 20, dev)
    monkeypatch.setattr(ops, "torch", _TorchProxy(arena))
    yield arena
    monkeypatch.undo()


def _lm_inputs(Tn, H, V, dev, dtype=torch.bfloat16, seed=5):
    g = torch.Generator(device=dev).manual_seed(seed)
    hidden = (torch.randn(Tn, H, device=dev, generator=g) * 0.5).to(dtype)
    weight = (torch.randn(V, H, device=dev, generator=g) * 0.05).to(dtype)
    labels = torch.randint(0, V, (Tn,), device=dev, generator=g)
    g_lp = torch.randn(Tn, device=dev, generator=g)
    g_ent = torch.randn(Tn, device=dev, generator=g)
    return hidden, weight, labels, g_lp, g_ent


@pytest.mark.parametrize("Tn,H,V", [(1000, 320, 5000), (257, 64, 264), (3, 128, 100003 // 8 * 8), (2049, 512, 1032)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_lm_head_writes_stay_inside_their_buffers(Tn, H, V, dtype, dev, guarded):
    from verl_b200 import ops

    hidden, weight, labels, g_lp, g_ent = _lm_inputs(Tn, H, V, dev, dtype)
    temp = 0.8
    plans = [dict(chunk_v=0), dict(chunk_v=256), dict(chunk_v=768), dict(chunk_t=256), dict(chunk_t=512)]
    n_before = len(guarded.regions)
    part = ops.lce_forward_partials(hidden, weight, labels, temp)
    lp, ent, lse, eb = ops.lce_finalize(part, temp)
    grads = [ops.lce_backward(hidden, weight, labels, lse, eb, g_lp, g_ent, temp, **kw) for kw in plans]
    torch.cuda.synchronize()
    assert len(guarded.regions) >= n_before + 4 + 2 * len(plans), "allocations did not go through the guard arena"
    guarded.check()
    # and the numbers are right (fp32 math of the same op)
    x = hidden.float() @ weight.float().t() / temp
    lse_r = torch.logsumexp(x, -1)
    lp_r = x.gather(1, labels[:, None]).squeeze(1) - lse_r
    p = torch.softmax(x, -1)
    ent_r = lse_r - (p * x).sum(-1)
    tol = 2e-3 if dtype == torch.bfloat16 else 2e-2  # tf32 recompute of fp32 operands
    torch.testing.assert_close(lp, lp_r, atol=tol, rtol=tol)
    torch.testing.assert_close(ent, ent_r, atol=tol, rtol=tol)
    d_x = (g_lp[:, None] * (torch.nn.functional.one_hot(labels, V) - p)
           + g_ent[:, None] * (-p * (x - (p * x).sum(-1, keepdim=True)))) / temp
    dh_r, dw_r = d_x @ weight.float(), d_x.t() @ hidden.float()
    for dh, dw in grads:
        torch.testing.assert_close(dh.float(), dh_r, atol=2e-2, rtol=4e-2)
        torch.testing.assert_close(dw.float(), dw_r, atol=2e-2, rtol=4e-2)


def test_split_k_and_plain_gemm_writes_stay_inside(dev, guarded):
    from verl_b200 import ops

    g = torch.Generator(device=dev).manual_seed(2)
    for (M, N, K) in [(19008, 520, 4160), (1000, 704, 1100 // 8 * 8), (130, 72, 64)]:
        A = (torch.rand(M, K, device=dev, generator=g) - 0.5).bfloat16()
        B = (torch.rand(N, K, device=dev, generator=g) - 0.5).bfloat16()
        ref = A.float() @ B.float().t()
        for split_k in (False, True):
            for od in (torch.bfloat16, torch.float32):
                out = ops.gemm_bf16(A, B, out_dtype=od, split_k=split_k)
                torch.testing.assert_close(out.float(), ref, atol=2e-2 * math.sqrt(K / 64), rtol=2e-2)
    torch.cuda.synchronize()
    assert guarded.regions
    guarded.check()


def test_rmpad_and_token_kernels_write_inside(dev, guarded):
    from verl_b200 import ops

    torch.manual_seed(11)
    for B, S, W in [(7, 1001, 1), (3, 4097, 5), (33, 130, 64), (2, 9, 3)]:
        mask = (torch.rand(B, S) > 0.4)
        mask[0] = False
        mask[-1, -1] = True
        mask_d = mask.long().to(dev)
        idx, seqlens, cu, mx = ops.unpad_indices(mask_d)
        want = mask.flatten().nonzero().flatten()
        assert torch.equal(idx.cpu(), want)
        src = torch.randn(B * S, W, device=dev)
        rows = ops.gather_rows(src, idx)
        assert torch.equal(rows, src[idx])
        for marked in (True, False):  # one-pass gap fill (ascending, provenance-marked) and memset + scatter
            i2 = idx if marked else idx.clone()
            back = ops.scatter_rows(rows, i2, B * S)
            ref = torch.zeros_like(src)
            ref[idx] = rows
            assert torch.equal(back, ref)
        old = -torch.rand(B, S, device=dev)
        lp = old + 0.1 * torch.randn(B, S, device=dev)
        adv = torch.randn(B, S, device=dev)
        out = ops.ppo_loss(old, lp, adv, mask_d, cliprange_low=0.2, cliprange_high=0.28, want_grad=True)
        assert out is not None
        rewards = torch.zeros(B, S, device=dev)
        rewards[:, -1] = 1.0
        ops.gae(rewards, torch.randn(B, S, device=dev), mask_d.clone().index_fill_(0, torch.tensor([0], device=dev), 1),
                0.99, 0.95)
        gid = (torch.arange(B, device=dev) % 2).int()
        ops.grpo_advantage(rewards, mask_d, gid, 2, 1e-6, True)
        ops.masked_sums(adv, mask_d)
    torch.cuda.synchronize()
    guarded.check()
