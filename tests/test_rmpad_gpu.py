"""GPU: rmpad index plumbing, bit-exact against flash_attn.bert_padding golden vectors and the oracle."""

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port

pytestmark = pytest.mark.gpu


def T(a, dev):
    return torch.from_numpy(np.asarray(a)).to(dev)


def test_unpad_pad_golden_bit_exact(dev):
    from verl_b200 import bert_padding as bp
    from verl_b200.utils import torch_functional as F

    g = load_golden("rmpad")
    mask, ids = T(g["mask"], dev), T(g["ids"], dev)
    x, indices, cu, max_s, seqused = bp.unpad_input(ids.unsqueeze(-1), mask)
    assert indices.dtype == torch.int64 and cu.dtype == torch.int32 and isinstance(max_s, int)
    assert torch.equal(indices.cpu(), T(g["indices"], "cpu"))
    assert torch.equal(cu.cpu(), T(g["cu_seqlens"], "cpu"))
    assert max_s == int(g["max_seqlen"])
    assert torch.equal(seqused.cpu(), T(g["seqused"], "cpu"))
    assert torch.equal(x.squeeze(-1).cpu(), T(g["ids_rmpad"], "cpu"))
    i2, cu2, m2 = F.get_unpad_data(mask)
    assert torch.equal(i2, indices) and torch.equal(cu2, cu) and m2 == max_s
    padded = bp.pad_input(T(g["vals"], dev), indices, *g["mask"].shape)
    assert torch.equal(padded.cpu(), T(g["padded"], "cpu"))
    hs = T(g["hs"], dev)
    hs_rmpad = bp.index_first_axis(hs.reshape(-1, hs.shape[-1]), indices)
    assert torch.equal(hs_rmpad.cpu(), T(g["hs_rmpad"], "cpu"))
    _, e_idx, e_cu, e_max, _ = bp.unpad_input(torch.zeros(4, 9, 1, device=dev), T(g["edge_mask"], dev))
    assert torch.equal(e_idx.cpu(), T(g["edge_indices"], "cpu")) and torch.equal(e_cu.cpu(), T(g["edge_cu"], "cpu"))
    assert e_max == int(g["edge_max"])


@pytest.mark.parametrize("B,S", [(1, 1), (3, 4097), (16, 16384), (128, 2048), (5, 0)])
@pytest.mark.parametrize("dtype", [torch.int64, torch.bool, torch.int32])
def test_unpad_vs_oracle_shapes(B, S, dtype, dev):
    from verl_b200 import ops

    torch.manual_seed(B * 131 + S)
    mask = (torch.rand(B, S) > 0.35)
    if S > 2:
        mask[0] = False  # an empty row
        mask[-1] = True  # a full row
    idx, seqlens, cu, mx = ops.unpad_indices(mask.to(dtype).to(dev))
    w_idx, w_len, w_cu, w_mx = port.unpad_indices(mask.long())
    assert torch.equal(idx.cpu(), w_idx) and torch.equal(seqlens.cpu(), w_len) and torch.equal(cu.cpu(), w_cu)
    assert mx == w_mx


@pytest.mark.parametrize("B,S", [(300, 4100), (5000, 40), (4500, 1000), (2, 1 << 20)])
def test_unpad_many_tiles_rows_and_groups(B, S, dev):
    """more than 128 tiles (two-level offsets), more than 2048 rows (several row-scan blocks), rows that are not a
    multiple of 32 (row boundaries inside a warp's 32 elements)"""
    from verl_b200 import ops

    g = torch.Generator().manual_seed(B + S)
    lens = torch.randint(0, S + 1, (B,), generator=g)
    mask = (torch.arange(S)[None] < lens[:, None])
    mask = torch.where(torch.rand(B, 1, generator=g) < 0.5, mask, mask.flip(1))  # left- and right-padded rows
    idx, seqlens, cu, mx = ops.unpad_indices(mask.long().to(dev))
    w_idx, w_len, w_cu, w_mx = port.unpad_indices(mask.long())
    assert torch.equal(idx.cpu(), w_idx) and torch.equal(seqlens.cpu(), w_len) and torch.equal(cu.cpu(), w_cu)
    assert mx == w_mx


@pytest.mark.parametrize("dtype", [torch.int64, torch.int32, torch.uint8])
def test_unpad_unaligned_base_and_non_binary_values(dtype, dev):
    """a mask view whose base pointer is not 16-byte aligned takes the guarded-load path; mask values other than 0/1
    count as valid tokens and are SUMMED into seqlens, as the reference's mask.sum(-1) does"""
    from verl_b200 import ops

    B, S = 37, 1111
    g = torch.Generator().manual_seed(5)
    vals = torch.randint(0, 4, (B * S + 3,), generator=g).to(dtype)
    vals[torch.rand(B * S + 3, generator=g) < 0.4] = 0
    buf = vals.to(dev)
    for shift in (0, 1, 3):
        m = buf[shift:shift + B * S].view(B, S)
        idx, seqlens, cu, mx = ops.unpad_indices(m)
        ref = m.cpu().long()
        assert torch.equal(idx.cpu(), ref.flatten().nonzero().flatten())
        assert torch.equal(seqlens.cpu().long(), ref.sum(-1))
        assert torch.equal(cu.cpu().long(), torch.nn.functional.pad(ref.sum(-1).cumsum(0), (1, 0)))
        assert mx == int(ref.sum(-1).max())


def test_all_masked_and_all_valid(dev):
    from verl_b200 import ops

    z = torch.zeros(4, 100, dtype=torch.long, device=dev)
    idx, seqlens, cu, mx = ops.unpad_indices(z)
    assert idx.numel() == 0 and mx == 0 and cu.tolist() == [0] * 5
    o = torch.ones(4, 100, dtype=torch.long, device=dev)
    idx, seqlens, cu, mx = ops.unpad_indices(o)
    assert torch.equal(idx, torch.arange(400, device=dev)) and mx == 100 and cu.tolist() == [0, 100, 200, 300, 400]


@pytest.mark.parametrize("row", [(), (1,), (3,), (8,), (3584,), (2, 5)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.int64, torch.uint8])
def test_gather_scatter_roundtrip(row, dtype, dev):
    """pad(unpad(x)) keeps valid rows and zero-fills the rest; unpad(pad(y)) == y (round-trip property)"""
    from verl_b200 import bert_padding as bp

    torch.manual_seed(0)
    B, S = 6, 50
    mask = (torch.rand(B, S, device=dev) > 0.4).long()
    x = (torch.rand(B, S, *row, device=dev) * 100).to(dtype)
    if len(row) == 0:
        x = x.unsqueeze(-1)
    xr, indices, *_ = bp.unpad_input(x, mask)
    assert torch.equal(xr, x.reshape(B * S, *x.shape[2:])[indices])
    back = bp.pad_input(xr, indices, B, S)
    keep = mask.bool().view(B, S, *([1] * (x.dim() - 2)))
    assert torch.equal(back, torch.where(keep, x, torch.zeros_like(x)))
    assert torch.equal(bp.unpad_input(back, mask)[0], xr)


def test_autograd_of_index_ops(dev):
    from verl_b200 import bert_padding as bp

    torch.manual_seed(1)
    B, S, D = 4, 33, 8
    mask = (torch.rand(B, S, device=dev) > 0.5).long()
    x = torch.randn(B, S, D, device=dev, requires_grad=True)
    xr, indices, *_ = bp.unpad_input(x, mask)
    w = torch.randn_like(xr)
    (xr * w).sum().backward()
    want = torch.zeros(B * S, D, device=dev)
    want[indices] = w
    assert torch.equal(x.grad.view(B * S, D), want)
    y = torch.randn(indices.numel(), D, device=dev, requires_grad=True)
    out = bp.pad_input(y, indices, B, S)
    w2 = torch.randn_like(out)
    (out * w2).sum().backward()
    assert torch.equal(y.grad, w2.view(B * S, D)[indices])


def test_response_rmpad_helpers(dev):
    """log_probs_from_logits_response_rmpad / _all_rmpad (verl/utils/torch_functional.py:438-490) agree with the
    padded computation, as the reference's tests/models/test_transformer.py:62-104 checks on real models"""
    from verl_b200 import bert_padding as bp
    from verl_b200.utils import torch_functional as F

    torch.manual_seed(2)
    B, S, V, R = 4, 24, 97, 8
    ids = torch.randint(0, V, (B, S), device=dev)
    mask = torch.ones(B, S, dtype=torch.long, device=dev)
    mask[0, :5] = 0  # left padding
    mask[2, :9] = 0
    logits = torch.randn(B, S, V, device=dev)
    logits_rmpad, indices, *_ = bp.unpad_input(logits, mask)
    got = F.log_probs_from_logits_response_rmpad(ids, mask, logits_rmpad, R)
    ids_rmpad = bp.unpad_input(ids.unsqueeze(-1), mask)[0].transpose(0, 1)
    got2 = F.log_probs_from_logits_all_rmpad(ids_rmpad, logits_rmpad, indices, B, S, R)
    want = port.logprobs_from_logits(logits[:, -R - 1:-1].cpu(), ids[:, -R:].cpu())
    torch.testing.assert_close(got.cpu(), want, atol=1e-5, rtol=1e-5)
    torch.testing.assert_close(got2.cpu(), want, atol=1e-5, rtol=1e-5)


@pytest.mark.parametrize("row", [1, 3, 8, 96])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.int64])
def test_pad_input_one_pass_gap_fill(row, dtype, dev):
    """pad_input with indices that come straight from unpad_input takes the one-pass gap-filling kernel (every output row
    written once, no memset); the result must be bit-identical to the memset + scatter path an unmarked copy of the same
    indices takes, and to flash-attn's definition (zeros elsewhere).  Masks with empty rows, a masked head and a masked
    tail, one valid token, everything valid."""
    from verl_b200 import bert_padding, ops

    g = torch.Generator().manual_seed(row)
    masks = [(torch.rand(7, 53, generator=g) < 0.6).long(), torch.ones(3, 17, dtype=torch.long)]
    m = torch.zeros(5, 40, dtype=torch.long)
    m[1, 5:9] = 1
    m[3, 39] = 1
    masks.append(m)
    one = torch.zeros(4, 9, dtype=torch.long)
    one[2, 4] = 1
    masks.append(one)
    big = (torch.rand(64, 4096, generator=g) < 0.63).long()
    big[5] = 0
    big[:, :100] = 0
    big[:, -300:] = 0
    masks.append(big)
    ragged = (torch.arange(1237)[None] < torch.randint(0, 1238, (129, 1), generator=g)).long()  # many narrow-row tiles
    masks.append(ragged)
    for mask in masks:
        B, S = mask.shape
        md = mask.to(dev)
        idx, _, _, _ = ops.unpad_indices(md)
        assert ops.is_marked_ascending(idx)
        n = idx.numel()
        vals = (torch.randn(n, row, generator=g) * 100).to(dtype).to(dev)
        fast = ops.scatter_rows(vals, idx, B * S)
        slow = ops.scatter_rows(vals, idx.clone(), B * S)
        assert torch.equal(fast, slow)
        ref = torch.zeros(B * S, row, dtype=dtype, device=dev)
        ref[idx] = vals
        assert torch.equal(fast, ref)
        out = bert_padding.pad_input(vals, idx, B, S)
        assert torch.equal(out.reshape(B * S, row), ref)


def test_pad_input_after_in_place_edit_of_marked_indices(dev):
    """A marked index tensor that is permuted in place afterwards must fall back to memset + scatter: the
    one-pass gap-fill kernel assumes strictly ascending indices and would drop rows silently."""
    from verl_b200 import ops

    g = torch.Generator().manual_seed(3)
    mask = (torch.rand(8, 257, generator=g) < 0.6).long().to(dev)
    idx, _, _, _ = ops.unpad_indices(mask)
    assert ops.is_marked_ascending(idx)
    idx.copy_(idx.flip(0))  # still a valid scatter target list, no longer ascending
    assert not ops.is_marked_ascending(idx)
    vals = torch.randn(idx.numel(), 24, generator=g).to(dev)
    out = ops.scatter_rows(vals, idx, mask.numel())
    ref = torch.zeros(mask.numel(), 24, device=dev)
    ref[idx] = vals
    assert torch.equal(out, ref)
