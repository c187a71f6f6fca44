"""GPU: dynamic micro-batches through the reference-named API (verl_b200.utils.seqlen_balancing) — row lengths and the
scheduled-order row gathers run on the device; indices bit-exact against the reference goldens and the oracle; the
round trip the reference's own tests check (tests/utils/test_seqlen_balancing.py:31-63)."""

import json

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port
from verl_b200 import ops
from verl_b200.utils import seqlen_balancing as sb

pytestmark = pytest.mark.gpu


def _attn(c):
    pos = torch.arange(c["S"])[None]
    left, lens = torch.tensor(c["left"]), torch.tensor(c["lens"])
    return ((pos >= left[:, None]) & (pos < (left + lens)[:, None])).long()


@pytest.mark.parametrize("mask_dtype", [torch.int64, torch.int32, torch.bool])
def test_row_lengths_exact(dev, mask_dtype):
    torch.manual_seed(0)
    for B, S in [(1, 1), (5, 31), (40, 512), (16, 16384), (3, 70001)]:
        m = (torch.rand(B, S) > 0.37).to(mask_dtype)
        got = ops.row_lengths(m.to(dev))
        assert got.dtype == torch.int64 and torch.equal(got.cpu(), m.long().sum(dim=1))


def test_rearrange_micro_batches_golden(dev):
    for c in json.loads(str(load_golden("seqlen")["micro_batches"])):
        attn = _attn(c)
        ids = torch.randint(0, 1000, attn.shape)
        feats = torch.randn(c["B"], 3, 5)
        batch = {"input_ids": ids.to(dev), "attention_mask": attn.to(dev), "feats": feats.to(dev)}
        mbs, idx = sb.rearrange_micro_batches(batch, c["max_token_len"], num_batches_divided_by=c["num_batches_divided_by"],
                                              min_num_micro_batch=c["min_num_micro_batch"],
                                              use_dynamic_bsz_balance=c["use_dynamic_bsz_balance"])
        assert idx == c["micro_bsz_idx"]
        for mb, rows in zip(mbs, idx):
            assert torch.equal(mb["input_ids"].cpu(), ids[rows])
            assert torch.equal(mb["attention_mask"].cpu(), attn[rows])
            assert torch.equal(mb["feats"].cpu(), feats[rows])
        flat = torch.cat([mb["input_ids"] for mb in mbs])
        assert torch.equal(sb.restore_dynamic_batch(flat, idx).cpu(), ids)
        assert sb.get_reverse_idx([i for p in idx for i in p]) == c["reverse_idx"]


def test_prepare_dynamic_batch_round_trip(dev):
    """the reference's test_dynamic_batch: 20 x 100, max_token_len 300"""
    torch.manual_seed(1)
    B, S = 20, 100
    lens = torch.randint(50, 91, (B,))
    left = torch.randint(0, 11, (B,))
    pos = torch.arange(S)[None]
    attn = ((pos >= left[:, None]) & (pos < (left + lens)[:, None])).long()
    ids = torch.randint(0, 10, (B, S))

    class Data:
        batch = {"input_ids": ids.to(dev), "attention_mask": attn.to(dev)}
        non_tensor_batch = {"uid": np.array([f"u{i}" for i in range(B)], dtype=object)}

    mbs, idx = sb.prepare_dynamic_batch(Data, max_token_len=300)
    assert idx == port.micro_batch_indices(attn, 300)
    assert sum(len(m) for m in mbs) == B
    for m, rows in zip(mbs, idx):
        assert list(m.non_tensor_batch["uid"]) == [f"u{i}" for i in rows]
        assert int(m.batch["attention_mask"].sum()) <= 300 + S  # balanced, not a hard cap (same in the reference)
    out = sb.restore_dynamic_batch(torch.cat([m.batch["input_ids"] for m in mbs], dim=0), idx)
    assert torch.equal(out.cpu(), ids)
    # per-token outputs (fp32 [B, L]) restore the same way
    lp = torch.randn(B, 17)
    perm = [i for p in idx for i in p]
    assert torch.equal(sb.restore_dynamic_batch(lp[perm].to(dev), idx).cpu(), lp)


def test_max_token_len_assertion(dev):
    attn = torch.ones(4, 64, dtype=torch.long, device=dev)
    with pytest.raises(AssertionError):
        sb.rearrange_micro_batches({"attention_mask": attn}, max_token_len=32)


def test_large_batch_properties(dev):
    """C5-like scheduling: 1024 sequences up to 16k tokens, 256k tokens per micro-batch"""
    torch.manual_seed(3)
    B, S = 1024, 16384
    lens = torch.randint(256, S + 1, (B,), device=dev)
    attn = (torch.arange(S, device=dev)[None] < lens[:, None]).to(torch.int32)
    resp = torch.arange(B, device=dev, dtype=torch.float32)[:, None].repeat(1, 4)
    mbs, idx = sb.rearrange_micro_batches({"attention_mask": attn, "resp": resp}, max_token_len=262144)
    flat = [i for p in idx for i in p]
    assert sorted(flat) == list(range(B)) and all(p == sorted(p) for p in idx)
    lens_h = lens.tolist()
    sums = [sum(lens_h[i] for i in p) for p in idx]
    assert len(idx) == -(sum(lens_h) // -262144)
    assert max(sums) - min(sums) <= S  # the differencing method leaves at most one sequence of spread
    sq = [sum(lens_h[i] ** 2 for i in p) for p in idx]
    assert sq == sorted(sq, reverse=True)
    assert torch.equal(torch.cat([m["resp"][:, 0] for m in mbs]).cpu(), torch.tensor(flat, dtype=torch.float32))
    assert torch.equal(sb.restore_dynamic_batch(torch.cat([m["resp"] for m in mbs]), idx), resp)
