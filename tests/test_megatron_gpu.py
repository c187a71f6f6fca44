"""GPU: Megatron-side equivalents through the reference-named API against the reference goldens and
the oracle — vocab-parallel entropy / log-probs on materialised logit shards (ranks emulated serially on one GPU with
the real kernels, plus the single-rank public functions), packed-sequence pre/post-processing for tp x cp layouts."""

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port
from verl_b200 import ops
from verl_b200.models.mcore import util as mutil
from verl_b200.utils.megatron import tensor_parallel as tpmod

pytestmark = pytest.mark.gpu


def T(a):
    return torch.from_numpy(np.asarray(a))


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_vocab_parallel_partials_emulated(world, dev):
    """every rank's shard through vpg_logits_partials, concatenated like the all-gather, merged by vpg_lce_finalize"""
    g = load_golden("megatron")
    logits, labels = T(g["vp_logits"]).to(dev), T(g["vp_labels"]).to(dev)
    v_local = logits.shape[1] // world
    parts = [ops.logits_partials(logits[:, r * v_local:(r + 1) * v_local], labels, r * v_local) for r in range(world)]
    lp, ent, lse, eb = ops.lce_finalize(torch.cat(parts, dim=1), 1.0)
    torch.testing.assert_close(ent.cpu(), T(g["vp_entropy"]), rtol=1e-5, atol=2e-5)
    torch.testing.assert_close(lp.cpu(), T(g["vp_logprobs"]), rtol=1e-5, atol=2e-5)
    # backward of each shard from the merged per-token scalars == the matching columns of the reference gradients
    g_ent, g_lp = T(g["vp_g_ent"]).to(dev), T(g["vp_g_lp"]).to(dev)
    for r in range(world):
        x = logits[:, r * v_local:(r + 1) * v_local].contiguous()
        d_ent = ops.logits_stats_backward(x, None, lse, ent, None, g_ent)
        d_lp = ops.logits_stats_backward(x, labels - r * v_local, lse, None, g_lp, None)
        torch.testing.assert_close(d_ent.cpu(), T(g["vp_entropy_dlogits"])[:, r * v_local:(r + 1) * v_local],
                                   rtol=1e-4, atol=1e-6)
        torch.testing.assert_close(d_lp.cpu(), T(g["vp_logprobs_dlogits"])[:, r * v_local:(r + 1) * v_local],
                                   rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_public_functions_single_rank(dev, dtype):
    g = load_golden("megatron")
    logits = T(g["vp_logits"]).to(dtype)
    labels = T(g["vp_labels"])
    x = logits.to(dev).requires_grad_(True)
    ent = tpmod.vocab_parallel_entropy(x)
    lp = tpmod.vocab_parallel_log_probs_from_logits(x, labels.to(dev))
    (ent * T(g["vp_g_ent"]).to(dev)).sum().backward(retain_graph=True)
    d_ent = x.grad.clone()
    x.grad = None
    (lp * T(g["vp_g_lp"]).to(dev)).sum().backward()
    xo = logits.float().requires_grad_(True)
    ent_o = port.vocab_parallel_entropy([xo])
    lp_o = port.vocab_parallel_log_probs([xo], labels)
    tol = dict(rtol=1e-5, atol=2e-5) if dtype == torch.float32 else dict(rtol=0, atol=2e-3)
    torch.testing.assert_close(ent.cpu(), ent_o.detach(), **tol)
    torch.testing.assert_close(lp.cpu(), lp_o.detach(), **tol)
    ent_o.backward(T(g["vp_g_ent"]), retain_graph=True)
    gtol = dict(rtol=1e-4, atol=1e-6) if dtype == torch.float32 else dict(rtol=2e-2, atol=2e-3)
    torch.testing.assert_close(d_ent.float().cpu(), xo.grad, **gtol)
    # 3-D logits keep their leading shape
    e3 = tpmod.vocab_parallel_entropy(logits[:20].view(4, 5, -1).to(dev))
    assert e3.shape == (4, 5)
    torch.testing.assert_close(e3.reshape(-1).cpu(), ent_o.detach()[:20], **tol)


def test_response_rmpad_composition(dev):
    """vocab_parallel_log_probs_from_logits_response_rmpad == pad back + response window of the oracle log-probs"""
    torch.manual_seed(4)
    B, S, V, R = 5, 24, 300, 7
    lens = torch.randint(R + 2, S + 1, (B,))
    attn = (torch.arange(S)[None] >= (S - lens)[:, None]).long()  # left-padded
    ids = torch.randint(0, V, (B, S))
    idx, _, _, _ = port.unpad_indices(attn)
    logits_rmpad = torch.randn(idx.numel(), V) * 2
    got = tpmod.vocab_parallel_log_probs_from_logits_response_rmpad(ids.to(dev), attn.to(dev), logits_rmpad.to(dev), R)
    rolled = torch.roll(ids.reshape(-1)[idx], -1, 0)
    lp = port.vocab_parallel_log_probs([logits_rmpad], rolled)
    want = port.pad_rows(lp[:, None], idx, B * S).view(B, S)[:, -R - 1:-1]
    torch.testing.assert_close(got.cpu(), want, rtol=1e-5, atol=2e-5)


@pytest.mark.parametrize("tp,cp", [(1, 1), (2, 1), (4, 1), (1, 2), (2, 2), (1, 4)])
def test_packed_seqs_golden(tp, cp, dev):
    g = load_golden("megatron")
    ids, attn, table = T(g["pk_ids"]), T(g["pk_attn"]), T(g["pk_table"])
    B, S = ids.shape
    per_rank, params = [], []
    for r in range(cp):
        packed, p = mutil.preprocess_packed_seqs(ids.to(dev), attn.bool().to(dev), tp_size=tp, cp_size=cp, cp_rank=r)
        assert torch.equal(packed.cpu(), T(g[f"pk|tp{tp}cp{cp}|r{r}|packed"]))
        assert torch.equal(p.cu_seqlens_q_padded.cpu(), T(g[f"pk|tp{tp}cp{cp}|cu_padded"]))
        assert p.max_seqlen_q == int(g[f"pk|tp{tp}cp{cp}|max_seqlen"]) and p.qkv_format == "thd"
        per_rank.append(table.to(dev)[packed[0]])
        params.append(p)
    # post-processing: the cp all-gather is emulated by handing the kernel the rank-major concatenation
    gathered = torch.cat(per_rank, dim=0)
    w = (torch.arange(B * S * table.shape[1], dtype=torch.float32).view(B, S, -1) / 100).to(dev)
    for r in range(cp):
        lay = params[r]._vpg
        padded = ops.permute_rows(gathered, lay["gathered_pos"], lay["indices"], B * S).view(B, S, -1)
        assert torch.equal(padded.cpu(), T(g[f"pk|tp{tp}cp{cp}|r{r}|padded"]))
        d_local = ops.permute_rows(w.view(B * S, -1), lay["indices"], lay["local_dst"], per_rank[r].shape[0])
        torch.testing.assert_close(d_local.cpu(), T(g[f"pk|tp{tp}cp{cp}|r{r}|d_output"])[0], rtol=0, atol=0)
    if cp == 1:  # the public round trip, with autograd
        o = per_rank[0].unsqueeze(0).clone().requires_grad_(True)
        padded = mutil.postprocess_packed_seqs(o, params[0], attn.bool().to(dev), B, S)
        assert torch.equal(padded.detach().cpu(), T(g[f"pk|tp{tp}cp{cp}|r0|padded"]))
        (padded * w).sum().backward()
        torch.testing.assert_close(o.grad.cpu(), T(g[f"pk|tp{tp}cp{cp}|r0|d_output"]), rtol=0, atol=0)
        # params from elsewhere (no cached maps): rebuilt from the mask
        bare = mutil.PackedSeqParams(cu_seqlens_q_padded=params[0].cu_seqlens_q_padded)
        again = mutil.postprocess_packed_seqs(o.detach(), bare, attn.to(dev), B, S, tp_size=tp, cp_size=1, cp_rank=0)
        assert torch.equal(again, padded.detach())
        assert mutil.postprocess_packed_seqs(o, params[0], attn, B, S, post_process=False) is o


def test_packed_seqs_large_round_trip(dev):
    """C5-like: 16 x 16384 tokens, hidden rows of 5120 bf16 — pack -> unpack is the identity on valid tokens"""
    torch.manual_seed(8)
    B, S = 16, 16384
    lens = torch.randint(S // 4, S + 1, (B,), device=dev)
    attn = torch.arange(S, device=dev)[None] < lens[:, None]
    ids = torch.randint(0, 152064, (B, S), device=dev)
    for tp, cp, r in [(8, 1, 0), (2, 4, 3)]:
        packed, p = mutil.preprocess_packed_seqs(ids, attn, tp_size=tp, cp_size=cp, cp_rank=r)
        want, cu_padded, mx = port.preprocess_packed_seqs(ids.cpu(), attn.cpu(), tp, cp, r) if cp == 1 else (None, None, None)
        if want is not None:
            assert torch.equal(packed.cpu(), want) and torch.equal(p.cu_seqlens_q.cpu(), cu_padded) and p.max_seqlen_q == mx
        lay = p._vpg
        total = lay["total_padded"]
        assert total % (tp * cp * (2 if cp > 1 else 1)) == 0 and packed.shape == (1, total // cp)
        owned = lay["local_dst"] >= 0
        assert torch.equal(packed[0][lay["local_dst"][owned]], ids.reshape(-1)[lay["indices"][owned]])
        # every slot of the gathered stream is hit at most once; ranks own near-equal shares (causal load balance)
        assert lay["gathered_pos"].unique().numel() == lay["gathered_pos"].numel()
        assert abs(int(owned.sum()) - lay["indices"].numel() / cp) <= B * tp * cp * 2
    h = torch.randn(B * S // 8, 5120, device=dev).bfloat16()  # hidden-sized rows through the same permute kernel
    perm = torch.randperm(h.shape[0], device=dev)
    assert torch.equal(ops.permute_rows(h, perm, None, h.shape[0]), h[perm])
    assert torch.equal(ops.permute_rows(h, None, perm, h.shape[0])[perm], h)
