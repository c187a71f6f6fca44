"""CPU: oracle restatements of the Megatron-side equivalents against golden vectors produced by the
reference's own class / function bodies (oracle/make_golden.py::gen_megatron)."""

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import port


def T(a):
    return torch.from_numpy(np.asarray(a))


@pytest.mark.parametrize("tp", [1, 2, 4, 8])
def test_vocab_parallel_entropy_and_logprobs_golden(tp):
    g = load_golden("megatron")
    logits, labels = T(g["vp_logits"]), T(g["vp_labels"])
    shards = list(logits.chunk(tp, dim=-1))
    torch.testing.assert_close(port.vocab_parallel_entropy(shards), T(g["vp_entropy"]), rtol=1e-5, atol=2e-5)
    torch.testing.assert_close(port.vocab_parallel_log_probs(shards, labels), T(g["vp_logprobs"]), rtol=1e-5, atol=2e-5)
    # gradients of the restatement (autograd) == the reference class's hand-written backward
    x = logits.clone().requires_grad_(True)
    port.vocab_parallel_entropy(list(x.chunk(tp, dim=-1))).backward(T(g["vp_g_ent"]))
    torch.testing.assert_close(x.grad, T(g["vp_entropy_dlogits"]), rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("tp,cp", [(1, 1), (2, 1), (4, 1), (1, 2), (2, 2), (1, 4)])
def test_packed_seqs_golden(tp, cp):
    g = load_golden("megatron")
    ids, attn, table = T(g["pk_ids"]), T(g["pk_attn"]), T(g["pk_table"])
    per_rank = []
    for r in range(cp):
        packed, cu_padded, max_len = port.preprocess_packed_seqs(ids, attn, tp, cp, r)
        assert torch.equal(packed, T(g[f"pk|tp{tp}cp{cp}|r{r}|packed"]))
        assert torch.equal(cu_padded, T(g[f"pk|tp{tp}cp{cp}|cu_padded"])) and max_len == int(g[f"pk|tp{tp}cp{cp}|max_seqlen"])
        per_rank.append(table[packed[0]].unsqueeze(0))
    padded = port.postprocess_packed_seqs(per_rank, attn, tp, cp)
    for r in range(cp):
        assert torch.equal(padded, T(g[f"pk|tp{tp}cp{cp}|r{r}|padded"]))
    # the packed round trip reproduces the row lookup at every valid position, zeros elsewhere
    want = table[ids] * attn.bool()[..., None]
    assert torch.equal(padded, want)
