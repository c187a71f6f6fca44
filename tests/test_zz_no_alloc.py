"""vpg_grpo_advantage_static: the GRPO advantage for hosts that cannot give the library scratch memory.

Who needs it: a C / ctypes host that runs without PyTorch's caching allocator (it could only get scratch from
cudaMalloc / cudaFree around the call, and those stall every stream on the GPU, including the rollout engine that shares
it), and a trainer whose HBM was handed to another pool up front, where an allocation in the middle of a step fails.
For them the library carries its own scratch: one fixed array per GPU (576.25 KiB) that is part of the loaded library
image, good for up to VPG_GRPO_STATIC_MAX_ROWS = 16384 rows.  The array is shared by all calls and is not wiped between
them, so this file checks the things a user has to be able to rely on:

  * called straight through the C ABI, the result equals vpg_grpo_advantage's (which gets a fresh torch workspace), on
    two calls in a row where the second call is smaller and differently grouped than the first, i.e. it runs on top of
    the first call's leftovers;
  * the call does not go to the device allocator behind the caller's back (torch's counters do not move, and the
    driver's free-memory figure does not drop);
  * a batch beyond the limit is refused with a message that names the limit, and nothing is launched.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must not hide the parity
tests of the other kernels.  It passes on its own in a fresh process (the first call then also does the one-time address
look-up on the device) and after the rest of the suite.
"""

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ROWS = 16384  # VPG_GRPO_STATIC_MAX_ROWS in include/vpg_b200.h


def _case(B, L, n, seed, dev):
    """Rewards on the last valid token, ragged response mask, shuffled group ids (groups of n, last one may be short)."""
    g = torch.Generator().manual_seed(seed)
    lens = torch.randint(max(1, L // 4), L + 1, (B,), generator=g)
    mask = (torch.arange(L)[None] < lens[:, None]).long()
    rewards = torch.zeros(B, L)
    rewards[torch.arange(B), lens - 1] = torch.randint(0, 5, (B,), generator=g).float() / 4
    gids = (torch.arange(B) // n)[torch.randperm(B, generator=g)].to(torch.int32)
    G = (B + n - 1) // n
    return rewards.to(dev), mask.to(dev), gids.to(dev), G


def _both(lib, check, ptr, stream_ptr, rewards, mask, gids, G, norm_by_std=1):
    """(static-scratch result, workspace result) of one GRPO advantage, both through the C ABI on torch's stream."""
    B, L = rewards.shape
    stream = stream_ptr(rewards.device)
    # outputs start as NaN: every element has to be written by the call
    got = torch.full_like(rewards, float("nan"))
    want = torch.full_like(rewards, float("nan"))
    ws_bytes = int(lib.vpg_grpo_workspace_bytes(B, G))
    ws = torch.empty((ws_bytes + 7) // 8, dtype=torch.float64, device=rewards.device)
    check(lib.vpg_grpo_advantage(ptr(rewards), ptr(mask), 0, ptr(gids), B, L, G, 1e-6, norm_by_std, ptr(want), ptr(ws),
                                 ws.numel() * 8, stream))
    torch.cuda.synchronize()
    before = (torch.cuda.memory_stats(rewards.device)["allocation.all.allocated"],
              torch.cuda.memory_reserved(rewards.device), torch.cuda.mem_get_info(rewards.device)[0])
    check(lib.vpg_grpo_advantage_static(ptr(rewards), ptr(mask), 0, ptr(gids), B, L, G, 1e-6, norm_by_std, ptr(got),
                                        stream))
    torch.cuda.synchronize()
    after = (torch.cuda.memory_stats(rewards.device)["allocation.all.allocated"],
             torch.cuda.memory_reserved(rewards.device), torch.cuda.mem_get_info(rewards.device)[0])
    assert after[0] == before[0] and after[1] == before[1], "the call went through torch's allocator"
    assert after[2] >= before[2], "free device memory dropped across the call"
    return got, want


def test_static_scratch_matches_workspace_entry_point_twice_in_a_row(dev):
    from verl_b200 import _lib

    lib = _lib.load()
    # first call: the largest batch the entry point takes, singleton-heavy tail, multi-block group kernel
    # second call: a small batch (C1 of BASELINE.json: 8 prompts x 5 rollouts x 512) on top of the first call's leftovers
    # third call: pairs, mean-only normalisation, odd L (scalar path)
    for B, L, n, seed, norm in [(MAX_ROWS, 64, 8, 1, 1), (40, 512, 5, 2, 1), (4097, 33, 2, 3, 0)]:
        rewards, mask, gids, G = _case(B, L, n, seed, dev)
        got, want = _both(lib, _lib.check, _lib.ptr, _lib.stream_ptr, rewards, mask, gids, G, norm)
        assert torch.isfinite(want).all()
        # tolerance of the GRPO parity tests (tests/test_token_ops_gpu.py): 1e-5 relative, 1e-6 absolute
        torch.testing.assert_close(got.cpu(), want.cpu(), rtol=1e-5, atol=1e-6)
        assert torch.equal(got == 0, want == 0)  # masked-out tokens are exactly zero in both


def test_static_scratch_same_batch_twice_gives_the_same_answer(dev):
    """The second of two identical calls sees the first one's sums in the array; it must clear them, not add to them."""
    from verl_b200 import _lib

    lib = _lib.load()
    rewards, mask, gids, G = _case(128, 1024, 8, 7, dev)
    first, want = _both(lib, _lib.check, _lib.ptr, _lib.stream_ptr, rewards, mask, gids, G)
    second, _ = _both(lib, _lib.check, _lib.ptr, _lib.stream_ptr, rewards, mask, gids, G)
    torch.testing.assert_close(first.cpu(), want.cpu(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(second.cpu(), want.cpu(), rtol=1e-5, atol=1e-6)


def test_static_scratch_refuses_batches_beyond_its_limit(dev):
    from verl_b200 import _lib

    lib = _lib.load()
    B, L = MAX_ROWS + 1, 4
    rewards, mask, gids, G = _case(B, L, 8, 11, dev)
    out = torch.full_like(rewards, float("nan"))
    launches = _lib.launch_count()
    rc = lib.vpg_grpo_advantage_static(_lib.ptr(rewards), _lib.ptr(mask), 0, _lib.ptr(gids), B, L, G, 1e-6, 1,
                                       _lib.ptr(out), _lib.stream_ptr(dev))
    assert rc == -3  # VPG_ERR_WORKSPACE
    msg = lib.vpg_last_error().decode()
    assert str(MAX_ROWS) in msg and "vpg_grpo_advantage" in msg
    assert _lib.launch_count() == launches
    torch.cuda.synchronize()
    assert torch.isnan(out).all()  # untouched
    with pytest.raises(RuntimeError):
        _lib.check(rc)


def test_header_limit_matches(dev):
    import os
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "include", "vpg_b200.h")).read()
    assert int(re.search(r"#define\s+VPG_GRPO_STATIC_MAX_ROWS\s+(\d+)", text).group(1)) == MAX_ROWS
