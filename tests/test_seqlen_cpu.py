"""CPU: dynamic micro-batch scheduling (verl/utils/seqlen_balancing.py) — the oracle restatement and the HOST C++
scheduler behind the C ABI (vpg_seqlen_balanced_partitions) against golden partitions produced by the reference's own
function bodies (oracle/make_golden.py::gen_seqlen), bit-exact including tie order; plus the Ulysses pad / gather
plumbing over gloo at world_size 2."""

import json
import os
import random
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import load_golden
from oracle import port
from verl_b200 import ops

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cases(key):
    return json.loads(str(load_golden("seqlen")[key]))


def test_oracle_karmarkar_karp_golden():
    n = 0
    for c in _cases("karmarkar_karp"):
        if c["partitions"] == "AssertionError":
            with pytest.raises(AssertionError):
                port.seqlen_balanced_partitions(c["lens"], c["k"], c["equal_size"])
            continue
        assert port.karmarkar_karp(c["lens"], c["k"], c["equal_size"]) == c["merge_order"]
        assert port.seqlen_balanced_partitions(c["lens"], c["k"], c["equal_size"]) == c["partitions"]
        n += 1
    assert n > 50


def test_scheduler_abi_golden():
    """the shipped scheduler (host C++, no GPU involved) == the reference, partition by partition"""
    for c in _cases("karmarkar_karp"):
        if c["partitions"] == "AssertionError":
            with pytest.raises(AssertionError):
                ops.seqlen_balanced_partitions(c["lens"], c["k"], c["equal_size"])
            continue
        assert ops.seqlen_balanced_partitions(c["lens"], c["k"], c["equal_size"]) == c["partitions"]


def test_scheduler_abi_vs_oracle_random_and_errors():
    rnd = random.Random(5)
    for _ in range(150):
        n = rnd.randint(1, 90)
        k = rnd.randint(1, min(n, 12))
        lens = [rnd.randint(1, rnd.choice([4, 50, 30000])) for _ in range(n)]
        for eq in (False, True):
            if eq and n % k:
                continue
            assert ops.seqlen_balanced_partitions(lens, k, eq) == port.seqlen_balanced_partitions(lens, k, eq)
    with pytest.raises(AssertionError):  # seqlen_balancing.py:175
        ops.seqlen_balanced_partitions([1, 2], 3, False)
    with pytest.raises(AssertionError):  # :102
        ops.seqlen_balanced_partitions([1, 2, 3], 2, True)
    # balance property at a production-like size: 4096 sequences into 64 micro-batches
    lens = [rnd.randint(100, 8000) for _ in range(4096)]
    parts = ops.seqlen_balanced_partitions(lens, 64, False)
    sums = [sum(lens[i] for i in p) for p in parts]
    assert sorted(i for p in parts for i in p) == list(range(4096))
    assert max(sums) - min(sums) <= 8000


def test_oracle_micro_batches_golden():
    for c in _cases("micro_batches"):
        pos = torch.arange(c["S"])[None]
        left, lens = torch.tensor(c["left"]), torch.tensor(c["lens"])
        attn = ((pos >= left[:, None]) & (pos < (left + lens)[:, None])).long()
        got = port.micro_batch_indices(attn, c["max_token_len"], c["min_num_micro_batch"], c["num_batches_divided_by"],
                                       c["use_dynamic_bsz_balance"])
        assert got == c["micro_bsz_idx"]
        flat = [i for p in got for i in p]
        assert port.reverse_idx(flat) == c["reverse_idx"]
        x = torch.arange(c["B"] * 3).view(c["B"], 3)
        assert torch.equal(port.restore_dynamic_batch(x[flat], got), x)


def test_host_helpers():
    from verl_b200.utils import seqlen_balancing as sb

    assert sb.ceildiv(7, 3) == 3 and sb.roundup_divisible(7, 4) == 8
    assert sb.get_reverse_idx([2, 0, 1]) == [1, 2, 0]
    lens = [5, 1, 9, 3, 7, 2]
    parts = sb.get_seqlen_balanced_partitions(lens, 2, True)
    stats = sb.log_seqlen_unbalance(lens, parts, "x")
    assert stats["x/balanced_max"] - stats["x/balanced_min"] <= stats["x/minmax_diff"]
    assert sb.greedy_partition([3, 1, 2], 3, True) == [[0], [1], [2]]


# ------------------------------------------------------------------------------------------------ ulysses over gloo
def _ulysses_worker(rank, world, port_no, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port_no)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from verl_b200.utils import ulysses as U

        U.set_ulysses_sequence_parallel_group(dist.group.WORLD)
        T = 37  # not a multiple of the group size
        ids = torch.arange(100, 100 + T).view(1, T)
        pos = torch.arange(T).view(1, T)
        mrope = pos.unsqueeze(0).repeat(3, 1, 1)
        ids_l, pos_l, pad = U.ulysses_pad_and_slice_inputs(ids, pos, sp_size=world)
        assert pad == (-T) % world and ids_l.shape == (1, (T + pad) // world) == pos_l.shape
        _, mrope_p, _ = U.ulysses_pad(ids, mrope, sp_size=world)
        assert mrope_p.shape == (3, 1, T + pad) and mrope_p[:, :, T:].tolist() == [[list(range(pad))]] * 3
        # per-token outputs of this rank's slice -> gathered and un-padded stream (dp_actor.py:206-219)
        local = (ids_l.squeeze(0).float() * 2).requires_grad_(True)
        full = U.gather_outputs_and_unpad(local, gather_dim=0, unpad_dim=0, padding_size=pad)
        assert torch.equal(full.detach(), ids.squeeze(0).float() * 2)
        w = torch.arange(1, T + 1).float()
        (full * w).sum().backward()
        per = (T + pad) // world
        want = torch.zeros(per)
        n_real = max(0, min(T - rank * per, per))
        want[:n_real] = w[rank * per:rank * per + n_real] * world  # grad_scaler multiplies by the group size
        assert torch.equal(local.grad, want)
        # gather along a non-leading axis
        x = torch.full((2, 3), float(rank))
        y = U.gather_outputs_and_unpad(x, gather_dim=1, grad_scaler=False)
        assert y.shape == (2, 3 * world) and y[0].tolist() == [float(r) for r in range(world) for _ in range(3)]
        # the plain axis-0 gather the reference also exports (verl/utils/ulysses.py:155-162)
        z = U.all_gather_tensor(torch.full((2, 3), float(rank)))
        assert z.shape == (2 * world, 3) and z[:, 0].tolist() == [float(r) for r in range(world) for _ in range(2)]
        q.put((rank, "ok"))
    except Exception as e:
        q.put((rank, f"{type(e).__name__}: {e}"))
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_ulysses_pad_slice_gather_world2():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port_no = 31500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_ulysses_worker, args=(r, world, port_no, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == "ok", f"rank {rank}: {msg}"


def test_ulysses_without_group_is_identity():
    from verl_b200.utils import ulysses as U

    U.set_ulysses_sequence_parallel_group(None)
    x = torch.arange(5.0)
    assert U.gather_outputs_and_unpad(x, gather_dim=0, unpad_dim=0, padding_size=2) is x
    ids = torch.arange(7).view(1, 7)
    assert U.ulysses_pad(ids, None, sp_size=1) == (ids, None, 0)
    with pytest.raises(AssertionError):
        U.validate_ulysses_config(28, 8)
    with pytest.raises(RuntimeError, match="gather_outputs_and_unpad"):
        U.gather_outpus_and_unpad(x, 0)
