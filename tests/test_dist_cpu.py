"""CPU, world_size 2 over gloo: the vocab-parallel exchange of the fused lm-head (verl_b200.utils.kernel.kernels.
exchange_partials) — each rank builds the partials of its vocab shard (oracle restatement of the kernel's format),
ONE all-gather packs them, and the merge reproduces the full-vocab result of the reference golden vectors."""

import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port_no, result_q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port_no)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import port  # noqa: I001
        from verl_b200.utils.kernel.kernels import exchange_partials

        g = np.load(os.path.join(ROOT, "tests", "golden", "lce.npz"))
        hidden = torch.from_numpy(g["c_hidden"]).bfloat16()
        weight = torch.from_numpy(g["c_weight"]).bfloat16()
        labels = torch.from_numpy(g["c_labels"])
        v_local = weight.shape[0] // world
        shard = weight[rank * v_local:(rank + 1) * v_local]
        part = port.lce_partials(hidden, shard, labels, 1.0, label_offset=rank * v_local, num_splits=2)
        merged = exchange_partials(part, dist.group.WORLD)
        assert merged.shape == (4, 2 * world, hidden.shape[0])
        lp, ent = port.lce_finalize(merged, 1.0)
        err = max((lp - torch.from_numpy(g["c_logprobs"])).abs().max().item(),
                  (ent - torch.from_numpy(g["c_entropy"])).abs().max().item())
        result_q.put((rank, err))
    except Exception as e:  # surface the failure instead of letting the parent wait for its timeout
        result_q.put((rank, f"{type(e).__name__}: {e}"))
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_vocab_parallel_exchange_world2():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port_no = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, world, port_no, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, err in results:
        assert not isinstance(err, str), f"rank {rank} failed: {err}"
        assert err < 5e-5, f"rank {rank}: {err}"
