"""Dynamic micro-batches: the scheduling that decides which sequences share an rmpad token stream
(mirror of verl/utils/seqlen_balancing.py).

    karmarkar_karp / get_seqlen_balanced_partitions   :26-190   HOST C++ scheduler behind the C ABI
                                                                (vpg_seqlen_balanced_partitions; bit-identical
                                                                partitions, heapq tie order included)
    rearrange_micro_batches                           :250-319  row lengths on the device (vpg_row_lengths), ONE
                                                                row gather per field into the scheduled order
                                                                (vpg_gather_rows); micro-batches are views of it
    get_reverse_idx / prepare_dynamic_batch / restore_dynamic_batch   :322-375

The reference slices the TensorDict row by row and `torch.cat`s every micro-batch (B small copies per field); here
every field is permuted once and split into views.  A "batch" is any mapping of name -> tensor whose first axis is
the sequence axis and that has an "attention_mask" [B,S] entry (TensorDict is not a dependency).
"""

from __future__ import annotations

from itertools import chain
from typing import Any, Mapping

import torch
from torch import distributed as dist

from .. import ops


def karmarkar_karp(seqlen_list: list[int], k_partitions: int, equal_size: bool):
    """seqlen_balancing.py:26-127.  The reference returns each partition in merge order; callers only ever see the
    sorted lists of get_seqlen_balanced_partitions, which is what the scheduler emits."""
    return ops.seqlen_balanced_partitions(seqlen_list, k_partitions, equal_size)


def greedy_partition(seqlen_list: list[int], k_partitions: int, equal_size: bool):
    """seqlen_balancing.py:130-147 (unused by the reference's own callers; kept for the name)."""
    bias = sum(seqlen_list) + 1 if equal_size else 0
    partitions = [[] for _ in range(k_partitions)]
    sums = [0] * k_partitions
    for i, seqlen in enumerate(seqlen_list):
        j = min(range(k_partitions), key=lambda p: (sums[p], p))
        partitions[j].append(i)
        sums[j] += seqlen + bias
    if equal_size:
        for p in partitions:
            assert len(p) * k_partitions == len(seqlen_list), f"{len(p)} * {k_partitions} != {len(seqlen_list)}"
    return partitions


def get_seqlen_balanced_partitions(seqlen_list: list[int], k_partitions: int, equal_size: bool):
    """seqlen_balancing.py:150-190: k index lists (each sorted) with balanced length sums; AssertionError when
    len(seqlen_list) < k_partitions, when equal_size does not divide, or when a partition would be empty."""
    return ops.seqlen_balanced_partitions(seqlen_list, k_partitions, equal_size)


def log_seqlen_unbalance(seqlen_list: list[int], partitions: list[list[int]], prefix):
    """seqlen_balancing.py:194-239: imbalance statistics of a partitioning against the naive contiguous split
    (consecutive runs of len // k sequences)."""
    k = len(partitions)
    run = len(seqlen_list) // k
    naive = [sum(seqlen_list[lo:lo + run]) for lo in range(0, len(seqlen_list), run)]
    scheduled = [sum(seqlen_list[i] for i in part) for part in partitions]
    lo, hi = min(naive), max(naive)
    return {
        f"{prefix}/min": lo,
        f"{prefix}/max": hi,
        f"{prefix}/minmax_diff": hi - lo,
        f"{prefix}/balanced_min": min(scheduled),
        f"{prefix}/balanced_max": max(scheduled),
        f"{prefix}/mean": sum(naive) / k,
    }


def ceildiv(a, b):
    return -(a // -b)


def roundup_divisible(a, b):
    return ceildiv(a, b) * b


def _fields(batch) -> dict:
    if hasattr(batch, "items"):
        return {k: v for k, v in batch.items() if isinstance(v, torch.Tensor)}
    raise TypeError("batch must be a mapping of name -> tensor with an 'attention_mask' entry")


def _micro_batch_count(lengths, max_token_len, device, dp_group, same_micro_num_in_dp, min_num_micro_batch,
                       num_batches_divided_by) -> int:
    """How many micro-batches the mini-batch is cut into (seqlen_balancing.py:281-296): enough for `max_token_len`
    tokens each, never more than sequences, at least what a pipeline-parallel caller asks for, the maximum over the
    data-parallel group (every rank must step the same number of times), rounded up to the caller's multiple."""
    count = min(len(lengths), ceildiv(sum(lengths), max_token_len))
    if min_num_micro_batch is not None:
        count = max(count, min_num_micro_batch)
    if same_micro_num_in_dp and dist.is_initialized():
        agreed = torch.tensor([count], device=device)
        dist.all_reduce(agreed, op=dist.ReduceOp.MAX, group=dp_group)
        count = int(agreed.cpu().item())
    if num_batches_divided_by is not None:
        count = roundup_divisible(count, num_batches_divided_by)
    assert count <= len(lengths)
    return count


def rearrange_micro_batches(
    batch,
    max_token_len,
    dp_group=None,
    num_batches_divided_by=None,
    same_micro_num_in_dp=True,
    min_num_micro_batch=None,
    use_dynamic_bsz_balance=True,
):
    """seqlen_balancing.py:250-319.  -> (micro_batches: list of {name: tensor view}, micro_bsz_idx: list of index lists).

    One device->host read (the row lengths; the reference reads the same numbers with `.item()` / `.tolist()`)."""
    fields = _fields(batch)
    attention_mask = fields["attention_mask"]
    max_seq_len = attention_mask.shape[-1]
    assert max_token_len >= max_seq_len, (
        f"max_token_len must be greater than the sequence length. Got {max_token_len=} and {max_seq_len=}"
    )
    lengths = ops.row_lengths(attention_mask).tolist()
    count = _micro_batch_count(lengths, max_token_len, attention_mask.device, dp_group, same_micro_num_in_dp,
                               min_num_micro_batch, num_batches_divided_by)
    micro_bsz_idx = get_seqlen_balanced_partitions(lengths, count, equal_size=False)
    if use_dynamic_bsz_balance:
        # heaviest micro-batch first, weight = sum of squared lengths ~ attention work (:298-306); ties by first index

        def work(part):
            return sum(lengths[i] * lengths[i] for i in part), (min(part) if part else 0)

        micro_bsz_idx.sort(key=work, reverse=True)
    order = torch.tensor(list(chain.from_iterable(micro_bsz_idx)), dtype=torch.int64).to(
        attention_mask.device, non_blocking=True)
    sizes = [len(p) for p in micro_bsz_idx]
    permuted = {k: ops.gather_rows(v, order).split(sizes) for k, v in fields.items()}
    micro_batches = [{k: permuted[k][i] for k in permuted} for i in range(len(sizes))]
    return micro_batches, micro_bsz_idx


def get_reverse_idx(idx_map):
    """seqlen_balancing.py:322-337: the inverse permutation, out[idx_map[i]] = i (a new list)."""
    out = list(idx_map)
    for position, source in enumerate(idx_map):
        out[source] = position
    return out


class MicroBatch:
    """What prepare_dynamic_batch hands to the worker loop: the `.batch` / `.non_tensor_batch` / `.meta_info` triple of
    the reference's DataProto (verl/protocol.py), nothing else of it (DataProto itself is out of scope)."""

    __slots__ = ("batch", "non_tensor_batch", "meta_info")

    def __init__(self, batch: dict, non_tensor_batch: dict | None = None, meta_info: dict | None = None):
        self.batch = batch
        self.non_tensor_batch = non_tensor_batch or {}
        self.meta_info = meta_info or {}

    def __len__(self):
        return next(iter(self.batch.values())).shape[0] if self.batch else 0


def prepare_dynamic_batch(data: Any, max_token_len: int):
    """seqlen_balancing.py:340-359: `data` has `.batch` (mapping of tensors) and `.non_tensor_batch` (mapping of numpy
    object arrays).  -> (list of MicroBatch, batch_idx_list)."""
    batch, batch_idx_list = rearrange_micro_batches(data.batch, max_token_len=max_token_len)
    non_tensor: Mapping = getattr(data, "non_tensor_batch", None) or {}
    micro_batches = []
    for i, batch_idx in enumerate(batch_idx_list):
        micro_batches.append(MicroBatch(dict(batch[i]), {k: v[batch_idx] for k, v in non_tensor.items()}))
    return micro_batches, batch_idx_list


def restore_dynamic_batch(data: torch.Tensor, batch_idx_list: list[list[int]]) -> torch.Tensor:
    """seqlen_balancing.py:362-375: undo the scheduled order, `data[revert_indices]` as one row gather."""
    indices = list(chain.from_iterable(batch_idx_list))
    revert = torch.tensor(get_reverse_idx(indices), dtype=torch.int64).to(data.device, non_blocking=True)
    return ops.gather_rows(data, revert)
