"""Multi-GPU plumbing: one process per GPU, torch.distributed for the exchange.

The ETL path shards by rows; only two things cross ranks:
  * small reductions - counts, (sum, count) pairs, label histograms - all-reduced;
  * the shuffle of the high-cardinality groupBy: rows are hash-partitioned by key
    on the device (etl_hash_partition), exchanged with ONE all-to-all of 16-byte
    {key, value} records (NCCL over NVLink), and merged locally
    (etl_groupby_accumulate reads the received records in place).
Every function here works on whatever device the tensors live on, so the same
code runs under NCCL on GPUs and under gloo in the CPU tests."""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def all_reduce_sum_count(sums: torch.Tensor, counts: torch.Tensor, group=None):
    """Merge of the Average partial states (k_means.py:45-48) across ranks."""
    if world()[1] > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    return sums, counts


def all_reduce_counts(counts: torch.Tensor, group=None) -> torch.Tensor:
    if world()[1] > 1:
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    return counts


def merge_label_tables(labels: Sequence[str], counts: Sequence[int], group=None):
    """StringIndexer.fit across ranks: every rank holds (label, count) pairs for
    its rows; the union with summed counts is what the driver would see."""
    rank, size = world()
    if size == 1:
        return list(labels), list(counts)
    gathered: List[Optional[list]] = [None] * size
    dist.all_gather_object(gathered, list(zip(labels, counts)), group=group)
    total = {}
    for part in gathered:
        for s, c in part:
            total[s] = total.get(s, 0) + int(c)
    names = sorted(total)
    return names, [total[s] for s in names]


def exchange_records(records: torch.Tensor, send_counts: torch.Tensor, group=None) -> torch.Tensor:
    """All-to-all-v of partitioned {key, value} records.  `records` is int64[n, 2]
    ordered by destination rank, send_counts int64[world] (on the records'
    device).  Returns the received records int64[m, 2]."""
    rank, size = world()
    if size == 1:
        return records
    recv_counts = torch.empty_like(send_counts)
    dist.all_to_all_single(recv_counts, send_counts, group=group)
    send = [int(x) for x in send_counts.cpu().tolist()]
    recv = [int(x) for x in recv_counts.cpu().tolist()]
    out = torch.empty((sum(recv), 2), dtype=records.dtype, device=records.device)
    dist.all_to_all_single(out, records.contiguous(), output_split_sizes=recv, input_split_sizes=send, group=group)
    return out


def shuffle_groupby_sum(eng, keys: torch.Tensor, vals: torch.Tensor, capacity: int, groups_hint: int = 1 << 30,
                        group=None):
    """Config 4: hash-partition -> all-to-all -> local final merge.  After the
    exchange every key lives on exactly one rank, so no second exchange is
    needed.  Returns the rank-local GroupTable."""
    rank, size = world()
    gt = eng.groupby_new(capacity, 1, minmax=False)
    if size == 1:
        eng.groupby_accumulate(gt, keys, None, [vals], groups_hint=groups_hint)
        return gt
    records, counts, _ = eng.hash_partition(keys, vals, size)
    received = exchange_records(records, counts, group=group)
    flat = received.view(-1)
    m = received.shape[0]
    if m:
        eng.groupby_accumulate(gt, flat, None, [flat[1:].view(torch.float64)], groups_hint=groups_hint, n=m, key_stride=2,
                               val_strides=[2])
    return gt
