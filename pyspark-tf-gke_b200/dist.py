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


def choose_buckets(capacity: int, world_size: int, entry_bytes: int = 32, slice_bytes: int = 32 << 20) -> int:
    """Buckets per rank so that the table slice being updated (capacity / buckets
    entries) stays well inside the 126 MB L2; ranks * buckets <= 256."""
    b = 1
    while capacity * entry_bytes // b > slice_bytes and b * 2 * world_size <= 256:
        b *= 2
    return b


def shuffle_groupby_sum(eng, keys: torch.Tensor, vals: torch.Tensor, capacity: int, groups_hint: int = 1 << 30,
                        buckets: Optional[int] = None, group=None):
    """Config 4: hash-partition -> all-to-all -> local final merge.

    One partition pass groups the rows by (destination rank, table bucket); the
    all-to-all moves each rank's range; the receiver then merges bucket by bucket
    (all sources' runs of bucket 0, then bucket 1, ...), so the 1/buckets slice of
    the aggregation table being updated stays resident in L2 instead of taking a
    DRAM round trip per row.  After the exchange every key lives on exactly one
    rank, so no second exchange is needed.  Returns the rank-local GroupTable."""
    rank, size = world()
    gt = eng.groupby_new(capacity, 1, minmax=False)
    B = choose_buckets(capacity, size) if buckets is None else buckets
    if size == 1 and B == 1:
        eng.groupby_accumulate(gt, keys, None, [vals], groups_hint=groups_hint)
        return gt
    records, counts, _ = eng.hash_partition(keys, vals, size, B)
    matrix = counts.view(size, B)
    if size == 1:
        received, recv_matrix = records, matrix.cpu()
    else:
        recv_matrix = torch.empty_like(matrix)
        dist.all_to_all_single(recv_matrix.view(-1), matrix.contiguous().view(-1), group=group)
        received = exchange_records(records, matrix.sum(dim=1), group=group)
        recv_matrix = recv_matrix.cpu()
    # run of (source s, bucket b) inside `received`: sources are concatenated, buckets ascend within a source
    lengths = recv_matrix.reshape(-1)
    starts = torch.cumsum(lengths, 0) - lengths
    starts, lengths = starts.view(size, B).tolist(), lengths.view(size, B).tolist()
    for b in range(B):
        for s in range(size):
            eng.groupby_accumulate_records(gt, received, starts[s][b], starts[s][b] + lengths[s][b], groups_hint)
    return gt
