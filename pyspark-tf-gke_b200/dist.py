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


LABEL_BLOB_BYTES = 16384


def merge_label_tables(labels: Sequence[str], counts: Sequence[int], group=None):
    """StringIndexer.fit across ranks: every rank holds (label, count) pairs for
    its rows; the union with summed counts is what the driver would see."""
    rank, size = world()
    if size == 1:
        return list(labels), list(counts)
    # one tensor all-gather of a fixed-size, length-prefixed JSON blob (a label table is a few hundred bytes);
    # all_gather_object - pickling plus a size exchange plus a payload exchange - only if some rank's table
    # does not fit
    import json
    blob = json.dumps([list(labels), [int(c) for c in counts]]).encode("utf-8")
    device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    buf = torch.zeros(LABEL_BLOB_BYTES, dtype=torch.uint8)
    fits = len(blob) <= LABEL_BLOB_BYTES - 8
    buf[:8] = torch.tensor(list((len(blob) if fits else 0xFFFFFFFF).to_bytes(8, "little")), dtype=torch.uint8)
    if fits:
        buf[8:8 + len(blob)] = torch.frombuffer(bytearray(blob), dtype=torch.uint8)
    out = torch.empty(size * LABEL_BLOB_BYTES, dtype=torch.uint8, device=device)
    dist.all_gather_into_tensor(out, buf.to(device), group=group)
    host = out.cpu().numpy().reshape(size, LABEL_BLOB_BYTES)
    lengths = [int.from_bytes(host[r, :8].tobytes(), "little") for r in range(size)]
    gathered: List[Optional[list]] = [None] * size
    if all(n != 0xFFFFFFFF for n in lengths):
        for r in range(size):
            names, cnts = json.loads(host[r, 8:8 + lengths[r]].tobytes().decode("utf-8"))
            gathered[r] = list(zip(names, cnts))
    else:
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


class PeerExchange:
    """The shuffle exchange as ONE kernel over NVLink peer memory: every rank owns
    a receive buffer in symmetric memory (mapped into all peers), and the
    partition's scatter pass stores each {key, value} record straight into its
    destination rank's buffer - no send buffer, no separate all-to-all.

    Per exchange: a count pass, one all-gather of world x world counts (so every
    rank knows where its runs start in every receive buffer and how much it will
    receive), the fused scatter kernel, and a symmetric-memory barrier before the
    merge reads.  A leading barrier keeps a fast rank from overwriting a buffer a
    slow rank is still merging."""

    def __init__(self, eng, capacity_records: int, group=None):
        import torch.distributed._symmetric_memory as symm

        self.eng = eng
        self.group = group or dist.group.WORLD
        self.rank, self.size = dist.get_rank(self.group), dist.get_world_size(self.group)
        self.capacity = int(capacity_records)
        self.buf = symm.empty((self.capacity, 2), dtype=torch.int64, device=eng.device)
        self.handle = symm.rendezvous(self.buf, self.group.group_name)
        self.peer_ptrs = [int(p) for p in self.handle.buffer_ptrs]

    def exchange(self, keys: torch.Tensor, vals: torch.Tensor, buckets: int = 1):
        """-> (records int64[m, 2] view of this rank's receive buffer, m, bucket_counts,
        bucket_base).  With buckets > 1 the records arrive grouped by table bucket
        (device int64[buckets] counts / first records): the receiver's first partition
        level is done by the senders' scatter.  world_size * buckets <= 256."""
        eng, size, rank = self.eng, self.size, self.rank
        counts = eng.hash_partition_count(keys, size, buckets)
        matrix = torch.empty((size, size * buckets), dtype=torch.int64, device=eng.device)   # [source, destination]
        dist.all_gather_into_tensor(matrix.view(-1), counts, group=self.group)
        offsets, tot, bucket_start = peer_layout(matrix, rank, size, buckets)
        inc = tot.sum(dim=1).cpu().tolist()
        if max(inc) > self.capacity:
            raise RuntimeError(f"PeerExchange: a rank would receive {max(inc)} records, buffer holds {self.capacity}")
        self.handle.barrier()            # everyone is done reading the previous exchange's records
        eng.hash_partition_scatter_peers(keys, vals, self.peer_ptrs, offsets, buckets=buckets)
        self.handle.barrier()            # every peer's stores into this buffer have landed
        m = int(inc[rank])
        return self.buf[:m], m, tot[rank].contiguous(), bucket_start[rank].contiguous()


def peer_layout(matrix: torch.Tensor, rank: int, size: int, buckets: int):
    """Layout of the receive buffers of a bucketed peer exchange.  matrix[s, r * buckets + b]
    = records source s sends to (rank r, bucket b).  A receive buffer holds its buckets one
    after another and, inside a bucket, the sources one after another.  -> (offsets
    int64[size * buckets]: first record of this rank's run for every (r, b); tot[r, b]:
    records of bucket b on rank r; bucket_start[r, b]: first record of that bucket)."""
    per = matrix.view(size, size, buckets)              # [source, rank, bucket]
    tot = per.sum(dim=0)
    bucket_start = torch.cumsum(tot, dim=1) - tot
    offsets = bucket_start + per[:rank].sum(dim=0)
    return offsets.contiguous().view(-1), tot, bucket_start


def choose_buckets(capacity: int, world_size: int, entry_bytes: int = 32, slice_bytes: int = 32 << 20) -> int:
    """Buckets per rank so that the table slice being updated (capacity / buckets
    entries) stays well inside the 126 MB L2; ranks * buckets <= 256."""
    b = 1
    while capacity * entry_bytes // b > slice_bytes and b * 2 * world_size <= 256:
        b *= 2
    return b


def shuffle_groupby_sum(eng, keys: torch.Tensor, vals: torch.Tensor, capacity: int, groups_hint: int = 1 << 30,
                        buckets: Optional[int] = None, group=None, mark=None, peer: Optional[PeerExchange] = None,
                        merge: str = "smem", exchange_buckets: Optional[int] = None):
    """Config 4: hash-partition -> all-to-all -> local final merge.

    Sender: one partition pass groups the rows by destination rank (few
    destinations = long contiguous runs).  The all-to-all moves each rank's range.
    Receiver: a second partition pass groups what arrived by table bucket (the
    bucket comes from the same hash word as a key's home slot), then the buckets
    are merged one after another, so the 1/buckets slice of the aggregation table
    being updated stays resident in L2 instead of taking a DRAM round trip per
    row.  After the exchange every key lives on exactly one rank, so no second
    exchange is needed.  With `peer` (a PeerExchange) partition and exchange are
    one kernel writing over NVLink into the receivers' buffers; without it the
    send buffer goes through NCCL's all-to-all.  `mark(name)` is called at the
    phase boundaries (for timing).

    merge="smem" (default) replaces the receiver's bucket-by-bucket table merge by
    Engine.groupby_partitioned_sum: two partition levels cut what arrived into
    pieces whose groups fit a shared-memory table, so no table update touches L2 or
    HBM (`capacity` is then the room for output groups; `groups_hint` = expected
    distinct keys over ALL ranks sizes the pieces).  merge="table" keeps the global
    table.  exchange_buckets > 1: the fused exchange also does the receiver's
    first partition level (destinations = ranks x buckets: one pass less, but
    shorter runs over NVLink - 385 instead of 555 GB/s per GPU); 1 = the exchange
    partitions by rank only and the receiver runs the first level itself; None
    picks the faster one as measured on config 4 (cut by bucket on 2 GPUs only).  Returns the rank-local result for Engine.groupby_result."""
    rank, size = world()
    mark = mark or (lambda name: None)
    if merge not in ("smem", "table"):
        raise ValueError(f"merge={merge!r}")
    gt = eng.groupby_new(capacity, 1, minmax=False) if merge == "table" else None
    B = choose_buckets(capacity, 1) if buckets is None else buckets
    mark("init")
    if size > 1:
        if exchange_buckets is None and size > 2:
            exchange_buckets = 1   # measured (config 4): short NVLink runs cost more than the receiver's own pass
        if peer is not None and merge == "smem" and exchange_buckets != 1:
            # the senders' scatter also does the receiver's first partition level
            n_local = keys.shape[0]
            hint = -(-groups_hint // size) + groups_hint // (8 * size) if groups_hint > 0 else 0
            B, sub = eng.merge_plan(n_local, hint, max_buckets=min(exchange_buckets or 256, max(1, 256 // size)))
            mark("partition")
            received, m, bcounts, bbase = peer.exchange(keys, vals, buckets=B)
            mark("exchange")
            mark("level1")
            # the receive buffer is overwritten by the next exchange: spilled records get their own room
            spill = torch.empty((max(m // 8, 1 << 16), 2), dtype=torch.int64, device=eng.device)
            pg = eng.groupby_merge_buckets(received, bcounts, bbase, B, sub, m, capacity, mark=mark, spill=spill)
            mark("merge")
            return pg
        if peer is not None:
            mark("partition")
            received, _, _, _ = peer.exchange(keys, vals)
        else:
            records, counts, _ = eng.hash_partition(keys, vals, size, 1)
            mark("partition")
            received = exchange_records(records, counts, group=group)
            del records
        mark("exchange")
        flat = received.view(-1)
        rk, rv, m, stride = flat, flat[1:].view(torch.float64), received.shape[0], 2
    else:
        rk, rv, m, stride = keys, vals, keys.shape[0], 1
        mark("partition")
        mark("exchange")
    if merge == "smem":
        local_hint = -(-groups_hint // size) + (groups_hint // (8 * size) if size > 1 else 0) if groups_hint > 0 else 0
        pg = eng.groupby_partitioned_sum(rk, rv, capacity, groups_hint=local_hint, n=m, key_stride=stride, val_stride=stride,
                                         mark=mark)
        mark("merge")
        return pg
    if B == 1 or m == 0:
        if m:
            eng.groupby_accumulate(gt, rk, None, [rv], groups_hint=groups_hint, n=m, key_stride=stride, val_strides=[stride])
    else:
        by_bucket, counts, base = eng.hash_partition(rk, rv, 1, B, n=m, key_stride=stride, val_stride=stride)
        starts, lengths = base.cpu().tolist(), counts.cpu().tolist()
        for b in range(B):
            eng.groupby_accumulate_records(gt, by_bucket, starts[b], starts[b] + lengths[b], groups_hint)
    mark("merge")
    return gt
