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


class SlabExchange:
    """The exchange without a count pass, streamed in chunks and overlapped with the
    receivers' first partition level.

    Every rank owns a receive buffer in symmetric memory cut into one fixed-capacity
    slab per SOURCE rank, plus a small table of arrival counts.  A sender scatters a
    chunk of its rows straight into its slab of every destination's buffer (NVLink
    peer stores, etl_slab_partition_rows with persistent cursors), publishes its
    cursors into the receivers' count tables (etl_peer_publish_counts) and joins a
    symmetric-memory barrier.  After the barrier of chunk c every receiver runs its
    first partition level over exactly the records that arrived with chunk c
    (etl_slab_partition_records, segments = sources) on a second stream while chunk
    c + 1 is crossing NVLink.  Nothing is read back by the host; no all-gather.

    A slab that overflows (skewed keys) only counts the surplus: the result object
    then repeats the exchange on the counted path (PeerExchange), see
    Engine.groupby_result."""

    def __init__(self, eng, rows_per_rank: int, groups_hint: int = 0, chunks: int = 4, group=None,
                 exchange_ctas: Optional[int] = None, level1_ctas: Optional[int] = None):
        import os
        import torch.distributed._symmetric_memory as symm

        self.eng = eng
        self.group = group or dist.group.WORLD
        self.rank, self.size = dist.get_rank(self.group), dist.get_world_size(self.group)
        self.rows_per_rank = int(rows_per_rank)
        total = self.rows_per_rank * self.size
        g = max(1, min(groups_hint if groups_hint > 0 else total, total))
        self.rows_per_key = max(1.0, total / g)
        self.slab = eng.slab_capacity(self.rows_per_rank / self.size, self.rows_per_key)
        self.chunks = max(1, int(os.environ.get("ETL_EXCHANGE_CHUNKS", chunks)))
        self.max_chunks = 64
        self.buf = symm.empty((self.size * self.slab, 2), dtype=torch.int64, device=eng.device)
        self.handle = symm.rendezvous(self.buf, self.group.group_name)
        self.peer_ptrs = [int(p) for p in self.handle.buffer_ptrs]
        self.counts = symm.empty((self.max_chunks, self.size), dtype=torch.int64, device=eng.device)
        self.counts_handle = symm.rendezvous(self.counts, self.group.group_name)
        self.count_ptrs = [int(p) for p in self.counts_handle.buffer_ptrs]
        self.side = torch.cuda.Stream(device=eng.device)
        sm = torch.cuda.get_device_properties(eng.device).multi_processor_count
        # the exchange scatter is NVLink-bound: capped at one 74 KB CTA per SM (a capped launch takes the 2048-row
        # tile) it leaves room on every SM for the 141 KB CTA of the receiver's level-1 scatter (uncapped: 0)
        self.exchange_ctas = int(os.environ.get("ETL_EXCHANGE_CTAS", exchange_ctas if exchange_ctas is not None else sm))
        self.level1_ctas = int(os.environ.get("ETL_LEVEL1_CTAS", level1_ctas if level1_ctas is not None else 0))
        self._counted: Optional[PeerExchange] = None

    def counted(self) -> "PeerExchange":
        """The exact-count exchange, created (collectively) the first time an overflow asks for it."""
        if self._counted is None:
            self._counted = PeerExchange(self.eng, int(self.rows_per_rank * 1.5) + (1 << 20), group=self.group)
        return self._counted

    def any_overflow(self, local: int) -> int:
        t = torch.tensor([int(local)], dtype=torch.int64, device=self.eng.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return int(t.item())

    def run(self, chunk_iter, buckets: int, cap1: int, cur1: torch.Tensor, level1: torch.Tensor, overflow: torch.Tensor,
            mark=None):
        """chunk_iter yields (keys, vals, n, ready_event or None) device chunks of this
        rank's rows.  On return (main stream) level1 / cur1 hold this rank's received
        records cut into `buckets` slabs of cap1 records."""
        eng, size, rank = self.eng, self.size, self.rank
        main = torch.cuda.current_stream(eng.device)
        xcur = torch.zeros(size, dtype=torch.int64, device=eng.device)
        self.handle.barrier()            # every rank has finished reading the previous exchange's records
        self.side.wait_stream(main)      # level1 / cur1 / overflow were set up on the main stream
        c = 0
        for keys, vals, n, ready in chunk_iter:
            if c >= self.max_chunks:
                raise RuntimeError(f"SlabExchange: more than {self.max_chunks} chunks")
            if ready is not None:
                main.wait_event(ready)
            eng.slab_partition_rows(keys, vals, n, size, 1, self.peer_ptrs, rank * self.slab, 0, self.slab, xcur, overflow,
                                    max_ctas=self.exchange_ctas)
            eng.peer_publish_counts(xcur, self.count_ptrs, c * size + rank)
            self.handle.barrier()        # chunk c of every source, and its counts, have landed here
            arrived = torch.cuda.Event()
            arrived.record(main)
            with torch.cuda.stream(self.side):
                self.side.wait_event(arrived)
                bound = min(self.slab, -(-n // size) * 2 + 4096)
                eng.slab_partition_records(self.buf, size, self.slab, self.counts[c], self.counts[c - 1] if c else None, bound,
                                           buckets, 0, level1, cap1, cur1, overflow, max_ctas=self.level1_ctas)
            c += 1
        if mark:
            mark("exchange")
        main.wait_stream(self.side)
        if mark:
            mark("level1")


def plan_waves(rows: int, waves: int, head: float = 0.25):
    """How BucketExchange cuts a rank's `rows` into `waves` parts: the last wave holds 1 - head of
    the rows, and the rows up to and including wave w are `head` of the rows up to and including
    wave w + 1 (so every wave is at least `head` of the next one).  -> (fractions, bounds) with wave w
    = rows [bounds[w], bounds[w + 1]); inner bounds are multiples of 4096 rows (the scatter tile,
    and 16-byte aligned in both columns), bounds[-1] = rows; a wave may be empty when rows is small."""
    waves = max(1, int(waves))
    head = min(0.5, max(0.01, float(head)))
    fracs = [head ** (waves - 1 - w) * (1.0 - head if w else 1.0) for w in range(waves)] if waves > 1 else [1.0]
    total = sum(fracs)
    fracs = [f / total for f in fracs]
    bounds, acc = [0], 0.0
    for w in range(waves):
        acc += fracs[w]
        end = rows if w == waves - 1 else min(rows, -(-int(acc * rows) // 4096) * 4096)
        bounds.append(max(end, bounds[-1]))
    return fracs, bounds


class BucketExchange:
    """The exchange as: partition locally by (destination rank, bucket), move the
    slabs bucket group by bucket group over NVLink, merge each group of buckets as
    soon as it has arrived - the transfer is what stays exposed.

    Sender: ONE slab scatter (etl_slab_partition_rows, 4096-row tiles, all SMs) cuts
    this rank's rows into ranks x Bx slabs of a local send buffer, destination-rank
    major - the receiver's first partition level is done here, at full HBM speed,
    with no NVLink store in the kernel.  The cursors go to the receivers' count
    tables (etl_peer_publish_counts).  Then, for each of G groups of buckets, the
    filled part of the group's slabs is stored into every receiver's buffer (layout
    [source][bucket], symmetric memory) by a small copy kernel (etl_peer_copy_slabs:
    16-byte stores over NVLink, no shared memory, so it runs beside the merge), followed
    by a symmetric-memory barrier.
    Receiver: after the barrier of group g it cuts the group's buckets into pieces
    (etl_slab_partition_records over the (bucket, source) slabs) and merges them in
    shared memory (etl_groupby_merge_slabs, append mode) on a compute stream while
    group g + 1 is still crossing NVLink.  No count pass, no all-gather, nothing
    read back by the host.
    Waves (optional; the default on 2 GPUs): the rows are cut in W parts with their
    own slabs, so that part w crosses NVLink while part w + 1 is being cut; a bucket
    group is merged once its last wave has landed (see __init__).
    Movers: "sm" (default, the store kernel above), "tma" (etl_peer_copy_slabs_bulk),
    "ce" (copy engines).

    A slab that overflows (skewed keys) only counts the surplus; the result object
    then repeats the exchange on the counted path (PeerExchange)."""

    def __init__(self, eng, rows_per_rank: int, groups_hint: int, bucket_groups: int = 8, group=None, copy_streams: int = 4,
                 waves: Optional[int] = None, mover: Optional[str] = None):
        import os
        import torch.distributed._symmetric_memory as symm

        self.eng = eng
        self.group = group or dist.group.WORLD
        self.rank, self.size = dist.get_rank(self.group), dist.get_world_size(self.group)
        size = self.size
        self.rows_per_rank = int(rows_per_rank)
        total = self.rows_per_rank * size
        g = max(1, min(groups_hint if groups_hint > 0 else total, total))
        self.rows_per_key = max(1.0, total / g)
        hint = -(-g // size) + g // (8 * size)                       # distinct keys this rank may end up owning
        fine = max(1, -(-hint // eng.MERGE_GROUPS_PER_PIECE))
        bx = 1
        while bx * 2 * size <= 256 and bx < fine:
            bx *= 2
        self.buckets = bx                                            # Bx buckets per rank (a power of two)
        self.sub = min(256, max(1, -(-fine // bx)))
        self.bucket_groups = max(1, min(int(os.environ.get("ETL_BUCKET_GROUPS", bucket_groups)), bx))
        while bx % self.bucket_groups:
            self.bucket_groups -= 1
        # The rows can go in `waves` parts, each with its own slabs: wave w crosses NVLink while wave w + 1 is still being
        # cut, so only the cut of the first wave is exposed.  The merge of a bucket group needs every wave's records of
        # the group, so the LAST wave is the large one (1 - head of the rows: its groups are sent one by one with the
        # merges right behind them) and everything before a wave is `head` of everything up to and including it, so every
        # wave is at least `head` of the next one - enough transfer to cover
        # the next cut (cut : transfer is about 1 : 4 on NVLink 5, hence head = 1/4).  Measured on config 4 (bench.py):
        # two waves 13.4 -> 12.4 ms at 2 GPUs, 7.85 -> 7.24 at 4, 4.60 -> 4.45 at 8 - less than the cut it hides, because
        # the cut and the transfer slow each other down; three waves gain nothing over two.
        default_waves = os.environ.get("ETL_EXCHANGE_WAVES", "2")
        W = self.waves = max(1, min(int(default_waves) if waves is None else int(waves), 6))
        head = float(os.environ.get("ETL_EXCHANGE_HEAD", "0.25"))
        fracs, bounds = plan_waves(self.rows_per_rank, W, head)
        self.wave_bounds = bounds                                      # wave w = rows [bounds[w], bounds[w + 1]) of this rank
        self.slabs = [eng.slab_capacity(max(bounds[w + 1] - bounds[w], 1) / (size * bx), self.rows_per_key) for w in range(W)]
        self.wave_groups = []
        for w in range(W):                                           # bucket groups a wave is sent in (powers of two divide bx)
            want = self.bucket_groups if w == W - 1 else max(1, int(self.bucket_groups * fracs[w] + 0.5))
            gw = 1
            while gw * 2 <= want and bx % (gw * 2) == 0:
                gw *= 2
            self.wave_groups.append(self.bucket_groups if w == W - 1 else gw)
        self.cap2 = eng.slab_capacity(self.rows_per_rank / (bx * self.sub), self.rows_per_key)
        dev = eng.device
        starts = [0]
        for w in range(W):
            starts.append(starts[-1] + size * bx * self.slabs[w])    # records of a wave's slabs, [rank or source][bucket]
        self.send_all = torch.empty((starts[-1], 2), dtype=torch.int64, device=dev)
        self.recv_all = symm.empty((starts[-1], 2), dtype=torch.int64, device=dev)
        self.handle = symm.rendezvous(self.recv_all, self.group.group_name)
        self.send = [self.send_all[starts[w]:starts[w + 1]] for w in range(W)]
        self.recv = [self.recv_all[starts[w]:starts[w + 1]] for w in range(W)]
        self.peer_ptrs = [[int(p) + starts[w] * 16 for p in self.handle.buffer_ptrs] for w in range(W)]
        self.counts_all = symm.empty(W * size * bx, dtype=torch.int64, device=dev)
        self.counts_handle = symm.rendezvous(self.counts_all, self.group.group_name)
        self.counts = self.counts_all.view(W, size * bx)
        self.count_ptrs = [[int(p) + w * size * bx * 8 for p in self.counts_handle.buffer_ptrs] for w in range(W)]
        self.pieces = torch.empty((bx * self.sub * self.cap2, 2), dtype=torch.int64, device=dev)
        self.copy_streams = [torch.cuda.Stream(device=dev) for _ in range(max(1, int(os.environ.get("ETL_COPY_STREAMS", copy_streams))))]
        self.sync_stream = torch.cuda.Stream(device=dev)
        self.compute = torch.cuda.Stream(device=dev)
        # how the slabs cross NVLink: "sm" = etl_peer_copy_slabs (16-byte stores from registers, runs beside the merge
        # kernels), "tma" = etl_peer_copy_slabs_bulk (bulk async copies through a shared-memory ring, one thread per CTA),
        # "ce" = stream-ordered peer copies through the copy engines (330-400 GB/s at 8 GPUs)
        self.mover = mover or os.environ.get("ETL_EXCHANGE_MOVER", "sm")
        if self.mover not in ("sm", "tma", "ce"):
            raise ValueError(f"BucketExchange: mover={self.mover!r}")
        self.copy_ctas = int(os.environ.get("ETL_COPY_CTAS", "0"))
        # measurement only (tools/measure_shuffle.py --skip; the result is then wrong): "copy" = leave the transfer out
        # (the receive buffers keep the previous exchange's records), "compute" = leave the receivers' second level and
        # merge out - what each side costs on its own
        self.skip = ""
        self.timing = False            # True: every shuffle leaves (first copy starts, last group landed) events in .window
        self.window = None
        self._counted: Optional[PeerExchange] = None

    chunks = 1   # device-resident rows need no chunking here (shuffle_groupby_sum_host brings its own chunks)

    def counted(self) -> "PeerExchange":
        if self._counted is None:
            self._counted = PeerExchange(self.eng, int(self.rows_per_rank * 1.5) + (1 << 20), group=self.group)
        return self._counted

    def any_overflow(self, local: int) -> int:
        t = torch.tensor([int(local)], dtype=torch.int64, device=self.eng.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return int(t.item())

    def shuffle(self, chunk_iter, capacity: int, minmax: bool = False, mark=None):
        """chunk_iter yields (keys, vals, n, ready_event or None) device chunks of this rank's rows
        -> PartitionedGroups of the keys this rank owns (enqueued; main stream waits for the merge)."""
        eng, size, rank, bx, W = self.eng, self.size, self.rank, self.buckets, self.waves
        mark = mark or (lambda name: None)
        main = torch.cuda.current_stream(eng.device)
        pg = eng.new_partitioned_groups(capacity, max(self.rows_per_rank // 8, 1 << 16), bx * self.sub, minmax=minmax)
        xcur = torch.zeros((W, size * bx), dtype=torch.int64, device=eng.device)
        cur2 = torch.zeros(bx * self.sub, dtype=torch.int64, device=eng.device)
        mark("init")
        self.handle.barrier()            # every rank has finished reading the previous exchange's records
        state = {"k": 0, "first": None, "landed": None}

        def send_wave(w):
            """Wave w has been cut into its slabs on the main stream: publish the counts, move the slabs group by group,
            and let the compute stream take every group as soon as it has landed."""
            slab, groups = self.slabs[w], self.wave_groups[w]
            bpg = bx // groups
            nbytes = bpg * slab * 16
            eng.peer_publish_counts(xcur[w], self.count_ptrs[w], rank * bx, per_rank=bx)
            cut = torch.cuda.Event()
            cut.record(main)
            movers = self.copy_streams if self.mover == "ce" else self.copy_streams[:1]
            for st in movers:
                st.wait_event(cut)
            if self.timing and state["first"] is None:
                state["first"] = torch.cuda.Event(enable_timing=True)
                state["first"].record(movers[0])
            send_ptr = self.send[w].data_ptr()
            for g in range(groups):
                copied = []
                if self.mover != "ce" and self.skip != "copy":    # one kernel per group over NVLink, only the filled part of the slabs
                    st = movers[0]
                    with torch.cuda.stream(st):
                        eng.peer_copy_slabs(self.send[w], self.peer_ptrs[w], rank, bx, g * bpg, bpg, slab, xcur[w], self.copy_ctas,
                                            bulk=self.mover == "tma")
                        ev = torch.cuda.Event()
                        ev.record(st)
                        copied.append(ev)
                for i in range(size if self.mover == "ce" else 0):
                    r = (rank + 1 + i) % size             # every rank starts with a different peer
                    st = movers[state["k"] % len(movers)]
                    state["k"] += 1
                    with torch.cuda.stream(st):
                        eng.copy_async(self.peer_ptrs[w][r] + ((rank * bx + g * bpg) * slab) * 16,
                                       send_ptr + ((r * bx + g * bpg) * slab) * 16, nbytes)
                        ev = torch.cuda.Event()
                        ev.record(st)
                        copied.append(ev)
                with torch.cuda.stream(self.sync_stream):
                    for ev in copied:
                        self.sync_stream.wait_event(ev)
                    self.handle.barrier()     # group g of every source (and the wave's count tables) has landed here
                    landed = torch.cuda.Event(enable_timing=self.timing)
                    landed.record(self.sync_stream)
                    state["landed"] = landed
                with torch.cuda.stream(self.compute):
                    self.compute.wait_event(landed)
                    if self.skip == "compute":
                        continue
                    eng.slab_partition_records(self.recv[w], bpg * size, slab, self.counts[w], None, slab, bx, self.sub, self.pieces,
                                               self.cap2, cur2, pg.counters[4:], sources=size, bucket0=g * bpg)
                    if w == W - 1:        # every wave's records of this group are in the pieces now
                        eng.merge_slab_pieces(self.pieces, cur2, self.cap2, g * bpg * self.sub, bpg * self.sub, g > 0, pg,
                                              shared_sms=True)

        w, pos = 0, 0
        for keys, vals, n, ready in chunk_iter:
            if ready is not None:
                main.wait_event(ready)
            a = 0
            while a < n:
                while w < W - 1 and pos >= self.wave_bounds[w + 1]:
                    send_wave(w)
                    w += 1
                take = n - a if w == W - 1 else min(n - a, self.wave_bounds[w + 1] - pos)
                eng.slab_partition_rows(keys[a:a + take], vals[a:a + take], take, size, bx, [self.send[w].data_ptr()] * size, 0,
                                        self.slabs[w], self.slabs[w], xcur[w], pg.counters[4:])
                a += take
                pos += take
        mark("partition")
        while w < W:                     # the last wave, and empty ones if this rank had fewer rows: every rank runs every barrier
            send_wave(w)
            w += 1
        main.wait_event(state["landed"])
        mark("exchange")
        mark("level1")
        main.wait_stream(self.compute)
        mark("level2")
        mark("merge")
        self.window = (state["first"], state["landed"]) if self.timing else None
        pg.any_overflow = self.any_overflow
        return pg


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
                        merge: str = "smem", exchange_buckets: Optional[int] = None, counted: bool = False,
                        minmax: bool = False):
    """Config 4: hash-partition -> all-to-all -> local final merge.

    Default path (merge="smem", groups_hint > 0, one GPU or peer = a SlabExchange):
    no count passes - every scatter writes into fixed-capacity slabs; at N > 1 the
    rows cross NVLink in chunks while the receivers already run their first
    partition level on what has arrived (_slab_shuffle_groupby_sum).  counted=True,
    peer = a PeerExchange, merge="table" or an unknown group count select the
    counted path described below, which sizes every destination exactly and is also
    what a slab overflow falls back to.

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
    if merge == "smem" and not counted and groups_hint > 0 and (size == 1 or isinstance(peer, (SlabExchange, BucketExchange))):
        return _slab_shuffle_groupby_sum(eng, keys, vals, capacity, groups_hint, peer, mark, group, minmax=minmax)
    if isinstance(peer, (SlabExchange, BucketExchange)):
        peer = peer.counted()
    gt = eng.groupby_new(capacity, 1, minmax=minmax) if merge == "table" else None
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
            pg = eng.groupby_merge_buckets(received, bcounts, bbase, B, sub, m, capacity, mark=mark, spill=spill, minmax=minmax)
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
                                         mark=mark, minmax=minmax)
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


def _device_chunks(keys: torch.Tensor, vals: torch.Tensor, chunks: int):
    n = keys.shape[0]
    step = max(2048, -(-(-(-n // chunks)) // 2048) * 2048)      # chunk starts stay 16-byte aligned in both columns
    for lo in range(0, max(n, 1), step):
        hi = min(n, lo + step)
        yield keys[lo:hi], vals[lo:hi], hi - lo, None


def _slab_shuffle_groupby_sum(eng, keys, vals, capacity: int, groups_hint: int, peer, mark, group, chunk_iter=None,
                              n_local: Optional[int] = None, minmax: bool = False):
    """The slab path of shuffle_groupby_sum (see there).  chunk_iter: this rank's rows
    as device chunks (shuffle_groupby_sum_host feeds them as they arrive from the host)."""
    rank, size = world()
    if size == 1:
        for name in ("init", "partition", "exchange"):
            mark(name)
        if chunk_iter is None:
            pg = eng.groupby_slab_sum(keys, vals, capacity, groups_hint, mark=mark, minmax=minmax)
            mark("merge")
            return pg
        n = n_local
        buckets, sub, rpk, cap1, cap2 = eng.slab_plan(n, groups_hint)
        counters = torch.zeros(5, dtype=torch.int64, device=eng.device)
        cur1 = torch.zeros(buckets, dtype=torch.int64, device=eng.device)
        level1 = torch.empty((buckets * cap1, 2), dtype=torch.int64, device=eng.device)
        main = torch.cuda.current_stream(eng.device)
        for k, v, m, ready in chunk_iter:
            if ready is not None:
                main.wait_event(ready)
            eng.slab_partition_rows(k, v, m, 1, buckets, [level1.data_ptr()], 0, cap1, cap1, cur1, counters[4:])
        mark("level1")
        pg = eng.groupby_merge_slab_buckets(level1, cur1, cap1, buckets, sub, cap2, capacity, counters, mark=mark, minmax=minmax)
        mark("merge")
        return pg
    if isinstance(peer, BucketExchange):
        pg = peer.shuffle(chunk_iter if chunk_iter is not None else _device_chunks(keys, vals, 1), capacity, minmax=minmax, mark=mark)
        if keys is not None:
            pg.rerun = lambda: shuffle_groupby_sum(eng, keys, vals, capacity, groups_hint, peer=peer.counted(), group=group,
                                                   counted=True, minmax=minmax)
        return pg
    n = keys.shape[0] if n_local is None else n_local
    m = n + n // 8 + (1 << 16)                      # records this rank expects to own (sizes the plan, not the buffers)
    hint = -(-groups_hint // size) + groups_hint // (8 * size)
    buckets, sub, rpk, cap1, cap2 = eng.slab_plan(m, hint)
    counters = torch.zeros(5, dtype=torch.int64, device=eng.device)
    cur1 = torch.zeros(buckets, dtype=torch.int64, device=eng.device)
    level1 = torch.empty((buckets * cap1, 2), dtype=torch.int64, device=eng.device)
    mark("init")
    mark("partition")
    peer.run(chunk_iter if chunk_iter is not None else _device_chunks(keys, vals, peer.chunks), buckets, cap1, cur1, level1,
             counters[4:], mark=mark)
    pg = eng.groupby_merge_slab_buckets(level1, cur1, cap1, buckets, sub, cap2, capacity, counters, mark=mark, minmax=minmax)
    mark("merge")
    pg.any_overflow = peer.any_overflow
    if keys is not None:
        pg.rerun = lambda: shuffle_groupby_sum(eng, keys, vals, capacity, groups_hint, peer=peer.counted(), group=group,
                                               counted=True, minmax=minmax)
    return pg


def shuffle_groupby_sum_host(eng, h_keys: torch.Tensor, h_vals: torch.Tensor, capacity: int, groups_hint: int = 1 << 30,
                             peer=None, out=None, group=None, chunks: int = 8):
    """shuffle_groupby_sum for callers whose rows live in HOST memory: `h_keys`
    int64[n] / `h_vals` float64[n] (pinned for full PCIe speed) are this rank's
    rows; the rank-local groups come back in host memory as out = (keys int64,
    count int64, sum float64) arrays of at least `capacity` entries (pinned tensors
    are filled in place; allocated when None).  Returns the number of groups.  The
    copies, the exchange and the merge all run inside this call: on the slab path
    the rows are copied in `chunks` pieces on a copy stream and every piece is
    scattered (first partition level, or the exchange over NVLink) as soon as it
    has landed, so the partitioning hides behind the PCIe transfer."""
    dev = eng.device
    rank, size = world()
    n = h_keys.shape[0]
    main = torch.cuda.current_stream(dev)
    slab_ok = groups_hint > 0 and n > 0 and (size == 1 or isinstance(peer, (SlabExchange, BucketExchange)))
    if not slab_ok:
        d_keys, d_vals = h_keys.to(dev, non_blocking=True), h_vals.to(dev, non_blocking=True)
        pg = shuffle_groupby_sum(eng, d_keys, d_vals, capacity, groups_hint=groups_hint, peer=peer, group=group)
    else:
        d_keys = torch.empty(n, dtype=torch.int64, device=dev)
        d_vals = torch.empty(n, dtype=torch.float64, device=dev)
        copy = _copy_stream(dev)
        copy.wait_stream(main)
        step = max(2048, -(-(-(-n // chunks)) // 2048) * 2048)

        def arriving():
            for lo in range(0, n, step):
                hi = min(n, lo + step)
                with torch.cuda.stream(copy):
                    d_keys[lo:hi].copy_(h_keys[lo:hi], non_blocking=True)
                    d_vals[lo:hi].copy_(h_vals[lo:hi], non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(copy)
                yield d_keys[lo:hi], d_vals[lo:hi], hi - lo, ev

        pg = _slab_shuffle_groupby_sum(eng, None, None, capacity, groups_hint, peer, lambda name: None, group,
                                       chunk_iter=arriving(), n_local=n)
        counted_peer = peer.counted if isinstance(peer, (SlabExchange, BucketExchange)) else (lambda: peer)
        pg.rerun = lambda: shuffle_groupby_sum(eng, d_keys, d_vals, capacity, groups_hint, peer=counted_peer(), group=group,
                                               counted=True)
    res = eng.groupby_result(pg, sort=False)
    g = int(res["keys"].shape[0])
    if out is None:
        out = tuple(torch.empty(max(g, 1), dtype=t, pin_memory=True) for t in (torch.int64, torch.int64, torch.float64))
    out[0][:g].copy_(res["keys"], non_blocking=True)
    out[1][:g].copy_(res["cols"][0]["count"], non_blocking=True)
    out[2][:g].copy_(res["cols"][0]["sum"], non_blocking=True)
    main.synchronize()
    return g


_COPY_STREAMS = {}


def _copy_stream(dev) -> "torch.cuda.Stream":
    key = (dev.type, dev.index)
    if key not in _COPY_STREAMS:
        _COPY_STREAMS[key] = torch.cuda.Stream(device=dev)
    return _COPY_STREAMS[key]
