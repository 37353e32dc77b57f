"""DeviceTable: the columnar, device-resident stand-in for the Spark DataFrame
the ETL stage receives (`input_df` of KMeansWorkload.k_means, k_means.py:17).

read_csv replaces spark.read.csv(path, header=True, inferSchema=True)
(spark_workload_to_cloud_k8s.py:48) / the JDBC read of the same table
(google_health_SQL.py:26-37): the file is copied to the GPU once and decoded
there; only the header line and the (small) string dictionaries come back to
the host."""
from __future__ import annotations

import csv
import io
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence
from urllib.request import urlopen

import numpy as np
import torch

from ._lib import EtlError
from .engine import Column, Engine

CSV_TRIM_STRINGS = 1
CSV_NAN_IS_NULL = 2
CSV_USE_INDEX_MASKS = 4   # the row scan may walk the masks the index passes left in the handle (same text, same engine)


@dataclass
class StringColumn:
    """Dictionary-encoded string column: `codes` index into `labels`
    (ascending by code point, the order Java/Python compare strings in), -1 / a
    cleared validity bit = null."""

    column: Column
    labels: List[str]


@dataclass
class DeviceTable:
    engine: Engine
    rows: int
    numeric: Dict[str, Column] = field(default_factory=dict)
    strings: Dict[str, StringColumn] = field(default_factory=dict)
    inferred_types: Dict[str, str] = field(default_factory=dict)
    header: List[str] = field(default_factory=list)

    @property
    def columns(self) -> List[str]:
        return list(self.header) if self.header else list(self.strings) + list(self.numeric)

    def count(self) -> int:
        return self.rows


def read_bytes(source) -> bytes:
    """Path, http(s) URL (open_text, train_tf_ps.py:53-72) or bytes."""
    if isinstance(source, (bytes, bytearray, memoryview)):
        return bytes(source)
    if source.startswith("http://") or source.startswith("https://"):
        return urlopen(source).read()
    with open(source, "rb") as fh:
        return fh.read()


def _padded(n: int) -> int:
    return (n + 127) // 64 * 64


class PinnedRing:
    """Host side of the bulk ingest (replaces the JDBC / file read of
    google_health_SQL.py:26-37 and open_text, train_tf_ps.py:53-72): a few
    page-locked slots that live as long as the engine.  A file (or a bytes object)
    is cut into slot-sized slices; worker threads fill the slots (readinto / memcpy
    release the GIL) while earlier slots are already crossing PCIe on a copy
    stream, so neither a per-call page-lock of the whole input nor a monolithic
    copy is paid."""

    def __init__(self, device, slots: int = 4, slot_bytes: int = 64 << 20):
        from concurrent.futures import ThreadPoolExecutor

        self.device = device
        self.slot_bytes = slot_bytes
        self.bufs = [torch.empty(slot_bytes, dtype=torch.uint8, pin_memory=True) for _ in range(slots)]
        self.views = [b.numpy() for b in self.bufs]
        self.events: List[Optional[torch.cuda.Event]] = [None] * slots
        self.stream = torch.cuda.Stream(device=device)
        self.pool = ThreadPoolExecutor(max_workers=slots)

    def upload(self, fill, nbytes: int, extra: int = 0) -> torch.Tensor:
        """fill(view: np.ndarray[uint8], offset, length) writes bytes [offset, offset + length) of the
        input into `view`.  -> device uint8 tensor, zero padded (see _padded), with `extra` more
        (uninitialised) bytes behind the padding."""
        dst = torch.empty(_padded(nbytes) + extra, dtype=torch.uint8, device=self.device)
        dst[nbytes:_padded(nbytes)].zero_()
        main = torch.cuda.current_stream(self.device)
        self.stream.wait_stream(main)
        slots, sb = len(self.bufs), self.slot_bytes
        slices = [(off, min(sb, nbytes - off)) for off in range(0, nbytes, sb)]

        def work(i):
            slot = i % slots
            if self.events[slot] is not None:
                self.events[slot].synchronize()        # the copy that last read this slot has finished
            off, ln = slices[i]
            fill(self.views[slot][:ln], off, ln)

        pending = {i: self.pool.submit(work, i) for i in range(min(slots, len(slices)))}
        for i, (off, ln) in enumerate(slices):
            pending.pop(i).result()
            slot = i % slots
            with torch.cuda.stream(self.stream):
                dst[off:off + ln].copy_(self.bufs[slot][:ln], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.stream)
            self.events[slot] = ev
            if i + slots < len(slices):
                pending[i + slots] = self.pool.submit(work, i + slots)
        main.wait_stream(self.stream)
        return dst


def _ring(eng: Engine) -> PinnedRing:
    ring = getattr(eng, "_pinned_ring", None)
    if ring is None:
        ring = eng._pinned_ring = PinnedRing(eng.device)
    return ring


def upload_text(eng: Engine, raw, extra: int = 0) -> torch.Tensor:
    """File bytes -> device uint8 tensor, zero padded to a multiple of 64 (+64)
    so the decoder's 8- and 16-byte window loads never leave the allocation;
    `extra` more bytes behind that (the Parquet reader inflates pages there)."""
    src = np.frombuffer(raw, dtype=np.uint8)

    def fill(view, off, ln):
        view[:] = src[off:off + ln]

    return _ring(eng).upload(fill, len(raw), extra)


def upload_file(eng: Engine, path: str):
    """Local file -> (device tensor, nbytes, first bytes of the file for the header):
    worker threads pread() slices of the file straight into the pinned ring while
    earlier slices are already on their way to the GPU."""
    import os

    n = os.path.getsize(path)
    fd = os.open(path, os.O_RDONLY)
    try:
        def fill(view, off, ln):
            got = 0
            while got < ln:
                k = os.preadv(fd, [memoryview(view)[got:ln]], off + got)
                if k <= 0:
                    raise IOError(f"{path}: short read at byte {off + got}")
                got += k

        text = _ring(eng).upload(fill, n)
        head = os.pread(fd, min(n, 1 << 20), 0)
    finally:
        os.close(fd)
    return text, n, head


def _dictionary_labels(text: torch.Tensor, hash_tab: torch.Tensor, rep_tab: torch.Tensor):
    """Used slots of a decode dictionary and their strings (one small gather)."""
    # the tables hold 2 x capacity x 8 bytes but only a handful of used slots: compact on the device, copy those
    used_d = torch.nonzero(hash_tab).flatten()
    used = used_d.cpu().numpy()
    rep = rep_tab[used_d].cpu().numpy().view(np.uint64)
    off = (rep >> np.uint64(24)).astype(np.int64)
    quoted = ((rep >> np.uint64(23)) & np.uint64(1)).astype(bool)
    length = (rep & np.uint64((1 << 23) - 1)).astype(np.int64)
    if used.size == 0:
        return used, []
    idx = np.concatenate([np.arange(o, o + l) for o, l in zip(off, length)])
    blob = text[torch.from_numpy(idx).to(text.device)].cpu().numpy().tobytes()
    labels, pos = [], 0
    for l, q in zip(length, quoted):
        s = blob[pos:pos + l].decode("utf-8")
        pos += l
        labels.append(s.replace('""', '"') if q else s)
    return used, labels


def byte_range_of(path: str, rank: int, world_size: int):
    """This rank's share of a CSV file when it is split by byte range at record
    boundaries, the way Spark's CSV source splits a file with its default
    multiLine=false (every newline ends a record): -> (header line bytes, first
    byte, end byte) with the range starting after the first newline at or after the
    nominal cut and ending after the first newline at or after the next cut.  The
    header line belongs to no range; every rank reads it separately."""
    import os

    size = os.path.getsize(path)
    with open(path, "rb") as fh:
        header = fh.readline()

        def cut(at: int) -> int:       # the first record start at or after byte `at`
            if at <= len(header):
                return len(header)
            if at >= size:
                return size
            fh.seek(at - 1)
            pos = at - 1
            while True:
                block = fh.read(1 << 16)
                if not block:
                    return size
                k = block.find(b"\n")
                if k >= 0:
                    return pos + k + 1
                pos += len(block)

        return header, cut(size * rank // world_size), cut(size * (rank + 1) // world_size)


def upload_file_range(eng: Engine, path: str, header: bytes, lo: int, hi: int):
    """header line + bytes [lo, hi) of the file -> device text (what read_csv decodes on this rank)."""
    import os

    if header and not header.endswith(b"\n"):
        header += b"\n"
    hl = len(header)
    n = hl + (hi - lo)
    hsrc = np.frombuffer(header, dtype=np.uint8)
    fd = os.open(path, os.O_RDONLY)
    try:
        def fill(view, off, ln):
            got = 0
            if off < hl:               # the slice starts inside the header line
                got = min(ln, hl - off)
                view[:got] = hsrc[off:off + got]
            while got < ln:
                k = os.preadv(fd, [memoryview(view)[got:ln]], lo + off + got - hl)
                if k <= 0:
                    raise IOError(f"{path}: short read at byte {lo + off + got - hl}")
                got += k

        text = _ring(eng).upload(fill, n)
    finally:
        os.close(fd)
    return text, n, header


def read_csv(eng: Engine, source, numeric: Sequence[str], strings: Sequence[str], trim_strings: bool = False,
             nan_is_null: bool = False, dict_capacity: int = 1 << 16, shard: Optional[tuple] = None,
             device_text: Optional[torch.Tensor] = None, byte_range: Optional[tuple] = None) -> DeviceTable:
    """shard=(rank, world): every rank indexes the whole text (the two index
    passes cost ~1 ms per GB) and decodes only its contiguous share of the records,
    so quoted newlines need no cross-rank protocol.  byte_range=(rank, world) (a
    local file): this rank reads, uploads and decodes ONLY its byte range of the
    file, cut at newlines (byte_range_of - Spark's split of a CSV file under its
    default multiLine=false; a file with newlines inside quoted cells needs `shard`
    instead).  device_text: the file already on this GPU (as returned by
    upload_text), to keep the copy out of a timed region."""
    head = None
    if byte_range is not None:
        if not isinstance(source, str) or device_text is not None or shard is not None:
            raise ValueError("read_csv: byte_range needs a local file path (and neither shard nor device_text)")
        text, n, head = upload_file_range(eng, source, *byte_range_of(source, *byte_range))
    elif isinstance(source, str) and not source.startswith(("http://", "https://")) and device_text is None:
        text, n, head = upload_file(eng, source)          # file -> pinned ring -> HBM
    else:
        raw = read_bytes(source)
        n = len(raw)
        text = upload_text(eng, raw) if device_text is None else device_text
    rows_total = int(eng.csv_index_rows(text, n).item())
    if rows_total == 0:
        return DeviceTable(eng, 0)
    row_start = eng.csv_row_offsets(text, n, rows_total)
    # header line -> field positions (DictReader: a repeated name keeps its last position)
    hdr_end = int(row_start[1].item()) if rows_total > 1 else n
    if head is not None:
        hdr_bytes = head[:hdr_end] if hdr_end <= len(head) else text[:hdr_end].cpu().numpy().tobytes()
    else:
        hdr_bytes = bytes(raw[:hdr_end])
    header = next(csv.reader(io.StringIO(hdr_bytes.decode("utf-8"), newline="")))
    pos = {name: j for j, name in enumerate(header)}
    rows = rows_total - 1
    first = 1
    if shard is not None:
        rank, world_size = shard
        lo, hi = rows * rank // world_size, rows * (rank + 1) // world_size
        first, rows = 1 + lo, hi - lo
    table = DeviceTable(eng, rows, header=header)
    wanted = [(c, 0) for c in numeric] + [(c, 1) for c in strings]
    present = [(c, k) for c, k in wanted if c in pos]
    # the index passes above ran on this very text through this engine's handle: the row scan may use their masks
    flags = (CSV_TRIM_STRINGS if trim_strings else 0) | (CSV_NAN_IS_NULL if nan_is_null else 0) | CSV_USE_INDEX_MASKS
    dicts = [None if k == 0 else (torch.zeros(dict_capacity, dtype=torch.int64, device=eng.device),
                                  torch.zeros(dict_capacity, dtype=torch.int64, device=eng.device)) for _, k in present]
    if present and rows:
        # the parse kernel takes up to 8 columns per launch: wider selections go in groups of 8 over the same row index
        # (the masks of the index passes stay valid: parsing does not touch the handle's arena)
        cols, type_bits, st = [], 0, np.zeros(4, dtype=np.int64)
        for g0 in range(0, len(present), 8):
            part = present[g0:g0 + 8]
            pcols, status, first_bad = eng.csv_parse_columns(text, n, row_start, first, rows, [pos[c] for c, _ in part],
                                                             [k for _, k in part], dicts[g0:g0 + 8], flags)
            pst = status.cpu().numpy().astype(np.int64)
            cols += pcols
            type_bits |= int(pst[2]) << g0
            st[:2] += pst[:2]
            if pst[0] or pst[1]:
                break
        st[2] = type_bits
        if st[0]:
            raise EtlError(-4, f"{int(st[0])} numeral(s) outside the exactly-converted domain (more than 60 significant "
                               f"digits); first at CSV record {int(first_bad.item())}")
        if st[1]:
            raise EtlError(-5, f"string dictionary full (capacity {dict_capacity}); first at CSV record "
                               f"{int(first_bad.item())}")
    else:
        cols, st = [], np.zeros(4, dtype=np.int32)
    for s, ((name, kind), col) in enumerate(zip(present, cols)):
        if kind == 0:
            table.numeric[name] = col
            if rows:
                table.inferred_types[name] = "double" if (int(st[2]) >> s) & 1 else "int"
        else:
            used, labels = _dictionary_labels(text, *dicts[s])
            order = sorted(range(len(labels)), key=lambda i: labels[i])
            remap = torch.full((dict_capacity,), -1, dtype=torch.int32)
            remap[torch.from_numpy(used[order].astype(np.int64))] = torch.arange(len(order), dtype=torch.int32)
            eng.remap_codes(col.data, remap.to(eng.device), -1)
            table.strings[name] = StringColumn(col, [labels[i] for i in order])
    # a requested column that the header lacks is all null (row.get(c, "") in load_csv)
    for name, kind in wanted:
        if name not in pos:
            from .engine import words_for

            nullmap = torch.zeros(words_for(rows), dtype=torch.int32, device=eng.device)
            if kind == 0:
                table.numeric[name] = Column(torch.zeros(rows, dtype=torch.float64, device=eng.device), nullmap)
            else:
                table.strings[name] = StringColumn(Column(torch.full((rows,), -1, dtype=torch.int32, device=eng.device),
                                                          nullmap), [])
    return table


def from_host_columns(eng: Engine, key: Sequence[Optional[str]], numerics: Dict[str, np.ndarray],
                      key_name: str = "measure_name") -> DeviceTable:
    """A table from already-decoded host columns (the JDBC path,
    google_health_SQL.py:26-37: numerics arrive as FLOAT with SQL NULLs).  NaN in
    a float array is a value; pass a numpy masked array to mark nulls."""
    labels = sorted({s for s in key if s is not None})
    code_of = {s: i for i, s in enumerate(labels)}
    codes = np.array([code_of[s] if s is not None else -1 for s in key], dtype=np.int32)
    kvalid = np.array([s is not None for s in key], dtype=bool)
    t = DeviceTable(eng, len(key))
    t.strings[key_name] = StringColumn(Column.from_host(codes, kvalid, eng.device), labels)
    for name, arr in numerics.items():
        if isinstance(arr, np.ma.MaskedArray):
            t.numeric[name] = Column.from_host(np.ascontiguousarray(arr.filled(0)), ~np.ma.getmaskarray(arr), eng.device)
        else:
            t.numeric[name] = Column.from_host(np.ascontiguousarray(arr), None, eng.device)
        t.inferred_types[name] = {"float64": "double", "float32": "float", "int32": "int"}[str(t.numeric[name].data.dtype).split(".")[-1]]
    return t
