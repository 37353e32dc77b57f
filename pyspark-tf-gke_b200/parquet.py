"""Parquet -> DeviceTable (north-star item 1, "CSV/Parquet decode"; the reference
itself reads CSV / JDBC only, so this is an extension whose oracle is pyarrow's
reader on the same file - parity unpinned by the reference).

Host side: the footer and page headers (Thrift compact protocol - a few hundred
bytes per megabyte of data) are parsed here; dictionaries of string columns
(a handful of labels) are decoded here.  Device side (csrc/parquet.cu): the
RLE / bit-packed hybrid streams (definition levels, dictionary indices) are
expanded, values are gathered to their rows, nulls become validity bits, numeric
types are cast to double, dictionary indices become label codes.

Supported: flat schemas; UNCOMPRESSED and SNAPPY column chunks (what Spark and
pyarrow write by default - the reference's model.save / pipeline_model.save,
spark_workload_to_cloud_k8s.py:149-154, produce *.snappy.parquet): compressed data
pages are inflated on the device (pq_snappy_kernel) into a region behind the file
bytes; dictionary pages (small) are inflated here; data pages v1 and v2; PLAIN for
DOUBLE / FLOAT / INT32 / INT64 / BYTE_ARRAY; RLE_DICTIONARY / PLAIN_DICTIONARY
for BYTE_ARRAY (strings) and the numeric types.  PLAIN string pages (a writer's
fallback when its dictionary grows too large) go through a device dictionary
like CSV cells.  Anything else raises ParquetUnsupported - never a silent wrong
answer."""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .engine import Column, Engine, words_for
from .table import DeviceTable, StringColumn, read_bytes, upload_text


class ParquetUnsupported(NotImplementedError):
    pass


# ---------------------------------------------------------------- Thrift compact
class _Thrift:
    """Minimal reader of the Thrift compact protocol: structs become
    {field_id: value} dicts, lists Python lists, binaries bytes."""

    def __init__(self, buf: bytes, pos: int = 0):
        self.b, self.p = buf, pos

    def varint(self) -> int:
        r = s = 0
        while True:
            c = self.b[self.p]
            self.p += 1
            r |= (c & 0x7F) << s
            if not c & 0x80:
                return r
            s += 7

    def zigzag(self) -> int:
        v = self.varint()
        return (v >> 1) ^ -(v & 1)

    def value(self, t: int):
        if t in (1, 2):
            return t == 1
        if t == 3:
            v = self.b[self.p]
            self.p += 1
            return v - 256 if v > 127 else v
        if t in (4, 5, 6):
            return self.zigzag()
        if t == 7:
            v = struct.unpack_from("<d", self.b, self.p)[0]
            self.p += 8
            return v
        if t == 8:
            n = self.varint()
            v = self.b[self.p:self.p + n]
            self.p += n
            return v
        if t in (9, 10):
            h = self.b[self.p]
            self.p += 1
            n, et = h >> 4, h & 0xF
            if n == 15:
                n = self.varint()
            if et in (1, 2):  # bools in a list are one byte each
                out = [self.b[self.p + i] == 1 for i in range(n)]
                self.p += n
                return out
            return [self.value(et) for _ in range(n)]
        if t == 12:
            return self.struct()
        raise ParquetUnsupported(f"thrift type {t}")

    def struct(self) -> dict:
        out, fid = {}, 0
        while True:
            h = self.b[self.p]
            self.p += 1
            if h == 0:
                return out
            delta, t = h >> 4, h & 0xF
            fid = fid + delta if delta else self.zigzag()
            out[fid] = self.value(t)


def snappy_decode(buf, pos: int, end: int, limit: Optional[int] = None) -> bytes:
    """Snappy block format (the raw format Parquet uses), bytes [pos, end) of buf ->
    the first `limit` uncompressed bytes (all of them when None).  Host side only:
    dictionary pages and the first bytes of a v1 data page (its definition-level
    length); the data pages themselves are inflated on the device."""
    n = shift = 0
    while True:
        c = buf[pos]
        pos += 1
        n |= (c & 0x7F) << shift
        if not c & 0x80:
            break
        shift += 7
    want = n if limit is None else min(n, limit)
    out = bytearray()
    while len(out) < want and pos < end:
        tag = buf[pos]
        pos += 1
        kind = tag & 3
        if kind == 0:
            ln = tag >> 2
            if ln >= 60:
                extra = ln - 59
                ln = int.from_bytes(buf[pos:pos + extra], "little")
                pos += extra
            ln += 1
            out += buf[pos:pos + ln]
            pos += ln
            continue
        if kind == 1:
            ln = 4 + ((tag >> 2) & 7)
            off = ((tag >> 5) << 8) | buf[pos]
            pos += 1
        elif kind == 2:
            ln = (tag >> 2) + 1
            off = buf[pos] | (buf[pos + 1] << 8)
            pos += 2
        else:
            ln = (tag >> 2) + 1
            off = int.from_bytes(buf[pos:pos + 4], "little")
            pos += 4
        if off == 0 or off > len(out):
            raise ParquetUnsupported("corrupt snappy stream (copy offset outside the output)")
        start = len(out) - off
        if off >= ln:
            out += out[start:start + ln]
        else:                      # overlapping copy: the last `off` bytes repeat
            for i in range(ln):
                out.append(out[start + i])
    if limit is None and len(out) != n:
        raise ParquetUnsupported(f"corrupt snappy stream ({len(out)} of {n} bytes)")
    return bytes(out[:want])


CODEC_NONE, CODEC_SNAPPY = 0, 1
T_BOOLEAN, T_INT32, T_INT64, T_INT96, T_FLOAT, T_DOUBLE, T_BYTE_ARRAY, T_FIXED = range(8)
ENC_PLAIN, ENC_PLAIN_DICT, ENC_RLE, ENC_BIT_PACKED, ENC_RLE_DICT = 0, 2, 3, 4, 8  # parquet.thrift Encoding
PAGE_DATA, PAGE_INDEX, PAGE_DICT, PAGE_DATA_V2 = range(4)
_TYPE_WIDTH = {T_INT32: 4, T_INT64: 8, T_FLOAT: 4, T_DOUBLE: 8}
# value kinds understood by the gather kernel
KIND_F64, KIND_F32, KIND_I32, KIND_I64, KIND_DICT = 0, 1, 2, 3, 4


@dataclass
class _Page:
    row_start: int          # first row (file-global) of the page
    num_values: int         # rows in the page (flat schema: values incl. nulls)
    def_offset: int         # byte offset of the definition-level hybrid stream, -1 = no nulls possible
    def_bytes: int
    val_offset: int         # byte offset of the value section
    val_bytes: int
    encoding: int           # ENC_PLAIN or a dictionary encoding
    dict_id: int            # index into the column's dictionaries, -1 if PLAIN
    bit_width: int = 0      # dictionary pages: width of the indices (first byte of the value section)


@dataclass
class _ColumnPlan:
    name: str
    ptype: int
    optional: bool
    pages: List[_Page] = field(default_factory=list)
    dictionaries: List[object] = field(default_factory=list)  # list[str] or np.ndarray per row group
    inflate: List[Tuple[int, int, int, int]] = field(default_factory=list)   # (src, src_bytes, dst, dst_bytes) per compressed page

    def page_table(self) -> np.ndarray:
        """int64[pages, 8]: row_start, num_values, def_offset, def_bytes, val_offset, val_bytes,
        dict_id, bit_width - what Engine.parquet_decode_columns vectorises over."""
        if getattr(self, "_table", None) is None or self._table.shape[0] != len(self.pages):
            self._table = np.array([[p.row_start, p.num_values, p.def_offset, p.def_bytes, p.val_offset, p.val_bytes,
                                     p.dict_id, p.bit_width] for p in self.pages], dtype=np.int64).reshape(-1, 8)
        return self._table


def _plain_strings(buf: bytes, n: int) -> List[str]:
    out, p = [], 0
    for _ in range(n):
        (ln,) = struct.unpack_from("<I", buf, p)
        out.append(buf[p + 4:p + 4 + ln].decode("utf-8"))
        p += 4 + ln
    return out


def _plan(raw: bytes, wanted: Sequence[str], inflated_base: Optional[int] = None) -> Tuple[int, Dict[str, _ColumnPlan], int]:
    """-> (rows, plans, end of the buffer).  Compressed data pages are given a place
    in the region that starts at `inflated_base` (default: right behind the padded
    file bytes) - 64-byte aligned, 64 bytes apart - and every page offset of such a
    page points there; the third result is where that region ends."""
    if raw[:4] != b"PAR1" or raw[-4:] != b"PAR1":
        raise ParquetUnsupported("not a Parquet file (or an encrypted footer)")
    (flen,) = struct.unpack_from("<I", raw, len(raw) - 8)
    meta = _Thrift(raw, len(raw) - 8 - flen).struct()
    schema = meta[2]
    root_children = schema[0].get(5, 0)
    leaves = schema[1:]
    if len(leaves) != root_children or any(5 in e and e[5] for e in leaves):
        raise ParquetUnsupported("nested schemas are not supported (flat columns only)")
    names = [e[4].decode() for e in leaves]
    plans: Dict[str, _ColumnPlan] = {}
    for name in wanted:
        if name not in names:
            continue
        e = leaves[names.index(name)]
        rep = e.get(3, 0)  # 0 REQUIRED, 1 OPTIONAL, 2 REPEATED
        if rep == 2:
            raise ParquetUnsupported(f"column {name}: repeated fields are not supported")
        plans[name] = _ColumnPlan(name, e[1], rep == 1)
    from .table import _padded
    cursor = _padded(len(raw)) if inflated_base is None else inflated_base

    def place(nbytes: int) -> int:
        nonlocal cursor
        at = cursor
        cursor += (nbytes + 127) // 64 * 64
        return at

    row0 = 0
    for rg in meta[4]:
        for cc in rg[1]:
            cm = cc[3]
            name = ".".join(x.decode() for x in cm[3])
            if name not in plans:
                continue
            plan = plans[name]
            codec = cm[4]
            if codec not in (CODEC_NONE, CODEC_SNAPPY):
                raise ParquetUnsupported(f"column {name}: compression codec {codec} is not supported (UNCOMPRESSED and SNAPPY are)")
            starts = [x for x in (cm.get(11), cm.get(9)) if x]  # 0 / absent = no such page
            if not starts:
                continue
            pos = min(starts)
            end = pos + cm[7]
            row = row0
            dict_id = -1
            while pos < end:
                th = _Thrift(raw, pos)
                ph = th.struct()
                body = th.p
                size, usize = ph[3], ph[2]          # compressed / uncompressed size of what follows the header
                if ph[1] == PAGE_DICT:
                    dh = ph[7]
                    if dh[2] not in (ENC_PLAIN, ENC_PLAIN_DICT):
                        raise ParquetUnsupported(f"column {name}: dictionary page encoding {dh[2]}")
                    blob = snappy_decode(raw, body, body + size) if codec == CODEC_SNAPPY else raw[body:body + size]
                    if plan.ptype == T_BYTE_ARRAY:
                        plan.dictionaries.append(_plain_strings(blob, dh[1]))
                    elif plan.ptype in _TYPE_WIDTH:
                        dt = {T_INT32: "<i4", T_INT64: "<i8", T_FLOAT: "<f4", T_DOUBLE: "<f8"}[plan.ptype]
                        plan.dictionaries.append(np.frombuffer(blob, dtype=dt, count=dh[1]).astype(np.float64))
                    else:
                        raise ParquetUnsupported(f"column {name}: physical type {plan.ptype}")
                    dict_id = len(plan.dictionaries) - 1
                elif ph[1] == PAGE_DATA:
                    dh = ph[5]
                    nv, enc = dh[1], dh[2]
                    if plan.optional and dh[3] != ENC_RLE:
                        raise ParquetUnsupported(f"column {name}: definition-level encoding {dh[3]}")
                    is_dict = enc in (ENC_PLAIN_DICT, ENC_RLE_DICT)
                    if codec == CODEC_SNAPPY:
                        # the whole page body is one Snappy block; it is inflated on the device.  The host only needs
                        # the 4-byte length of the definition levels in front of it: the block's first bytes
                        p = place(usize)
                        plan.inflate.append((body, size, p, usize))
                        p_end, bw = p + usize, -1
                        first = snappy_decode(raw, body, body + size, 4) if plan.optional else b""
                    else:
                        p, p_end, first = body, body + size, raw[body:body + 4]
                    def_off, def_bytes = -1, 0
                    if plan.optional:
                        (def_bytes,) = struct.unpack("<I", first)
                        def_off = p + 4
                        p += 4 + def_bytes
                    if codec != CODEC_SNAPPY:
                        bw = raw[p] if is_dict and p_end > p else 0
                    plan.pages.append(_Page(row, nv, def_off, def_bytes, p, p_end - p, enc, dict_id if is_dict else -1, bw))
                    row += nv
                elif ph[1] == PAGE_DATA_V2:
                    dh = ph[8]
                    nv, enc = dh[1], dh[4]
                    dl, rl = dh[5], dh[6]
                    if rl:
                        raise ParquetUnsupported(f"column {name}: repetition levels present")
                    def_off = body if (plan.optional and dl) else -1      # v2 levels are never compressed
                    is_dict = enc in (ENC_PLAIN_DICT, ENC_RLE_DICT)
                    vo, vbytes = body + dl + rl, size - dl - rl
                    if codec == CODEC_SNAPPY and dh.get(7, True) and vbytes > 0:
                        ubytes = usize - dl - rl
                        at = place(ubytes)
                        plan.inflate.append((vo, vbytes, at, ubytes))
                        vo, vbytes, bw = at, ubytes, -1
                    else:
                        bw = raw[vo] if is_dict and vbytes > 0 else 0
                    plan.pages.append(_Page(row, nv, def_off, dl, vo, vbytes, enc, dict_id if is_dict else -1, bw))
                    row += nv
                else:
                    pass  # index pages carry no values
                pos = body + size
            if row - row0 != rg[3]:
                raise ParquetUnsupported(f"column {name}: page rows {row - row0} != row group rows {rg[3]}")
        row0 += rg[3]
    return int(meta[3]), plans, cursor + 64


def read_parquet(eng: Engine, source, numeric: Sequence[str], strings: Sequence[str], dict_capacity: int = 1 << 16) -> DeviceTable:
    """The Parquet counterpart of read_csv: numeric columns -> f64 + validity,
    string columns -> dictionary codes (labels sorted ascending) + validity."""
    raw = read_bytes(source)
    rows, plans, buffer_end = _plan(raw, list(numeric) + list(strings))
    # the file bytes, then room for the pages the device inflates
    data = upload_text(eng, raw, extra=max(0, buffer_end - _padded_len(len(raw))))
    jobs_inflate = [j for plan in plans.values() for j in plan.inflate]
    status = eng.parquet_inflate(data, jobs_inflate)
    table = DeviceTable(eng, rows, header=list(plans))
    jobs = []   # (name, want_string, plan): decoded together, one launch for all their streams
    for name in list(numeric) + list(strings):
        want_string = name in strings
        if name not in plans:  # absent column: all null, as read_csv does
            nullmap = torch.zeros(words_for(rows), dtype=torch.int32, device=eng.device)
            if want_string:
                table.strings[name] = StringColumn(Column(torch.full((rows,), -1, dtype=torch.int32, device=eng.device), nullmap), [])
            else:
                table.numeric[name] = Column(torch.zeros(rows, dtype=torch.float64, device=eng.device), nullmap)
            continue
        plan = plans[name]
        if want_string != (plan.ptype == T_BYTE_ARRAY):
            raise ParquetUnsupported(f"column {name}: physical type {plan.ptype} does not match the requested kind")
        if not want_string:
            if plan.ptype not in _TYPE_WIDTH:
                raise ParquetUnsupported(f"column {name}: physical type {plan.ptype}")
        for pg in plan.pages:
            if pg.dict_id < 0 and pg.encoding != ENC_PLAIN:
                raise ParquetUnsupported(f"column {name}: value encoding {pg.encoding}")
        jobs.append((name, want_string, plan))
    cols, label_lists = eng.parquet_decode_columns(data, rows, [column_job(plan, want_string) for _, want_string, plan in jobs],
                                                   status=status, dict_capacity=dict_capacity)
    for (name, want_string, plan), col, labels in zip(jobs, cols, label_lists):
        if want_string:
            table.strings[name] = StringColumn(col, labels)
        else:
            table.numeric[name] = col
            table.inferred_types[name] = {T_DOUBLE: "double", T_FLOAT: "float", T_INT32: "int", T_INT64: "long"}[plan.ptype]
    return table


def _padded_len(n: int) -> int:
    from .table import _padded
    return _padded(n)


def column_job(plan: _ColumnPlan, want_string: bool) -> dict:
    """One column's work order for Engine.parquet_decode_columns: the page table, the PLAIN value
    kind and the per-row-group dictionaries (label lists for strings, value arrays for numerics)."""
    kind = {T_DOUBLE: KIND_F64, T_FLOAT: KIND_F32, T_INT32: KIND_I32, T_INT64: KIND_I64}.get(plan.ptype, KIND_DICT)
    return {"pages": plan.page_table(), "kind_plain": kind, "string": want_string, "dictionaries": plan.dictionaries}


# ---------------------------------------------------------------- small nested files (Spark ML model data)
# model.save / pipeline_model.save (spark_workload_to_cloud_k8s.py:149-154) write a few hundred bytes of
# Parquet per model: one row of labelsArray list<list<string>>, one of categorySizes list<int>, k rows of
# {clusterIdx int, clusterCenter struct<type, size, indices list<int>, values list<double>>}.  Spark reads them
# on the driver; so does this reader - with the footer / page-header parser, the Snappy decoder and the level
# decoding of this module (no other Parquet library), assembling records from repetition and definition levels.

def _hybrid_host(buf: bytes, pos: int, end: int, bit_width: int, count: int) -> List[int]:
    """RLE / bit-packed hybrid stream -> the first `count` values."""
    out: List[int] = []
    if bit_width == 0:
        return [0] * count
    vbytes = (bit_width + 7) // 8
    while len(out) < count and pos < end:
        h = shift = 0
        while True:
            c = buf[pos]
            pos += 1
            h |= (c & 0x7F) << shift
            shift += 7
            if not c & 0x80:
                break
        if h & 1:                                   # bit-packed groups of 8
            n = (h >> 1) * 8
            nbytes = (h >> 1) * bit_width
            bits = int.from_bytes(buf[pos:pos + nbytes], "little")
            mask = (1 << bit_width) - 1
            out.extend((bits >> (i * bit_width)) & mask for i in range(n))
            pos += nbytes
        else:                                       # a run
            v = int.from_bytes(buf[pos:pos + vbytes], "little")
            pos += vbytes
            out.extend([v] * (h >> 1))
    return out[:count]


def _plain_host(ptype: int, buf: bytes, pos: int, n: int) -> list:
    if ptype == T_BYTE_ARRAY:
        out = []
        for _ in range(n):
            (ln,) = struct.unpack_from("<I", buf, pos)
            out.append(buf[pos + 4:pos + 4 + ln])
            pos += 4 + ln
        return out
    if ptype == T_BOOLEAN:
        return [bool((buf[pos + i // 8] >> (i % 8)) & 1) for i in range(n)]
    fmt = {T_INT32: "<i4", T_INT64: "<i8", T_FLOAT: "<f4", T_DOUBLE: "<f8"}.get(ptype)
    if fmt is None:
        raise ParquetUnsupported(f"physical type {ptype}")
    return np.frombuffer(buf, dtype=fmt, count=n, offset=pos).tolist()


@dataclass
class _Node:
    name: str
    repetition: int                 # 0 required, 1 optional, 2 repeated
    ptype: Optional[int]            # None for groups
    converted: Optional[int]        # ConvertedType (0 = UTF8, 3 = LIST, 15 INT_8 ...)
    children: List["_Node"] = field(default_factory=list)


def _schema_tree(elements: List[dict]) -> _Node:
    it = iter(elements)

    def take() -> _Node:
        e = next(it)
        node = _Node(e[4].decode(), e.get(3, 0), e.get(1) if not e.get(5) else None, e.get(6))
        for _ in range(e.get(5, 0) or 0):
            node.children.append(take())
        return node

    return take()


def _read_leaf(raw: bytes, cm: dict, ptype: int, max_rep: int, max_def: int, string: bool):
    """One column chunk -> (rep levels, def levels, non-null values), v1 / v2 pages, PLAIN or dictionary."""
    codec = cm[4]
    if codec not in (CODEC_NONE, CODEC_SNAPPY):
        raise ParquetUnsupported(f"compression codec {codec}")
    starts = [x for x in (cm.get(11), cm.get(9)) if x]
    pos, end = min(starts), min(starts) + cm[7]
    reps: List[int] = []
    defs: List[int] = []
    vals: list = []
    dictionary: Optional[list] = None
    bits = lambda m: max(m, 0).bit_length()       # noqa: E731
    while pos < end:
        th = _Thrift(raw, pos)
        ph = th.struct()
        body, size = th.p, ph[3]
        pos = body + size
        if ph[1] == PAGE_DICT:
            blob = snappy_decode(raw, body, body + size) if codec == CODEC_SNAPPY else raw[body:body + size]
            dictionary = _plain_host(ptype, blob, 0, ph[7][1])
            continue
        if ph[1] == PAGE_DATA:
            page = snappy_decode(raw, body, body + size) if codec == CODEC_SNAPPY else raw[body:body + size]
            nv, enc = ph[5][1], ph[5][2]
            p = 0
            r = [0] * nv
            d = [max_def] * nv
            if max_rep:
                (ln,) = struct.unpack_from("<I", page, p)
                r = _hybrid_host(page, p + 4, p + 4 + ln, bits(max_rep), nv)
                p += 4 + ln
            if max_def:
                (ln,) = struct.unpack_from("<I", page, p)
                d = _hybrid_host(page, p + 4, p + 4 + ln, bits(max_def), nv)
                p += 4 + ln
        elif ph[1] == PAGE_DATA_V2:
            dh = ph[8]
            nv, enc, dl, rl = dh[1], dh[4], dh[5], dh[6]
            r = _hybrid_host(raw, body, body + rl, bits(max_rep), nv) if max_rep else [0] * nv
            d = _hybrid_host(raw, body + rl, body + rl + dl, bits(max_def), nv) if max_def else [max_def] * nv
            vb = body + rl + dl
            page = snappy_decode(raw, vb, body + size) if (codec == CODEC_SNAPPY and dh.get(7, True) and body + size > vb) \
                else raw[vb:body + size]
            p = 0
        else:
            continue
        present = sum(1 for x in d if x == max_def)
        if enc in (ENC_PLAIN_DICT, ENC_RLE_DICT):
            if dictionary is None:
                raise ParquetUnsupported("dictionary-encoded page without a dictionary page")
            idx = _hybrid_host(page, p + 1, len(page), page[p], present) if present else []
            v = [dictionary[i] for i in idx]
        elif enc == ENC_PLAIN:
            v = _plain_host(ptype, page, p, present)
        else:
            raise ParquetUnsupported(f"value encoding {enc}")
        reps += r
        defs += d
        vals += [x.decode("utf-8") for x in v] if string else v
    return reps, defs, vals


class _Null:
    """A null met while descending a leaf's path: `level` = index of the path node that was null."""

    def __init__(self, level: int):
        self.level = level


def _build_leaf(nodes: List[_Node], reps: List[int], defs: List[int], vals: list) -> list:
    """Record assembly for ONE leaf column: per row, the nested value along the leaf's path - a list per
    repeated node, _Null(level) where an optional node is null, the leaf value at the bottom.  Groups that
    are neither null nor repeated are transparent (their fields are zipped by the caller)."""
    max_def = sum(1 for n in nodes if n.repetition != 0)
    vidx, k = [], 0
    for d in defs:
        vidx.append(k if d == max_def else -1)
        k += d == max_def

    def build(level: int, lo: int, hi: int, rep_level: int, def_level: int):
        node = nodes[level]
        if node.repetition == 1:
            def_level += 1
            if defs[lo] < def_level:
                return _Null(level)
        elif node.repetition == 2:
            def_level += 1
            rep_level += 1
            if defs[lo] < def_level:
                return []
            starts = [lo] + [i for i in range(lo + 1, hi) if reps[i] == rep_level]
            return [down(level, a, b, rep_level, def_level) for a, b in zip(starts, starts[1:] + [hi])]
        return down(level, lo, hi, rep_level, def_level)

    def down(level: int, lo: int, hi: int, rep_level: int, def_level: int):
        if level == len(nodes) - 1:
            return vals[vidx[lo]]
        v = build(level + 1, lo, hi, rep_level, def_level)
        return v

    row_starts = [i for i, r in enumerate(reps) if r == 0] + [len(reps)]
    return [build(0, a, b, 0, 0) for a, b in zip(row_starts[:-1], row_starts[1:])]


def _denull(v):
    if isinstance(v, _Null):
        return None
    if isinstance(v, list):
        return [_denull(x) for x in v]
    return v


def read_nested_small(source) -> Dict[str, list]:
    """A (small) Parquet file with nested columns -> {column: [one Python value per row]}: lists for
    LIST-annotated groups, dicts for structs, None for nulls - what pyarrow's Table.to_pydict() gives.
    Host-side record assembly from repetition / definition levels; meant for the Spark ML model files, not
    for bulk data (read_parquet).  Structs inside lists are not supported."""
    raw = read_bytes(source)
    if raw[:4] != b"PAR1" or raw[-4:] != b"PAR1":
        raise ParquetUnsupported("not a Parquet file (or an encrypted footer)")
    (flen,) = struct.unpack_from("<I", raw, len(raw) - 8)
    meta = _Thrift(raw, len(raw) - 8 - flen).struct()
    root = _schema_tree(meta[2])
    rows_total = int(meta[3])

    def leaves(node: _Node, path, nodes):
        for ch in node.children:
            if ch.ptype is None:
                yield from leaves(ch, path + (ch.name,), nodes + [ch])
            else:
                yield path + (ch.name,), nodes + [ch]

    info = dict(leaves(root, (), []))
    streams: Dict[Tuple[str, ...], list] = {p: [[], [], []] for p in info}
    for rg in meta[4]:
        for cc in rg[1]:
            cm = cc[3]
            path = tuple(x.decode() for x in cm[3])
            nodes = info[path]
            leaf = nodes[-1]
            max_rep = sum(1 for n in nodes if n.repetition == 2)
            max_def = sum(1 for n in nodes if n.repetition != 0)
            r, d, v = _read_leaf(raw, cm, leaf.ptype, max_rep, max_def, leaf.ptype == T_BYTE_ARRAY and leaf.converted in (0, None))
            for acc, part in zip(streams[path], (r, d, v)):
                acc += part
    built = {}
    for path, (r, d, v) in streams.items():
        built[path] = _build_leaf(info[path], r, d, v)
        if len(built[path]) != rows_total:
            raise ParquetUnsupported(f"column {'.'.join(path)}: {len(built[path])} records for {rows_total} rows")

    def value_of(node: _Node, path, depth: int, row: int):
        below = [p for p in info if p[:len(path)] == path]
        if node.ptype is not None or len(below) == 1:
            return _denull(built[below[0]][row])
        if any(n.repetition == 2 for p in below for n in info[p][:depth + 1]) or node.repetition == 2:
            raise ParquetUnsupported(f"column {'.'.join(path)}: a struct inside a list is not supported")
        states = [built[p][row] for p in below]
        if all(isinstance(x, _Null) and x.level <= depth for x in states):
            return None                                    # the struct itself is null
        return {ch.name: value_of(ch, path + (ch.name,), depth + 1, row) for ch in node.children}

    return {ch.name: [value_of(ch, (ch.name,), 0, row) for row in range(rows_total)] for ch in root.children}
