"""Per-GPU engine: owns the C-ABI handle and wraps each entry point for torch
device buffers.  torch is used for device memory and streams only; all compute
is in libetl_b200.so.  No CPU path exists here: every method needs CUDA."""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import ETL_F32, ETL_F64, ETL_I32, ETL_I64, F64x3, I32x3, VP3

_DTYPE_CODE = {torch.float64: ETL_F64, torch.float32: ETL_F32, torch.int32: ETL_I32, torch.int64: ETL_I64}


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def words_for(n: int) -> int:
    return (n + 31) // 32


@dataclass
class Column:
    """A device column: `data` (1-D tensor) + optional validity bitmap (`valid`,
    int32 words, Arrow bit order, 1 = non-null)."""

    data: torch.Tensor
    valid: Optional[torch.Tensor] = None

    def __len__(self):
        return self.data.shape[0]

    @property
    def dtype_code(self) -> int:
        return _DTYPE_CODE[self.data.dtype]

    @staticmethod
    def from_host(values: np.ndarray, valid: Optional[np.ndarray], device) -> "Column":
        data = torch.from_numpy(np.ascontiguousarray(values)).to(device)
        bitmap = None
        if valid is not None:
            bitmap = torch.from_numpy(pack_validity_host(valid).view(np.int32)).to(device)
        return Column(data, bitmap)

    def valid_host(self) -> np.ndarray:
        n = len(self)
        if self.valid is None:
            return np.ones(n, dtype=bool)
        w = self.valid.cpu().numpy().view(np.uint8)
        return np.unpackbits(w, bitorder="little")[:n].astype(bool)


def pack_validity_host(valid: np.ndarray) -> np.ndarray:
    valid = np.asarray(valid, dtype=bool)
    n = valid.shape[0]
    padded = np.zeros(words_for(n) * 32, dtype=np.uint8)
    padded[:n] = valid
    return np.packbits(padded, bitorder="little").view(np.uint32).copy()


@dataclass
class GroupTable:
    """Caller-owned hash-aggregation table (see etl_groupby_* in etl_b200.h)."""

    table: torch.Tensor
    capacity: int
    ncols: int
    minmax: bool
    status: torch.Tensor


@dataclass
class PartitionedGroups:
    """Result of the shared-memory merge (etl_groupby_merge_partitioned): dense
    columns in no particular order + device counters [groups, spilled, -, reserved]
    + the spilled records; Engine.groupby_result turns it into the result dict."""

    keys: torch.Tensor
    rows: torch.Tensor
    count: torch.Tensor
    sum: torch.Tensor
    counters: torch.Tensor
    spill: torch.Tensor
    capacity: int
    pieces: int
    # slab path only: counters[4] counts records that found their slab full.  If it (any_overflow: over all
    # ranks) is non-zero the groups above are incomplete and rerun() repeats the work on the counted path.
    rerun: Optional[Callable[[], "PartitionedGroups"]] = None
    any_overflow: Optional[Callable[[int], int]] = None
    min: Optional[torch.Tensor] = None      # min(value) / max(value) per group when asked for (minmax=True)
    max: Optional[torch.Tensor] = None


class Engine:
    """One per GPU (replaces the per-executor SparkSession state,
    spark_session.py:37-88)."""

    def __init__(self, device: Optional[int] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("pyspark_tf_gke_b200 needs a CUDA device: there is no CPU fallback")
        self.lib = _lib.load()
        self.device_index = torch.cuda.current_device() if device is None else int(device)
        self.device = torch.device("cuda", self.device_index)
        h = C.c_void_p()
        _lib.check(self.lib.etl_create(self.device_index, C.byref(h)))
        self._h = h
        self.launches = 0  # kernels of ours launched through this engine

    def close(self):
        if getattr(self, "_h", None):
            self.lib.etl_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------
    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def reserve(self, nbytes: int):
        _lib.check(self.lib.etl_reserve(self._h, nbytes))

    def empty(self, n: int, dtype) -> torch.Tensor:
        return torch.empty(n, dtype=dtype, device=self.device)

    # ---------------------------------------------------------------- a2
    def null_count(self, col: Column, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """filter(col.isNull()).count() (k_means.py:23) -> device int64 scalar."""
        out = self.empty(1, torch.int64) if out is None else out
        _lib.check(self.lib.etl_null_count(self._h, _ptr(col.valid), len(col), out.data_ptr(), self._stream()))
        self.launches += 1 if col.valid is not None and len(col) else 0
        return out

    # ---------------------------------------------------------------- a4
    def sum_count(self, cols: Sequence[Column], row_valid: Optional[torch.Tensor] = None,
                  out_sum: Optional[torch.Tensor] = None, out_count: Optional[torch.Tensor] = None):
        """(sum, count) over ~isnan & isNotNull for up to 3 columns
        (k_means.py:45-48) -> device tensors f64[ncols], i64[ncols]."""
        nc = len(cols)
        n = len(cols[0])
        out_sum = self.empty(3, torch.float64) if out_sum is None else out_sum
        out_count = self.empty(3, torch.int64) if out_count is None else out_count
        val = VP3(*[_ptr(c.data) for c in cols], *([None] * (3 - nc)))
        vv = VP3(*[_ptr(c.valid) for c in cols], *([None] * (3 - nc)))
        dt = I32x3(*[c.dtype_code for c in cols], *([0] * (3 - nc)))
        _lib.check(self.lib.etl_column_sum_count(self._h, nc, C.byref(val), C.byref(vv), C.byref(dt), _ptr(row_valid),
                                                 n, out_sum.data_ptr(), out_count.data_ptr(), self._stream()))
        self.launches += 1
        return out_sum[:nc], out_count[:nc]

    def column_moments(self, cols: Sequence[Column], row_valid: Optional[torch.Tensor] = None):
        """StandardScaler.fit's summary of up to 3 columns -> device tensors (count
        i64[nc], mean f64[nc], m2 f64[nc]): m2 = sum of squared deviations from the
        mean over the non-null, non-NaN cells (two passes, nothing read back)."""
        nc, n = len(cols), len(cols[0])
        count, mean, m2 = self.empty(3, torch.int64), self.empty(3, torch.float64), self.empty(3, torch.float64)
        val = VP3(*[_ptr(c.data) for c in cols], *([None] * (3 - nc)))
        vv = VP3(*[_ptr(c.valid) for c in cols], *([None] * (3 - nc)))
        dt = I32x3(*[c.dtype_code for c in cols], *([0] * (3 - nc)))
        _lib.check(self.lib.etl_column_moments(self._h, nc, C.byref(val), C.byref(vv), C.byref(dt), _ptr(row_valid), n,
                                               count.data_ptr(), mean.data_ptr(), m2.data_ptr(), self._stream()))
        self.launches += 2
        return count[:nc], mean[:nc], m2[:nc]

    def standard_scaler_fit(self, cols: Sequence[Column], row_valid: Optional[torch.Tensor] = None,
                            with_mean: bool = False, with_std: bool = True):
        """Spark ML StandardScaler(withMean, withStd).fit on up to 3 columns ->
        (shift, scale) lists for assemble_dense / assemble_sparse / pack_tf: shift =
        mean or 0, scale = 1 / std (unbiased; 0 where std == 0, as Spark does) or 1."""
        count, mean, m2 = (t.cpu().tolist() for t in self.column_moments(cols, row_valid))
        shift = [m if with_mean else 0.0 for m in mean]
        scale = []
        for k, q in zip(count, m2):
            std = float(np.sqrt(q / (k - 1))) if k > 1 else 0.0
            scale.append((1.0 / std if std != 0.0 else 0.0) if with_std else 1.0)
        return shift, scale

    # ------------------------------------------------------------- a3/a5
    def project_filter_impute(self, key: Column, vals: Sequence[Column], fill: Sequence[float],
                              out_code: Optional[torch.Tensor] = None,
                              out_vals: Optional[Sequence[torch.Tensor]] = None,
                              out_rows: Optional[torch.Tensor] = None):
        """filter(key.isNotNull()) + 3x when(isNull|isnan, mean).otherwise(c) as
        double, compacted in input order (k_means.py:27-28,49-51).  Returns
        (out_code i32[n], [3 x f64[n]], rows i64[1]) - the first `rows` entries
        of each output are the kept rows."""
        n = len(key)
        assert len(vals) == 3 and len(fill) == 3
        out_code = self.empty(n, torch.int32) if out_code is None else out_code
        out_vals = [self.empty(n, torch.float64) for _ in range(3)] if out_vals is None else list(out_vals)
        out_rows = self.empty(1, torch.int64) if out_rows is None else out_rows
        val = VP3(*[_ptr(c.data) for c in vals])
        vv = VP3(*[_ptr(c.valid) for c in vals])
        dt = I32x3(*[c.dtype_code for c in vals])
        fl = F64x3(*[float(f) for f in fill])
        ov = VP3(*[t.data_ptr() for t in out_vals])
        _lib.check(self.lib.etl_project_filter_impute(self._h, _ptr(key.data), _ptr(key.valid), C.byref(val),
                                                      C.byref(vv), C.byref(dt), C.byref(fl), n, out_code.data_ptr(),
                                                      C.byref(ov), out_rows.data_ptr(), self._stream()))
        self.launches += 1 if n else 0
        return out_code, out_vals, out_rows

    def project_filter_impute_host(self, key_code: np.ndarray, key_valid: Optional[np.ndarray],
                                   vals: Sequence[np.ndarray], vals_valid: Sequence[Optional[np.ndarray]],
                                   fill: Sequence[float], out_code: np.ndarray, out_vals: Sequence[np.ndarray]) -> int:
        """Host-buffer entry (numpy arrays; bitmaps as uint32 words).  Returns the
        kept-row count; outputs are written into the given host arrays."""
        n = key_code.shape[0]
        np_code = {np.dtype(np.float64): ETL_F64, np.dtype(np.float32): ETL_F32, np.dtype(np.int32): ETL_I32}

        def hp(a):
            return None if a is None else a.ctypes.data

        val = VP3(*[hp(v) for v in vals])
        vv = VP3(*[hp(v) for v in vals_valid])
        dt = I32x3(*[np_code[v.dtype] for v in vals])
        fl = F64x3(*[float(f) for f in fill])
        ov = VP3(*[hp(v) for v in out_vals])
        rows = C.c_int64(0)
        _lib.check(self.lib.etl_project_filter_impute_host(self._h, hp(key_code), hp(key_valid), C.byref(val),
                                                           C.byref(vv), C.byref(dt), C.byref(fl), n, hp(out_code),
                                                           C.byref(ov), C.byref(rows)))
        self.launches += int(self.lib.etl_project_filter_impute_host_launches(n))
        return rows.value

    # -------------------------------------------------------------- a1 csv
    def csv_index_rows(self, text: torch.Tensor, nbytes: int) -> torch.Tensor:
        out = self.empty(1, torch.int64)
        _lib.check(self.lib.etl_csv_index_rows(self._h, text.data_ptr(), nbytes, out.data_ptr(), self._stream()))
        self.launches += 2 if nbytes else 0
        return out

    def csv_row_offsets(self, text: torch.Tensor, nbytes: int, rows: int) -> torch.Tensor:
        row_start = self.empty(max(rows, 1), torch.int64)
        _lib.check(self.lib.etl_csv_row_offsets(self._h, text.data_ptr(), nbytes, row_start.data_ptr(), self._stream()))
        self.launches += 1 if nbytes else 0
        return row_start[:rows]

    def csv_parse_columns(self, text: torch.Tensor, nbytes: int, row_start: torch.Tensor, first_row: int, rows: int,
                          fields: Sequence[int], kinds: Sequence[int], dicts: Sequence[Optional[Tuple[torch.Tensor, torch.Tensor]]],
                          flags: int = 0):
        """-> (columns, status int32[4] on host, first_bad_row).  kinds: 0 numeric
        (f64), 1 string (int32 dictionary slot).  dicts[s] = (hash, rep) uint64
        tables (as int64 tensors, power-of-two length) for string columns."""
        nc = len(fields)
        outs, valids = [], []
        for k in kinds:
            outs.append(self.empty(rows, torch.float64 if k == 0 else torch.int32))
            valids.append(torch.zeros(words_for(rows), dtype=torch.int32, device=self.device))
        status = torch.zeros(4, dtype=torch.int32, device=self.device)
        first_bad = torch.full((1,), torch.iinfo(torch.int64).max, dtype=torch.int64, device=self.device)
        I = C.c_int32 * nc
        P = C.c_void_p * nc
        dh = P(*[(d[0].data_ptr() if d is not None else None) for d in dicts])
        dr = P(*[(d[1].data_ptr() if d is not None else None) for d in dicts])
        dc = I(*[(d[0].shape[0] if d is not None else 0) for d in dicts])
        _lib.check(self.lib.etl_csv_parse_columns(
            self._h, text.data_ptr(), nbytes, _ptr(row_start), first_row, rows, nc, I(*fields), I(*kinds),
            P(*[t.data_ptr() for t in outs]), P(*[t.data_ptr() for t in valids]), dh, dr, dc, flags,
            status.data_ptr(), first_bad.data_ptr(), self._stream()))
        self.launches += 1 if rows else 0
        cols = [Column(o, v) for o, v in zip(outs, valids)]
        return cols, status, first_bad

    def remap_codes(self, codes: torch.Tensor, table: torch.Tensor, missing: int) -> torch.Tensor:
        """StringIndexerModel.transform on codes, in place (k_means.py:74)."""
        _lib.check(self.lib.etl_remap_codes(self._h, codes.data_ptr(), codes.shape[0], table.data_ptr(), table.shape[0],
                                            missing, self._stream()))
        self.launches += 1 if codes.shape[0] else 0
        return codes

    # ---------------------------------------------------------------- a6
    def code_histogram(self, col: Column, bins: int):
        """group-by-count of StringIndexer.fit (k_means.py:34,73) -> (int64[bins], out_of_range int64[1])."""
        counts = self.empty(bins, torch.int64)
        oor = self.empty(1, torch.int64)
        _lib.check(self.lib.etl_code_histogram(self._h, _ptr(col.data), _ptr(col.valid), len(col), bins,
                                               counts.data_ptr(), oor.data_ptr(), self._stream()))
        self.launches += 1 if len(col) else 0
        return counts, oor

    # ------------------------------------------------------------ a8-a10
    def assemble_dense(self, index: torch.Tensor, nums: Sequence[torch.Tensor], num_labels: int, repeats: int,
                       dtype=torch.float64, shift=None, scale=None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        n = index.shape[0]
        D = repeats * num_labels + 3
        out = torch.empty((n, D), dtype=dtype, device=self.device) if out is None else out
        sh = F64x3(*shift) if shift is not None else None
        sc = F64x3(*scale) if scale is not None else None
        _lib.check(self.lib.etl_assemble_features_dense(
            self._h, index.data_ptr(), C.byref(VP3(*[t.data_ptr() for t in nums])),
            C.cast(sh, C.c_void_p) if sh is not None else None, C.cast(sc, C.c_void_p) if sc is not None else None,
            n, num_labels, repeats, _DTYPE_CODE[dtype], out.data_ptr(), self._stream()))
        self.launches += 1 if n else 0
        return out

    def assemble_sparse(self, index: torch.Tensor, nums: Sequence[torch.Tensor], num_labels: int, repeats: int,
                        shift=None, scale=None):
        n = index.shape[0]
        W = repeats + 3
        oi = torch.empty((n, W), dtype=torch.int32, device=self.device)
        ov = torch.empty((n, W), dtype=torch.float64, device=self.device)
        sh = F64x3(*shift) if shift is not None else None
        sc = F64x3(*scale) if scale is not None else None
        _lib.check(self.lib.etl_assemble_features_sparse(
            self._h, index.data_ptr(), C.byref(VP3(*[t.data_ptr() for t in nums])),
            C.cast(sh, C.c_void_p) if sh is not None else None, C.cast(sc, C.c_void_p) if sc is not None else None,
            n, num_labels, repeats, oi.data_ptr(), ov.data_ptr(), self._stream()))
        self.launches += 1 if n else 0
        return oi, ov

    # ----------------------------------------------------------- a11/a12
    def pack_tf(self, label: Column, nums: Sequence[Column], remap: Optional[torch.Tensor] = None, shift=None,
                scale=None):
        """Row filter + X float32[N,3] / y int32[N] of load_csv (train_tf_ps.py:120-149)."""
        n = len(label)
        X = torch.empty((n, 3), dtype=torch.float32, device=self.device)
        y = self.empty(n, torch.int32)
        rows = self.empty(1, torch.int64)
        sh = F64x3(*shift) if shift is not None else None
        sc = F64x3(*scale) if scale is not None else None
        _lib.check(self.lib.etl_pack_tf_features(
            self._h, _ptr(label.data), _ptr(label.valid), C.byref(VP3(*[_ptr(c.data) for c in nums])),
            C.byref(VP3(*[_ptr(c.valid) for c in nums])), C.cast(sh, C.c_void_p) if sh is not None else None,
            C.cast(sc, C.c_void_p) if sc is not None else None, _ptr(remap), 0 if remap is None else remap.shape[0], n,
            X.data_ptr(), y.data_ptr(), rows.data_ptr(), self._stream()))
        self.launches += 3 if n else 0
        return X, y, rows

    def nan_to_null(self, col: Column) -> Column:
        n = len(col)
        if col.valid is None:
            col.valid = torch.full((words_for(n),), -1, dtype=torch.int32, device=self.device)
        _lib.check(self.lib.etl_nan_to_null(self._h, _ptr(col.data), _ptr(col.valid), n, self._stream()))
        self.launches += 1 if n else 0
        return col

    def bucketize(self, col: Column, splits: Sequence[float]):
        n = len(col)
        d_splits = torch.tensor(list(splits), dtype=torch.float64, device=self.device)
        out = self.empty(n, torch.int32)
        oor = self.empty(1, torch.int64)
        _lib.check(self.lib.etl_bucketize(self._h, _ptr(col.data), _ptr(col.valid), n, d_splits.data_ptr(),
                                          d_splits.shape[0], out.data_ptr(), oor.data_ptr(), self._stream()))
        self.launches += 1 if n else 0
        return out, oor

    # ------------------------------------------------- groupBy (extension)
    def groupby_new(self, capacity: int, ncols: int, minmax: bool = False) -> "GroupTable":
        words = C.c_int64()
        _lib.check(self.lib.etl_groupby_table_words(ncols, int(minmax), capacity, C.byref(words)))
        table = self.empty(words.value, torch.int64)
        _lib.check(self.lib.etl_groupby_init(self._h, table.data_ptr(), capacity, ncols, int(minmax), self._stream()))
        self.launches += 1
        status = torch.zeros(2, dtype=torch.int64, device=self.device)
        return GroupTable(table, capacity, ncols, bool(minmax), status)

    def groupby_accumulate(self, gt: "GroupTable", keys: torch.Tensor, key_valid: Optional[torch.Tensor],
                           vals: Sequence[torch.Tensor], vals_valid: Sequence[Optional[torch.Tensor]] = (),
                           groups_hint: int = 0, n: Optional[int] = None, key_stride: int = 1,
                           val_strides: Optional[Sequence[int]] = None, key_range: int = 0) -> None:
        """One batch of rows into the table.  keys int32/int64; vals f64.  With
        strides it reads the {key, value} records of the exchange in place."""
        nc = gt.ncols
        assert len(vals) == nc
        n = keys.shape[0] if n is None else n
        P = C.c_void_p * max(nc, 1)
        L = C.c_int64 * max(nc, 1)
        vv = list(vals_valid) + [None] * (nc - len(vals_valid))
        strides = list(val_strides) if val_strides is not None else [1] * nc
        _lib.check(self.lib.etl_groupby_accumulate(
            self._h, keys.data_ptr(), _DTYPE_CODE[keys.dtype], key_stride, _ptr(key_valid), nc,
            P(*[v.data_ptr() for v in vals]) if nc else None, L(*strides) if nc else None,
            P(*[_ptr(v) for v in vv]) if nc else None, n, gt.table.data_ptr(), gt.capacity, int(gt.minmax),
            groups_hint, key_range, gt.status.data_ptr(), self._stream()))
        self.launches += 1 if n else 0

    def groupby_result(self, gt: "GroupTable", max_groups: Optional[int] = None, sort: bool = False) -> dict:
        """Dense result: keys, rows, per column count/sum/mean[/min/max] (device
        tensors), plus the null-key group.  Raises if rows were dropped.  The groups
        come as the kernels left them - Spark's own groupBy output is unordered too -
        and no library kernel runs; sort=True orders them by key (a radix sort of the
        key column by torch: a convenience for callers that compare results, which is
        what the tests do)."""
        if isinstance(gt, PartitionedGroups):
            return self._partitioned_result(gt, sort)
        cap = gt.capacity if max_groups is None else max_groups
        nc = gt.ncols
        keys = self.empty(cap, torch.int64)
        rows = self.empty(cap, torch.int64)
        cnt = torch.empty((max(nc, 1), cap), dtype=torch.int64, device=self.device)
        sm = torch.empty((max(nc, 1), cap), dtype=torch.float64, device=self.device)
        mn = torch.empty((max(nc, 1), cap), dtype=torch.float64, device=self.device) if gt.minmax else None
        mx = torch.empty((max(nc, 1), cap), dtype=torch.float64, device=self.device) if gt.minmax else None
        groups = self.empty(1, torch.int64)
        _lib.check(self.lib.etl_groupby_compact(self._h, gt.table.data_ptr(), gt.capacity, nc, int(gt.minmax), cap,
                                                keys.data_ptr(), rows.data_ptr(), cnt.data_ptr(), sm.data_ptr(),
                                                _ptr(mn), _ptr(mx), groups.data_ptr(), self._stream()))
        self.launches += 1
        st = gt.status.cpu().tolist()
        if st[0] or st[1]:
            raise _lib.EtlError(-5, f"groupBy: {st[0]} row(s) dropped (table full), {st[1]} row(s) with the reserved key")
        g = int(groups.item())
        if g > cap:
            raise _lib.EtlError(-5, f"groupBy: {g} groups exceed the output capacity {cap}")
        order = torch.argsort(keys[:g]) if sort else slice(None)
        out = {"keys": keys[:g][order], "rows": rows[:g][order], "cols": []}
        for c in range(nc):
            col = {"count": cnt[c, :g][order], "sum": sm[c, :g][order]}
            col["mean"] = col["sum"] / col["count"].to(torch.float64)
            if gt.minmax:
                col["min"], col["max"] = mn[c, :g][order], mx[c, :g][order]
            out["cols"].append(col)
        stride = 2 + 2 * nc + (2 * nc if gt.minmax else 0)
        null_entry = gt.table[gt.capacity * stride:(gt.capacity + 1) * stride].cpu()
        out["null_group_rows"] = int(null_entry[1].item())
        return out

    def hash_partition(self, keys: torch.Tensor, vals: torch.Tensor, ranks: int, buckets: int = 1,
                       n: Optional[int] = None, key_stride: int = 1, val_stride: int = 1):
        """Map side of the exchange -> (records int64[n, 2], counts int64[ranks*buckets],
        base int64[ranks*buckets]); destination = rank * buckets + bucket."""
        n = keys.shape[0] if n is None else n
        parts = ranks * buckets
        rec = torch.empty((max(n, 1), 2), dtype=torch.int64, device=self.device)
        counts = self.empty(parts, torch.int64)
        base = self.empty(parts, torch.int64)
        _lib.check(self.lib.etl_hash_partition(self._h, _ptr(keys), key_stride, _ptr(vals), val_stride, n, ranks, buckets,
                                               counts.data_ptr(), base.data_ptr(), rec.data_ptr(), self._stream()))
        self.launches += 3 if n else 0
        return rec[:n], counts, base

    def hash_partition_count(self, keys: torch.Tensor, ranks: int, buckets: int = 1, n: Optional[int] = None,
                             key_stride: int = 1) -> torch.Tensor:
        n = keys.shape[0] if n is None else n
        counts = self.empty(ranks * buckets, torch.int64)
        _lib.check(self.lib.etl_hash_partition_count(self._h, _ptr(keys), key_stride, n, ranks, buckets, counts.data_ptr(),
                                                     self._stream()))
        self.launches += 1 if n else 0
        return counts

    def hash_partition_scatter_peers(self, keys: torch.Tensor, vals: torch.Tensor, peer_ptrs: Sequence[int],
                                     peer_offset: torch.Tensor, n: Optional[int] = None, key_stride: int = 1,
                                     val_stride: int = 1, buckets: int = 1) -> None:
        """Partition + exchange in one kernel: records go straight into the peers'
        receive buffers (peer_ptrs[r] = rank r's buffer mapped into this process);
        peer_offset int64[ranks * buckets] = first record of every (rank, bucket) run."""
        n = keys.shape[0] if n is None else n
        ranks = len(peer_ptrs)
        P = C.c_void_p * ranks
        _lib.check(self.lib.etl_hash_partition_scatter_peers(self._h, _ptr(keys), key_stride, _ptr(vals), val_stride, n,
                                                             ranks, buckets, P(*peer_ptrs), peer_offset.data_ptr(),
                                                             self._stream()))
        self.launches += 1 if n else 0

    def groupby_accumulate_records(self, gt: "GroupTable", records: torch.Tensor, lo: int, hi: int,
                                   groups_hint: int = 1 << 30) -> None:
        """Rows [lo, hi) of a {key, value} record buffer into a 1-column table."""
        if hi <= lo:
            return
        flat = records.view(-1)
        part = flat[2 * lo:]
        self.groupby_accumulate(gt, part, None, [part[1:].view(torch.float64)], groups_hint=groups_hint, n=hi - lo,
                                key_stride=2, val_strides=[2])

    # ------------------------------------------- shared-memory merge of the exchange
    # shared-table slots per CTA (4096 = 80 KB: two CTAs per SM) and target groups per piece (load factor ~0.37);
    # the environment overrides are tuning knobs for tools/measure_shuffle.py
    MERGE_SLOTS = int(os.environ.get("ETL_MERGE_SLOTS", "4096"))
    MERGE_GROUPS_PER_PIECE = int(os.environ.get("ETL_MERGE_GROUPS", "1536"))

    @staticmethod
    def merge_plan(records: int, groups_hint: int, groups_per_piece: int = MERGE_GROUPS_PER_PIECE, max_buckets: int = 256):
        """-> (buckets, sub): how finely to cut `records` records holding about
        `groups_hint` distinct keys so one piece's groups fit a shared table.
        max_buckets bounds the first level (256 / ranks when the exchange does it)."""
        g = max(1, min(int(groups_hint) if groups_hint > 0 else records, records))
        fine = min(65536, max(1, -(-g // groups_per_piece)))
        if fine <= max_buckets:
            return fine, 1
        b = min(max_buckets, int(np.ceil(np.sqrt(fine))))
        return b, min(256, -(-fine // b))

    def groupby_partitioned_sum(self, keys: torch.Tensor, vals: torch.Tensor, out_capacity: int, groups_hint: int = 0,
                                n: Optional[int] = None, key_stride: int = 1, val_stride: int = 1,
                                plan: Optional[Tuple[int, int]] = None, slots: Optional[int] = None,
                                mark=None, minmax: bool = False) -> "PartitionedGroups":
        """groupBy(key).agg(count, sum(value)) of (key int64, value f64) rows through
        two partition levels and a shared-memory merge (csrc/merge.cu).  Everything is
        enqueued on the current stream; nothing is read back until groupby_result."""
        n = keys.shape[0] if n is None else n
        buckets, sub = plan or self.merge_plan(n, groups_hint)
        level1, counts, base = self.hash_partition(keys, vals, 1, buckets, n=n, key_stride=key_stride, val_stride=val_stride)
        if mark:
            mark("level1")
        return self.groupby_merge_buckets(level1, counts, base, buckets, sub, n, out_capacity, slots=slots, mark=mark, minmax=minmax)

    def groupby_merge_buckets(self, level1: torch.Tensor, counts: torch.Tensor, base: torch.Tensor, buckets: int, sub: int,
                              n: int, out_capacity: int, slots: Optional[int] = None, mark=None,
                              spill: Optional[torch.Tensor] = None, minmax: bool = False) -> "PartitionedGroups":
        """The part after the first partition level: `level1` int64[>= n, 2] holds n
        records grouped into `buckets` buckets (device counts / base), as written by
        hash_partition(ranks=1) or by a bucketed peer exchange.  sub > 1 cuts every
        bucket into `sub` pieces first.  `spill` (int64[k, 2]) takes records of
        over-full pieces; by default the level-1 buffer when the second level has made
        it redundant, else a new buffer of n records."""
        slots = slots or self.MERGE_SLOTS
        fine, pieces = buckets, level1
        if sub > 1 and n:
            fine = buckets * sub
            fcounts, fbase = self.empty(fine, torch.int64), self.empty(fine, torch.int64)
            pieces = torch.empty((n, 2), dtype=torch.int64, device=self.device)
            _lib.check(self.lib.etl_hash_subpartition(self._h, level1.data_ptr(), buckets, sub, counts.data_ptr(),
                                                      base.data_ptr(), n, fcounts.data_ptr(), fbase.data_ptr(),
                                                      pieces.data_ptr(), self._stream()))
            self.launches += 3
            counts, base = fcounts, fbase
            if spill is None:
                spill = level1[:n]     # dead from here on
        if mark:
            mark("level2")
        if spill is None:
            spill = torch.empty((max(n, 1), 2), dtype=torch.int64, device=self.device)
        out_keys, out_rows = self.empty(out_capacity, torch.int64), self.empty(out_capacity, torch.int64)
        out_cnt, out_sum = self.empty(out_capacity, torch.int64), self.empty(out_capacity, torch.float64)
        counters = self.empty(4, torch.int64)
        out_min = self.empty(out_capacity, torch.float64) if minmax else None
        out_max = self.empty(out_capacity, torch.float64) if minmax else None
        _lib.check(self.lib.etl_groupby_merge_partitioned(
            self._h, pieces.data_ptr() if n else None, fine, counts.data_ptr(), base.data_ptr(), min(slots, 4096) if minmax else slots,
            out_capacity, out_keys.data_ptr(), out_rows.data_ptr(), out_cnt.data_ptr(), out_sum.data_ptr(), _ptr(out_min),
            _ptr(out_max), spill.data_ptr(), spill.shape[0], counters.data_ptr(), self._stream()))
        self.launches += 1 if n else 0
        return PartitionedGroups(out_keys, out_rows, out_cnt, out_sum, counters, spill, out_capacity, fine, min=out_min, max=out_max)

    # ---------------------------------------------- the same, without count passes (slabs)
    SLAB_SIGMAS = float(os.environ.get("ETL_SLAB_SIGMAS", "6"))

    @classmethod
    def slab_capacity(cls, expected: float, rows_per_key: float) -> int:
        """Room for a destination expected to receive `expected` records of keys that
        carry ~rows_per_key rows each (all rows of a key land together, so the
        spread is that of a compound Poisson count): mean + SLAB_SIGMAS sigma + 64, a
        multiple of 8 records."""
        sigma = float(np.sqrt(max(expected, 1.0) * (max(rows_per_key, 1.0) + 1.0)))
        return int(-(-(expected + cls.SLAB_SIGMAS * sigma + 64) // 8) * 8)

    def slab_partition_rows(self, keys: torch.Tensor, vals: torch.Tensor, n: int, ranks: int, buckets: int,
                            dest_ptrs: Sequence[int], slab_origin: int, slab_stride: int, slab_capacity: int,
                            cursor: torch.Tensor, overflow: torch.Tensor, key_stride: int = 1, val_stride: int = 1,
                            max_ctas: int = 0) -> None:
        P = C.c_void_p * ranks
        _lib.check(self.lib.etl_slab_partition_rows(self._h, _ptr(keys), key_stride, _ptr(vals), val_stride, n, ranks, buckets,
                                                    P(*dest_ptrs), slab_origin, slab_stride, slab_capacity,
                                                    cursor.data_ptr(), overflow.data_ptr(), max_ctas, self._stream()))
        self.launches += 1 if n else 0

    def slab_partition_records(self, records: torch.Tensor, segments: int, seg_capacity: int, seg_counts: torch.Tensor,
                               seg_prev: Optional[torch.Tensor], max_seg_records: int, buckets: int, sub: int,
                               out_records: torch.Tensor, slab_capacity: int, cursor: torch.Tensor, overflow: torch.Tensor,
                               max_ctas: int = 0, sources: int = 0, bucket0: int = 0) -> None:
        _lib.check(self.lib.etl_slab_partition_records(self._h, records.data_ptr(), segments, seg_capacity,
                                                       seg_counts.data_ptr(), _ptr(seg_prev), max_seg_records, buckets, sub,
                                                       sources, bucket0, out_records.data_ptr(), slab_capacity,
                                                       cursor.data_ptr(), overflow.data_ptr(), max_ctas, self._stream()))
        self.launches += 1 if max_seg_records else 0

    def peer_publish_counts(self, cursor: torch.Tensor, table_ptrs: Sequence[int], slot: int, per_rank: int = 1) -> None:
        ranks = len(table_ptrs)
        P = C.c_void_p * ranks
        _lib.check(self.lib.etl_peer_publish_counts(self._h, cursor.data_ptr(), ranks, per_rank, P(*table_ptrs), slot,
                                                    self._stream()))
        self.launches += 1

    def peer_copy_slabs(self, send: torch.Tensor, peer_ptrs: Sequence[int], my_rank: int, buckets: int, bucket0: int,
                        nbuckets: int, slab_capacity: int, cursor: torch.Tensor, ctas_per_slab: int = 0, bulk: bool = False) -> None:
        """The filled part of this rank's slabs of a bucket group -> every owner's receive buffer over NVLink:
        16-byte stores from registers, or (bulk) TMA bulk copies through a shared-memory ring."""
        ranks = len(peer_ptrs)
        P = C.c_void_p * ranks
        fn = self.lib.etl_peer_copy_slabs_bulk if bulk else self.lib.etl_peer_copy_slabs
        _lib.check(fn(self._h, send.data_ptr(), P(*peer_ptrs), ranks, my_rank, buckets, bucket0, nbuckets,
                      slab_capacity, cursor.data_ptr(), ctas_per_slab, self._stream()))
        self.launches += 1

    def copy_async(self, dst_ptr: int, src_ptr: int, nbytes: int) -> None:
        """Stream-ordered device-to-device copy on the current stream (copy engine); dst may be peer memory."""
        _lib.check(self.lib.etl_copy_async(self._h, dst_ptr, src_ptr, nbytes, self._stream()))

    def merge_slab_pieces(self, pieces: torch.Tensor, cursor: torch.Tensor, piece_capacity: int, piece0: int, count: int,
                          append: bool, pg: "PartitionedGroups", slots: Optional[int] = None, shared_sms: bool = False) -> None:
        """Shared-memory merge of the piece slabs piece0 .. piece0 + count - 1 into pg's output columns
        (append: after the groups earlier calls left there; shared_sms: the launch runs beside the exchange's
        copy kernel and keeps to 512-thread CTAs so that the copy CTAs fit on the same SMs)."""
        slots = slots or self.MERGE_SLOTS
        minmax = pg.min is not None
        _lib.check(self.lib.etl_groupby_merge_slabs(
            self._h, pieces.data_ptr(), piece0, count, int(bool(append)) | (2 if shared_sms else 0), cursor.data_ptr(), piece_capacity,
            min(slots, 4096) if minmax else slots, pg.capacity, pg.keys.data_ptr(), pg.rows.data_ptr(), pg.count.data_ptr(),
            pg.sum.data_ptr(), _ptr(pg.min), _ptr(pg.max), pg.spill.data_ptr(), pg.spill.shape[0], pg.counters.data_ptr(),
            self._stream()))
        self.launches += 1

    def new_partitioned_groups(self, out_capacity: int, spill_records: int, pieces: int, minmax: bool = False,
                               counters: Optional[torch.Tensor] = None) -> "PartitionedGroups":
        """Empty output of the slab merge: dense columns for out_capacity groups, a spill buffer, counters
        int64[5] ([4] = slab overflow) zeroed."""
        e = self.empty
        counters = torch.zeros(5, dtype=torch.int64, device=self.device) if counters is None else counters
        spill = torch.empty((max(spill_records, 1), 2), dtype=torch.int64, device=self.device)
        return PartitionedGroups(e(out_capacity, torch.int64), e(out_capacity, torch.int64), e(out_capacity, torch.int64),
                                 e(out_capacity, torch.float64), counters, spill, out_capacity, pieces,
                                 min=e(out_capacity, torch.float64) if minmax else None,
                                 max=e(out_capacity, torch.float64) if minmax else None)

    def slab_plan(self, n: int, groups_hint: int, max_buckets: int = 256):
        """-> (buckets, sub, rows_per_key, cap1, cap2) for n records holding ~groups_hint distinct keys."""
        buckets, sub = self.merge_plan(n, groups_hint, max_buckets=max_buckets)
        g = max(1, min(int(groups_hint) if groups_hint > 0 else n, max(n, 1)))
        rpk = max(1.0, n / g)
        cap1 = self.slab_capacity(n / buckets, rpk)
        cap2 = self.slab_capacity(n / (buckets * sub), rpk) if sub > 1 else cap1
        return buckets, sub, rpk, cap1, cap2

    def groupby_slab_sum(self, keys: torch.Tensor, vals: torch.Tensor, out_capacity: int, groups_hint: int,
                         n: Optional[int] = None, key_stride: int = 1, val_stride: int = 1, slots: Optional[int] = None,
                         mark=None, minmax: bool = False) -> "PartitionedGroups":
        """groupby_partitioned_sum without the two count passes: both partition levels
        scatter into fixed-capacity slabs (csrc/slabs.cu).  groups_hint (> 0) sizes the
        slabs; if a slab overflows (skewed keys, a hint far off) groupby_result repeats
        the work with groupby_partitioned_sum, so the result is always complete."""
        n = keys.shape[0] if n is None else n
        if n == 0 or groups_hint <= 0:
            return self.groupby_partitioned_sum(keys, vals, out_capacity, groups_hint, n, key_stride, val_stride, slots=slots,
                                                mark=mark, minmax=minmax)
        buckets, sub, rpk, cap1, cap2 = self.slab_plan(n, groups_hint)
        counters = torch.zeros(5, dtype=torch.int64, device=self.device)
        cur1 = torch.zeros(buckets, dtype=torch.int64, device=self.device)
        level1 = torch.empty((buckets * cap1, 2), dtype=torch.int64, device=self.device)
        self.slab_partition_rows(keys, vals, n, 1, buckets, [level1.data_ptr()], 0, cap1, cap1, cur1, counters[4:], key_stride, val_stride)
        if mark:
            mark("level1")
        pg = self.groupby_merge_slab_buckets(level1, cur1, cap1, buckets, sub, cap2, out_capacity, counters, slots=slots, mark=mark,
                                             minmax=minmax)
        pg.rerun = lambda: self.groupby_partitioned_sum(keys, vals, out_capacity, groups_hint, n, key_stride, val_stride,
                                                        slots=slots, minmax=minmax)
        return pg

    def groupby_merge_slab_buckets(self, level1: torch.Tensor, cur1: torch.Tensor, cap1: int, buckets: int, sub: int,
                                   cap2: int, out_capacity: int, counters: torch.Tensor, slots: Optional[int] = None,
                                   mark=None, minmax: bool = False) -> "PartitionedGroups":
        """After the first level (local, or done on what the exchange delivered): the second
        level into piece slabs and the shared-memory merge.  counters = int64[5], [4] = overflow."""
        slots = slots or self.MERGE_SLOTS
        fine, pieces, cur2 = buckets, level1, cur1
        spill = None
        if sub > 1:
            fine = buckets * sub
            cur2 = torch.zeros(fine, dtype=torch.int64, device=self.device)
            pieces = torch.empty((fine * cap2, 2), dtype=torch.int64, device=self.device)
            self.slab_partition_records(level1, buckets, cap1, cur1, None, cap1, buckets, sub, pieces, cap2, cur2, counters[4:])
            spill = level1             # dead from here on
        if mark:
            mark("level2")
        if spill is None:
            spill = torch.empty((max(buckets * cap1 // 8, 1 << 16), 2), dtype=torch.int64, device=self.device)
        out_keys, out_rows = self.empty(out_capacity, torch.int64), self.empty(out_capacity, torch.int64)
        out_cnt, out_sum = self.empty(out_capacity, torch.int64), self.empty(out_capacity, torch.float64)
        out_min = self.empty(out_capacity, torch.float64) if minmax else None
        out_max = self.empty(out_capacity, torch.float64) if minmax else None
        _lib.check(self.lib.etl_groupby_merge_slabs(
            self._h, pieces.data_ptr(), 0, fine, 0, cur2.data_ptr(), cap2, min(slots, 4096) if minmax else slots, out_capacity,
            out_keys.data_ptr(), out_rows.data_ptr(), out_cnt.data_ptr(), out_sum.data_ptr(), _ptr(out_min), _ptr(out_max),
            spill.data_ptr(), spill.shape[0], counters.data_ptr(), self._stream()))
        self.launches += 1
        return PartitionedGroups(out_keys, out_rows, out_cnt, out_sum, counters, spill, out_capacity, fine, min=out_min, max=out_max)

    def _partitioned_result(self, pg: "PartitionedGroups", sort: bool = False) -> dict:
        ctr = pg.counters.cpu().tolist()
        groups, spilled, _, reserved = ctr[:4]
        overflow = ctr[4] if len(ctr) > 4 else 0
        if pg.any_overflow is not None:
            overflow = pg.any_overflow(overflow)       # a collective: every rank of the exchange calls it
        if overflow:
            if pg.rerun is None:
                raise _lib.EtlError(-5, f"groupBy: {overflow} record(s) found their slab full and no counted rerun was given")
            return self._partitioned_result(pg.rerun(), sort)
        if reserved:
            raise _lib.EtlError(-5, f"groupBy: 0 row(s) dropped (table full), {reserved} row(s) with the reserved key")
        if groups > pg.capacity:
            raise _lib.EtlError(-5, f"groupBy: {groups} groups exceed the output capacity {pg.capacity}")
        if spilled > pg.spill.shape[0]:
            raise _lib.EtlError(-5, f"groupBy: {spilled} spilled records exceed the spill buffer {pg.spill.shape[0]}")
        keys, rows, cnt, sm = pg.keys[:groups], pg.rows[:groups], pg.count[:groups], pg.sum[:groups]
        minmax = pg.min is not None
        mn, mx = (pg.min[:groups], pg.max[:groups]) if minmax else (None, None)
        if spilled:   # pieces with more groups than the shared table holds: the global-table path takes them
            cap = 1 << max(4, int(np.ceil(np.log2(2 * spilled))))
            gt = self.groupby_new(cap, 1, minmax=minmax)
            self.groupby_accumulate_records(gt, pg.spill, 0, spilled, groups_hint=spilled)
            extra = self.groupby_result(gt, max_groups=spilled, sort=False)
            keys, rows = torch.cat([keys, extra["keys"]]), torch.cat([rows, extra["rows"]])
            cnt, sm = torch.cat([cnt, extra["cols"][0]["count"]]), torch.cat([sm, extra["cols"][0]["sum"]])
            if minmax:
                mn, mx = torch.cat([mn, extra["cols"][0]["min"]]), torch.cat([mx, extra["cols"][0]["max"]])
        order = torch.argsort(keys) if sort else slice(None)
        col = {"count": cnt[order], "sum": sm[order]}
        if minmax:
            col["min"], col["max"] = mn[order], mx[order]
        col["mean"] = col["sum"] / col["count"].to(torch.float64)
        return {"keys": keys[order], "rows": rows[order], "cols": [col], "null_group_rows": 0, "spilled_records": spilled}

    # ------------------------------------------------------- Parquet (extension)
    _PQ_STREAM = np.dtype([("src", "<i8"), ("bytes", "<i4"), ("bw", "<i4"), ("out", "<i8"), ("maxv", "<i4"), ("exact", "<i4"),
                           ("run", "<i8"), ("cap", "<i8")])
    _PQ_PAGE = np.dtype([("row", "<i8"), ("n", "<i4"), ("kind", "<i4"), ("lvl", "<i8"), ("val", "<i8"), ("dict", "<i4"),
                         ("pad", "<i4")])

    _PQ_INFLATE = np.dtype([("src", "<i8"), ("src_bytes", "<i8"), ("dst", "<i8"), ("dst_bytes", "<i8")])
    _PQ_STRPAGE = np.dtype([("val", "<i8"), ("bytes", "<i4"), ("maxv", "<i4"), ("out", "<i8")])

    def parquet_inflate(self, data: torch.Tensor, jobs: Sequence[Tuple[int, int, int, int]]) -> torch.Tensor:
        """Snappy pages (src, src_bytes, dst, dst_bytes: byte offsets in `data`) -> inflated
        in place behind the file bytes.  Returns the int32[4] status words every later
        Parquet call adds to (read back once, by parquet_decode_columns)."""
        status = torch.zeros(4, dtype=torch.int32, device=self.device)
        if jobs:
            tab = np.zeros(len(jobs), dtype=self._PQ_INFLATE)
            arr = np.asarray(jobs, dtype=np.int64).reshape(-1, 4)
            tab["src"], tab["src_bytes"], tab["dst"], tab["dst_bytes"] = arr[:, 0], arr[:, 1], arr[:, 2], arr[:, 3]
            if int((arr[:, 2] + arr[:, 3]).max()) > data.shape[0]:
                raise _lib.EtlError(-1, "parquet: the buffer has no room for the inflated pages")
            d_tab = torch.from_numpy(tab.view(np.uint8)).to(self.device)
            _lib.check(self.lib.etl_parquet_inflate_pages(self._h, data.data_ptr(), d_tab.data_ptr(), len(jobs),
                                                          status.data_ptr(), self._stream()))
            self.launches += 1
        return status

    def parquet_decode_columns(self, data: torch.Tensor, rows: int, jobs: Sequence[dict], status: Optional[torch.Tensor] = None,
                               dict_capacity: int = 1 << 16):
        """Decode several columns of one file (device buffer `data`, compressed pages
        already inflated by parquet_inflate) from the page tables of parquet.py.
        jobs[i] = {"pages": int64[p, 8] (row_start, num_values, def_offset, def_bytes,
        val_offset, val_bytes, dict_id, bit_width; bit_width -1 = read it on the
        device), "kind_plain", "string", "dictionaries": per row group a label list
        (strings) or a float64 array (numerics)}.  All hybrid streams of all columns
        (definition levels and dictionary indices) expand in ONE launch into one int32
        buffer; PLAIN string pages are entered into a device dictionary; then one
        gather launch per column.  -> (columns, label lists: sorted labels of the
        string columns, [] for numeric ones)."""
        if not jobs:
            return [], []
        from .table import _dictionary_labels   # (shared with the CSV decoder's dictionary read-back)

        status = torch.zeros(4, dtype=torch.int32, device=self.device) if status is None else status
        streams, str_pages, out_total, run_total = [], [], 0, 0
        plans = []
        for job in jobs:
            P = np.asarray(job["pages"], dtype=np.int64).reshape(-1, 8)
            n_pages = P.shape[0]
            row, nv, def_off, def_bytes, val_off, val_bytes, dict_id, bw = (P[:, i] for i in range(8))
            has_def, is_dict = def_off >= 0, dict_id >= 0
            is_plain_str = (~is_dict) & bool(job["string"])
            if (bw[is_dict] > 32).any():
                raise _lib.EtlError(-4, f"parquet: dictionary index width {int(bw[is_dict].max())}")
            # output offsets: levels of the pages that have them, then the indices of the dictionary pages and the
            # dictionary slots of the PLAIN string pages
            span = nv + 8
            lvl_out = np.full(n_pages, -1, dtype=np.int64)
            lvl_out[has_def] = out_total + np.cumsum(span[has_def]) - span[has_def]
            out_total += int(span[has_def].sum())
            idx_out = np.zeros(n_pages, dtype=np.int64)
            indexed = is_dict | is_plain_str
            idx_out[indexed] = out_total + np.cumsum(span[indexed]) - span[indexed]
            out_total += int(span[indexed].sum())
            for mask, src, nbytes, width, out, exact in ((has_def, def_off, def_bytes, np.ones_like(bw), lvl_out, 1),
                                                         (is_dict, val_off + 1, val_bytes - 1, bw, idx_out, 0)):
                k = int(mask.sum())
                if not k:
                    continue
                st = np.zeros(k, dtype=self._PQ_STREAM)
                cap = nbytes[mask] // 2 + 2
                st["src"], st["bytes"], st["bw"], st["out"] = src[mask], nbytes[mask], width[mask], out[mask]
                st["maxv"], st["exact"], st["cap"] = nv[mask], exact, cap
                st["run"] = run_total + np.cumsum(cap) - cap
                run_total += int(cap.sum())
                streams.append(st)
            sp = None
            if is_plain_str.any():
                sp = np.zeros(int(is_plain_str.sum()), dtype=self._PQ_STRPAGE)
                sp["val"], sp["bytes"], sp["maxv"], sp["out"] = (val_off[is_plain_str], val_bytes[is_plain_str], nv[is_plain_str],
                                                                 idx_out[is_plain_str])
            plans.append((P, lvl_out, idx_out, is_dict, is_plain_str, sp))
        decoded = self.empty(max(out_total, 1), torch.int32)
        nstreams = sum(x.shape[0] for x in streams)
        if nstreams:
            d_streams = torch.from_numpy(np.concatenate(streams).view(np.uint8)).to(self.device)
            _lib.check(self.lib.etl_parquet_expand_streams(self._h, data.data_ptr(), d_streams.data_ptr(), nstreams, None,
                                                           decoded.data_ptr(), status.data_ptr(), self._stream()))
            self.launches += 1
        # PLAIN string pages -> device dictionaries (one per column), read back for their labels
        device_dicts = []
        for job, (P, lvl_out, idx_out, is_dict, is_plain_str, sp) in zip(jobs, plans):
            if sp is None:
                device_dicts.append(None)
                continue
            hash_tab = torch.zeros(dict_capacity, dtype=torch.int64, device=self.device)
            rep_tab = torch.zeros(dict_capacity, dtype=torch.int64, device=self.device)
            d_sp = torch.from_numpy(sp.view(np.uint8)).to(self.device)
            _lib.check(self.lib.etl_parquet_plain_strings(self._h, data.data_ptr(), d_sp.data_ptr(), sp.shape[0],
                                                          hash_tab.data_ptr(), rep_tab.data_ptr(), dict_capacity,
                                                          decoded.data_ptr(), status.data_ptr(), self._stream()))
            self.launches += 1
            device_dicts.append((hash_tab, rep_tab))
        cols, label_lists, keep = [], [], []
        for job, (P, lvl_out, idx_out, is_dict, is_plain_str, sp), dd in zip(jobs, plans, device_dicts):
            n_pages = P.shape[0]
            row, nv, val_off, dict_id = P[:, 0], P[:, 1], P[:, 4], P[:, 6]
            # lookup tables of this column: one per row-group dictionary (+ one for the device dictionary's slots)
            tables, labels = [], []
            if job["string"]:
                used, dev_labels = _dictionary_labels(data, *dd) if dd is not None else (np.zeros(0, np.int64), [])
                labels = sorted({s for d in job["dictionaries"] for s in d} | set(dev_labels))
                pos = {s: i for i, s in enumerate(labels)}
                tables = [torch.tensor([pos[s] for s in d] or [0], dtype=torch.int32, device=self.device) for d in job["dictionaries"]]
                if dd is not None:
                    slot_codes = np.full(dict_capacity, -1, dtype=np.int32)
                    slot_codes[used] = [pos[s] for s in dev_labels]
                    tables.append(torch.from_numpy(slot_codes).to(self.device))
            else:
                tables = [torch.from_numpy(np.ascontiguousarray(d)).to(self.device) for d in job["dictionaries"]]
            keep += tables
            pg = np.zeros(n_pages, dtype=self._PQ_PAGE)
            pg["row"], pg["n"], pg["lvl"] = row, nv, lvl_out
            indexed = is_dict | is_plain_str
            pg["kind"] = np.where(indexed, 4, job["kind_plain"])
            pg["val"] = np.where(indexed, idx_out, val_off)
            pg["dict"] = np.where(is_dict, dict_id, np.where(is_plain_str, len(tables) - 1, -1))
            out = self.empty(rows, torch.int32 if job["string"] else torch.float64)
            valid = torch.zeros(words_for(rows), dtype=torch.int32, device=self.device)
            if n_pages:
                d_pg = torch.from_numpy(pg.view(np.uint8)).to(self.device)
                d_ptrs = torch.tensor([t.data_ptr() for t in tables] or [0], dtype=torch.int64, device=self.device)
                d_sizes = torch.tensor([t.shape[0] for t in tables] or [0], dtype=torch.int32, device=self.device)
                keep += [d_pg, d_ptrs, d_sizes]
                _lib.check(self.lib.etl_parquet_gather_pages(
                    self._h, data.data_ptr(), d_pg.data_ptr(), n_pages, decoded.data_ptr(), decoded.data_ptr(), d_ptrs.data_ptr(),
                    d_sizes.data_ptr(), int(job["string"]), out.data_ptr(), valid.data_ptr(), status.data_ptr(), self._stream()))
                self.launches += 1
            cols.append(Column(out, valid))
            label_lists.append(labels)
        st = status.cpu().tolist()
        if any(st):
            raise _lib.EtlError(-4, f"parquet decode: string dictionary full {st[0]}, short or malformed streams {st[1]}, "
                                    f"dictionary indices out of range {st[2]}, corrupt Snappy pages {st[3]}")
        del keep
        return cols, label_lists

    # ------------------------------------------------- KMeans passes (8f rank 2)
    def kmeans_prepare(self, centers: torch.Tensor, num_labels: int, repeats: int):
        k = centers.shape[0]
        T = torch.empty((k, num_labels + 1), dtype=torch.float64, device=self.device)
        cn = torch.empty((k, 3), dtype=torch.float64, device=self.device)
        n2 = self.empty(k, torch.float64)
        _lib.check(self.lib.etl_kmeans_prepare(self._h, centers.data_ptr(), k, num_labels, repeats, T.data_ptr(),
                                               cn.data_ptr(), n2.data_ptr(), self._stream()))
        self.launches += 1
        return T, cn, n2

    def kmeans_assign(self, index: torch.Tensor, nums: Sequence[torch.Tensor], centers: torch.Tensor, num_labels: int,
                      repeats: int, prev_cost: Optional[torch.Tensor] = None, cluster: Optional[torch.Tensor] = None,
                      cluster_offset: int = 0, want_cluster: bool = True):
        """Closest centre per row -> (cluster int32[n] or None, cost f64[n])."""
        n, k = index.shape[0], centers.shape[0]
        T, cn, n2 = self.kmeans_prepare(centers.contiguous(), num_labels, repeats)
        if want_cluster and cluster is None:
            cluster = torch.zeros(n, dtype=torch.int32, device=self.device)
        cost = self.empty(n, torch.float64)
        _lib.check(self.lib.etl_kmeans_assign(self._h, _ptr(index), C.byref(VP3(*[t.data_ptr() for t in nums])), n, k,
                                              num_labels, repeats, T.data_ptr(), cn.data_ptr(), n2.data_ptr(),
                                              _ptr(prev_cost), cluster_offset, _ptr(cluster) if want_cluster else None,
                                              cost.data_ptr(), self._stream()))
        self.launches += 1 if n else 0
        return cluster, cost

    def kmeans_accumulate(self, index: torch.Tensor, nums: Sequence[torch.Tensor], cluster: torch.Tensor, k: int,
                          num_labels: int):
        n = index.shape[0]
        count = self.empty(k, torch.int64)
        hist = torch.empty((k, num_labels + 1), dtype=torch.int64, device=self.device)
        numsum = torch.empty((k, 3), dtype=torch.float64, device=self.device)
        _lib.check(self.lib.etl_kmeans_accumulate(self._h, _ptr(index), C.byref(VP3(*[t.data_ptr() for t in nums])),
                                                  _ptr(cluster), n, k, num_labels, count.data_ptr(), hist.data_ptr(),
                                                  numsum.data_ptr(), self._stream()))
        self.launches += 1 if n else 0
        return count, hist, numsum

    def kmeans_update(self, count: torch.Tensor, hist: torch.Tensor, numsum: torch.Tensor, centers: torch.Tensor,
                      num_labels: int, repeats: int) -> None:
        """centers[c] <- mean of cluster c's rows, in place (empty clusters keep theirs)."""
        _lib.check(self.lib.etl_kmeans_update(self._h, count.data_ptr(), hist.data_ptr(), numsum.data_ptr(), centers.shape[0],
                                              num_labels, repeats, centers.data_ptr(), self._stream()))
        self.launches += 1

    def kmeans_row_norms(self, index: torch.Tensor, nums: Sequence[torch.Tensor], num_labels: int, repeats: int) -> torch.Tensor:
        """|x_i|^2 of every structured feature row -> f64[n]."""
        n = index.shape[0]
        out = self.empty(n, torch.float64)
        _lib.check(self.lib.etl_kmeans_row_norms(self._h, _ptr(index), C.byref(VP3(*[t.data_ptr() for t in nums])), n,
                                                 num_labels, repeats, out.data_ptr(), self._stream()))
        self.launches += 1 if n else 0
        return out

    def kmeans_silhouette(self, index: torch.Tensor, nums: Sequence[torch.Tensor], cluster: torch.Tensor, label_sum: torch.Tensor,
                          num_sum: torch.Tensor, psi: torch.Tensor, count: torch.Tensor, num_labels: int,
                          repeats: int) -> torch.Tensor:
        """Silhouette coefficient (squared Euclidean) of every row from the per-cluster tables -> f64[n]."""
        n, k = index.shape[0], count.shape[0]
        out = self.empty(n, torch.float64)
        _lib.check(self.lib.etl_kmeans_silhouette(self._h, _ptr(index), C.byref(VP3(*[t.data_ptr() for t in nums])),
                                                  _ptr(cluster), n, k, num_labels, repeats, label_sum.data_ptr(),
                                                  num_sum.data_ptr(), psi.data_ptr(), count.data_ptr(), out.data_ptr(),
                                                  self._stream()))
        self.launches += 1 if n else 0
        return out

    # --------------------------------------------- trainer hand-off (8f rank 1)
    def mlp_param_count(self, classes: int) -> int:
        return int(self.lib.etl_mlp_param_count(classes))

    def mlp_train_steps(self, X: torch.Tensor, y: torch.Tensor, order: torch.Tensor, batch: int, classes: int,
                        params: torch.Tensor, adam_m: torch.Tensor, adam_v: torch.Tensor, step0: int = 0,
                        lr: float = 1e-3, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-7,
                        step_rows: Optional[torch.Tensor] = None):
        """`order` int32[steps * batch] row indices -> (loss f32[steps], acc f32[steps]);
        params / moments are updated in place.  step_rows int32[steps] (optional): how
        many of a step's `batch` slots count - the short batch at the end of a pass."""
        steps = order.numel() // batch
        loss = self.empty(max(steps, 1), torch.float32)
        acc = self.empty(max(steps, 1), torch.float32)
        _lib.check(self.lib.etl_mlp_train_steps_ragged(self._h, X.data_ptr(), y.data_ptr(), X.shape[0], order.data_ptr(),
                                                       _ptr(step_rows), batch, steps, classes, params.data_ptr(),
                                                       adam_m.data_ptr(), adam_v.data_ptr(), step0, lr, beta1, beta2, eps,
                                                       loss.data_ptr(), acc.data_ptr(), self._stream()))
        self.launches += 1 if steps else 0
        return loss[:steps], acc[:steps]

    def mlp_evaluate(self, X: torch.Tensor, y: Optional[torch.Tensor], params: torch.Tensor, classes: int):
        n = X.shape[0]
        loss = self.empty(n, torch.float32) if y is not None else None
        pred = self.empty(n, torch.int32)
        _lib.check(self.lib.etl_mlp_evaluate(self._h, X.data_ptr(), _ptr(y), n, params.data_ptr(), classes, _ptr(loss),
                                             pred.data_ptr(), self._stream()))
        self.launches += 1 if n else 0
        return loss, pred

    # ------------------------------------------------------------ synth
    def synth_column(self, seed: int, stream_id: int, kind: int, n: int, row0: int = 0, modulus: int = 1,
                     null_rate: float = 0.0, nan_rate: float = 0.0, with_valid: bool = True) -> Column:
        dtype = {0: torch.float64, 1: torch.float64, 2: torch.int32, 3: torch.int64}[kind]
        data = self.empty(n, dtype)
        valid = torch.zeros(words_for(n), dtype=torch.int32, device=self.device) if with_valid else None
        _lib.check(self.lib.etl_synth_column(self._h, seed, stream_id, kind, modulus, null_rate, nan_rate, row0, n,
                                             data.data_ptr(), _ptr(valid), self._stream()))
        self.launches += 1 if n else 0
        return Column(data, valid)


def project_filter_impute_noalloc(key: Column, vals: Sequence[Column], fill: Sequence[float], out_code: torch.Tensor,
                                  out_vals: Sequence[torch.Tensor], out_rows: torch.Tensor) -> None:
    """The projection for callers that must not allocate device memory (for
    instance because a pool allocator owns the whole GPU, or because the
    device-wide sync inside cudaMalloc is unaffordable): no Engine / handle is
    involved and every buffer, outputs included, is passed in.  Runs on torch's
    current device and stream.  Limits and sharing rules are those of
    etl_project_filter_impute_noalloc in include/etl_b200.h: at most 2^31 rows,
    first call per device outside stream capture, calls that overlap on the GPU
    corrupt each other because they share one fixed scratch array."""
    lib = _lib.load()
    n = len(key)
    val = VP3(*[_ptr(c.data) for c in vals])
    vv = VP3(*[_ptr(c.valid) for c in vals])
    dt = I32x3(*[c.dtype_code for c in vals])
    fl = F64x3(*[float(f) for f in fill])
    ov = VP3(*[t.data_ptr() for t in out_vals])
    stream = torch.cuda.current_stream(key.data.device).cuda_stream
    _lib.check(lib.etl_project_filter_impute_noalloc(_ptr(key.data), _ptr(key.valid), C.byref(val), C.byref(vv),
                                                     C.byref(dt), C.byref(fl), n, out_code.data_ptr(), C.byref(ov),
                                                     out_rows.data_ptr(), stream))
