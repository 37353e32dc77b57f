"""CPU oracle for the raw-spark ETL hot path.  TEST INFRASTRUCTURE ONLY.

This file restates, in NumPy / stdlib Python, what the reference computes on the
path SURVEY.md §8 scopes.  It is the checker for the CUDA path; it is never the
product.  Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline /
`--impl reference` legs may import it.  The product package
(pyspark-tf-gke_b200/) must not, and fails loudly when its CUDA library is
missing.

Pinning status:
  * load_csv(): PINNED - checked in tests/test_oracle_golden.py against outputs
    of the reference's own load_csv (workloads/raw-tf/train_tf_ps.py:75-149)
    captured by tests/golden/make_golden.py on both health.csv copies and on 19
    edge-case CSV texts.
  * Spark-side functions (k_means_etl and the columnar primitives): PARITY
    UNPINNED - the arithmetic lives in Apache Spark (pyspark, unpinned:
    infra/local/local_spark/requirements.txt:9), which is absent from
    /root/reference and cannot run in this image (no JVM).  They restate the
    published Spark >= 3.0 semantics at the reference's call sites
    (workloads/raw-spark/k_means.py:22-77) and are checked against the exact-
    rational known answers in tests/golden/spark_known_answers.json.

All `file:line` citations are relative to /root/reference/.
"""
from __future__ import annotations

import csv
import io
import math
from collections import Counter
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

# ----------------------------------------------------------------------------
# validity bitmaps: Arrow layout, bit i of the column lives in bit (i % 32) of
# little-endian uint32 word i // 32; 1 = valid (non-null).
# ----------------------------------------------------------------------------


def pack_validity(valid: np.ndarray) -> np.ndarray:
    valid = np.asarray(valid, dtype=bool)
    n = valid.shape[0]
    words = (n + 31) // 32
    padded = np.zeros(words * 32, dtype=np.uint8)
    padded[:n] = valid
    return np.packbits(padded, bitorder="little").view(np.uint32).copy()


def unpack_validity(bitmap: Optional[np.ndarray], n: int) -> np.ndarray:
    if bitmap is None:
        return np.ones(n, dtype=bool)
    bits = np.unpackbits(np.ascontiguousarray(bitmap).view(np.uint8), bitorder="little")
    return bits[:n].astype(bool)


# ----------------------------------------------------------------------------
# a2/a3  null count + isNotNull filter        k_means.py:23, :27-28
# a4     avg over ~isnan & isNotNull          k_means.py:45-48
# a5     when(isNull | isnan, mean).otherwise k_means.py:49-51
# ----------------------------------------------------------------------------


def null_count(valid: np.ndarray) -> int:
    """filter(col.isNull()).count()  (k_means.py:23)."""
    return int((~np.asarray(valid, dtype=bool)).sum())


def sum_count(values: np.ndarray, valid: np.ndarray) -> Tuple[float, int]:
    """Partial state of Spark's Average over rows passing ~isnan & isNotNull
    (k_means.py:45-48): (fp64 sum, int64 count).  math.fsum gives the correctly
    rounded sum, the reference point the 1e-6 tolerance is measured from."""
    v = np.asarray(values).astype(np.float64)
    m = np.asarray(valid, dtype=bool) & ~np.isnan(v)
    return math.fsum(v[m].tolist()) if m.any() else 0.0, int(m.sum())


def mean_of(values: np.ndarray, valid: np.ndarray) -> Optional[float]:
    """agg({c: "avg"}) -> sum / count as fp64; None (SQL NULL) if no rows."""
    s, c = sum_count(values, valid)
    return s / c if c else None


def project_filter_impute(
    code: np.ndarray,
    code_valid: np.ndarray,
    values: Sequence[np.ndarray],
    values_valid: Sequence[np.ndarray],
    fill: Sequence[float],
) -> Tuple[np.ndarray, List[np.ndarray]]:
    """The map-side projection of k_means.py:27 + :49-51 on columnar input.

    keep   = key isNotNull                                        (:27)
    out[c] = fill[c] if (null or NaN) else double(value)          (:49-51; the
             literal double coerces int/float columns to double - the one cast
             on the path)
    Kept rows stay in input order (a Spark filter is order preserving inside a
    partition)."""
    keep = np.asarray(code_valid, dtype=bool)
    out_code = np.asarray(code)[keep].astype(np.int32)
    outs = []
    for v, vv, f in zip(values, values_valid, fill):
        d = np.asarray(v).astype(np.float64)
        bad = ~np.asarray(vv, dtype=bool) | np.isnan(d)
        outs.append(np.where(bad, np.float64(f), d)[keep])
    return out_code, outs


# ----------------------------------------------------------------------------
# a6  StringIndexer(handleInvalid="keep").fit    k_means.py:34,73
# a7  StringIndexerModel.transform                k_means.py:74
# ----------------------------------------------------------------------------


def string_indexer_fit(labels: Sequence[Optional[str]]) -> Tuple[List[str], List[int]]:
    """group-by-count on the string key, ordered frequencyDesc with ties broken
    by label ascending (Spark >= 3.0 StringIndexer doc; nulls are not labels)."""
    counts = Counter(s for s in labels if s is not None)
    order = sorted(counts, key=lambda s: (-counts[s], s))
    return order, [counts[s] for s in order]


def string_indexer_order_from_counts(labels: Sequence[str], counts: Sequence[int]) -> List[int]:
    """Same ordering rule applied to an already-aggregated (label, count) table:
    returns the permutation p such that labels[p[0]] is index 0.  Ties compare
    the labels as Java strings (UTF-16 code units); for BMP-only text that is
    code-point order, which is what Python str comparison gives."""
    return sorted(range(len(labels)), key=lambda i: (-int(counts[i]), labels[i]))


def string_indexer_transform(labels: Sequence[Optional[str]], vocab: Sequence[str]) -> np.ndarray:
    """label -> index as double; unseen or null -> len(vocab) (handleInvalid=keep)."""
    pos = {s: i for i, s in enumerate(vocab)}
    return np.array([pos.get(s, len(vocab)) if s is not None else len(vocab) for s in labels], dtype=np.float64)


# ----------------------------------------------------------------------------
# a8  OneHotEncoder (dropLast=True, handleInvalid="error")  k_means.py:38
# a9  VectorAssembler([vec]*repeats + numeric)              k_means.py:64-68
# ----------------------------------------------------------------------------


def assemble_features_dense(index: np.ndarray, num_labels: int, numerics: Sequence[np.ndarray], repeats: int,
                            dtype=np.float64) -> np.ndarray:
    """features[i] = [onehot(index_i)] * repeats ++ numerics[i].

    The indexer (keep) declares num_labels + 1 categories; dropLast removes the
    last, so the one-hot block has size num_labels and index == num_labels maps
    to all zeros."""
    n = len(index)
    L = int(num_labels)
    D = repeats * L + len(numerics)
    out = np.zeros((n, D), dtype=dtype)
    idx = np.asarray(index).astype(np.int64)
    rows = np.nonzero(idx < L)[0]
    for r in range(repeats):
        out[rows, r * L + idx[rows]] = 1
    for j, col in enumerate(numerics):
        out[:, repeats * L + j] = np.asarray(col).astype(dtype)
    return out


def assemble_features_sparse(index: np.ndarray, num_labels: int, numerics: Sequence[np.ndarray], repeats: int):
    """Spark's SparseVector view of the same rows: per row, (indices, values)
    of the non-zero entries in ascending index order (VectorAssembler drops
    explicit zeros; `.compressed` keeps the sparse form when it is smaller)."""
    L = int(num_labels)
    rows = []
    for i, ix in enumerate(np.asarray(index).astype(np.int64)):
        ind, val = [], []
        if ix < L:
            for r in range(repeats):
                ind.append(r * L + int(ix))
                val.append(1.0)
        for j, col in enumerate(numerics):
            v = float(col[i])
            if v != 0.0:
                ind.append(repeats * L + j)
                val.append(v)
        rows.append((ind, val))
    return rows


# ----------------------------------------------------------------------------
# a1  spark.read.csv(header=True, inferSchema=True)   spark_workload_to_cloud_k8s.py:48
# ----------------------------------------------------------------------------


def read_csv_columns(text: str) -> Tuple[List[str], Dict[str, List[Optional[str]]]]:
    """CSV text -> {column: [str | None]} with Spark's default null rule for
    this data: an empty field is null (nullValue = "").  A QUOTED empty cell
    (`a,"",b`) is null as well under the reference's default options
    (spark_workload_to_cloud_k8s.py:48 sets none): Spark 2.4-4.0 hand univocity
    nullValue = "" for an unquoted empty field and emptyValue = "" (the read-side
    default) for a quoted one, so both arrive as "", and UnivocityParser
    .nullSafeDatum turns any datum equal to nullValue into null - for string
    columns too.  Only a non-default nullValue / emptyValue tells the two apart,
    which is outside this path.  The parse
    itself is RFC-4180 quoting (quote ", doubled-quote escape), which is what
    both Spark's univocity settings and Python's csv agree on for input that has
    no backslash escapes - the documented input domain."""
    reader = csv.reader(io.StringIO(text, newline=""))
    header = next(reader)
    cols: Dict[str, List[Optional[str]]] = {h: [] for h in header}
    for row in reader:
        if not row:  # Spark and csv.reader both drop blank lines
            continue
        for j, h in enumerate(header):
            cell = row[j] if j < len(row) else None
            cols[h].append(None if cell is None or cell == "" else cell)
    return header, cols


def infer_numeric_type(cells: Sequence[Optional[str]]) -> str:
    """Spark's inferSchema ladder restricted to what the hot columns can hit:
    int (all fit int32) -> long -> double -> string."""
    kind = "int"
    for c in cells:
        if c is None:
            continue
        s = c.strip()
        try:
            iv = int(s)
            if not (-2**31 <= iv < 2**31) and kind == "int":
                kind = "long"
            if not (-2**63 <= iv < 2**63):
                kind = "double"
            continue
        except ValueError:
            pass
        try:
            float(s)
            kind = "double"
        except ValueError:
            return "string"
    return kind


def to_numeric_column(cells: Sequence[Optional[str]]) -> Tuple[np.ndarray, np.ndarray]:
    """text cells -> (fp64 values, validity).  Decimal -> binary64 is Python
    float(), i.e. correctly rounded, as Java's Double.parseDouble is."""
    n = len(cells)
    vals = np.zeros(n, dtype=np.float64)
    valid = np.zeros(n, dtype=bool)
    for i, c in enumerate(cells):
        if c is None:
            continue
        vals[i] = float(c)
        valid[i] = True
    return vals, valid


def k_means_etl(text: str, repeats: int = 5, numeric_cols=("value", "lower_ci", "upper_ci"),
                key_col: str = "measure_name") -> dict:
    """The whole Spark-side ETL stage, k_means.py:22-77, on CSV text.

    Order of operations follows the reference: null-key count (:23), filter
    (:27), count (:28), per-column mean over ~isnan & isNotNull (:45-48) and
    impute (:49-51), StringIndexer fit/transform (:34,:73-74), OneHotEncoder
    (:38), VectorAssembler with the one-hot block repeated (:64-68), select
    features (:77)."""
    header, cols = read_csv_columns(text)
    key = cols[key_col]
    key_valid = np.array([k is not None for k in key], dtype=bool)
    nulls = null_count(key_valid)
    kept_rows = int(key_valid.sum())
    means, imputed = {}, []
    for c in numeric_cols:
        v, vv = to_numeric_column(cols[c])
        v, vv = v[key_valid], vv[key_valid]
        m = mean_of(v, vv)
        means[c] = m
        bad = ~vv | np.isnan(v)
        imputed.append(np.where(bad, np.float64(m) if m is not None else np.nan, v))
    kept_keys = [k for k in key if k is not None]
    labels, counts = string_indexer_fit(kept_keys)
    index = string_indexer_transform(kept_keys, labels)
    feats = assemble_features_dense(index, len(labels), imputed, repeats)
    return {
        "null_measure_name": nulls,
        "rows_after_filter": kept_rows,
        "means": means,
        "labels": labels,
        "label_counts": counts,
        "index": index,
        "numerics": imputed,
        "features": feats,
        "inferred_types": {c: infer_numeric_type(cols[c]) for c in numeric_cols},
    }


# ----------------------------------------------------------------------------
# a11  load_csv                                  train_tf_ps.py:75-149
# ----------------------------------------------------------------------------


def load_csv(text: str, numeric_features: Optional[List[str]] = None,
             label_col: str = "subpopulation") -> Tuple[np.ndarray, np.ndarray, List[str]]:
    """Restatement of the TF-side ETL (train_tf_ps.py:75-149).

    Per data row: label = stripped label cell, skip if empty (:120-123); for each
    numeric feature in order, stripped cell empty or case-insensitive "nan" ->
    skip the row (:128-131); float(cell) (:132); any exception (a short row
    whose cell is None, an unparsable number) skips the row (:135-137).  Then
    vocab = sorted(set(labels)) (:143), y = index in vocab as int32 (:145), X as
    float32 (:147).  Empty result raises RuntimeError (:139-140)."""
    if numeric_features is None:
        numeric_features = ["value", "lower_ci", "upper_ci"]
    reader = csv.reader(io.StringIO(text, newline=""))
    try:
        header = next(reader)
    except StopIteration:
        header = []
    # DictReader: duplicate header names -> the last one wins
    pos = {h: j for j, h in enumerate(header)}
    X: List[List[float]] = []
    y_raw: List[str] = []
    for row in reader:
        if not row:  # DictReader skips rows that are completely empty
            continue

        def cell(name):
            j = pos.get(name)
            if j is None:
                return ""  # row.get(c, "") default
            return row[j] if j < len(row) else None  # restval=None for short rows

        lab = cell(label_col)
        if lab is None:
            continue  # None.strip() raises -> row skipped
        lab = lab.strip()
        if not lab:
            continue
        feats, ok = [], True
        for c in numeric_features:
            v = cell(c)
            if v is None:
                ok = False
                break
            v = v.strip()
            if v == "" or v.lower() == "nan":
                ok = False
                break
            try:
                feats.append(float(v))
            except ValueError:
                ok = False
                break
        if not ok:
            continue
        X.append(feats)
        y_raw.append(lab)
    if not X:
        raise RuntimeError("No valid rows were parsed from the dataset.")
    vocab = sorted(set(y_raw))
    index_map = {s: i for i, s in enumerate(vocab)}
    y_idx = np.array([index_map[s] for s in y_raw], dtype=np.int32)
    with np.errstate(over="ignore"):
        X_arr = np.asarray(X, dtype=np.float32)
    return X_arr, y_idx, vocab


# ----------------------------------------------------------------------------
# north-star extensions with no reference site (SURVEY.md §8a "ext"):
# explicit groupBy aggregates, standardisation, bucketing, hash partitioning.
# Oracle = restated Spark semantics; parity unpinned.
# ----------------------------------------------------------------------------


def groupby_agg(keys: np.ndarray, keys_valid: Optional[np.ndarray], values: Sequence[np.ndarray],
                values_valid: Sequence[Optional[np.ndarray]]) -> dict:
    """df.groupBy(k).agg(count(*), count(c), sum(c), avg(c), min(c), max(c)).

    Spark semantics: a null key is its own group (reported here under
    `null_group`, not in `keys`); count(c) counts non-null cells (NaN included);
    sum/avg skip nulls and propagate NaN; min/max skip nulls and order NaN above
    every other double; aggregates of an all-null group are null (returned as
    NaN with count 0).  Output sorted by key ascending (Spark's own output order
    is unspecified; callers sort before comparing)."""
    keys = np.asarray(keys)
    n = keys.shape[0]
    kv = np.ones(n, dtype=bool) if keys_valid is None else np.asarray(keys_valid, dtype=bool)
    uniq, inv = np.unique(keys[kv], return_inverse=True)
    G = uniq.shape[0]
    out = {"keys": uniq, "rows": np.bincount(inv, minlength=G).astype(np.int64), "cols": []}
    null_rows = int((~kv).sum())
    out["null_group_rows"] = null_rows
    for v, vv in zip(values, values_valid):
        d = np.asarray(v).astype(np.float64)
        valid = np.ones(n, dtype=bool) if vv is None else np.asarray(vv, dtype=bool)
        d_k, valid_k = d[kv], valid[kv]
        cnt = np.bincount(inv[valid_k], minlength=G).astype(np.int64)
        sums = np.zeros(G, dtype=np.float64)
        order = np.argsort(inv[valid_k], kind="stable")
        dv = d_k[valid_k][order]
        bounds = np.concatenate([[0], np.cumsum(cnt)])
        mins = np.full(G, np.nan)
        maxs = np.full(G, np.nan)
        for g in range(G):
            seg = dv[bounds[g]:bounds[g + 1]]
            if seg.size:
                sums[g] = math.fsum(seg.tolist())
                nn = seg[~np.isnan(seg)]
                mins[g] = nn.min() if nn.size else np.nan
                maxs[g] = np.nan if nn.size < seg.size else nn.max()
        with np.errstate(invalid="ignore", divide="ignore"):
            mean = np.where(cnt > 0, sums / np.maximum(cnt, 1), np.nan)
        out["cols"].append({"count": cnt, "sum": sums, "mean": mean, "min": mins, "max": maxs})
    return out


def groupby_agg_fast(keys: np.ndarray, values: Sequence[np.ndarray],
                     values_valid: Sequence[Optional[np.ndarray]], num_groups: Optional[int] = None) -> dict:
    """Vectorised variant for large inputs (no per-group Python loop): count /
    sum / mean only, np.add.at-free via bincount with weights (pairwise fp64
    sums; good to ~1e-13 rel on the synthetic data, far inside the 1e-6 bar)."""
    keys = np.asarray(keys)
    uniq, inv = np.unique(keys, return_inverse=True)
    G = uniq.shape[0]
    out = {"keys": uniq, "rows": np.bincount(inv, minlength=G).astype(np.int64), "cols": []}
    for v, vv in zip(values, values_valid):
        d = np.asarray(v).astype(np.float64)
        valid = np.ones(d.shape[0], dtype=bool) if vv is None else np.asarray(vv, dtype=bool)
        cnt = np.bincount(inv[valid], minlength=G).astype(np.int64)
        sums = np.bincount(inv[valid], weights=d[valid], minlength=G)
        with np.errstate(invalid="ignore", divide="ignore"):
            mean = np.where(cnt > 0, sums / np.maximum(cnt, 1), np.nan)
        out["cols"].append({"count": cnt, "sum": sums, "mean": mean})
    return out


def column_stats(values: np.ndarray) -> Tuple[float, float]:
    """StandardScaler.fit's summary of one (already imputed, non-null) column:
    mean and the corrected sample standard deviation sqrt(M2 / (n - 1))."""
    d = np.asarray(values, dtype=np.float64)
    n = d.shape[0]
    mean = math.fsum(d.tolist()) / n
    m2 = math.fsum(((d - mean) ** 2).tolist())
    return mean, math.sqrt(m2 / (n - 1)) if n > 1 else 0.0


def standardize(values: np.ndarray, mean: float, std: float, with_mean: bool = False,
                with_std: bool = True) -> np.ndarray:
    """Spark ML StandardScalerModel.transform on one column: subtract the mean if
    withMean, multiply by 1/std if withStd (a zero std maps to factor 0.0, as
    Spark does).  Spark's defaults are withMean=False, withStd=True."""
    d = np.asarray(values, dtype=np.float64)
    if with_mean:
        d = d - np.float64(mean)
    if with_std:
        scale = np.float64(1.0 / std) if std != 0.0 else np.float64(0.0)
        d = d * scale
    return d


def bucketize(values: np.ndarray, splits: Sequence[float]) -> np.ndarray:
    """Spark ML Bucketizer(handleInvalid="keep"): x in [s_i, s_{i+1}) -> i, the
    last bucket also includes its upper bound; NaN -> len(splits) - 1 (the extra
    bucket); values outside [s_0, s_last] are an error in Spark and are reported
    here as -1."""
    s = np.asarray(splits, dtype=np.float64)
    d = np.asarray(values, dtype=np.float64)
    nb = s.shape[0] - 1
    out = np.searchsorted(s, d, side="right").astype(np.int64) - 1
    out[d == s[-1]] = nb - 1
    out[(d < s[0]) | (d > s[-1])] = -1
    out[np.isnan(d)] = nb
    return out.astype(np.float64)


MIX_C1 = np.uint64(0x9E3779B97F4A7C15)
MIX_C2 = np.uint64(0xBF58476D1CE4E5B9)
MIX_C3 = np.uint64(0x94D049BB133111EB)


def mix64(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser; the key hash used for slot choice and for the
    destination rank of the exchange (dest = mix64(key) >> 32 scaled to P)."""
    z = np.asarray(x).astype(np.uint64)
    z = (z ^ (z >> np.uint64(30))) * MIX_C2
    z = (z ^ (z >> np.uint64(27))) * MIX_C3
    return z ^ (z >> np.uint64(31))


def partition_dest(keys: np.ndarray, ranks: int, buckets: int = 1) -> np.ndarray:
    """Destination of each key in the hash exchange, two levels from one hash h =
    mix64(key): rank = (hi32(h) * ranks) >> 32 (Lemire's range reduction) picks
    the GPU - every rank owns a disjoint key set - and bucket = (lo32(h) * buckets)
    >> 32 a slice of that GPU's table.  dest = rank * buckets + bucket."""
    h = mix64(np.asarray(keys).astype(np.int64).view(np.uint64))
    rank = ((h >> np.uint64(32)) * np.uint64(ranks)) >> np.uint64(32)
    bucket = ((h & np.uint64(0xFFFFFFFF)) * np.uint64(buckets)) >> np.uint64(32)
    return (rank * np.uint64(buckets) + bucket).astype(np.int32)
