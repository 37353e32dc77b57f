"""Parquet decode (north-star extension, no reference site): the CUDA page decoder
against pyarrow's reader on the same file - values bit for bit, nulls exactly,
string codes through the label table.  Files are written here with pyarrow,
uncompressed and with Snappy (Spark's and pyarrow's default), dictionary-encoded
and PLAIN."""
import io

import numpy as np
import pyarrow as pa
import pyarrow.csv as pacsv
import pyarrow.parquet as pq
import pytest

import pyspark_tf_gke_b200 as pkg
from tests.conftest import health_text

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    e = pkg.Engine(0)
    yield e
    e.close()


def _table(n, seed, null_rate=0.07):
    rng = np.random.default_rng(seed)
    names = np.array(["alpha", "beta", "gamma δ", "x" * 40, ""])
    return pa.table({
        "measure_name": pa.array(rng.choice(names, n), mask=rng.random(n) < 0.02),
        "value": pa.array(rng.random(n) * 1e5, mask=rng.random(n) < null_rate),
        "lower_ci": pa.array((rng.random(n) * 100).astype(np.float32), mask=rng.random(n) < null_rate),
        "upper_ci": pa.array(rng.integers(-1000, 1000, n).astype(np.int32), mask=rng.random(n) < null_rate),
        "id": pa.array(np.arange(n, dtype=np.int64)),
        "few": pa.array(rng.integers(0, 7, n).astype(np.float64)),
    })


def _write(t, **kw):
    buf = io.BytesIO()
    opts = dict(compression="NONE", use_dictionary=["measure_name"], data_page_version="1.0")
    opts.update(kw)
    pq.write_table(t, buf, **opts)
    return buf.getvalue()


def _check(eng, raw, t, numeric, strings):
    got = pkg.read_parquet(eng, raw, numeric, strings)
    assert got.rows == t.num_rows
    for c in numeric:
        exp = t.column(c).combine_chunks()
        valid = np.array(exp.is_valid())
        col = got.numeric[c]
        assert (col.valid_host() == valid).all(), c
        ev = exp.to_numpy(zero_copy_only=False).astype(np.float64)
        assert col.data.cpu().numpy()[valid].tobytes() == ev[valid].tobytes(), c
    for c in strings:
        exp = t.column(c).to_pylist()
        sc = got.strings[c]
        valid = sc.column.valid_host()
        codes = sc.column.data.cpu().numpy()
        assert [sc.labels[k] if ok else None for k, ok in zip(codes, valid)] == exp
        assert sc.labels == sorted(set(s for s in exp if s is not None))


@pytest.mark.parametrize("n", [0, 1, 31, 1000, 250_003])
@pytest.mark.parametrize("version", ["1.0", "2.0"])
@pytest.mark.parametrize("codec", ["NONE", "SNAPPY"])
def test_parquet_matches_pyarrow(eng, n, version, codec):
    t = _table(n, seed=n + 1)
    raw = _write(t, data_page_version=version, row_group_size=100_000, data_page_size=64 * 1024, compression=codec)
    _check(eng, raw, t, ["value", "lower_ci", "upper_ci", "id"], ["measure_name"])


def test_parquet_as_spark_and_pyarrow_write_it_by_default(eng):
    # no options at all: Snappy, dictionary encoding for every column, 1 MB pages, v1 data pages - and a column
    # whose dictionary outgrows its limit, so later pages fall back to PLAIN inside a dictionary-encoded chunk
    n = 600_000
    t = _table(n, seed=77)
    rng = np.random.default_rng(78)
    t = t.append_column("wide", pa.array(rng.random(n) * 1e9, mask=rng.random(n) < 0.05))
    t = t.append_column("runs", pa.array(np.repeat(np.arange(n // 1000, dtype=np.float64), 1000)))     # compresses to almost nothing
    buf = io.BytesIO()
    pq.write_table(t, buf)
    raw = buf.getvalue()
    md = pq.ParquetFile(io.BytesIO(raw)).metadata
    assert md.row_group(0).column(0).compression == "SNAPPY"
    _check(eng, raw, t, ["value", "lower_ci", "upper_ci", "id", "few", "wide", "runs"], ["measure_name"])


@pytest.mark.parametrize("codec", ["NONE", "SNAPPY"])
def test_parquet_plain_string_pages(eng, codec):
    # PLAIN BYTE_ARRAY pages (no dictionary, or a writer's fallback when the dictionary grew too large): the
    # strings go through a device dictionary like CSV cells
    n = 40_003
    rng = np.random.default_rng(11)
    labels = np.array([f"label {i:04d} " + "é" * (i % 7) for i in range(300)] + ["", "x"])
    t = pa.table({"measure_name": pa.array(rng.choice(labels, n), mask=rng.random(n) < 0.03),
                  "value": pa.array(rng.random(n), mask=rng.random(n) < 0.07)})
    raw = _write(t, use_dictionary=False, compression=codec, data_page_size=32 * 1024)
    _check(eng, raw, t, ["value"], ["measure_name"])
    # a tiny dictionary page limit: the first pages of each chunk are dictionary-encoded, the rest PLAIN
    raw = _write(t, use_dictionary=["measure_name"], dictionary_pagesize_limit=1024, compression=codec, data_page_size=8 * 1024,
                 row_group_size=15_000)
    _check(eng, raw, t, ["value"], ["measure_name"])


def test_parquet_dictionary_encoded_numerics_and_required_columns(eng):
    t = _table(120_000, seed=3)
    raw = _write(t, use_dictionary=True, row_group_size=50_000)
    _check(eng, raw, t, ["few", "value", "upper_ci"], ["measure_name"])
    # required (non-nullable) columns carry no definition levels
    schema = pa.schema([pa.field("value", pa.float64(), nullable=False), pa.field("measure_name", pa.string(), nullable=False)])
    rng = np.random.default_rng(5)
    t2 = pa.table({"value": rng.random(70_001), "measure_name": rng.choice(["p", "q"], 70_001)}, schema=schema)
    _check(eng, _write(t2), t2, ["value"], ["measure_name"])


def test_parquet_unsupported_inputs_fail_loudly(eng):
    t = _table(1000, seed=9)
    with pytest.raises(pkg.ParquetUnsupported, match="codec"):
        pkg.read_parquet(eng, _write(t, compression="GZIP"), ["value"], ["measure_name"])
    with pytest.raises(pkg.ParquetUnsupported, match="not a Parquet file"):
        pkg.read_parquet(eng, b"PAR1 nonsense", ["value"], [])
    with pytest.raises(pkg.ParquetUnsupported, match="value encoding"):
        pkg.read_parquet(eng, _write(t, use_dictionary=False, column_encoding={"value": "BYTE_STREAM_SPLIT"}), ["value"], [])
    # a corrupted Snappy page is reported, not decoded into garbage
    raw = bytearray(_write(_table(50_000, seed=10), compression="SNAPPY", use_dictionary=False))
    from pyspark_tf_gke_b200 import parquet as P
    _, plans, _ = P._plan(bytes(raw), ["value"])
    src, nbytes, _, _ = plans["value"].inflate[0]
    raw[src] ^= 0x55                                   # the block's length preamble no longer matches the page header
    with pytest.raises(pkg.EtlError, match="corrupt Snappy"):
        pkg.read_parquet(eng, bytes(raw), ["value"], [])
    missing = pkg.read_parquet(eng, _write(t), ["nope"], ["also_nope"])
    assert int(eng.null_count(missing.numeric["nope"]).item()) == 1000


@pytest.mark.parametrize("which", ["workloads", "infra"])
def test_health_parquet_gives_the_same_stage_output_as_health_csv(eng, which, monkeypatch):
    text = health_text(which)
    monkeypatch.setenv("MEASURE_NAME_WEIGHT", "5")
    monkeypatch.setenv("ETL_SKIP_KMEANS", "1")
    arrow = pacsv.read_csv(io.BytesIO(text.encode()), convert_options=pacsv.ConvertOptions(strings_can_be_null=True))
    raw = _write(arrow, compression="SNAPPY" if which == "infra" else "NONE")
    cols = (["value", "lower_ci", "upper_ci"], ["measure_name"])
    from_csv, _ = pkg.KMeansWorkload(eng).k_means(pkg.read_csv(eng, text.encode(), *cols))
    from_pq, _ = pkg.KMeansWorkload(eng).k_means(pkg.read_parquet(eng, raw, *cols))
    assert from_pq.labels == from_csv.labels and from_pq.label_counts == from_csv.label_counts
    assert from_pq.means == from_csv.means == [1738.2895095771808, 1724.1278848441589, 1762.6607185343803]
    assert from_pq.features().cpu().numpy().tobytes() == from_csv.features().cpu().numpy().tobytes()
