"""Property tests (hypothesis) of the CPU-side pieces: the oracle's invariants, the
product's numeral parser (host build of csrc/decimal.cuh) and Thrift reader, the
StringIndexer ordering.  CPU only."""
import ctypes as C
import io
import os
import struct
import subprocess

import numpy as np
import pytest
from hypothesis import given, settings
from hypothesis import strategies as st

from oracle import etl_oracle as O
from pyspark_tf_gke_b200.k_means import string_indexer_order
from tests.conftest import ROOT


@pytest.fixture(scope="module")
def dec(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("dec") / "libdec.so")
    subprocess.run(["g++", "-O2", "-shared", "-fPIC", "-o", so, os.path.join(ROOT, "tests", "native", "decimal_harness.cpp")],
                   check=True)
    lib = C.CDLL(so)

    def parse(s: str):
        b = s.encode()
        out, fl = C.c_double(), C.c_int()
        return lib.dec_parse_host(b, len(b), C.byref(out), C.byref(fl)), out.value

    return parse


@settings(max_examples=300, deadline=None)
@given(st.floats(allow_nan=False, allow_infinity=False, width=64))
def test_repr_of_any_double_parses_back_exactly_or_is_refused(dec, x):
    status, v = dec(repr(x))
    assert status in (0, 3)
    if status == 0:
        assert struct.pack("<d", v) == struct.pack("<d", x)


@settings(max_examples=300, deadline=None)
@given(st.integers(min_value=-10**18, max_value=10**18), st.integers(min_value=0, max_value=18))
def test_fixed_point_numerals_are_correctly_rounded(dec, m, frac):
    digits = str(abs(m)).rjust(frac + 1, "0")
    s = ("-" if m < 0 else "") + digits[:len(digits) - frac] + ("." + digits[len(digits) - frac:] if frac else "")
    status, v = dec(s)
    assert status == 0 and struct.pack("<d", v) == struct.pack("<d", float(s))


@settings(max_examples=200, deadline=None)
@given(st.text(alphabet="0123456789.eE+- naNifty", max_size=12))
def test_parser_never_accepts_what_float_rejects(dec, s):
    status, v = dec(s)
    try:
        ref = float(s)
    except ValueError:
        ref = None
    if status == 0:
        assert ref is not None and (struct.pack("<d", v) == struct.pack("<d", ref) or (v != v and ref != ref))
    if ref is None:
        assert status in (1, 2)


@settings(max_examples=100, deadline=None)
@given(st.lists(st.tuples(st.sampled_from(["a", "b", "c", "Ab", "é", ""]), st.booleans()), max_size=60))
def test_string_indexer_order_is_count_desc_then_label_asc(rows):
    labels = [s if ok else None for s, ok in rows]
    vocab, counts = O.string_indexer_fit(labels)
    assert counts == sorted(counts, reverse=True)
    for (a, ca), (b, cb) in zip(zip(vocab, counts), zip(vocab[1:], counts[1:])):
        assert ca > cb or a < b
    uniq = sorted(set(s for s in labels if s is not None))
    cnt = [sum(1 for s in labels if s == u) for u in uniq]
    assert [uniq[i] for i in string_indexer_order(uniq, cnt)] == vocab
    idx = O.string_indexer_transform(labels, vocab)
    assert all((i == len(vocab)) == (s is None) for i, s in zip(idx, labels))


@settings(max_examples=100, deadline=None)
@given(st.integers(0, 2000), st.integers(0, 2**31))
def test_projection_oracle_invariants(n, seed):
    rng = np.random.default_rng(seed)
    code = rng.integers(0, 30, n).astype(np.int32)
    kv = rng.random(n) < 0.9
    vals = [np.where(rng.random(n) < 0.05, np.nan, rng.random(n) * 100) for _ in range(3)]
    valid = [rng.random(n) < 0.9 for _ in range(3)]
    fill = [1.0, 2.0, 3.0]
    out_code, outs = O.project_filter_impute(code, kv, vals, valid, fill)
    assert len(out_code) == int(kv.sum()) and (out_code == code[kv]).all()
    for o, v, vv, f in zip(outs, vals, valid, fill):
        assert not np.isnan(o).any()
        keep_val = (vv & ~np.isnan(v))[kv]
        assert (o[keep_val] == v[kv][keep_val]).all() and (o[~keep_val] == f).all()
    # idempotent
    again_code, again = O.project_filter_impute(out_code, np.ones(len(out_code), bool), outs, [np.ones(len(out_code), bool)] * 3, fill)
    assert (again_code == out_code).all() and all((a == b).all() for a, b in zip(again, outs))


@settings(max_examples=60, deadline=None)
@given(st.integers(1, 64), st.integers(1, 64), st.integers(0, 2**31))
def test_partition_destinations_are_a_function_of_the_key_and_in_range(ranks, buckets, seed):
    rng = np.random.default_rng(seed)
    keys = rng.integers(-2**62, 2**62, 500)
    d = O.partition_dest(keys, ranks, buckets)
    assert ((d >= 0) & (d < ranks * buckets)).all()
    assert (O.partition_dest(keys, ranks, buckets) == d).all()
    assert (d // buckets == O.partition_dest(keys, ranks, 1)).all()   # the rank does not depend on the bucket count


def test_thrift_reader_round_trips_a_pyarrow_footer():
    import pyarrow as pa
    import pyarrow.parquet as pq

    from pyspark_tf_gke_b200 import parquet as P

    t = pa.table({"value": pa.array([1.0, None, 3.0]), "measure_name": pa.array(["a", "b", None])})
    buf = io.BytesIO()
    pq.write_table(t, buf, compression="NONE")
    rows, plans, end = P._plan(buf.getvalue(), ["value", "measure_name", "absent"])
    assert rows == 3 and set(plans) == {"value", "measure_name"}
    assert plans["measure_name"].dictionaries == [["a", "b"]]
    assert sum(pg.num_values for pg in plans["value"].pages) == 3 and plans["value"].optional
    assert not plans["value"].inflate
    # the same table as Spark / pyarrow write it by default: snappy pages get a place behind the file bytes
    buf = io.BytesIO()
    pq.write_table(t, buf, compression="snappy")
    raw = buf.getvalue()
    rows, plans, end = P._plan(raw, ["value", "measure_name"])
    assert rows == 3 and plans["measure_name"].dictionaries == [["a", "b"]]       # dictionary pages are inflated on the host
    for plan in plans.values():
        assert len(plan.inflate) == len(plan.pages) == 1
        src, src_bytes, dst, dst_bytes = plan.inflate[0]
        assert dst >= len(raw) and dst % 64 == 0 and dst + dst_bytes <= end
        page = P.snappy_decode(raw, src, src + src_bytes)
        assert len(page) == dst_bytes
        pg = plan.pages[0]
        assert pg.def_offset == dst + 4 and pg.def_bytes == int.from_bytes(page[:4], "little") and pg.bit_width == -1
        assert pg.val_offset == pg.def_offset + pg.def_bytes and pg.val_offset + pg.val_bytes == dst + dst_bytes


def test_snappy_decoder_against_pyarrow():
    import pyarrow as pa

    from pyspark_tf_gke_b200.parquet import snappy_decode

    rng = np.random.default_rng(0)
    for data in (b"", b"a", b"abcd" * 1000, bytes(rng.integers(0, 4, 100000, dtype=np.uint8)),
                 bytes(rng.integers(0, 256, 70000, dtype=np.uint8)), b"\x00" * 100000, b"ab" * 70000 + b"xyz"):
        c = pa.compress(data, codec="snappy", asbytes=True)
        assert snappy_decode(c, 0, len(c)) == data
        assert snappy_decode(c, 0, len(c), 5) == data[:5]
