"""The compiled oracle (oracle/etl_oracle.c) against the NumPy restatement."""
import numpy as np
import pytest

from oracle import c_oracle as CO
from oracle import etl_oracle as O
import pyspark_tf_gke_b200.synth as S


def _table(n, seed=7, dtype=np.float64, kind=S.KIND_REAL, nan_rate=0.01):
    code, kv = S.host_column(seed, 0, S.KIND_KEY32, n, modulus=30, null_rate=0.05)
    vals, valid = [], []
    for c in range(3):
        v, vv = S.host_column(seed, 1 + c, kind, n, null_rate=0.07, nan_rate=nan_rate)
        vals.append(v.astype(dtype) if dtype != np.int32 else np.nan_to_num(v).astype(np.int32))
        valid.append(vv)
    return code, kv, vals, valid


@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 1000, 100_003])
@pytest.mark.parametrize("dtype", [np.float64, np.float32, np.int32])
def test_project_matches_numpy(n, dtype):
    code, kv, vals, valid = _table(n, dtype=dtype)
    fill = [1.5, -2.25, 1e9]
    exp_code, exp_vals = O.project_filter_impute(code, kv, vals, valid, fill)
    rows, out_code, out_vals = CO.project_filter_impute(
        code, O.pack_validity(kv), vals, [O.pack_validity(v) for v in valid], fill)
    assert rows == exp_code.shape[0]
    assert (out_code[:rows] == exp_code).all()
    for a, b in zip(out_vals, exp_vals):
        assert a[:rows].tobytes() == b.tobytes()


@pytest.mark.parametrize("n", [0, 5, 64, 100_003])
def test_sum_count_and_null_count(n):
    code, kv, vals, valid = _table(n, kind=S.KIND_EXACT)
    assert CO.null_count(O.pack_validity(kv), n) == O.null_count(kv)
    for v, vv in zip(vals, valid):
        s, c = CO.sum_count(v, O.pack_validity(vv), O.pack_validity(kv))
        es, ec = O.sum_count(v, vv & kv)
        assert (s, c) == (es, ec)  # integer-valued data: exact
    v, vv = S.host_column(3, 9, S.KIND_REAL, n, null_rate=0.1, nan_rate=0.02)
    s, c = CO.sum_count(v, O.pack_validity(vv))
    es, ec = O.sum_count(v, vv)
    assert c == ec and abs(s - es) <= 1e-12 * max(1.0, abs(es))


def test_no_bitmaps_means_no_nulls():
    code, kv, vals, valid = _table(1000, nan_rate=0.0)
    rows, out_code, out_vals = CO.project_filter_impute(code, None, vals, [None] * 3, [0, 0, 0])
    assert rows == 1000 and (out_code == code).all()
    for a, b in zip(out_vals, vals):
        assert a.tobytes() == b.tobytes()


@pytest.mark.parametrize("n,groups", [(0, 5), (1, 5), (1000, 7), (200_003, 20_000), (300_000, 1)])
def test_groupby_sum_and_synth_match_numpy(n, groups):
    # the compiled groupBy (bench.py's CPU arm for config 4) and the compiled generator against NumPy
    keys = CO.synth_column(3, 0, S.KIND_KEY64, n, row0=96, modulus=groups)
    vals = CO.synth_column(3, 1, S.KIND_EXACT, n, row0=96)
    hk, _ = S.host_column(3, 0, S.KIND_KEY64, n, row0=96, modulus=groups)
    hv, _ = S.host_column(3, 1, S.KIND_EXACT, n, row0=96)
    assert (keys == hk).all() and vals.tobytes() == hv.tobytes()
    gk, gc, gs = CO.groupby_sum(keys, vals)
    want = O.groupby_agg_fast(keys, [vals], [None])
    order = np.argsort(gk)
    assert (gk[order] == want["keys"]).all() and (gc[order] == want["rows"]).all()
    assert gs[order].tobytes() == want["cols"][0]["sum"].tobytes()     # integer-valued data: exact in any order
