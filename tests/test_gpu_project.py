"""Parity of the CUDA projection / aggregate kernels against the oracle, through
the C ABI (ctypes -> libetl_b200.so).  Integer/byte results are compared bit for
bit; fp64 sums of non-integer data within 1e-6 relative (the north star's
tolerance; the observed error is ~1e-15)."""
import numpy as np
import pytest
import torch

import pyspark_tf_gke_b200 as pkg
import pyspark_tf_gke_b200.synth as S
from oracle import etl_oracle as O

pytestmark = pytest.mark.gpu
REL_TOL = 1e-6


@pytest.fixture(scope="module")
def eng():
    e = pkg.Engine(0)
    yield e
    e.close()


def _host_table(n, seed, kind=S.KIND_REAL, dtype=np.float64, key_null=0.05, nan_rate=0.01, val_null=0.07):
    code, kv = S.host_column(seed, 0, S.KIND_KEY32, n, modulus=30, null_rate=key_null)
    vals, valid = [], []
    for c in range(3):
        v, vv = S.host_column(seed, 1 + c, kind, n, null_rate=val_null, nan_rate=nan_rate if dtype != np.int32 else 0.0)
        vals.append(v.astype(dtype))
        valid.append(vv)
    return code, kv, vals, valid


def _to_device(eng, code, kv, vals, valid, bitmaps=True):
    key = pkg.Column.from_host(code, kv if bitmaps else None, eng.device)
    cols = [pkg.Column.from_host(v, vv if bitmaps else None, eng.device) for v, vv in zip(vals, valid)]
    return key, cols


@pytest.mark.parametrize("n", [0, 1, 3, 4, 31, 32, 33, 1023, 1024, 1025, 4096, 65_537, 1_000_003])
def test_project_matches_oracle_sizes(eng, n):
    code, kv, vals, valid = _host_table(n, seed=11)
    fill = [1738.25, -3.5, 1e300]
    key, cols = _to_device(eng, code, kv, vals, valid)
    out_code, out_vals, rows = eng.project_filter_impute(key, cols, fill)
    exp_code, exp_vals = O.project_filter_impute(code, kv, vals, valid, fill)
    r = int(rows.item())
    assert r == exp_code.shape[0]
    assert out_code[:r].cpu().numpy().tobytes() == exp_code.tobytes()
    for a, b in zip(out_vals, exp_vals):
        assert a[:r].cpu().numpy().tobytes() == b.tobytes()


@pytest.mark.parametrize("dtype", [np.float32, np.int32])
def test_project_casts_to_double(eng, dtype):
    n = 200_001
    code, kv, vals, valid = _host_table(n, seed=5, kind=S.KIND_EXACT, dtype=dtype)
    fill = [0.5, 0.25, 0.125]
    key, cols = _to_device(eng, code, kv, vals, valid)
    out_code, out_vals, rows = eng.project_filter_impute(key, cols, fill)
    exp_code, exp_vals = O.project_filter_impute(code, kv, vals, valid, fill)
    r = int(rows.item())
    assert r == exp_code.shape[0] and out_vals[0].dtype == torch.float64
    assert out_code[:r].cpu().numpy().tobytes() == exp_code.tobytes()
    for a, b in zip(out_vals, exp_vals):
        assert a[:r].cpu().numpy().tobytes() == b.tobytes()


def test_project_mixed_dtypes_and_missing_bitmaps(eng):
    n = 50_000
    code, kv, vals, valid = _host_table(n, seed=2, nan_rate=0.02)
    vals[1] = np.nan_to_num(vals[1]).astype(np.float32)
    vals[2] = np.nan_to_num(vals[2]).astype(np.int32)
    fill = [7.0, 8.0, 9.0]
    key = pkg.Column.from_host(code, None, eng.device)  # no key nulls: nothing filtered
    cols = [pkg.Column.from_host(vals[0], None, eng.device), pkg.Column.from_host(vals[1], valid[1], eng.device),
            pkg.Column.from_host(vals[2], None, eng.device)]
    out_code, out_vals, rows = eng.project_filter_impute(key, cols, fill)
    ones = np.ones(n, dtype=bool)
    exp_code, exp_vals = O.project_filter_impute(code, ones, vals, [ones, valid[1], ones], fill)
    assert int(rows.item()) == n
    assert out_code.cpu().numpy().tobytes() == exp_code.tobytes()
    for a, b in zip(out_vals, exp_vals):
        assert a.cpu().numpy().tobytes() == b.tobytes()


@pytest.mark.parametrize("pattern", ["all_null", "none_null", "alternating", "first_tile_only"])
def test_project_edge_patterns(eng, pattern):
    n = 10_000
    code, kv, vals, valid = _host_table(n, seed=9)
    if pattern == "all_null":
        kv[:] = False
    elif pattern == "none_null":
        kv[:] = True
    elif pattern == "alternating":
        kv[:] = np.arange(n) % 2 == 0
    else:
        kv[:] = np.arange(n) < 1024
    key, cols = _to_device(eng, code, kv, vals, valid)
    out_code, out_vals, rows = eng.project_filter_impute(key, cols, [0.0, 0.0, 0.0])
    exp_code, exp_vals = O.project_filter_impute(code, kv, vals, valid, [0.0, 0.0, 0.0])
    r = int(rows.item())
    assert r == exp_code.shape[0]
    assert out_code[:r].cpu().numpy().tobytes() == exp_code.tobytes()
    for a, b in zip(out_vals, exp_vals):
        assert a[:r].cpu().numpy().tobytes() == b.tobytes()


def test_project_repeated_launches_reuse_tile_state(eng):
    # consecutive launches of different sizes share the look-back words
    for n, seed in [(300_000, 1), (5_000, 2), (1_000_000, 3), (5_000, 4)]:
        code, kv, vals, valid = _host_table(n, seed=seed)
        key, cols = _to_device(eng, code, kv, vals, valid)
        out_code, out_vals, rows = eng.project_filter_impute(key, cols, [1.0, 2.0, 3.0])
        exp_code, _ = O.project_filter_impute(code, kv, vals, valid, [1.0, 2.0, 3.0])
        r = int(rows.item())
        assert r == exp_code.shape[0] and out_code[:r].cpu().numpy().tobytes() == exp_code.tobytes()


def test_project_host_entry_matches_device_entry(eng):
    n = 2_500_000  # three 1 Mi-row chunks, last one ragged
    code, kv, vals, valid = _host_table(n, seed=21)
    fill = [1.0, 2.0, 3.0]
    out_code = np.empty(n, dtype=np.int32)
    out_vals = [np.empty(n, dtype=np.float64) for _ in range(3)]
    r = eng.project_filter_impute_host(code, O.pack_validity(kv), vals, [O.pack_validity(v) for v in valid], fill,
                                       out_code, out_vals)
    exp_code, exp_vals = O.project_filter_impute(code, kv, vals, valid, fill)
    assert r == exp_code.shape[0]
    assert out_code[:r].tobytes() == exp_code.tobytes()
    for a, b in zip(out_vals, exp_vals):
        assert a[:r].tobytes() == b.tobytes()


def test_project_config2_full_size_properties(eng):
    """BASELINE.json configs[1] at full size (10 M rows): size-independent
    properties - kept count = popcount of the key bitmap, order preserved (key
    codes of kept rows equal the filtered input), idempotence (projecting the
    output again changes nothing), and column sums consistent with
    sum(valid) + nulls * fill."""
    cfg = S.CONFIG2
    n = cfg["rows"]
    key = eng.synth_column(cfg["seed"], 0, S.KIND_KEY32, n, modulus=cfg["categories"], null_rate=cfg["key_null"])
    cols = [eng.synth_column(cfg["seed"], 1 + c, S.KIND_EXACT, n, null_rate=cfg["val_null"][c]) for c in range(3)]
    fill = [3.0, 5.0, 7.0]
    out_code, out_vals, rows = eng.project_filter_impute(key, cols, fill)
    r = int(rows.item())
    nulls = int(eng.null_count(key).item())
    assert r == n - nulls
    kv = torch.from_numpy(key.valid_host()).to(eng.device)
    assert torch.equal(out_code[:r], key.data[kv])
    # sums: integer-valued, exact
    sums, counts = eng.sum_count(cols, row_valid=key.valid)
    out_cols = [pkg.Column(v[:r]) for v in out_vals]
    sums2, counts2 = eng.sum_count(out_cols)
    for c in range(3):
        assert int(counts2[c]) == r
        assert float(sums2[c]) == float(sums[c]) + (r - int(counts[c])) * fill[c]
    # idempotence
    again_code, again_vals, rows2 = eng.project_filter_impute(pkg.Column(out_code[:r]), out_cols, fill)
    assert int(rows2.item()) == r and torch.equal(again_code[:r], out_code[:r])
    for a, b in zip(again_vals, out_vals):
        assert torch.equal(a[:r], b[:r])


@pytest.mark.parametrize("n", [0, 1, 33, 1024, 70_001, 3_000_001])
def test_null_count_and_sum_count_exact(eng, n):
    code, kv, vals, valid = _host_table(n, seed=13, kind=S.KIND_EXACT)
    key, cols = _to_device(eng, code, kv, vals, valid)
    assert int(eng.null_count(key).item()) == O.null_count(kv)
    sums, counts = eng.sum_count(cols, row_valid=key.valid)
    for c in range(3):
        es, ec = O.sum_count(vals[c], valid[c] & kv)
        assert int(counts[c]) == ec
        assert float(sums[c]) == es  # integer-valued: bit exact


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_sum_count_real_within_tolerance_and_reproducible(eng, dtype):
    n = 2_000_003
    code, kv, vals, valid = _host_table(n, seed=17, kind=S.KIND_REAL, dtype=dtype, nan_rate=0.02)
    key, cols = _to_device(eng, code, kv, vals, valid)
    sums, counts = eng.sum_count(cols)
    first = sums.cpu().numpy().tobytes()
    for c in range(3):
        es, ec = O.sum_count(vals[c], valid[c])
        assert int(counts[c]) == ec
        assert abs(float(sums[c]) - es) <= REL_TOL * abs(es)
    sums2, _ = eng.sum_count(cols)
    assert sums2.cpu().numpy().tobytes() == first  # fixed-shape tree: same bits every run
    s1, c1 = eng.sum_count(cols[:1])
    assert float(s1[0]) == float(sums[0]) and int(c1[0]) == int(counts[0])


def test_synth_device_matches_host(eng):
    n = 100_003
    for kind, kw in [(S.KIND_EXACT, dict(null_rate=0.07, nan_rate=0.01)), (S.KIND_REAL, dict(null_rate=0.07)),
                     (S.KIND_KEY32, dict(modulus=30, null_rate=0.01)), (S.KIND_KEY64, dict(modulus=10**8))]:
        col = eng.synth_column(3, 4, kind, n, **kw)
        v, vv = S.host_column(3, 4, kind, n, **kw)
        assert col.data.cpu().numpy().tobytes() == v.tobytes()
        assert (col.valid_host() == vv).all()
    # slices are reproducible: rows [64, 64+1000) of the same column
    col = eng.synth_column(3, 4, S.KIND_REAL, 1000, row0=64)
    v, _ = S.host_column(3, 4, S.KIND_REAL, 1000, row0=64)
    assert col.data.cpu().numpy().tobytes() == v.tobytes()


def test_errors_are_loud(eng):
    key = pkg.Column(torch.zeros(16, dtype=torch.int32, device=eng.device))
    bad = pkg.Column(torch.zeros(17, dtype=torch.float64, device=eng.device)[1:])  # 8-byte aligned only
    good = pkg.Column(torch.zeros(16, dtype=torch.float64, device=eng.device))
    with pytest.raises(pkg.EtlError, match="misaligned"):
        eng.project_filter_impute(key, [bad, good, good], [0, 0, 0])
