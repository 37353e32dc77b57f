"""StandardScaler.fit on the device (etl_column_moments) and its transform through
the feature writers, against oracle.column_stats / oracle.standardize.  count is
exact; mean, m2 and the standardised features are floating point: <= 1e-6
relative (north_star's bar; observed ~1e-15)."""
import math

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


@pytest.mark.parametrize("n", [1, 2, 33, 1000, 100_003, 2_000_001])
@pytest.mark.parametrize("kind", [S.KIND_EXACT, S.KIND_REAL])
def test_column_moments_match_oracle(eng, n, kind):
    _, kv = S.host_column(9, 0, S.KIND_KEY32, n, modulus=30, null_rate=0.05)
    vals, valid = [], []
    for c in range(3):
        v, vv = S.host_column(9, 1 + c, kind, n, null_rate=0.07, nan_rate=0.01 if c == 1 else 0.0)
        vals.append(v + 1e6 * c)            # a large offset: the naive sum-of-squares formula would lose digits here
        valid.append(vv)
    key_valid = torch.from_numpy(pkg.engine.pack_validity_host(kv).view(np.int32)).to(eng.device)
    cols = [pkg.Column.from_host(v, vv, eng.device) for v, vv in zip(vals, valid)]
    count, mean, m2 = (t.cpu().numpy() for t in eng.column_moments(cols, row_valid=key_valid))
    for c in range(3):
        sel = kv & valid[c] & ~np.isnan(vals[c])
        assert int(count[c]) == int(sel.sum())
        if not sel.any():
            assert math.isnan(mean[c])
            continue
        want_mean, want_std = O.column_stats(vals[c][sel])
        assert abs(mean[c] - want_mean) <= REL_TOL * abs(want_mean)
        want_m2 = want_std ** 2 * (sel.sum() - 1)
        assert abs(m2[c] - want_m2) <= REL_TOL * max(want_m2, 1e-300)


def test_moments_are_reproducible_and_f32_i32_inputs_cast_like_the_sum_pass(eng):
    n = 300_007
    rng = np.random.default_rng(4)
    a = (rng.random(n) * 1000).astype(np.float32)
    b = rng.integers(-500, 500, n).astype(np.int32)
    cols = [pkg.Column.from_host(a, None, eng.device), pkg.Column.from_host(b, None, eng.device)]
    r1 = [t.clone() for t in eng.column_moments(cols)]
    r2 = eng.column_moments(cols)
    for x, y in zip(r1, r2):
        assert torch.equal(x, y)                       # fixed-shape tree: same bits every run
    for c, host in enumerate((a.astype(np.float64), b.astype(np.float64))):
        want_mean, want_std = O.column_stats(host)
        assert abs(float(r2[1][c]) - want_mean) <= REL_TOL * abs(want_mean)
        assert abs(math.sqrt(float(r2[2][c]) / (n - 1)) - want_std) <= REL_TOL * want_std


@pytest.mark.parametrize("with_mean", [False, True])
def test_standard_scaler_fit_then_assemble(eng, with_mean):
    # fit on the device, transform inside the feature writers (dense f64 / f32, sparse, TF layout)
    n, L, repeats = 50_001, 7, 2
    rng = np.random.default_rng(12)
    nums = [rng.random(n) * 100 + 50 * c for c in range(3)]
    nums[2][:] = 3.25                                   # a constant column: std 0 -> factor 0, as Spark does
    index = rng.integers(0, L + 1, n).astype(np.int32)
    cols = [pkg.Column.from_host(v, None, eng.device) for v in nums]
    shift, scale = eng.standard_scaler_fit(cols, with_mean=with_mean)
    stats = [O.column_stats(v) for v in nums]
    want = [O.standardize(v, m, s, with_mean=with_mean) for v, (m, s) in zip(nums, stats)]
    assert scale[2] == 0.0
    d_index = torch.from_numpy(index).to(eng.device)
    d_nums = [c.data for c in cols]
    for dtype, tol in ((torch.float64, REL_TOL), (torch.float32, 1e-5)):
        F = eng.assemble_dense(d_index, d_nums, L, repeats, dtype=dtype, shift=shift, scale=scale).cpu().numpy()
        exp = O.assemble_features_dense(index, L, want, repeats)
        np.testing.assert_allclose(F[:, :repeats * L], exp[:, :repeats * L], rtol=0, atol=0)       # the one-hot part: exact
        np.testing.assert_allclose(F[:, repeats * L:], exp[:, repeats * L:], rtol=tol, atol=1e-9)
    oi, ov = eng.assemble_sparse(d_index, d_nums, L, repeats, shift=shift, scale=scale)
    np.testing.assert_allclose(ov.cpu().numpy()[:, -3:], np.stack(want, axis=1), rtol=REL_TOL, atol=1e-9)
