"""groupBy aggregates and the hash partition of the exchange against the oracle.
Keys, row counts and counts are exact; sums are exact on integer-valued data and
within 1e-6 relative otherwise (atomic accumulation order is not fixed);
min / max are exact."""
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


def _check(res, exp, exact_sums, minmax):
    assert res["keys"].cpu().numpy().tolist() == exp["keys"].tolist()
    assert res["rows"].cpu().numpy().tolist() == exp["rows"].tolist()
    assert res["null_group_rows"] == exp.get("null_group_rows", 0)
    for got, want in zip(res["cols"], exp["cols"]):
        assert got["count"].cpu().numpy().tolist() == want["count"].tolist()
        gs, ws = got["sum"].cpu().numpy(), want["sum"]
        if exact_sums:
            assert gs.tobytes() == ws.tobytes()
        else:
            np.testing.assert_allclose(gs, ws, rtol=REL_TOL, atol=0, equal_nan=True)
            np.testing.assert_allclose(got["mean"].cpu().numpy(), want["mean"], rtol=REL_TOL, atol=0, equal_nan=True)
        if minmax:
            assert np.array_equal(got["min"].cpu().numpy(), want["min"], equal_nan=True)
            assert np.array_equal(got["max"].cpu().numpy(), want["max"], equal_nan=True)


@pytest.mark.parametrize("n,groups,hint", [(0, 5, 0), (1, 1, 0), (1000, 7, 0), (200_003, 1000, 1000), (200_003, 30, 30),
                                           (300_000, 5000, 5000), (300_000, 5000, 100)])
@pytest.mark.parametrize("key_dtype", [np.int32, np.int64])
def test_groupby_matches_oracle(eng, n, groups, hint, key_dtype):
    keys, kv = S.host_column(5, 0, S.KIND_KEY64, n, modulus=groups, null_rate=0.02)
    keys = (keys - groups // 2).astype(key_dtype)  # negative keys too
    vals, valid = [], []
    for c in range(3):
        v, vv = S.host_column(5, 1 + c, S.KIND_REAL, n, null_rate=0.07, nan_rate=0.001 if c == 2 else 0.0)
        vals.append(v - 5e4)
        valid.append(vv)
    gt = eng.groupby_new(1 << 14, 3, minmax=True)
    kcol = pkg.Column.from_host(keys, kv, eng.device)
    cols = [pkg.Column.from_host(v, vv, eng.device) for v, vv in zip(vals, valid)]
    eng.groupby_accumulate(gt, kcol.data, kcol.valid, [c.data for c in cols], [c.valid for c in cols], groups_hint=hint)
    res = eng.groupby_result(gt, sort=True)
    exp = O.groupby_agg(keys, kv, vals, valid)
    _check(res, exp, exact_sums=False, minmax=True)


def test_groupby_exact_sums_on_integer_valued_data_in_batches(eng):
    n, groups = 500_000, 1000
    keys, _ = S.host_column(2, 0, S.KIND_KEY32, n, modulus=groups)
    vals, valid = [], []
    for c in range(3):
        v, vv = S.host_column(2, 1 + c, S.KIND_EXACT, n, null_rate=0.07)
        vals.append(v)
        valid.append(vv)
    gt = eng.groupby_new(4096, 3)
    half = (n // 2) // 32 * 32
    for lo, hi in ((0, half), (half, n)):  # two batches into one table
        k = pkg.Column.from_host(keys[lo:hi], None, eng.device)
        cols = [pkg.Column.from_host(v[lo:hi], vv[lo:hi], eng.device) for v, vv in zip(vals, valid)]
        eng.groupby_accumulate(gt, k.data, None, [c.data for c in cols], [c.valid for c in cols], groups_hint=groups)
    res = eng.groupby_result(gt, sort=True)
    exp = O.groupby_agg(keys, None, vals, valid)
    _check(res, exp, exact_sums=True, minmax=False)


@pytest.mark.parametrize("ncols", [1, 2, 3])
def test_groupby_direct_indexed_path(eng, ncols):
    # key_range promises int32 keys in [0, range): direct-indexed shared table; keys
    # outside the promise and null keys must still come out right
    n, groups = 300_001, 1000
    keys, kv = S.host_column(8, 0, S.KIND_KEY32, n, modulus=groups, null_rate=0.01)
    keys = keys.copy()
    keys[::5003] = 5000 + (np.arange(len(keys[::5003])) % 7)   # beyond the promised range
    keys[7::9001] = -3
    vals, valid = [], []
    for c in range(ncols):
        v, vv = S.host_column(8, 1 + c, S.KIND_EXACT, n, null_rate=0.07, nan_rate=0.0)
        vals.append(v)
        valid.append(vv)
    gt = eng.groupby_new(4096, ncols)
    k = pkg.Column.from_host(keys, kv, eng.device)
    cols = [pkg.Column.from_host(v, vv, eng.device) for v, vv in zip(vals, valid)]
    eng.groupby_accumulate(gt, k.data, k.valid, [c.data for c in cols], [c.valid for c in cols], groups_hint=groups,
                           key_range=groups)
    res = eng.groupby_result(gt, sort=True)
    exp = O.groupby_agg(keys, kv, vals, valid)
    _check(res, exp, exact_sums=True, minmax=False)


def test_groupby_all_null_group_and_full_table(eng):
    keys = np.array([1, 1, 2, 3], dtype=np.int64)
    v = np.array([1.0, 2.0, 5.0, 7.0])
    vv = np.array([True, True, False, True])
    gt = eng.groupby_new(8, 1, minmax=True)
    c = pkg.Column.from_host(v, vv, eng.device)
    eng.groupby_accumulate(gt, torch.from_numpy(keys).to(eng.device), None, [c.data], [c.valid])
    res = eng.groupby_result(gt, sort=True)
    assert res["keys"].cpu().tolist() == [1, 2, 3] and res["rows"].cpu().tolist() == [2, 1, 1]
    col = res["cols"][0]
    assert col["count"].cpu().tolist() == [2, 0, 1] and col["sum"].cpu().tolist() == [3.0, 0.0, 7.0]
    assert np.array_equal(col["min"].cpu().numpy(), np.array([1.0, np.nan, 7.0]), equal_nan=True)
    assert np.isnan(col["mean"].cpu().numpy()[1])
    # 100 distinct keys cannot fit 8 slots: loud failure, not a silent drop
    gt = eng.groupby_new(8, 0)
    eng.groupby_accumulate(gt, torch.arange(100, dtype=torch.int64, device=eng.device), None, [], groups_hint=4096)
    with pytest.raises(pkg.EtlError, match="dropped"):
        eng.groupby_result(gt, sort=True)


@pytest.mark.parametrize("ranks,buckets", [(1, 1), (2, 1), (8, 1), (1, 16), (8, 32)])
def test_hash_partition_matches_oracle_destinations(eng, ranks, buckets):
    n = 300_007
    parts = ranks * buckets
    keys, _ = S.host_column(3, 0, S.KIND_KEY64, n, modulus=10**8)
    vals, _ = S.host_column(3, 1, S.KIND_REAL, n)
    rec, counts, base = eng.hash_partition(torch.from_numpy(keys).to(eng.device), torch.from_numpy(vals).to(eng.device),
                                           ranks, buckets)
    rec = rec.cpu().numpy()
    dest = O.partition_dest(keys, ranks, buckets)
    assert counts.cpu().numpy().tolist() == np.bincount(dest, minlength=parts).tolist()
    assert base.cpu().numpy().tolist() == np.concatenate([[0], np.cumsum(np.bincount(dest, minlength=parts))[:-1]]).tolist()
    # every destination range holds exactly its rows (as a multiset of records)
    b = base.cpu().numpy().tolist() + [n]
    for d in range(parts):
        got = rec[b[d]:b[d + 1]]
        assert (O.partition_dest(got[:, 0], ranks, buckets) == d).all()
        want = np.stack([keys[dest == d], vals[dest == d].view(np.int64)], axis=1)
        assert np.array_equal(got[np.lexsort((got[:, 1], got[:, 0]))], want[np.lexsort((want[:, 1], want[:, 0]))])


@pytest.mark.parametrize("buckets", [1, 8])
def test_partition_then_strided_merge_equals_direct_groupby(eng, buckets):
    # single-process emulation of the exchange: every rank's range is merged into
    # its own table (bucket by bucket); together they must equal the direct aggregation
    n, groups, ranks = 400_000, 50_000, 4
    keys, _ = S.host_column(9, 0, S.KIND_KEY64, n, modulus=groups)
    vals, _ = S.host_column(9, 1, S.KIND_EXACT, n)
    dk, dv = torch.from_numpy(keys).to(eng.device), torch.from_numpy(vals).to(eng.device)
    rec, counts, base = eng.hash_partition(dk, dv, ranks, buckets)
    b = base.cpu().tolist() + [n]
    all_keys, all_sums, all_rows = [], [], []
    for r in range(ranks):
        gt = eng.groupby_new(1 << 17, 1)
        for k in range(buckets):
            d = r * buckets + k
            eng.groupby_accumulate_records(gt, rec, b[d], b[d + 1], groups)
        res = eng.groupby_result(gt, sort=True)
        all_keys.append(res["keys"]); all_sums.append(res["cols"][0]["sum"]); all_rows.append(res["rows"])
    k = torch.cat(all_keys).cpu().numpy()
    order = np.argsort(k)
    exp = O.groupby_agg_fast(keys, [vals], [None])
    assert k[order].tolist() == exp["keys"].tolist()  # disjoint ownership: no key on two ranks
    assert torch.cat(all_rows).cpu().numpy()[order].tolist() == exp["rows"].tolist()
    assert torch.cat(all_sums).cpu().numpy()[order].tobytes() == exp["cols"][0]["sum"].tobytes()


def test_bucketed_single_gpu_shuffle_path(eng):
    import pyspark_tf_gke_b200.dist as D

    n, groups = 500_000, 100_000
    keys, _ = S.host_column(4, 0, S.KIND_KEY64, n, modulus=groups)
    vals, _ = S.host_column(4, 1, S.KIND_EXACT, n)
    dk, dv = torch.from_numpy(keys).to(eng.device), torch.from_numpy(vals).to(eng.device)
    exp = O.groupby_agg_fast(keys, [vals], [None])
    for buckets, merge in ((1, "table"), (16, "table"), (None, "smem")):
        gt = D.shuffle_groupby_sum(eng, dk, dv, 1 << 18, groups_hint=groups, buckets=buckets, merge=merge)
        res = eng.groupby_result(gt, sort=True)
        assert res["keys"].cpu().numpy().tolist() == exp["keys"].tolist()
        assert res["cols"][0]["sum"].cpu().numpy().tobytes() == exp["cols"][0]["sum"].tobytes()
    assert D.choose_buckets(1 << 28, 1) == 256 and D.choose_buckets(1 << 25, 8) == 32 and D.choose_buckets(1 << 10, 8) == 1


# ---- shared-memory merge (csrc/merge.cu): partition levels + per-piece tables + spill list
@pytest.mark.parametrize("n,groups,plan,slots", [
    (0, 10, None, None),                    # nothing to do
    (1, 1, None, None),
    (300_007, 1000, None, None),            # one piece
    (300_007, 100_000, None, None),         # one level (66 pieces)
    (1_000_003, 600_000, None, None),       # two levels (20 x 20 pieces)
    (400_000, 200_000, (3, 5), 4096),       # far too few pieces: most keys spill to the global-table path
    (400_000, 50_000, (7, 1), 64),          # tiny tables: nearly everything spills
    (500_000, 70_000, (256, 256), 256),     # the finest cut: 65536 pieces, most of them tiny or empty
])
def test_smem_merge_matches_oracle(eng, n, groups, plan, slots):
    keys, _ = S.host_column(21, 0, S.KIND_KEY64, n, modulus=groups)
    keys = keys - groups // 2                                   # negative keys too
    vals, _ = S.host_column(21, 1, S.KIND_EXACT, n)
    dk, dv = torch.from_numpy(keys).to(eng.device), torch.from_numpy(vals).to(eng.device)
    pg = eng.groupby_partitioned_sum(dk, dv, max(n, 1), groups_hint=groups, plan=plan, slots=slots)
    res = eng.groupby_result(pg, sort=True)
    exp = O.groupby_agg_fast(keys, [vals], [None])
    _check(res, exp, exact_sums=True, minmax=False)
    if plan in ((3, 5), (7, 1)):
        assert res["spilled_records"] > n // 2
    elif plan is None:
        assert res["spilled_records"] == 0


def test_smem_merge_skew_real_values_and_strided_records(eng):
    # 90% of the rows on one key (one piece takes nearly everything), real-valued sums, and the
    # {key, value} record layout of the exchange as input
    n, groups = 700_001, 40_000
    keys, _ = S.host_column(22, 0, S.KIND_KEY64, n, modulus=groups)
    hot, _ = S.host_column(22, 2, S.KIND_KEY64, n, modulus=10)
    keys = np.where(hot != 0, np.int64(123456789012345), keys)
    vals, _ = S.host_column(22, 1, S.KIND_REAL, n)
    rec = torch.from_numpy(np.stack([keys, vals.view(np.int64)], axis=1)).to(eng.device)
    flat = rec.view(-1)
    pg = eng.groupby_partitioned_sum(flat, flat[1:].view(torch.float64), n, groups_hint=groups, n=n, key_stride=2, val_stride=2)
    res = eng.groupby_result(pg, sort=True)
    _check(res, O.groupby_agg_fast(keys, [vals], [None]), exact_sums=False, minmax=False)


def test_smem_merge_reports_reserved_key_and_small_output(eng):
    keys = np.arange(1000, dtype=np.int64)
    vals = np.ones(1000)
    dk, dv = torch.from_numpy(keys).to(eng.device), torch.from_numpy(vals).to(eng.device)
    with pytest.raises(pkg.EtlError, match="exceed the output capacity"):
        eng.groupby_result(eng.groupby_partitioned_sum(dk, dv, 999, groups_hint=1000), sort=True)
    keys[17] = np.iinfo(np.int64).min
    with pytest.raises(pkg.EtlError, match="reserved key"):
        eng.groupby_result(eng.groupby_partitioned_sum(torch.from_numpy(keys).to(eng.device), dv, 1000, groups_hint=1000), sort=True)
    assert pkg.Engine.merge_plan(10**9, 10**8) == (256, 255) and pkg.Engine.merge_plan(1000, 0) == (1, 1)
    assert pkg.Engine.merge_plan(125_000_000, 14_000_000) == (96, 95)


def test_hash_partition_unaligned_and_strided_inputs_take_the_register_staged_kernel(eng):
    # bulk async copies need 16-byte aligned sources: a column view that starts 8 bytes into an
    # allocation, and a stride-3 layout, must give the same partition through the other kernel
    n, ranks, buckets = 100_001, 2, 4
    keys, _ = S.host_column(5, 0, S.KIND_KEY64, n + 1, modulus=10**6)
    vals, _ = S.host_column(5, 1, S.KIND_REAL, n + 1)
    dk, dv = torch.from_numpy(keys).to(eng.device), torch.from_numpy(vals).to(eng.device)
    wide = torch.zeros(3 * n, dtype=torch.int64, device=eng.device)
    wide[0::3], wide[1::3] = dk[1:], dv[1:].view(torch.int64)
    dest = O.partition_dest(keys[1:], ranks, buckets)
    want_counts = np.bincount(dest, minlength=ranks * buckets).tolist()
    for k_in, v_in, ks, vs in ((dk[1:], dv[1:], 1, 1), (wide, wide[1:].view(torch.float64), 3, 3)):
        rec, counts, base = eng.hash_partition(k_in, v_in, ranks, buckets, n=n, key_stride=ks, val_stride=vs)
        assert counts.cpu().tolist() == want_counts
        rec = rec.cpu().numpy()
        b = base.cpu().tolist() + [n]
        for d in range(ranks * buckets):
            got = rec[b[d]:b[d + 1]]
            want = np.stack([keys[1:][dest == d], vals[1:][dest == d].view(np.int64)], axis=1)
            assert np.array_equal(got[np.lexsort((got[:, 1], got[:, 0]))], want[np.lexsort((want[:, 1], want[:, 0]))])
