"""BASELINE.json configs at (or near) full size: size-independent properties -
conservation of rows and of the grand total, sortedness / uniqueness of keys,
idempotence - plus a per-group diff against the oracle on a sample of the keys
(every row of ~1 % of the keys is pulled to the host and aggregated there), as
the oracle cannot finish the whole tables in seconds."""
import numpy as np
import pytest
import torch

from oracle import etl_oracle as O

import pyspark_tf_gke_b200 as pkg
import pyspark_tf_gke_b200.dist as D
import pyspark_tf_gke_b200.synth as S

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    e = pkg.Engine(0)
    yield e
    e.close()


def test_config3_groupby_100m_rows_1k_groups(eng):
    cfg = S.CONFIG3
    n, G = cfg["rows"], cfg["groups"]
    key = eng.synth_column(cfg["seed"], 0, S.KIND_KEY32, n, modulus=G, with_valid=False)
    cols = [eng.synth_column(cfg["seed"], 1 + c, S.KIND_EXACT, n, null_rate=cfg["val_null"][c]) for c in range(3)]
    results = []
    for key_range in (G, 0):  # direct-indexed kernel, then the hashed shared-memory kernel
        gt = eng.groupby_new(4096, 3)
        eng.groupby_accumulate(gt, key.data, None, [c.data for c in cols], [c.valid for c in cols], groups_hint=G,
                               key_range=key_range)
        results.append(eng.groupby_result(gt, sort=True))
    sums, counts = eng.sum_count(cols)
    for res in results:
        assert res["keys"].tolist() == list(range(G))           # every key present, once, sorted
        assert int(res["rows"].sum()) == n and res["null_group_rows"] == 0
        for c in range(3):
            assert int(res["cols"][c]["count"].sum()) == int(counts[c])
            assert float(res["cols"][c]["sum"].sum()) == float(sums[c])   # integer-valued data: exact
    # per-group diff against the oracle on 1 % of the keys (every row of keys 0..9)
    idx = torch.nonzero(key.data < 10).flatten()
    hk = key.data[idx].cpu().numpy()
    hv = [c.data[idx].cpu().numpy() for c in cols]
    hvv = [(((c.valid[idx >> 5] >> (idx & 31).to(torch.int32)) & 1) != 0).cpu().numpy() for c in cols]
    want = O.groupby_agg_fast(hk, hv, hvv)
    for res in results:
        assert res["rows"][:10].tolist() == want["rows"].tolist()
        for c in range(3):
            assert res["cols"][c]["count"][:10].tolist() == want["cols"][c]["count"].tolist()
            assert res["cols"][c]["sum"][:10].cpu().numpy().tobytes() == want["cols"][c]["sum"].tobytes()
    for c in range(3):  # the two kernels agree group by group
        assert torch.equal(results[0]["cols"][c]["count"], results[1]["cols"][c]["count"])
        assert torch.equal(results[0]["cols"][c]["sum"], results[1]["cols"][c]["sum"])
    assert torch.equal(results[0]["rows"], results[1]["rows"])


def test_config4_shape_groupby_quarter_size_single_gpu(eng):
    cfg = S.CONFIG4
    n, G = cfg["rows"] // 4, cfg["groups"] // 4
    key = eng.synth_column(cfg["seed"], 0, S.KIND_KEY64, n, modulus=G, with_valid=False).data
    val = eng.synth_column(cfg["seed"], 1, S.KIND_EXACT, n, with_valid=False).data
    cap = 1 << 26
    direct = eng.groupby_new(cap, 1)
    eng.groupby_accumulate(direct, key, None, [val], groups_hint=G)
    pieces = D.shuffle_groupby_sum(eng, key, val, cap, groups_hint=G)     # two partition levels + shared-memory merge
    b = eng.groupby_result(pieces, sort=True)
    assert b["spilled_records"] == 0
    del pieces
    bucketed = D.shuffle_groupby_sum(eng, key, val, cap, groups_hint=G, merge="table")   # bucket-by-bucket table merge
    a, c = eng.groupby_result(direct, max_groups=G, sort=True), eng.groupby_result(bucketed, max_groups=G, sort=True)
    assert torch.equal(a["keys"], c["keys"]) and torch.equal(a["rows"], c["rows"])
    assert torch.equal(a["cols"][0]["sum"], c["cols"][0]["sum"])
    del c, bucketed
    s, _ = eng.sum_count([pkg.Column(val)])
    for res in (a, b):
        k = res["keys"]
        assert bool((k[1:] > k[:-1]).all()) and int(k[0]) >= 0 and int(k[-1]) < G   # unique, in range
        assert int(res["rows"].sum()) == n
        assert float(res["cols"][0]["sum"].sum()) == float(s[0])
    assert torch.equal(a["keys"], b["keys"]) and torch.equal(a["rows"], b["rows"])
    assert torch.equal(a["cols"][0]["sum"], b["cols"][0]["sum"])


def test_config4_full_size_single_gpu(eng):
    # BASELINE configs[3] on one GPU at its full size: 1 B rows, 100 M groups through the two partition
    # levels and the shared-memory merge; conservation of rows, of the grand total and of the key set
    cfg = S.CONFIG4
    n, G = cfg["rows"], cfg["groups"]
    key = eng.synth_column(cfg["seed"], 0, S.KIND_KEY64, n, modulus=G, with_valid=False).data
    val = eng.synth_column(cfg["seed"], 1, S.KIND_EXACT, n, with_valid=False).data
    pieces = D.shuffle_groupby_sum(eng, key, val, G + G // 8, groups_hint=G)
    res = eng.groupby_result(pieces, sort=True)
    del pieces
    s, _ = eng.sum_count([pkg.Column(val)])
    k = res["keys"]
    assert res["spilled_records"] == 0
    assert bool((k[1:] > k[:-1]).all()) and int(k[0]) >= 0 and int(k[-1]) < G          # unique, sorted, in range
    assert G - k.shape[0] < G // 1000                                                   # ~e^-10 of the keys never drawn
    assert int(res["rows"].sum()) == n and int(res["cols"][0]["count"].sum()) == n
    assert float(res["cols"][0]["sum"].sum()) == float(s[0])                            # integer-valued data: exact
    # a key's row count is what a direct count of that key gives (spot check on the busiest group)
    top = int(torch.argmax(res["rows"]))
    assert int((key == k[top]).sum()) == int(res["rows"][top])
    # per-group diff against the oracle on 1 % of the keys: every row whose key is below G / 100
    sel = key < G // 100
    want = O.groupby_agg_fast(key[sel].cpu().numpy(), [val[sel].cpu().numpy()], [None])
    m = int((k < G // 100).sum())
    assert k[:m].cpu().numpy().tolist() == want["keys"].tolist()
    assert res["rows"][:m].cpu().numpy().tolist() == want["rows"].tolist()
    assert res["cols"][0]["sum"][:m].cpu().numpy().tobytes() == want["cols"][0]["sum"].tobytes()


def test_config5_shape_csv_decode_10m_records(eng):
    # the default input repeated to ~10 M records (1.4 GB of text): every statistic of the stage must be
    # the default input's, scaled - label table, null counts, means
    import gzip
    import os

    from tests.conftest import ROOT
    with gzip.open(os.path.join(ROOT, "tests", "golden", "health_workloads.csv.gz"), "rb") as fh:
        raw = fh.read()
    head, body = raw.split(b"\n", 1)
    reps = 550
    numeric, strings = ["value", "lower_ci", "upper_ci"], ["measure_name"]
    small = pkg.read_csv(eng, raw, numeric, strings)
    big = pkg.read_csv(eng, head + b"\n" + body * reps, numeric, strings)
    assert big.rows == small.rows * reps == 18155 * reps
    assert big.strings["measure_name"].labels == small.strings["measure_name"].labels
    hist_small = torch.bincount(small.strings["measure_name"].column.data.to(torch.int64), minlength=30)
    hist_big = torch.bincount(big.strings["measure_name"].column.data.to(torch.int64), minlength=30)
    assert torch.equal(hist_big, hist_small * reps)
    for c in numeric:
        (s1, c1), (s2, c2) = eng.sum_count([small.numeric[c]]), eng.sum_count([big.numeric[c]])
        assert int(c2[0]) == int(c1[0]) * reps
        assert abs(float(s2[0]) / float(c2[0]) - float(s1[0]) / float(c1[0])) <= 1e-9 * abs(float(s1[0]) / float(c1[0]))
        # the first and the last copy of the block decode to the same bits as the single one
        ok = torch.from_numpy(small.numeric[c].valid_host()).to(eng.device)
        want = small.numeric[c].data.view(torch.int64)[ok]
        assert torch.equal(big.numeric[c].data[:small.rows].view(torch.int64)[ok], want)
        assert torch.equal(big.numeric[c].data[-small.rows:].view(torch.int64)[ok], want)
