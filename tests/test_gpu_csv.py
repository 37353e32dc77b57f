"""Config 1 (BASELINE.json configs[0]): the default input end to end, plus the
CSV decoder's edge cases - CUDA path against the golden vectors captured from
the reference's own load_csv and against the oracle's restatement of the Spark
stage.  Everything is compared bit for bit (integer / byte / index work, and
fp64 values that are parsed, copied or imputed; the means are exact on this
input because every sum is an integer < 2^53)."""
import json
import os

import numpy as np
import pytest
import torch

import pyspark_tf_gke_b200 as pkg
from oracle import etl_oracle as O
from tests.conftest import GOLDEN, health_text

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    e = pkg.Engine(0)
    yield e
    e.close()


@pytest.fixture(autouse=True)
def _etl_stage_only(monkeypatch):
    monkeypatch.setenv("ETL_SKIP_KMEANS", "1")  # these tests cover the ETL stage; KMeans has its own file


def _cases():
    with open(os.path.join(GOLDEN, "load_csv_cases.json")) as fh:
        return json.load(fh)


@pytest.mark.parametrize("which", ["workloads", "infra"])
def test_load_csv_health_matches_reference_golden(eng, which):
    g = np.load(os.path.join(GOLDEN, "load_csv_health.npz"))
    X, y, vocab = pkg.load_csv(health_text(which).encode())
    assert X.dtype == np.float32 and y.dtype == np.int32 and X.flags["C_CONTIGUOUS"]
    assert X.shape == (15355, 3) and y.shape == (15355,)
    assert X.tobytes() == g["X"].tobytes()
    assert y.tobytes() == g["y"].tobytes()
    assert vocab == [str(s) for s in g["vocab"]]
    assert np.bincount(y).tolist() == [1133, 1015, 1225, 1040, 1196, 1040, 1301, 1040, 1196, 521, 1075, 475, 707, 1040, 1351]


@pytest.mark.parametrize("name", sorted(_cases()))
def test_load_csv_cases_match_reference_golden(eng, name):
    case = _cases()[name]
    raw = case["text"].encode()
    if "raises" in case:
        with pytest.raises(RuntimeError, match="No valid rows"):
            pkg.load_csv(raw)
        return
    X, y, vocab = pkg.load_csv(raw)
    assert [[float(v).hex() for v in row] for row in X] == case["X_hex"]
    assert y.tolist() == case["y"]
    assert vocab == case["vocab"]


def test_load_csv_numerals_across_the_double_range(eng):
    text = ("subpopulation,value,lower_ci,upper_ci\n"
            "A,123456789012345,0.1,1e-5\nB,16777217,9007199254740993,1e22\n"
            "C,0.30000000000000004,12345678901234567890,-0\nD,1234567890123456789e-19,.5,5.\n"
            "E,3.4e38,1e23,1.7976931348623157e308\nF,4.9e-324,2.2250738585072011e-308,1e-400\nG,1e400,-1e309,6.02214076e23\n"
            "H,3.14159265358979323846264338327950288,0.12345678901234567891,9007199254740992.999999999999999999999\n")
    X, y, vocab = pkg.load_csv(text.encode())
    eX, ey, ev = O.load_csv(text)
    assert X.tobytes() == eX.tobytes() and y.tobytes() == ey.tobytes() and vocab == ev
    t = pkg.read_csv(eng, text.encode(), ["value", "lower_ci", "upper_ci"], ["subpopulation"])
    header, cols = O.read_csv_columns(text)
    for c in ("value", "lower_ci", "upper_ci"):   # fp64 before the float32 cast: bit for bit
        v, vv = O.to_numeric_column(cols[c])
        assert t.numeric[c].data.cpu().numpy().tobytes() == v.tobytes()
    # more than 60 significant digits with a non-zero tail stay outside the exactly-converted domain: loud failure
    with pytest.raises(pkg.EtlError, match="outside the exactly-converted domain"):
        pkg.load_csv(("subpopulation,value,lower_ci,upper_ci\nA,0.%s,1,2\n" % ("7" * 70)).encode())


def test_load_csv_other_columns_and_label_map(eng, tmp_path):
    text = health_text("workloads")
    X, y, vocab = pkg.load_csv(text.encode(), ["upper_ci", "value", "lower_ci"], "state_name")
    eX, ey, ev = O.load_csv(text, ["upper_ci", "value", "lower_ci"], "state_name")
    assert X.tobytes() == eX.tobytes() and y.tobytes() == ey.tobytes() and vocab == ev and len(vocab) == 52
    path = tmp_path / "label_map.json"
    pkg.write_label_map(vocab, str(path))
    with open(path) as fh:
        raw = fh.read()
    assert json.loads(raw) == {str(i): s for i, s in enumerate(vocab)}
    assert raw.startswith('{\n  "0": "Alabama"')  # indent=2, int keys (train_tf_ps.py:582-583)


@pytest.mark.parametrize("which", ["workloads", "infra"])
@pytest.mark.parametrize("repeats", [5, 1])
def test_k_means_etl_matches_oracle(eng, which, repeats, monkeypatch):
    text = health_text(which)
    monkeypatch.setenv("MEASURE_NAME_WEIGHT", str(repeats))
    table = pkg.read_csv(eng, text.encode(), ["value", "lower_ci", "upper_ci"], ["measure_name"])
    assert table.rows == 18155
    assert table.inferred_types["value"] == ("int" if which == "workloads" else "double")
    model, kmeans = pkg.KMeansWorkload(eng).k_means(table)
    exp = O.k_means_etl(text, repeats=repeats)
    assert model.null_key_count == exp["null_measure_name"] == 0
    assert model.rows_after_filter == exp["rows_after_filter"] == 18155
    assert [float(m).hex() for m in model.means] == [float(exp["means"][c]).hex() for c in ("value", "lower_ci", "upper_ci")]
    assert model.means == [1738.2895095771808, 1724.1278848441589, 1762.6607185343803]
    assert model.labels == exp["labels"] and model.label_counts == exp["label_counts"]
    assert model.index.cpu().numpy().tolist() == exp["index"].astype(np.int32).tolist()
    for a, b in zip(model.numerics, exp["numerics"]):
        assert a.cpu().numpy().tobytes() == b.tobytes()
    F = model.features().cpu().numpy()
    assert F.shape == (18155, repeats * 30 + 3) == exp["features"].shape
    assert F.tobytes() == exp["features"].tobytes()
    F32 = model.features(torch.float32).cpu().numpy()
    assert F32.tobytes() == exp["features"].astype(np.float32).tobytes()
    # the SparseVector view
    oi, ov = (t.cpu().numpy() for t in model.features_sparse())
    sparse = O.assemble_features_sparse(exp["index"], 30, exp["numerics"], repeats)
    for i in (0, 1, 17, 9000, 18154):
        ind = [int(k) for k, v in zip(oi[i], ov[i]) if k >= 0 and v != 0.0]
        val = [float(v) for k, v in zip(oi[i], ov[i]) if k >= 0 and v != 0.0]
        assert (ind, val) == sparse[i]


def test_k_means_logs_and_reraises(eng, caplog):
    wl = pkg.KMeansWorkload(eng)
    table = pkg.read_csv(eng, b"measure_name,value\nA,1\n", ["value", "lower_ci", "upper_ci"], ["measure_name"])
    del table.numeric["lower_ci"]
    with caplog.at_level("ERROR"):
        with pytest.raises(KeyError):
            wl.k_means(table)
    assert "Error training K-Means model" in caplog.text


def test_unseen_and_null_keys_go_to_the_unknown_index(eng):
    train = b"measure_name,value,lower_ci,upper_ci\nb,1,2,3\na,4,5,6\nb,7,,9\n,1,1,1\n"
    table = pkg.read_csv(eng, train, ["value", "lower_ci", "upper_ci"], ["measure_name"])
    model, _ = pkg.KMeansWorkload(eng).k_means(table)
    assert model.null_key_count == 1 and model.rows_after_filter == 3
    assert model.labels == ["b", "a"] and model.label_counts == [2, 1]
    assert model.means == [4.0, 3.5, 6.0]
    assert model.numerics[1].cpu().tolist() == [2.0, 5.0, 3.5]
    other = pkg.read_csv(eng, b"measure_name,value,lower_ci,upper_ci\nzzz,1,2,3\na,,5,6\n", ["value", "lower_ci", "upper_ci"],
                         ["measure_name"])
    index, nums = model.transform(other)
    assert index.cpu().tolist() == [2, 1]          # unseen -> numLabels (handleInvalid="keep")
    assert nums[0].cpu().tolist() == [1.0, 4.0]    # imputed with the training mean
    F = model.features(index=index, numerics=nums).cpu().numpy()
    assert F.shape == (2, 5 * 2 + 3) and F[0, :10].sum() == 0 and F[1, :10].tolist() == [0, 1] * 5


def test_dense_features_wide_and_edge_dimensions(eng):
    rng = np.random.default_rng(3)
    n = 5003
    nums = [rng.random(n) * 100 for _ in range(3)]
    dn = [torch.from_numpy(v).to(eng.device) for v in nums]
    for L, repeats in [(1, 1), (29, 1), (30, 5), (84, 3), (85, 3), (300, 2), (7, 40)]:
        idx = rng.integers(0, L + 1, n).astype(np.int32)  # L itself = unknown label
        F = eng.assemble_dense(torch.from_numpy(idx).to(eng.device), dn, L, repeats).cpu().numpy()
        exp = O.assemble_features_dense(idx.astype(np.float64), L, nums, repeats)
        assert F.shape == exp.shape and F.tobytes() == exp.tobytes(), (L, repeats)


def _synthetic_csv(rows, seed):
    rng = np.random.default_rng(seed)
    names = ["alpha", "be,ta", 'ga"mma', "delta\nline", "  eps  ", ""]
    lines = ["id,measure_name,filler,value,lower_ci,upper_ci"]
    for i in range(rows):
        k = names[rng.integers(len(names))]
        if any(ch in k for ch in ',"\n') or rng.random() < 0.1:
            k = '"' + k.replace('"', '""') + '"'
        filler = "x" * int(rng.integers(0, 400)) if rng.random() < 0.05 else "f"
        cells = []
        for _ in range(3):
            u = rng.random()
            cells.append("" if u < 0.07 else ("NaN" if u < 0.09 else (str(int(rng.integers(0, 100000))) if u < 0.6
                                                                       else repr(float(np.round(rng.random() * 1e5, 3))))))
        eol = "\r\n" if rng.random() < 0.1 else "\n"
        lines.append(f"{i},{k},{filler},{','.join(cells)}" + ("\n" if rng.random() < 0.02 else "") + eol.rstrip("\n"))
    return "\n".join(lines) + "\n"


@pytest.mark.parametrize("rows,seed", [(1, 0), (37, 1), (5000, 2), (120_000, 3)])
def test_decoder_on_synthetic_csv_with_quotes_newlines_and_nulls(eng, rows, seed):
    text = _synthetic_csv(rows, seed)
    table = pkg.read_csv(eng, text.encode(), ["value", "lower_ci", "upper_ci"], ["measure_name"])
    header, cols = O.read_csv_columns(text)
    assert table.rows == len(cols["value"])
    for c in ("value", "lower_ci", "upper_ci"):
        v, vv = O.to_numeric_column(cols[c])
        got = table.numeric[c]
        assert (got.valid_host() == vv).all()
        g = got.data.cpu().numpy()
        assert g[vv].tobytes() == v[vv].tobytes()  # NaN text -> NaN value, same bits
    key = table.strings["measure_name"]
    exp_keys = cols["measure_name"]
    kv = key.column.valid_host()
    assert (kv == np.array([k is not None for k in exp_keys])).all()
    codes = key.column.data.cpu().numpy()
    assert [key.labels[c] if ok else None for c, ok in zip(codes, kv)] == exp_keys
    # and the TF-side reader on the same text
    try:
        eX, ey, ev = O.load_csv(text, label_col="measure_name")
    except RuntimeError:
        with pytest.raises(RuntimeError):
            pkg.load_csv(text.encode(), label_col="measure_name")
        return
    X, y, vocab = pkg.load_csv(text.encode(), label_col="measure_name")
    assert X.tobytes() == eX.tobytes() and y.tobytes() == ey.tobytes() and vocab == ev


def test_quoted_empty_cell_is_null_in_the_decoder_too(eng):
    # Spark's defaults (nullValue = "" and read-side emptyValue = "") make `""` and an empty field the same
    # null (oracle.read_csv_columns states the rule): not an empty-string label, not a zero
    text = 'measure_name,value,lower_ci,upper_ci\n"",1,2,3\n,4,5,6\nx,"",8,9\n"x",10,"",12\n" ",1,1,1\n'
    table = pkg.read_csv(eng, text.encode(), ["value", "lower_ci", "upper_ci"], ["measure_name"])
    header, cols = O.read_csv_columns(text)
    key = table.strings["measure_name"]
    kv = key.column.valid_host()
    codes = key.column.data.cpu().numpy()
    assert [key.labels[c] if ok else None for c, ok in zip(codes, kv)] == cols["measure_name"] == [None, None, "x", "x", " "]
    for c in ("value", "lower_ci", "upper_ci"):
        v, vv = O.to_numeric_column(cols[c])
        assert (table.numeric[c].valid_host() == vv).all()
        assert table.numeric[c].data.cpu().numpy()[vv].tobytes() == v[vv].tobytes()


def test_row_index_across_block_boundaries(eng):
    # quoted fields with newlines and commas straddling the 16 KiB scan blocks
    parts = ["a,b"]
    for i in range(300):
        parts.append(f'{i},"' + ("line\n" * (i % 50)) + ',"",' + "z" * (i * 7 % 997) + '"')
        if i % 17 == 0:
            parts.append("")  # blank line
    text = "\r\n".join(parts)
    raw = text.encode()
    t = pkg.table.upload_text(eng, raw)
    rows = int(eng.csv_index_rows(t, len(raw)).item())
    import csv, io
    exp = [r for r in csv.reader(io.StringIO(text, newline="")) if r]
    assert rows == len(exp) == 301
    starts = eng.csv_row_offsets(t, len(raw), rows).cpu().numpy()
    assert starts[0] == 0
    for r in (1, 2, 51, 299, 300):
        assert raw[starts[r]:].decode().startswith(f"{r - 1},\"")


def test_empty_and_header_only_inputs(eng):
    assert pkg.read_csv(eng, b"", ["value"], ["measure_name"]).rows == 0
    t = pkg.read_csv(eng, b"measure_name,value\n", ["value"], ["measure_name"])
    assert t.rows == 0
    t = pkg.read_csv(eng, b"\n\n\r\n", ["value"], ["measure_name"])
    assert t.rows == 0


def test_bucketize_standardise_and_histogram(eng):
    rng = np.random.default_rng(5)
    n = 100_001
    x = rng.random(n) * 10
    x[::97] = np.nan
    valid = rng.random(n) > 0.05
    col = pkg.Column.from_host(x, valid, eng.device)
    splits = [0.0, 1.0, 2.5, 5.0, 10.0]
    out, oor = eng.bucketize(col, splits)
    exp = O.bucketize(np.where(valid, x, np.nan), splits)
    assert out.cpu().numpy().tolist() == exp.astype(np.int32).tolist() and int(oor.item()) == 0
    out, oor = eng.bucketize(pkg.Column.from_host(np.array([-1.0, 0.0, 10.0, 11.0]), None, eng.device), splits)
    assert out.cpu().tolist() == [-1, 0, 3, -1] and int(oor.item()) == 2
    # histogram: shared-memory path and the global path (bins > 8192)
    for bins in (30, 20_000):
        codes = rng.integers(0, bins, n).astype(np.int32)
        counts, bad = eng.code_histogram(pkg.Column.from_host(codes, valid, eng.device), bins)
        assert counts.cpu().numpy().tolist() == np.bincount(codes[valid], minlength=bins).tolist() and int(bad.item()) == 0
    # standardised numerics in the TF layout: (x - mean) * (1 / std), float32
    nums = [rng.random(n) * 100 for _ in range(3)]
    stats = [O.column_stats(v) for v in nums]
    lab = pkg.Column.from_host(rng.integers(0, 7, n).astype(np.int32), valid, eng.device)
    cols = [pkg.Column.from_host(v, None, eng.device) for v in nums]
    X, y, rows = eng.pack_tf(lab, cols, shift=[m for m, _ in stats], scale=[1.0 / s for _, s in stats])
    r = int(rows.item())
    assert r == int(valid.sum())
    expX = np.stack([O.standardize(v[valid], m, s, with_mean=True) for v, (m, s) in zip(nums, stats)], axis=1)
    np.testing.assert_allclose(X[:r].cpu().numpy(), expX.astype(np.float32), rtol=1e-6, atol=0)
    assert y[:r].cpu().numpy().tolist() == lab.data.cpu().numpy()[valid].tolist()


def test_nan_to_null_then_pack(eng):
    n = 5000
    rng = np.random.default_rng(9)
    nums = [rng.random(n) for _ in range(3)]
    nums[1][::13] = np.nan
    lab = pkg.Column.from_host(rng.integers(0, 3, n).astype(np.int32), None, eng.device)
    cols = [eng.nan_to_null(pkg.Column.from_host(v, None, eng.device)) for v in nums]
    X, y, rows = eng.pack_tf(lab, cols)
    keep = ~np.isnan(nums[1])
    assert int(rows.item()) == int(keep.sum())
    assert X[:int(rows.item())].cpu().numpy().tobytes() == np.stack([v[keep] for v in nums], axis=1).astype(np.float32).tobytes()


def test_read_csv_from_a_path_and_from_host_columns(eng, tmp_path):
    text = health_text("workloads")
    path = tmp_path / "health.csv"
    path.write_text(text)
    cols = (["value", "lower_ci", "upper_ci"], ["measure_name"])
    a, _ = pkg.KMeansWorkload(eng).k_means(pkg.read_csv(eng, str(path), *cols))      # pinned-file path
    b, _ = pkg.KMeansWorkload(eng).k_means(pkg.read_csv(eng, text.encode(), *cols))  # bytes path
    assert a.labels == b.labels and a.means == b.means
    assert a.features().cpu().numpy().tobytes() == b.features().cpu().numpy().tobytes()
    X1, y1, v1 = pkg.load_csv(str(path))
    X2, y2, v2 = pkg.load_csv(text.encode())
    assert X1.tobytes() == X2.tobytes() and y1.tobytes() == y2.tobytes() and v1 == v2
    # the JDBC shape (google_health_SQL.py:26-37): already-decoded host columns, FLOAT numerics with SQL NULLs
    header, c = O.read_csv_columns(text)
    nums = {}
    for name in cols[0]:
        v, vv = O.to_numeric_column(c[name])
        nums[name] = np.ma.masked_array(v.astype(np.float32), mask=~vv)
    t = pkg.table.from_host_columns(eng, c["measure_name"], nums)
    m, _ = pkg.KMeansWorkload(eng).k_means(t)
    assert m.labels == a.labels and m.label_counts == a.label_counts
    assert m.means == a.means  # every value is an integer < 2^24: exact in float32 too
    assert m.features().cpu().numpy().tobytes() == a.features().cpu().numpy().tobytes()


@pytest.mark.parametrize("rows,seed", [(37, 11), (60_000, 12)])
def test_row_scan_over_index_masks_equals_row_scan_over_text(eng, rows, seed, monkeypatch):
    # read_csv lets the field parse walk the quote / comma / newline masks the index pass kept; without
    # the flag the same kernel scans the text bytes - both must decode every cell identically
    import pyspark_tf_gke_b200.table as T
    text = _synthetic_csv(rows, seed).encode()
    cols_n, cols_s = ["value", "lower_ci", "upper_ci"], ["measure_name", "filler"]
    with_masks = pkg.read_csv(eng, text, cols_n, cols_s)
    monkeypatch.setattr(T, "CSV_USE_INDEX_MASKS", 0)
    from_text = pkg.read_csv(eng, text, cols_n, cols_s)
    assert with_masks.rows == from_text.rows
    for c in cols_n:
        a, b = with_masks.numeric[c], from_text.numeric[c]
        assert torch.equal(a.valid, b.valid) and a.data.cpu().numpy().tobytes() == b.data.cpu().numpy().tobytes()
    for c in cols_s:
        a, b = with_masks.strings[c], from_text.strings[c]
        assert torch.equal(a.column.valid, b.column.valid)
        la = [a.labels[i] if i >= 0 else None for i in a.column.data.cpu().tolist()]
        lb = [b.labels[i] if i >= 0 else None for i in b.column.data.cpu().tolist()]
        assert la == lb
    # any other use of the handle's arena drops the masks: the flag is then ignored, not trusted
    n = len(text)
    d_text = T.upload_text(eng, text)
    total = int(eng.csv_index_rows(d_text, n).item())
    row_start = eng.csv_row_offsets(d_text, n, total)
    eng.reserve(1 << 20)
    cols, status, _ = eng.csv_parse_columns(d_text, n, row_start, 1, total - 1, [3], [0], [None], flags=T.CSV_NAN_IS_NULL | 4)
    assert cols[0].data.cpu().numpy().tobytes() == from_text.numeric["value"].data.cpu().numpy().tobytes()


def test_pinned_ring_upload_of_bytes_and_files(eng, tmp_path):
    import pyspark_tf_gke_b200.table as T

    # the ingest path: bytes / a file cut into slot-sized slices through the engine's pinned ring; small slots so
    # that a few MB exercise slot reuse, the tail slice and the zero padding
    from pyspark_tf_gke_b200.table import PinnedRing, _padded

    rng = np.random.default_rng(3)
    raw = rng.integers(0, 256, 5_000_123, dtype=np.uint8).tobytes()
    eng._pinned_ring = PinnedRing(eng.device, slots=3, slot_bytes=1 << 20)
    try:
        for data in (raw, raw[:1 << 20], raw[:17], b""):
            d = T.upload_text(eng, data)
            assert d.shape[0] == _padded(len(data))
            host = d.cpu().numpy()
            assert host[:len(data)].tobytes() == data and not host[len(data):].any()
        path = tmp_path / "blob.bin"
        path.write_bytes(raw)
        d, n, head = T.upload_file(eng, str(path))
        assert n == len(raw) and d.cpu().numpy()[:n].tobytes() == raw and head == raw[:1 << 20]
    finally:
        del eng._pinned_ring


def test_byte_range_split_of_a_file_covers_every_record_once(eng, tmp_path):
    # config 5's input split: each "rank" reads, uploads and decodes only its byte range of the file (cut at
    # newlines, the header line read by everyone); together the ranges hold every record exactly once, in order
    import pyspark_tf_gke_b200.table as T

    text = health_text("workloads")
    path = tmp_path / "health.csv"
    path.write_text(text)
    whole = pkg.read_csv(eng, str(path), ["value", "lower_ci", "upper_ci"], ["measure_name"])
    for world in (1, 3, 8):
        parts = [pkg.read_csv(eng, str(path), ["value", "lower_ci", "upper_ci"], ["measure_name"], byte_range=(r, world))
                 for r in range(world)]
        assert sum(p.rows for p in parts) == whole.rows == 18155
        ranges = [T.byte_range_of(str(path), r, world) for r in range(world)]
        assert ranges[0][1] == len(ranges[0][0]) and ranges[-1][2] == len(text.encode())
        assert all(a[2] == b[1] for a, b in zip(ranges[:-1], ranges[1:]))
        for c in ("value", "lower_ci", "upper_ci"):
            got = np.concatenate([p.numeric[c].data.cpu().numpy() for p in parts])
            ok = np.concatenate([p.numeric[c].valid_host() for p in parts])
            assert (ok == whole.numeric[c].valid_host()).all()
            assert got[ok].tobytes() == whole.numeric[c].data.cpu().numpy()[ok].tobytes()
        labels = np.concatenate([np.array(p.strings["measure_name"].labels, dtype=object)[p.strings["measure_name"].column.data.cpu().numpy()]
                                 for p in parts if p.rows])
        want = np.array(whole.strings["measure_name"].labels, dtype=object)[whole.strings["measure_name"].column.data.cpu().numpy()]
        assert labels.tolist() == want.tolist()
    # more ranks than records: the surplus ranks get empty tables
    tiny = tmp_path / "tiny.csv"
    tiny.write_text("measure_name,value,lower_ci,upper_ci\nx,1,2,3\n")
    rows = [pkg.read_csv(eng, str(tiny), ["value"], ["measure_name"], byte_range=(r, 4)).rows for r in range(4)]
    assert sum(rows) == 1


def test_read_csv_more_than_eight_columns(eng):
    # the parse kernel takes 8 columns per launch; a wider selection is decoded in groups over one row index
    rng = np.random.default_rng(2)
    names = [f"c{i}" for i in range(13)]
    rows = 3001
    data = rng.integers(-1000, 1000, (rows, 11))
    lines = [",".join(names)]
    for r in range(rows):
        cells = [str(v) if (r + i) % 17 else "" for i, v in enumerate(data[r])]
        cells[3] = f"{data[r][3]}.5"                    # a double column among the integer ones
        lines.append(",".join(cells + [f"lab{r % 5}", f'"q,{r % 3}"']))
    text = "\n".join(lines) + "\n"
    table = pkg.read_csv(eng, text.encode(), names[:11], names[11:])
    header, cols = O.read_csv_columns(text)
    for i, name in enumerate(names[:11]):
        v, vv = O.to_numeric_column(cols[name])
        assert (table.numeric[name].valid_host() == vv).all()
        assert table.numeric[name].data.cpu().numpy()[vv].tobytes() == v[vv].tobytes()
        assert table.inferred_types[name] == ("double" if i == 3 else "int")
    for name in names[11:]:
        sc = table.strings[name]
        codes = sc.column.data.cpu().numpy()
        assert [sc.labels[c] for c in codes] == cols[name]
