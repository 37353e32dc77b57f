"""The oracle against the golden vectors (CPU only).

load_csv is pinned against outputs of the reference's own load_csv
(tests/golden/make_golden.py); the Spark side against the exact-rational known
answers of SURVEY.md §2.3."""
import json
import os

import numpy as np
import pytest

from oracle import etl_oracle as O
from tests.conftest import GOLDEN


def _cases():
    with open(os.path.join(GOLDEN, "load_csv_cases.json")) as fh:
        return json.load(fh)


@pytest.mark.parametrize("which", ["workloads", "infra"])
def test_load_csv_health_matches_reference(which, health_workloads_text, health_infra_text):
    text = health_workloads_text if which == "workloads" else health_infra_text
    g = np.load(os.path.join(GOLDEN, "load_csv_health.npz"))
    X, y, vocab = O.load_csv(text)
    assert X.dtype == np.float32 and y.dtype == np.int32
    assert X.shape == (15355, 3) and X.flags["C_CONTIGUOUS"]
    assert X.tobytes() == g["X"].tobytes()
    assert y.tobytes() == g["y"].tobytes()
    assert vocab == [str(s) for s in g["vocab"]]


@pytest.mark.parametrize("name", sorted(_cases()))
def test_load_csv_cases_match_reference(name):
    case = _cases()[name]
    if "raises" in case:
        with pytest.raises(RuntimeError, match="No valid rows"):
            O.load_csv(case["text"])
        return
    X, y, vocab = O.load_csv(case["text"])
    assert [[float(v).hex() for v in row] for row in X] == case["X_hex"]
    assert y.tolist() == case["y"]
    assert vocab == case["vocab"]


@pytest.mark.parametrize("which", ["workloads", "infra"])
def test_spark_side_known_answers(which, health_workloads_text, health_infra_text):
    text = health_workloads_text if which == "workloads" else health_infra_text
    with open(os.path.join(GOLDEN, "spark_known_answers.json")) as fh:
        ka = json.load(fh)[which]
    r = O.k_means_etl(text, repeats=5)
    assert r["null_measure_name"] == ka["null_measure_name"] == 0
    assert r["rows_after_filter"] == ka["rows_after_filter"] == 18155
    for c in ("value", "lower_ci", "upper_ci"):
        assert float(r["means"][c]).hex() == ka["means"][c]["mean_hex"]
    # SURVEY.md §2.3 literal values
    assert r["means"]["value"] == 1738.2895095771808
    assert r["means"]["lower_ci"] == 1724.1278848441589
    assert r["means"]["upper_ci"] == 1762.6607185343803
    assert r["labels"] == ka["labels"] and r["label_counts"] == ka["label_counts"]
    assert r["labels"][:5] == ["Able-Bodied", "Per Capita Income", "Poverty", "Unemployment", "Uninsured"]
    assert r["labels"][-1] == "Maternal Mortality"
    F = r["features"]
    assert F.shape == (18155, 153)
    # five ones per row in the one-hot blocks, at r*30 + idx
    assert (F[:, :150].sum(axis=1) == 5).all()
    idx = r["index"].astype(int)
    for rep in range(5):
        assert (F[np.arange(len(idx)), rep * 30 + idx] == 1).all()
    assert int((F[:, 150:] == 0).sum()) == 4
    assert r["inferred_types"]["value"] == ("int" if which == "workloads" else "double")
    assert O.k_means_etl(text, repeats=1)["features"].shape == (18155, 33)


def test_sparse_and_dense_agree():
    idx = np.array([0, 2, 3, 1], dtype=np.float64)  # 3 == numLabels -> zero vector
    nums = [np.array([1.0, 0.0, 2.0, 3.0]), np.array([0.0, 0.0, 5.0, 6.0])]
    dense = O.assemble_features_dense(idx, 3, nums, repeats=2)
    sparse = O.assemble_features_sparse(idx, 3, nums, repeats=2)
    for i, (ind, val) in enumerate(sparse):
        row = np.zeros(dense.shape[1])
        row[ind] = val
        assert (row == dense[i]).all() and ind == sorted(ind) and 0.0 not in val


def test_validity_roundtrip():
    rng = np.random.default_rng(0)
    for n in (0, 1, 31, 32, 33, 1000):
        v = rng.random(n) < 0.7
        assert (O.unpack_validity(O.pack_validity(v), n) == v).all()


def test_string_indexer_ties_and_unseen():
    labels, counts = O.string_indexer_fit(["b", "a", "c", "c", None, "a", "b"])
    assert labels == ["a", "b", "c"] and counts == [2, 2, 2]
    assert O.string_indexer_transform(["c", "zzz", None], labels).tolist() == [2.0, 3.0, 3.0]
    assert O.string_indexer_order_from_counts(["x", "b", "a"], [1, 5, 5]) == [2, 1, 0]


def test_bucketize_and_standardize():
    s = [0.0, 1.0, 5.0, 10.0]
    x = np.array([0.0, 0.5, 1.0, 4.9, 5.0, 10.0, np.nan, -1.0, 11.0])
    assert O.bucketize(x, s).tolist() == [0, 0, 1, 1, 2, 2, 3, -1, -1]
    m, sd = O.column_stats(np.array([1.0, 2.0, 3.0, 4.0]))
    assert m == 2.5 and abs(sd - 1.2909944487358056) < 1e-15
    assert O.standardize(np.array([2.0]), m, sd).tolist() == [2.0 * (1.0 / sd)]
    assert O.standardize(np.array([2.0]), m, 0.0).tolist() == [0.0]


def test_quoted_empty_cell_is_null_like_an_unquoted_one():
    # the rule stated in read_csv_columns: with Spark's default nullValue = "" / emptyValue = "" both forms
    # reach nullSafeDatum as "" and become null - a quoted empty cell is NOT an empty-string label
    text = 'measure_name,value,lower_ci,upper_ci\n"",1,2,3\n,4,5,6\nx,"",8,9\n"x",10,"",12\n'
    header, cols = O.read_csv_columns(text)
    assert cols["measure_name"] == [None, None, "x", "x"]
    assert cols["value"] == ["1", "4", None, "10"] and cols["lower_ci"] == ["2", "5", "8", None]
    out = O.k_means_etl(text, repeats=1)
    assert out["null_measure_name"] == 2 and out["rows_after_filter"] == 2 and out["labels"] == ["x"]
