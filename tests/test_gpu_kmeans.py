"""KMeans on the structured features (k_means.py:83-87): the
CUDA passes + host driver against an independent dense NumPy restatement of the
same algorithm with the same counter-based random numbers.  Data are integer
valued, so per-cluster sums are exact and the two trajectories coincide; centres
are compared to 1e-12 relative, assignments and cluster sizes exactly."""
import numpy as np
import pytest
import torch

import pyspark_tf_gke_b200 as pkg
from oracle import kmeans_oracle as KO
from pyspark_tf_gke_b200.kmeans import dense_rows
from tests.conftest import health_text

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    e = pkg.Engine(0)
    yield e
    e.close()


class _Dataset:  # what KMeans.fit needs from an EtlPipelineModel
    def __init__(self, eng, index, nums, L, repeats):
        self.engine, self.labels, self.repeats = eng, ["l%d" % i for i in range(L)], repeats
        self.index = torch.from_numpy(index).to(eng.device)
        self.numerics = [torch.from_numpy(v).to(eng.device) for v in nums]


def _blobs(n, L, seed):
    rng = np.random.default_rng(seed)
    blob = rng.integers(0, 6, n)
    index = ((blob * 3 + rng.integers(0, 2, n)) % (L + 1)).astype(np.int32)      # L itself = unknown label
    centres = rng.integers(0, 2000, (6, 3)).astype(np.float64)
    nums = [np.floor(centres[blob, j] + rng.integers(-20, 21, n)) for j in range(3)]
    return index, nums


def test_assign_and_accumulate_match_dense_numpy(eng):
    n, L, repeats, k = 50_000, 12, 3, 9
    index, nums = _blobs(n, L, 1)
    rng = np.random.default_rng(2)
    centers = rng.random((k, repeats * L + 3)) * np.array([1.0] * (repeats * L) + [2000.0] * 3)
    X = dense_rows(index, nums, L, repeats)
    d = KO.sq_dist(X, index, L, repeats, centers)
    ds = _Dataset(eng, index, nums, L, repeats)
    cluster, cost = eng.kmeans_assign(ds.index, ds.numerics, torch.from_numpy(centers).to(eng.device), L, repeats)
    assert cluster.cpu().numpy().tolist() == d.argmin(axis=1).tolist()
    assert cost.cpu().numpy().tobytes() == d.min(axis=1).tobytes()          # same operations, same bits
    direct = ((X[:, None, :] - centers[None, :, :]) ** 2).sum(axis=2)       # and it is the Euclidean distance
    np.testing.assert_allclose(cost.cpu().numpy(), direct.min(axis=1), rtol=1e-9, atol=1e-6)
    count, hist, numsum = eng.kmeans_accumulate(ds.index, ds.numerics, cluster, k, L)
    a = d.argmin(axis=1)
    assert count.cpu().numpy().tolist() == np.bincount(a, minlength=k).tolist()
    for c in range(k):
        idx = np.where((index >= 0) & (index < L), index, L)[a == c]
        assert hist[c].cpu().numpy().tolist() == np.bincount(idx, minlength=L + 1).tolist()
        assert numsum[c].cpu().numpy().tolist() == [float(v[a == c].sum()) for v in nums]
    # k-means|| step: min-combine with a previous cost, cluster ids offset, untouched where not beaten
    prev = torch.from_numpy(np.where(np.arange(n) % 2 == 0, 0.0, np.inf)).to(eng.device)
    cl = torch.full((n,), -5, dtype=torch.int32, device=eng.device)
    cl2, cost2 = eng.kmeans_assign(ds.index, ds.numerics, torch.from_numpy(centers).to(eng.device), L, repeats,
                                   prev_cost=prev, cluster=cl, cluster_offset=100)
    assert (cost2.cpu().numpy()[::2] == 0).all() and (cl2.cpu().numpy()[::2] == -5).all()
    assert (cl2.cpu().numpy()[1::2] == a[1::2] + 100).all()


@pytest.mark.parametrize("n,L,repeats,k,seed", [(20_000, 12, 3, 6, 1), (5_000, 30, 5, 25, 1), (300, 4, 1, 3, 7)])
def test_fit_matches_the_numpy_restatement(eng, n, L, repeats, k, seed):
    index, nums = _blobs(n, L, seed + 10)
    ds = _Dataset(eng, index, nums, L, repeats)
    model = pkg.KMeans().setK(k).setSeed(seed).setMaxIter(40).fit(ds)
    X = dense_rows(index, nums, L, repeats)
    centers, cost, iters, sizes = KO.kmeans_fit(X, index, L, repeats, k, seed, max_iter=40)
    assert model.centers.shape == centers.shape
    np.testing.assert_allclose(model.centers, centers, rtol=1e-12, atol=1e-12)
    assert model.summary.numIter == iters and model.summary.clusterSizes == sizes.tolist()
    assert abs(model.summary.trainingCost - cost) <= 1e-9 * max(1.0, cost)
    pred = model.predict_structured(ds.index, ds.numerics).cpu().numpy()
    assert pred.tolist() == KO.sq_dist(X, index, L, repeats, model.centers).argmin(axis=1).tolist()
    assert len(model.clusterCenters()) == centers.shape[0]


def test_kmeans_workload_returns_both_models_and_infers_single_rows(eng, monkeypatch, tmp_path):
    monkeypatch.delenv("ETL_SKIP_KMEANS", raising=False)
    monkeypatch.setenv("MEASURE_NAME_WEIGHT", "5")
    monkeypatch.setenv("SAVE_MODELS", "true")                      # the block k_means.py:94-104 keeps commented out
    monkeypatch.setenv("MODEL_OUTPUT_DIR", str(tmp_path))
    table = pkg.read_csv(eng, health_text("workloads").encode(), ["value", "lower_ci", "upper_ci"], ["measure_name"])
    wl = pkg.KMeansWorkload(eng)
    pipeline_model, model = wl.k_means(table)                      # k_means.py:17,107: (pipeline_model, model)
    assert model is not None and model.centers.shape == (25, 153)  # setK(25) on the 153-dim features
    assert sum(model.summary.clusterSizes) == 18155 and model.summary.numIter <= 1000
    # the cost of the returned centres is a local optimum of Lloyd: one more pass does not increase it
    d_centers = torch.from_numpy(model.centers).to(eng.device)
    _, cost = eng.kmeans_assign(pipeline_model.index, pipeline_model.numerics, d_centers, 30, 5)
    assert float(cost.sum()) <= model.summary.trainingCost * (1 + 1e-9)
    pkg.KMeansWorkload.pipeline_model, pkg.KMeansWorkload.kmeans_model = pipeline_model, model
    pred, feats = wl.infer_single_row(None, "Asthma", 10)          # k_means.py:138: row (label, 10, 17, 15)
    f = feats.cpu().numpy()[0]
    assert 0 <= pred < 25 and f.shape == (153,) and f[150:].tolist() == [10.0, 17.0, 15.0]
    assert f[:150].sum() == 5 and f[pipeline_model.labels.index("Asthma")] == 1
    pred2, feats2 = wl.infer_single_row(None, "no such measure", 0)  # unseen label -> all-zero one-hot blocks
    assert feats2.cpu().numpy()[0, :150].sum() == 0 and 0 <= pred2 < 25
    # models saved in Spark ML's layout and loaded back (k_means.py:118-135) give the same answers
    loaded_pipeline, loaded_model = wl.load_models()
    assert loaded_pipeline.labels == pipeline_model.labels and loaded_pipeline.means == pipeline_model.means
    assert loaded_model.centers.tobytes() == model.centers.tobytes()
    index2, nums2 = loaded_pipeline.transform(table)
    assert torch.equal(index2, pipeline_model.index) and all(torch.equal(a, b) for a, b in zip(nums2, pipeline_model.numerics))
    assert torch.equal(loaded_model.transform(loaded_pipeline, index2, nums2), model.transform(pipeline_model))
    pkg.KMeansWorkload.pipeline_model, pkg.KMeansWorkload.kmeans_model = loaded_pipeline, loaded_model
    assert wl.infer_single_row(None, "Asthma", 10)[0] == pred


@pytest.mark.parametrize("k", [1, 25, 32, 40])
def test_accumulate_owner_and_atomic_kernels_match_numpy(eng, k):
    # k <= 32: lane c of every warp owns cluster c (no floating-point atomics); above: shared atomics
    n, L = 70_001, 30
    index, nums = _blobs(n, L, 3)
    rng = np.random.default_rng(k)
    cluster = rng.integers(-1, k + 1, n).astype(np.int32)         # -1 and k are out of range: skipped
    cluster[rng.random(n) < 0.5] = 0                              # a hot cluster
    ds = _Dataset(eng, index, nums, L, 5)
    count, hist, numsum = eng.kmeans_accumulate(ds.index, ds.numerics, torch.from_numpy(cluster).to(eng.device), k, L)
    ok = (cluster >= 0) & (cluster < k)
    assert count.cpu().numpy().tolist() == np.bincount(cluster[ok], minlength=k).tolist()
    idx = np.where((index >= 0) & (index < L), index, L)
    for c in range(k):
        sel = cluster == c
        assert hist[c].cpu().numpy().tolist() == np.bincount(idx[sel], minlength=L + 1).tolist()
        assert numsum[c].cpu().numpy().tolist() == [float(v[sel].sum()) for v in nums]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (run by hand with gpurun --gpus 2)")
def test_sharded_fit_equals_single_gpu_fit():
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29533", os.path.join(root, "tools", "check_dist_kmeans.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert '"identical": true' in r.stdout


def test_silhouette_matches_spark_formula_and_textbook_definition(eng):
    # ClusteringEvaluator (silhouette, squared Euclidean): spark_workload_to_cloud_k8s.py:139-144
    n, L, repeats, k = 3000, 12, 3, 7
    index, nums = _blobs(n, L, 5)
    rng = np.random.default_rng(9)
    cluster = rng.integers(0, k - 1, n).astype(np.int32)
    cluster[11] = k - 1                                   # a cluster of one row: coefficient 0
    ds = _Dataset(eng, index, nums, L, repeats)
    d_cluster = torch.from_numpy(cluster).to(eng.device)
    X = dense_rows(index, nums, L, repeats)
    want = KO.silhouette(X, index, L, repeats, cluster)
    # per row, through the C ABI pieces
    xi = eng.kmeans_row_norms(ds.index, ds.numerics, L, repeats)
    count, hist, numsum = eng.kmeans_accumulate(ds.index, ds.numerics, d_cluster, k, L)
    psi = eng.kmeans_accumulate(ds.index, [xi, xi, xi], d_cluster, k, L)[2][:, 0].contiguous()
    label_sum = (hist.to(torch.float64) * float(repeats)).contiguous()
    label_sum[:, L] = 0.0
    got = eng.kmeans_silhouette(ds.index, ds.numerics, d_cluster, label_sum, numsum, psi, count.to(torch.float64), L, repeats)
    assert got.cpu().numpy().tobytes() == want.tobytes()              # integer-valued data: same operations, same bits
    assert want[11] == 0.0 and -1.0 <= want.min() and want.max() <= 1.0
    sub = slice(0, 400)                                               # the O(n^2) definition on a slice with its own clusters
    np.testing.assert_allclose(KO.silhouette(X[sub], index[sub], L, repeats, cluster[sub]), KO.silhouette_direct(X[sub], cluster[sub]),
                               rtol=0, atol=1e-10)
    value = pkg.ClusteringEvaluator().evaluate(ds, d_cluster)
    assert abs(value - float(want.mean())) <= 1e-12
    with pytest.raises(ValueError):
        pkg.ClusteringEvaluator(metricName="other")


def test_silhouette_of_a_fitted_model_is_high_on_separated_blobs(eng):
    index, nums = _blobs(20_000, 12, 4)
    ds = _Dataset(eng, index, nums, 12, 3)
    model = pkg.KMeans().setK(6).setSeed(1).fit(ds)
    predictions = model.transform(ds)
    value = pkg.ClusteringEvaluator().evaluate(ds, predictions)
    assert 0.5 < value <= 1.0


def test_csv_twin_script_end_to_end(eng, tmp_path, capsys):
    # spark_workload_to_cloud_k8s.py:48-154 on the default input: un-repeated one-hot block, k = 5, seed 1,
    # silhouette, both models saved and loadable
    from pyspark_tf_gke_b200 import workload_to_cloud as W
    pipeline_model, model, sil = W.run(health_text("workloads").encode(), model_dir=str(tmp_path), engine=eng)
    out = capsys.readouterr().out
    assert "Total number of rows: 18155" in out and "Column 'measure_name' has 0 missing values" in out
    assert "Silhouette with squared Euclidean distance = " in out and "Saving K-Means pipeline to" in out
    assert pipeline_model.repeats == 1 and model.centers.shape == (5, 33) and -1.0 <= sil <= 1.0
    # the evaluator against the O(n k) restatement on the same predictions
    pred = model.transform(pipeline_model).cpu().numpy()
    index = pipeline_model.index.cpu().numpy()
    nums = [t.cpu().numpy() for t in pipeline_model.numerics]
    want = KO.silhouette(dense_rows(index, nums, 30, 1), index, 30, 1, pred)
    assert abs(sil - float(want.mean())) <= 1e-9
    back = pkg.KMeansModel.load(str(tmp_path / "health_kmeans_model"))
    assert back.centers.tobytes() == model.centers.tobytes() and back.repeats == 1
    assert pkg.EtlPipelineModel.load(str(tmp_path / "health_kmeans_pipeline")).labels == pipeline_model.labels
