"""KMeans on the ETL stage's `features` - the call that follows the stage in the
reference: `KMeans().setK(25).setSeed(1).setMaxIter(1000).fit(dataset)`
(workloads/raw-spark/k_means.py:83-87), SURVEY.md §8f rank 2.

Same estimator surface (setK / setSeed / setMaxIter / setTol / setInitSteps, fit,
model.clusterCenters(), model.transform / predict, summary.trainingCost) and the
same algorithm as Spark MLlib's KMeans: k-means|| initialisation (one uniformly
chosen centre, initSteps=2 rounds that sample every point with probability
2k * cost / sumCosts, candidates weighted by the points closest to them, a local
weighted k-means++ + Lloyd on the candidates) followed by Lloyd iterations until
every centre moves less than tol (1e-4) or maxIter.  The passes over the rows
(closest centre, per-cluster sums, cost sums) are CUDA kernels on the structured
features - the dense [N, D] matrix is never built; the driver-side steps on the
<= few hundred candidates run in NumPy, as they run on the driver in Spark.

Parity: Spark draws its random numbers from XORShiftRandom / java.util.Random
seeded per partition; those streams cannot be reproduced without Spark, so the
chosen initial centres differ from a Spark run with the same seed - parity
unpinned.  The algorithm itself is checked against an
independent NumPy restatement using the same counter-based random numbers
(tests/test_gpu_kmeans.py)."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import synth
from .engine import Column, Engine

# random streams (synth.bits(seed, stream, row)): first centre, then one per init step
STREAM_FIRST = 2000
STREAM_STEP0 = 2001


def dense_rows(index: np.ndarray, nums: Sequence[np.ndarray], L: int, repeats: int) -> np.ndarray:
    """Structured rows -> dense float64 [m, repeats * L + 3] (host, small m only)."""
    m = len(index)
    out = np.zeros((m, repeats * L + 3), dtype=np.float64)
    for i, ix in enumerate(index):
        if 0 <= ix < L:
            out[i, [r * L + ix for r in range(repeats)]] = 1.0
    for j in range(3):
        out[:, repeats * L + j] = nums[j]
    return out


def local_kmeans_pp(rng: np.random.Generator, points: np.ndarray, weights: np.ndarray, k: int, max_iter: int = 30):
    """Spark's LocalKMeans.kMeansPlusPlus on the k-means|| candidates: weighted
    k-means++ seeding, then weighted Lloyd for at most max_iter rounds; an empty
    cluster is re-seeded with a random candidate."""
    m = points.shape[0]
    centers = np.zeros((k, points.shape[1]), dtype=np.float64)

    def pick(cost):
        r = rng.random() * float(cost.sum())
        j = int(np.searchsorted(np.cumsum(cost), r, side="right"))
        return min(j, m - 1)

    centers[0] = points[pick(weights.astype(np.float64))]
    cost = ((points - centers[0]) ** 2).sum(axis=1)
    for i in range(1, k):
        centers[i] = points[pick(cost * weights)]
        cost = np.minimum(cost, ((points - centers[i]) ** 2).sum(axis=1))
    assign = np.full(m, -1)
    for _ in range(max_iter):
        d = ((points[:, None, :] - centers[None, :, :]) ** 2).sum(axis=2)
        new_assign = d.argmin(axis=1)
        moved = bool((new_assign != assign).any())
        assign = new_assign
        for c in range(k):
            w = weights * (assign == c)
            if w.sum() > 0:
                centers[c] = (points * w[:, None]).sum(axis=0) / w.sum()
            else:
                centers[c] = points[int(rng.integers(m))]
        if not moved:
            break
    return centers


@dataclass
class KMeansSummary:
    trainingCost: float
    numIter: int
    clusterSizes: List[int]
    k: int


@dataclass
class KMeansModel:
    engine: Engine
    centers: np.ndarray              # [k, D] float64
    num_labels: int
    repeats: int
    summary: Optional[KMeansSummary] = None

    def clusterCenters(self) -> List[np.ndarray]:
        return [c.copy() for c in self.centers]

    def save(self, path: str, params: Optional[dict] = None) -> None:
        """KMeansModel.save(path): Spark ML's directory layout (mlformat.py)."""
        from . import mlformat
        mlformat.save_kmeans_model(self, path, params)

    @staticmethod
    def load(path: str, engine: Optional[Engine] = None) -> "KMeansModel":
        from . import mlformat
        return mlformat.load_kmeans_model(engine, path)

    def predict_structured(self, index: torch.Tensor, nums: Sequence[torch.Tensor]) -> torch.Tensor:
        c = torch.from_numpy(self.centers).to(self.engine.device)
        cluster, _ = self.engine.kmeans_assign(index, nums, c, self.num_labels, self.repeats)
        return cluster

    def transform(self, pipeline_model, index=None, numerics=None) -> torch.Tensor:
        """predictions for the rows a pipeline model produced (k_means.py:153)."""
        index = pipeline_model.index if index is None else index
        numerics = pipeline_model.numerics if numerics is None else numerics
        return self.predict_structured(index, numerics)


class KMeans:
    """pyspark.ml.clustering.KMeans' fluent setters, as the reference uses them."""

    def __init__(self, k: int = 2, seed: int = 0, maxIter: int = 20, tol: float = 1e-4, initSteps: int = 2):
        self.k, self.seed, self.maxIter, self.tol, self.initSteps = k, seed, maxIter, tol, initSteps

    def setK(self, k):
        self.k = int(k)
        return self

    def setSeed(self, seed):
        self.seed = int(seed)
        return self

    def setMaxIter(self, n):
        self.maxIter = int(n)
        return self

    def setTol(self, tol):
        self.tol = float(tol)
        return self

    def setInitSteps(self, n):
        self.initSteps = int(n)
        return self

    # ------------------------------------------------------------------
    def _gather_dense(self, index, nums, rows: np.ndarray, L, repeats) -> np.ndarray:
        sel = torch.from_numpy(rows.astype(np.int64)).to(index.device)
        return dense_rows(index[sel].cpu().numpy(), [t[sel].cpu().numpy() for t in nums], L, repeats)

    # With torch.distributed initialised the dataset is this rank's shard of the rows (rank order = row
    # order): per-cluster sums, costs and candidate weights are all-reduced and the sampled candidate
    # rows all-gathered, so every rank ends with the model a single GPU would have fitted on all rows
    # (bit for bit when the sums are exact).  Random draws are keyed by the GLOBAL row number.
    @staticmethod
    def _all_sum(t: torch.Tensor) -> torch.Tensor:
        from . import dist as D
        if D.world()[1] > 1:
            import torch.distributed as td
            td.all_reduce(t)
        return t

    @staticmethod
    def _all_max(t: torch.Tensor) -> torch.Tensor:
        from . import dist as D
        if D.world()[1] > 1:
            import torch.distributed as td
            td.all_reduce(t, op=td.ReduceOp.MAX)
        return t

    @staticmethod
    def _all_rows(local: np.ndarray) -> np.ndarray:
        """concatenation over ranks (rank order) of a few dense rows"""
        from . import dist as D
        if D.world()[1] == 1:
            return local
        import torch.distributed as td
        parts = [None] * D.world()[1]
        td.all_gather_object(parts, local)
        return np.concatenate(parts)

    def _init_parallel(self, eng: Engine, index, nums, L, repeats) -> np.ndarray:
        from . import dist as D
        rank, size = D.world()
        n, k = index.shape[0], self.k
        counts = torch.zeros(size, dtype=torch.int64, device=eng.device)
        counts[rank] = n
        counts = self._all_sum(counts).cpu().numpy()
        row0, total = int(counts[:rank].sum()), int(counts.sum())
        D_feat = repeats * L + 3
        first = int(synth.bits(self.seed, STREAM_FIRST, 0, 1)[0] % np.uint64(total))
        mine = row0 <= first < row0 + n
        candidates = self._all_rows(self._gather_dense(index, nums, np.array([first - row0]), L, repeats) if mine
                                    else np.zeros((0, D_feat)))
        new = candidates
        cost = torch.full((n,), float("inf"), dtype=torch.float64, device=eng.device)
        for step in range(self.initSteps):
            _, cost = eng.kmeans_assign(index, nums, torch.from_numpy(new).to(eng.device), L, repeats, prev_cost=cost,
                                        want_cluster=False)
            s, _ = eng.sum_count([Column(cost)])
            sum_costs = float(self._all_sum(s[:1].clone())[0])
            if not sum_costs > 0.0:
                break
            # u * 1e5 < 1e5 * 2k * cost / sumCosts  (the synthetic generator hands out u * 1e5)
            lead = row0 % 32      # the generator starts at multiples of 32 rows
            u = eng.synth_column(self.seed, STREAM_STEP0 + step, synth.KIND_REAL, n + lead, row0=row0 - lead,
                                 with_valid=False).data[lead:]
            chosen = torch.nonzero(u < cost * (100000.0 * 2.0 * k / sum_costs)).flatten().cpu().numpy()
            picked = self._all_rows(self._gather_dense(index, nums, chosen, L, repeats) if chosen.size else np.zeros((0, D_feat)))
            if picked.shape[0] == 0:
                continue
            new = picked
            candidates = np.concatenate([candidates, new])
        distinct = np.unique(candidates, axis=0)
        if distinct.shape[0] <= k:
            return distinct
        cluster, _ = eng.kmeans_assign(index, nums, torch.from_numpy(distinct).to(eng.device), L, repeats)
        weights = self._all_sum(eng.code_histogram(Column(cluster), distinct.shape[0])[0])     # rows per candidate
        return local_kmeans_pp(np.random.default_rng(self.seed), distinct, weights.cpu().numpy().astype(np.float64), k)

    def fit(self, dataset) -> KMeansModel:
        """dataset: the EtlPipelineModel returned by KMeansWorkload.k_means (its
        `index`, `numerics`, label count and repeat factor describe `features`)."""
        eng, index, nums = dataset.engine, dataset.index, dataset.numerics
        L, repeats = len(dataset.labels), dataset.repeats
        n = index.shape[0]
        if n == 0:
            raise ValueError("KMeans.fit: empty dataset")
        centers = self._init_parallel(eng, index, nums, L, repeats)
        k = centers.shape[0]
        it, cost_sum, count = 0, 0.0, None
        d_centers = torch.from_numpy(np.ascontiguousarray(centers)).to(eng.device)
        while it < self.maxIter:
            # one Lloyd iteration, all on the device: assign, per-cluster sums, new centres; a single
            # read-back (the new centres + the cost) decides about convergence
            cluster, cost = eng.kmeans_assign(index, nums, d_centers, L, repeats)
            count, hist, numsum = eng.kmeans_accumulate(index, nums, cluster, k, L)
            s, _ = eng.sum_count([Column(cost)])
            for t in (count, hist, numsum, s):   # rows sharded over ranks: the sums of all of them
                self._all_sum(t)
            eng.kmeans_update(count, hist, numsum, d_centers, L, repeats)
            new = d_centers.cpu().numpy()
            cost_sum = float(s[0])
            moved = ((new - centers) ** 2).sum(axis=1)
            centers = new
            it += 1
            if bool((moved <= self.tol * self.tol).all()):
                break
        sizes = count.cpu().tolist() if count is not None else []
        return KMeansModel(eng, centers, L, repeats, KMeansSummary(cost_sum, it, [int(x) for x in sizes], k))


class ClusteringEvaluator:
    """pyspark.ml.evaluation.ClusteringEvaluator as spark_workload_to_cloud_k8s.py:139-144 uses it:
    default metric "silhouette", default distance "squaredEuclidean".  evaluate() takes what
    model.transform produced: the dataset (structured features) and its predictions.  With
    torch.distributed initialised the rows are this rank's shard: the per-cluster statistics and the
    final (sum, count) are all-reduced."""

    def __init__(self, metricName: str = "silhouette", distanceMeasure: str = "squaredEuclidean"):
        if metricName != "silhouette" or distanceMeasure != "squaredEuclidean":
            raise ValueError("ClusteringEvaluator: only silhouette / squaredEuclidean (the reference's defaults) are implemented")

    def evaluate(self, dataset, predictions: torch.Tensor, k: Optional[int] = None) -> float:
        eng, index, nums = dataset.engine, dataset.index, dataset.numerics
        L, repeats = len(dataset.labels), dataset.repeats
        if k is None:
            kk = predictions.max().reshape(1).to(torch.int64) + 1 if predictions.numel() else torch.ones(1, dtype=torch.int64, device=eng.device)
            k = int(KMeans._all_max(kk)[0])
        xi = eng.kmeans_row_norms(index, nums, L, repeats)
        count, hist, numsum = eng.kmeans_accumulate(index, nums, predictions, k, L)
        _, _, psi3 = eng.kmeans_accumulate(index, [xi, xi, xi], predictions, k, L)
        psi = psi3[:, 0].contiguous()
        for t in (count, hist, numsum, psi):
            KMeans._all_sum(t)
        label_sum = (hist.to(torch.float64) * float(repeats)).contiguous()
        label_sum[:, L] = 0.0                               # unknown label: the all-zero one-hot block
        sil = eng.kmeans_silhouette(index, nums, predictions, label_sum, numsum, psi, count.to(torch.float64), L, repeats)
        s, c = eng.sum_count([Column(sil)])
        tot = torch.stack([s[0], c[0].to(torch.float64)])
        KMeans._all_sum(tot)
        return float(tot[0]) / float(tot[1]) if float(tot[1]) else float("nan")
