"""CPU oracle of KMeans on the dense feature matrix.  TEST INFRASTRUCTURE ONLY.

Restates Spark MLlib's algorithm at the reference's call site
(workloads/raw-spark/k_means.py:83-87): k-means|| initialisation
(KMeans.initKMeansParallel: one random centre, initSteps rounds sampling each
point with probability 2k * cost / sumCosts, distinct candidates weighted by their
closest points, LocalKMeans.kMeansPlusPlus) and Lloyd iterations with the
tol = 1e-4 stopping rule.  It works on the DENSE rows [N, D], unlike the product,
and takes its random numbers from the repo's counter-based generator so that both
sides see the same draws.  Parity with a Spark run is unpinned: Spark's
XORShiftRandom streams are not reproducible here."""
from __future__ import annotations

import numpy as np

import pyspark_tf_gke_b200.synth as S

STREAM_FIRST, STREAM_STEP0 = 2000, 2001


def _norm2_tree(row: np.ndarray) -> float:
    """|c|^2 summed the way the prepare kernel does: 256 partials, pairwise tree."""
    part = np.zeros(256, dtype=np.float64)
    sq = row * row
    for t in range(256):
        acc = 0.0
        for d in range(t, len(row), 256):
            acc = acc + sq[d]
        part[t] = acc
    o = 128
    while o > 0:
        part[:o] = part[:o] + part[o:2 * o]
        o //= 2
    return float(part[0])


def sq_dist(X: np.ndarray, index: np.ndarray, L: int, repeats: int, centers: np.ndarray) -> np.ndarray:
    """|x|^2 - 2 x.c + |c|^2 with the product's operation order (no FMA), clamped at 0."""
    onehot = repeats * L
    a, b, c = X[:, onehot], X[:, onehot + 1], X[:, onehot + 2]
    known = (index >= 0) & (index < L)
    x2 = ((np.where(known, float(repeats), 0.0) + a * a) + b * b) + c * c
    out = np.empty((X.shape[0], centers.shape[0]))
    idx = np.where(known, index, L)
    for j, ctr in enumerate(centers):
        T = np.zeros(L + 1)
        for r in range(repeats):
            T[:L] = T[:L] + ctr[r * L:(r + 1) * L]
        dot = ((T[idx] + a * ctr[onehot]) + b * ctr[onehot + 1]) + c * ctr[onehot + 2]
        d2 = (x2 + -(2.0 * dot)) + _norm2_tree(ctr)
        out[:, j] = np.maximum(d2, 0.0)
    return out


def local_kmeans_pp(rng, points, weights, k, max_iter=30):
    m = points.shape[0]
    centers = np.zeros((k, points.shape[1]))

    def pick(cost):
        r = rng.random() * float(cost.sum())
        return min(int(np.searchsorted(np.cumsum(cost), r, side="right")), m - 1)

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
            centers[c] = (points * w[:, None]).sum(axis=0) / w.sum() if w.sum() > 0 else points[int(rng.integers(m))]
        if not moved:
            break
    return centers


def kmeans_fit(X: np.ndarray, index: np.ndarray, L: int, repeats: int, k: int, seed: int, max_iter: int = 20,
               tol: float = 1e-4, init_steps: int = 2):
    n = X.shape[0]
    first = int(S.bits(seed, STREAM_FIRST, 0, 1)[0] % np.uint64(n))
    candidates = X[[first]]
    new = candidates
    cost = np.full(n, np.inf)
    for step in range(init_steps):
        cost = np.minimum(cost, sq_dist(X, index, L, repeats, new).min(axis=1))
        sum_costs = float(cost.sum())
        if not sum_costs > 0:
            break
        u = S.unit(S.bits(seed, 2 * (STREAM_STEP0 + step), 0, n)) * 100000.0
        chosen = np.nonzero(u < cost * (100000.0 * 2.0 * k / sum_costs))[0]
        if chosen.size == 0:
            continue
        new = X[chosen]
        candidates = np.concatenate([candidates, new])
    distinct = np.unique(candidates, axis=0)
    if distinct.shape[0] <= k:
        centers = distinct
    else:
        closest = sq_dist(X, index, L, repeats, distinct).argmin(axis=1)
        weights = np.bincount(closest, minlength=distinct.shape[0]).astype(np.float64)
        centers = local_kmeans_pp(np.random.default_rng(seed), distinct, weights, k)
    kk = centers.shape[0]
    it, cost_sum, sizes = 0, 0.0, None
    while it < max_iter:
        d = sq_dist(X, index, L, repeats, centers)
        assign = d.argmin(axis=1)
        cost_sum = float(d.min(axis=1).sum())
        sizes = np.bincount(assign, minlength=kk)
        new_c = centers.copy()
        for c in range(kk):
            if sizes[c] > 0:
                new_c[c] = X[assign == c].sum(axis=0) / float(sizes[c])
        moved = ((new_c - centers) ** 2).sum(axis=1)
        centers = new_c
        it += 1
        if (moved <= tol * tol).all():
            break
    return centers, cost_sum, it, sizes


def silhouette(X: np.ndarray, index: np.ndarray, L: int, repeats: int, cluster: np.ndarray) -> np.ndarray:
    """Per-row silhouette coefficient with squared Euclidean distance, as Spark's
    ClusteringEvaluator computes it (org.apache.spark.ml.evaluation.SquaredEuclideanSilhouette;
    called at spark_workload_to_cloud_k8s.py:139-144 - Spark itself is an unpinned third-party
    dependency, so this follows its published algorithm): per cluster G the count N_G, Psi_G =
    sum |y|^2 and Y_G = sum y; mean squared distance x -> G = |x|^2 + Psi_G / N_G - 2 x.Y_G / N_G;
    a = own-cluster value * N / (N - 1); b = min over the other clusters; 1 - a/b, b/a - 1 or 0;
    0 for a cluster of one.  Operation order = the CUDA kernel's (structured dot product, no FMA)."""
    n = X.shape[0]
    onehot = repeats * L
    known = (index >= 0) & (index < L)
    a, b, c = X[:, onehot], X[:, onehot + 1], X[:, onehot + 2]
    x2 = ((np.where(known, float(repeats), 0.0) + a * a) + b * b) + c * c
    ks = np.unique(cluster)
    out = np.zeros(n)
    stats = {}
    for g in ks:
        sel = cluster == g
        lab = np.bincount(np.where(known, index, L)[sel], minlength=L + 1).astype(np.float64) * float(repeats)
        lab[L] = 0.0
        stats[g] = (float(sel.sum()), lab, [float(X[sel, onehot + j].sum()) for j in range(3)], float(_seq_sum(x2[sel])))
    idx = np.where(known, index, L)
    dist = {}
    for g, (ng, lab, ns, psi) in stats.items():
        dot = ((lab[idx] + a * ns[0]) + b * ns[1]) + c * ns[2]
        dist[g] = (x2 + psi / ng) + -((2.0 * dot) / ng)
    for i in range(n):
        own = cluster[i]
        n_own = stats[own][0]
        others = [dist[g][i] for g in ks if g != own]
        if n_own <= 1.0 or not others:
            continue
        cur = (dist[own][i] * n_own) / (n_own - 1.0)
        nb = min(others)
        out[i] = 1.0 - cur / nb if cur < nb else (nb / cur - 1.0 if cur > nb else 0.0)
    return out


def _seq_sum(v: np.ndarray) -> float:
    return float(np.sum(v))


def silhouette_direct(X: np.ndarray, cluster: np.ndarray) -> np.ndarray:
    """The textbook O(n^2) definition (mean squared distance to every other row of the own
    cluster / nearest other cluster) - used on small inputs to check the O(n k) formula."""
    n = X.shape[0]
    d = ((X[:, None, :] - X[None, :, :]) ** 2).sum(axis=2)
    out = np.zeros(n)
    ks = np.unique(cluster)
    for i in range(n):
        own = cluster == cluster[i]
        if own.sum() <= 1 or len(ks) < 2:
            continue
        a = d[i, own].sum() / (own.sum() - 1)
        b = min(d[i, cluster == g].mean() for g in ks if g != cluster[i])
        out[i] = 1 - a / b if a < b else (b / a - 1 if a > b else 0.0)
    return out
