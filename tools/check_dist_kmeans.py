#!/usr/bin/env python
"""KMeans.fit with the rows sharded over N GPUs must give the model one GPU fits on all rows.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port P tools/check_dist_kmeans.py

Every rank first fits the whole (integer-valued, so exactly summable) dataset alone, then the process
group is initialised and each rank fits its contiguous shard; centres, sizes, iterations and cost
must be identical.  Prints one JSON line on rank 0; exits non-zero on a mismatch."""
from __future__ import annotations

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import pyspark_tf_gke_b200 as pkg  # noqa: E402


class Shard:
    def __init__(self, eng, index, nums, L, repeats):
        self.engine, self.labels, self.repeats = eng, ["x"] * L, repeats
        self.index = torch.from_numpy(index).to(eng.device)
        self.numerics = [torch.from_numpy(v).to(eng.device) for v in nums]


def main():
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    eng = pkg.Engine(local)
    n, L, repeats, k = 400_003, 30, 5, 25
    rng = np.random.default_rng(5)
    blob = rng.integers(0, k, n)
    index = ((blob * 7 + rng.integers(0, 3, n)) % (L + 1)).astype(np.int32)          # L = an unseen label
    nums = [np.floor(blob * 500.0 + rng.random(n) * 300.0) for _ in range(3)]
    whole = pkg.KMeans().setK(k).setSeed(1).setMaxIter(50).fit(Shard(eng, index, nums, L, repeats))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lo, hi = rank * n // world, (rank + 1) * n // world
    part = pkg.KMeans().setK(k).setSeed(1).setMaxIter(50).fit(Shard(eng, index[lo:hi], [v[lo:hi] for v in nums], L, repeats))
    same = (part.centers.tobytes() == whole.centers.tobytes() and part.summary.clusterSizes == whole.summary.clusterSizes
            and part.summary.numIter == whole.summary.numIter and part.summary.trainingCost == whole.summary.trainingCost)
    ok = torch.tensor([int(same)], device=eng.device)
    if world > 1:
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(json.dumps({"check": "sharded KMeans.fit == single-GPU fit", "n_gpus": world, "rows": n, "k": k,
                          "iterations": whole.summary.numIter, "identical": bool(ok.item()),
                          "max_abs_centre_diff": float(np.abs(part.centers - whole.centers).max())}), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    sys.exit(0 if bool(ok.item()) else 1)


if __name__ == "__main__":
    main()
