#!/usr/bin/env python
"""Config 1 / config 5 (BASELINE.json configs[0], [4]): the full stage on CSV text -
decode -> null count -> means -> filter/impute -> StringIndexer -> features - timed
per step with CUDA events (text already in HBM; the H2D copy is timed separately),
on 1 GPU or sharded by rows over N GPUs (torchrun).  (CPU figures for the default
input come from the test suite's oracle, never from here: only tests/, smoke() and
bench.py may touch oracle/.)

    python tools/measure_pipeline.py [--rows 10000000] [--dtype f32]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import pyspark_tf_gke_b200 as pkg  # noqa: E402
from tools.measure_kernels import health_like_text  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64", "none"])
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--split-text", action="store_true",
                    help="N GPUs: every rank holds only its own 1/N of the text (config 5: the file split by byte range "
                         "at record boundaries) instead of the whole text sharded by row index")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = pkg.Engine(local)
    split = args.split_text and world > 1
    raw, rows = health_like_text(args.rows // world if split else args.rows)
    t0 = time.perf_counter()
    text = pkg.table.upload_text(eng, raw)
    torch.cuda.synchronize()
    h2d_s = time.perf_counter() - t0
    os.environ.setdefault("ETL_SKIP_KMEANS", "1")  # this tool times the ETL stage; KMeans is measured in measure_kernels
    wl = pkg.KMeansWorkload(eng)
    wl.logger.setLevel("ERROR")
    shard = (rank, world) if (world > 1 and not split) else None
    dtype = {"f32": torch.float32, "f64": torch.float64, "none": None}[args.dtype]

    def run():
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record()
        table = pkg.read_csv(eng, raw, ["value", "lower_ci", "upper_ci"], ["measure_name"], shard=shard, device_text=text)
        ev[1].record()
        model, _ = wl.k_means(table)
        ev[2].record()
        F = model.features(dtype) if dtype is not None else None
        ev[3].record()
        torch.cuda.synchronize()
        return model, F, [ev[i].elapsed_time(ev[i + 1]) for i in range(3)]

    run()
    from tools.clocks import ClockSampler
    sampler = ClockSampler(local)
    sampler.start()
    times = []
    for _ in range(args.steps):
        if world > 1:
            dist.barrier()
        model, F, t = run()
        times.append(t)
    clocks = sampler.stop()
    best = min(times, key=sum)
    tt = torch.tensor(best, dtype=torch.float64, device=eng.device)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    if rank == 0:
        total = float(tt.sum().item())
        if split:
            rows, nbytes = rows * world, len(raw) * world
        else:
            nbytes = len(raw)
        line = {"workload": f"full stage on {rows} CSV records ({nbytes / 1e9:.2f} GB of text)", "n_gpus": world,
                "text": "one byte range per rank" if split else "whole text on every rank, rows sharded",
                "ms": {"decode": float(tt[0]), "aggregate+project+index": float(tt[1]), "features": float(tt[2]),
                       "total": total},
                "rows_per_s": rows / (total * 1e-3), "text_gb_per_s": nbytes / 1e9 / (total * 1e-3),
                "features": None if F is None else [int(x) for x in F.shape] + [str(F.dtype)],
                "h2d_copy_s": h2d_s, "labels": len(model.labels), "means": model.means,
                "rows_after_filter": model.rows_after_filter, "null_keys": model.null_key_count, "clocks": clocks}
        print(json.dumps(line), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
