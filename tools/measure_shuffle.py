#!/usr/bin/env python
"""Config 4 (BASELINE.json configs[3]): 1B-row groupBy on a high-cardinality key
(100M groups), strong scaling over 1/2/4/8 GPUs: hash-partition -> NCCL
all-to-all over NVLink -> local final merge.

    python tools/measure_shuffle.py                       # 1 GPU (direct aggregation, no exchange)
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/measure_shuffle.py [--rows R --groups G --steps K]

Total rows are fixed; rank r generates rows [r*R/N, (r+1)*R/N) of the synthetic
table on its own GPU.  Timed with CUDA events between barriers, max over ranks;
phases (partition / exchange / merge) are timed separately the same way.  Checks:
row conservation, disjoint key ownership (sum of per-rank groups == distinct
keys), and the grand total of the sums against a plain column sum."""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import pyspark_tf_gke_b200 as pkg  # noqa: E402
import pyspark_tf_gke_b200.dist as D  # noqa: E402
import pyspark_tf_gke_b200.synth as S  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=S.CONFIG4["rows"])
    ap.add_argument("--groups", type=int, default=S.CONFIG4["groups"])
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--buckets", type=int, default=-1, help="table buckets per rank (-1 = choose from the table size)")
    ap.add_argument("--exchange", default="bucketed", choices=["bucketed", "peer", "nccl"],
                    help="bucketed = local (rank, bucket) partition in waves + slab copies over NVLink, merge behind the transfer; "
                         "peer = partition fused with NVLink peer stores (chunked slabs, or counted with --counted); "
                         "nccl = send buffer + all_to_all_single")
    ap.add_argument("--bucket-groups", type=int, default=8)
    ap.add_argument("--waves", type=int, default=None, help="bucketed exchange: parts the rows are cut and sent in (default 2)")
    ap.add_argument("--skip", default="", choices=["", "copy", "compute"],
                    help="bucketed exchange, timed steps only: leave out the NVLink transfer or the receivers' work, to time "
                         "the other side alone (the result is then wrong by design: no checks)")
    ap.add_argument("--out", default=None, help="append the JSON line to this file as well (rank 0)")
    ap.add_argument("--merge", default="smem", choices=["smem", "table"],
                    help="smem = two partition levels + shared-memory tables; table = bucket-by-bucket global table")
    ap.add_argument("--exchange-buckets", type=int, default=0,
                    help="peer + smem: 0 = the exchange also cuts by table bucket; 1 = by rank only, receiver partitions")
    ap.add_argument("--out-capacity", type=int, default=0, help="room for output groups per rank (0 = table capacity)")
    ap.add_argument("--counted", action="store_true", help="the counted path (exact count passes) instead of the slab path")
    ap.add_argument("--chunks", type=int, default=4, help="slab exchange: chunks the rows cross NVLink in")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = pkg.Engine(local)
    n = args.rows // world // 32 * 32
    row0 = rank * n
    key = eng.synth_column(S.CONFIG4["seed"], 0, S.KIND_KEY64, n, row0=row0, modulus=args.groups, with_valid=False).data
    val = eng.synth_column(S.CONFIG4["seed"], 1, S.KIND_EXACT, n, row0=row0, with_valid=False).data
    groups_local = args.groups // world + args.groups // (world * 4) + 1024
    cap = 1 << int(np.ceil(np.log2(2 * groups_local)))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def ev():
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def max_ms(a, b):
        t = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device=eng.device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    phases = {"init": [], "partition": [], "exchange": [], "merge": [], "total": [], "level1": [], "level2": []}
    if args.out_capacity:
        cap = args.out_capacity
    gt = None
    sent = n - n // world if world > 1 else 0  # expected rows leaving this GPU (uniform hash)
    B = D.choose_buckets(cap, 1) if args.buckets < 0 else args.buckets
    peer = None
    if world > 1 and args.exchange == "bucketed" and not args.counted:
        peer = D.BucketExchange(eng, n, groups_hint=args.groups, bucket_groups=args.bucket_groups, waves=args.waves)
    elif world > 1 and args.exchange in ("peer", "bucketed"):
        peer = (D.PeerExchange(eng, int(n * 1.05) + (1 << 20)) if args.counted
                else D.SlabExchange(eng, n, groups_hint=args.groups, chunks=args.chunks))
    from tools.clocks import ClockSampler
    sampler = ClockSampler(local)
    sampler.start()
    transfer = []
    if isinstance(peer, D.BucketExchange):
        peer.timing = True
    for it in range(args.warmup + args.steps):
        gt = None
        if it == args.warmup and args.skip:
            peer.skip = args.skip
        barrier()
        marks = {"start": ev()}
        gt = D.shuffle_groupby_sum(eng, key, val, cap, groups_hint=args.groups, buckets=B, peer=peer, merge=args.merge,
                                   exchange_buckets=args.exchange_buckets or None, counted=args.counted,
                                   mark=lambda name: marks.__setitem__(name, ev()))
        torch.cuda.synchronize()
        if it >= args.warmup:
            order = ["start", "init", "partition", "exchange", "merge"]
            for a, b in zip(order[:-1], order[1:]):
                phases[b].append(max_ms(marks[a], marks[b]))
            phases["total"].append(max_ms(marks["start"], marks["merge"]))
            if getattr(peer, "window", None):   # first copy kernel starts .. last bucket group landed
                transfer.append(max_ms(*peer.window))
            if "level1" in marks:   # inside "merge": the two partition levels of the shared-memory path
                phases["level1"].append(max_ms(marks["exchange"], marks["level1"]))
                phases["level2"].append(max_ms(marks["level1"], marks["level2"]))
    clocks = sampler.stop()
    # ---- checks
    if args.skip:                    # a timing-only run: there is no result to check
        res = {"rows": torch.zeros(1, dtype=torch.int64, device=eng.device), "keys": torch.zeros(0, device=eng.device),
               "cols": [{"sum": torch.zeros(1, dtype=torch.float64, device=eng.device)}]}
    else:
        res = eng.groupby_result(gt, max_groups=cap, sort=True)
    tot = torch.stack([res["rows"].sum(), torch.tensor(res["keys"].shape[0], device=eng.device)])
    gsum = res["cols"][0]["sum"].sum().reshape(1)
    s, c = eng.sum_count([pkg.Column(val)])
    direct = s[:1].clone()
    if world > 1:
        dist.all_reduce(tot)
        dist.all_reduce(gsum)
        dist.all_reduce(direct)
    ok_rows = int(tot[0].item()) == n * world
    ok_sum = float(gsum.item()) == float(direct.item())  # integer-valued data: exact
    if rank == 0:
        mean = {k: float(np.mean(v)) for k, v in phases.items() if v}
        if transfer:
            mean["transfer"] = float(np.mean(transfer))
        line = {"workload": "config4 shuffle+aggregate", "n_gpus": world, "rows_total": n * world, "groups": args.groups,
                "distinct_keys_found": int(tot[1].item()), "rows_conserved": ok_rows, "sum_matches_column_sum": ok_sum,
                "ms": mean, "rows_per_s": n * world / (mean["total"] * 1e-3),
                "buckets_per_rank": B if args.merge == "table" else None, "merge": args.merge, "exchange_buckets": args.exchange_buckets or "auto",
                "pieces_per_rank": getattr(gt, "pieces", None), "exchange": args.exchange if world > 1 else "none", "nvlink_sent_gb_per_gpu": sent * 16 / 1e9,
                "nvlink_gbs_per_gpu": (sent * 16 / 1e9) / (mean.get("transfer", mean["exchange"]) * 1e-3) if mean["exchange"] > 0 else None,
                "waves": getattr(peer, "waves", None), "skip": args.skip,
                "steps": args.steps, "path": "counted" if args.counted else "slabs", "clocks": clocks,
                "knobs": {k: v for k, v in os.environ.items() if k.startswith("ETL_")}}
        print(json.dumps(line), flush=True)
        if args.out:
            with open(args.out, "a") as fh:
                fh.write(json.dumps(line) + "\n")
    assert args.skip or (ok_rows and ok_sum)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
