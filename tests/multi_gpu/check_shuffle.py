#!/usr/bin/env python
"""Full oracle diff of the multi-GPU shuffle + aggregate (test infrastructure; run by
tests/test_gpu_multi.py when the box has >= 2 GPUs, or by hand):

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port P \
        tests/multi_gpu/check_shuffle.py [--rows 2000000 --groups 150000]

Every exchange variant of dist.shuffle_groupby_sum - the bucketed exchange (the default), the chunked slab exchange, NCCL all-to-all, the counted fused peer scatter (by rank, and by rank x
bucket), the host-buffer entry - runs on this rank's shard of a seeded
table; every rank's groups are gathered on rank 0 and compared with
oracle.groupby_agg_fast on ALL rows: same keys, counts exact, sums bit-exact
(integer-valued data), each key on the rank oracle.partition_dest gives it.
Prints one JSON line on rank 0; exits non-zero on a mismatch."""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import pyspark_tf_gke_b200 as pkg  # noqa: E402
import pyspark_tf_gke_b200.dist as D  # noqa: E402
import pyspark_tf_gke_b200.synth as S  # noqa: E402
from oracle import etl_oracle as O  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=2_000_000)
    ap.add_argument("--groups", type=int, default=150_000)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = pkg.Engine(local)
    n = args.rows // world // 32 * 32 + (0 if rank else 0)
    hk, _ = S.host_column(3, 0, S.KIND_KEY64, n, row0=rank * n, modulus=args.groups)
    hv, _ = S.host_column(3, 1, S.KIND_EXACT, n, row0=rank * n)
    if rank == world - 1:            # an unaligned, shorter last shard
        hk, hv = hk[:n - 77], hv[:n - 77]
    key, val = torch.from_numpy(hk).to(eng.device), torch.from_numpy(hv).to(eng.device)
    cap = args.groups // world * 2 + 4096
    peer = D.PeerExchange(eng, int(n * 1.3) + 4096)
    slab = D.SlabExchange(eng, n, groups_hint=args.groups, chunks=3)
    bucketed = D.BucketExchange(eng, n, groups_hint=args.groups, bucket_groups=4)
    one_wave = D.BucketExchange(eng, n, groups_hint=args.groups, bucket_groups=2, waves=1)
    three_waves = D.BucketExchange(eng, n, groups_hint=args.groups, bucket_groups=8, waves=3, mover="tma")
    variants = {"bucketed": dict(peer=bucketed), "bucketed_minmax": dict(peer=bucketed, minmax=True),
                "bucketed_one_wave": dict(peer=one_wave), "bucketed_three_waves_tma": dict(peer=three_waves),
                "slab_chunked": dict(peer=slab), "nccl": dict(peer=None, counted=True),
                "peer_by_rank": dict(peer=peer, exchange_buckets=1),
                "peer_by_rank_and_bucket": dict(peer=peer, exchange_buckets=16), "host_entry": None}
    results = {}
    rows_all = [None] * world
    dist.all_gather_object(rows_all, (hk, hv))
    want = O.groupby_agg_fast(np.concatenate([k for k, _ in rows_all]), [np.concatenate([v for _, v in rows_all])], [None]) \
        if rank == 0 else None
    ok_all = True
    for name, kw in variants.items():
        if kw is None:               # pinned host rows in, host group columns out, over the slab exchange
            hk_p, hv_p = torch.from_numpy(hk).pin_memory(), torch.from_numpy(hv).pin_memory()
            out = tuple(torch.empty(cap, dtype=t, pin_memory=True) for t in (torch.int64, torch.int64, torch.float64))
            g = D.shuffle_groupby_sum_host(eng, hk_p, hv_p, cap, groups_hint=args.groups, peer=bucketed, out=out, chunks=3)
            mine = tuple(o[:g].numpy().copy() for o in out)
        else:
            for rep in range(2):     # twice: the receive buffer is reused
                pg = D.shuffle_groupby_sum(eng, key, val, cap, groups_hint=args.groups, **kw)
                res = eng.groupby_result(pg, sort=False)
            mine = (res["keys"].cpu().numpy(), res["rows"].cpu().numpy(), res["cols"][0]["sum"].cpu().numpy())
        parts = [None] * world
        dist.all_gather_object(parts, mine)
        if rank == 0:
            owner_ok = all(bool((O.partition_dest(parts[r][0], world) == r).all()) for r in range(world))
            k = np.concatenate([p[0] for p in parts])
            o = np.argsort(k, kind="stable")
            same = k.shape == want["keys"].shape and bool((k[o] == want["keys"]).all())
            cnt_ok = same and bool((np.concatenate([p[1] for p in parts])[o] == want["rows"]).all())
            sum_ok = same and np.concatenate([p[2] for p in parts])[o].tobytes() == want["cols"][0]["sum"].tobytes()
            results[name] = {"keys": same, "counts": cnt_ok, "sums_bit_exact": sum_ok, "owner_rank": owner_ok,
                             "groups": int(k.shape[0])}
            ok_all = ok_all and same and cnt_ok and sum_ok and owner_ok
    flag = torch.tensor([1 if ok_all else 0], device=eng.device)
    dist.broadcast(flag, 0)
    if rank == 0:
        print(json.dumps({"check": "dist shuffle vs oracle", "n_gpus": world, "rows": int(sum(k.shape[0] for k, _ in rows_all)),
                          "groups": args.groups, "variants": results, "ok": ok_all}), flush=True)
    eng.close()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
