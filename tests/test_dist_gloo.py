"""world_size-2 gloo tests (CPU) of the N>1 host logic in pyspark_tf_gke_b200.dist:
the all-to-all-v of partitioned {key, value} records, the merge of (sum, count)
partials and of label tables.  Per-rank compute is done by the oracle here (no
GPU); on a GPU box the same functions run over NCCL with the CUDA kernels."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import pyspark_tf_gke_b200.dist as D
import pyspark_tf_gke_b200.synth as S
from oracle import etl_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, groups, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # this rank's shard of the synthetic table (weak scaling: rows [rank*n, (rank+1)*n))
        keys, _ = S.host_column(3, 0, S.KIND_KEY64, n, row0=rank * n, modulus=groups)
        vals, vv = S.host_column(3, 1, S.KIND_EXACT, n, row0=rank * n, null_rate=0.05)
        # map side: partition by destination rank (oracle stand-in for etl_hash_partition)
        dest = O.partition_dest(keys, world)
        order = np.argsort(dest, kind="stable")
        rec = torch.from_numpy(np.stack([keys[order], vals[order].view(np.int64)], axis=1).copy())
        counts = torch.from_numpy(np.bincount(dest, minlength=world).astype(np.int64))
        got = D.exchange_records(rec, counts).numpy()
        assert (O.partition_dest(got[:, 0], world) == rank).all()  # every received key is ours
        # final merge on the receiver
        local = O.groupby_agg_fast(got[:, 0], [got[:, 1].copy().view(np.float64)], [None])
        # small reductions
        s, c = O.sum_count(vals, vv)
        sums, cnts = D.all_reduce_sum_count(torch.tensor([s], dtype=torch.float64), torch.tensor([c], dtype=torch.int64))
        names, lc = D.merge_label_tables(["b", f"r{rank}"], [2 + rank, 1])
        # a table too large for the fixed-size blob on ONE rank: every rank must take the object path together
        big = [f"label {rank} {i:05d} " + "x" * 40 for i in range(600 if rank == 1 else 3)]
        bn, bc = D.merge_label_tables(big + ["é shared"], [1] * len(big) + [5])
        assert len(bn) == 604 and bc[bn.index("é shared")] == 10 and sum(bc) == 613
        torch.save({"keys": local["keys"], "rows": local["rows"], "sum": local["cols"][0]["sum"], "sums": sums,
                    "cnts": cnts, "names": names, "lc": lc}, f"{out}/r{rank}.pt")
    finally:
        dist.destroy_process_group()


def test_exchange_and_merges_world_size_2(tmp_path):
    world, n, groups = 2, 50_000, 7_000
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n, groups, str(tmp_path)), nprocs=world, join=True)
    parts = [torch.load(f"{tmp_path}/r{r}.pt", weights_only=False) for r in range(world)]
    keys, _ = S.host_column(3, 0, S.KIND_KEY64, world * n, modulus=groups)
    vals, vv = S.host_column(3, 1, S.KIND_EXACT, world * n, null_rate=0.05)
    exp = O.groupby_agg_fast(keys, [vals], [None])
    k = np.concatenate([p["keys"] for p in parts])
    assert len(set(k.tolist())) == len(k)  # disjoint ownership after the exchange
    order = np.argsort(k)
    assert k[order].tolist() == exp["keys"].tolist()
    assert np.concatenate([p["rows"] for p in parts])[order].tolist() == exp["rows"].tolist()
    assert np.concatenate([p["sum"] for p in parts])[order].tobytes() == exp["cols"][0]["sum"].tobytes()
    s, c = O.sum_count(vals, vv)
    for p in parts:
        assert float(p["sums"][0]) == s and int(p["cnts"][0]) == c
        assert p["names"] == ["b", "r0", "r1"] and p["lc"] == [5, 1, 1]


def test_single_process_paths_are_identity():
    rec = torch.arange(10, dtype=torch.int64).view(5, 2)
    assert D.exchange_records(rec, torch.tensor([5])) is rec
    assert D.merge_label_tables(["a"], [3]) == (["a"], [3])
    assert D.world() == (0, 1)


def test_peer_layout_places_every_run_exactly_once():
    # bucketed peer exchange: receive buffers hold their buckets one after another, inside a
    # bucket the sources one after another; every rank derives its own runs from the gathered
    # count matrix - together the runs must tile each buffer without gap or overlap
    rng = np.random.default_rng(7)
    for size, buckets in ((1, 1), (2, 1), (2, 5), (4, 64), (8, 32)):
        m = rng.integers(0, 50, size=(size, size * buckets)).astype(np.int64)
        m[rng.random(m.shape) < 0.2] = 0
        matrix = torch.from_numpy(m)
        runs = {r: [] for r in range(size)}
        for rank in range(size):
            off, tot, start = D.peer_layout(matrix, rank, size, buckets)
            assert tot.shape == (size, buckets) and tot.numpy().tolist() == m.reshape(size, size, buckets).sum(0).tolist()
            for r in range(size):
                for b in range(buckets):
                    d = r * buckets + b
                    runs[r].append((int(off[d]), int(m[rank, d]), b, rank))
                    assert int(start[r, b]) <= int(off[d])
        for r in range(size):
            pos, last = 0, (-1, -1)
            for first, length, b, src in sorted(runs[r], key=lambda x: (x[2], x[3])):   # bucket-major, then source
                assert first == pos
                pos += length
            assert pos == int(m.reshape(size, size, buckets)[:, r, :].sum())


def test_plan_waves_covers_every_row_once_with_aligned_cuts():
    # BucketExchange's wave plan: contiguous, aligned, last wave the large one, tiny shards leave later waves empty
    for rows in (0, 1, 1000, 4096, 4097, 1_000_003, 125_000_000, 500_000_000):
        for waves in (1, 2, 3, 6):
            for head in (0.25, 0.1, 0.5):
                fracs, bounds = D.plan_waves(rows, waves, head)
                assert len(fracs) == waves and abs(sum(fracs) - 1.0) < 1e-12
                assert len(bounds) == waves + 1 and bounds[0] == 0 and bounds[-1] == rows
                assert all(a <= b for a, b in zip(bounds, bounds[1:]))
                assert all(b % 4096 == 0 or b == rows for b in bounds[1:-1])
                if waves > 1:
                    assert fracs[-1] == max(fracs) and abs(fracs[-1] - (1.0 - head)) < 1e-12
                    # the rows up to and including wave w are `head` of the rows up to and including wave w + 1
                    assert all(abs(sum(fracs[:w + 1]) - head * sum(fracs[:w + 2])) < 1e-12 for w in range(waves - 1))
    fr, b = D.plan_waves(125_000_000, 2)
    assert fr == [0.25, 0.75] and b == [0, 31_252_480, 125_000_000]
    assert D.plan_waves(1000, 3)[1] == [0, 1000, 1000, 1000]
    assert D.plan_waves(10, 0)[1] == [0, 10] and D.plan_waves(10, 2, head=7.0)[0] == [0.5, 0.5]
