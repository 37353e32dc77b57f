#!/usr/bin/env python
"""Headline benchmark of the B200-native raw-spark ETL hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

Headline workload = BASELINE.json configs[3]: groupBy(key)
.agg(count, sum) of 1 B rows (int64 key uniform over 100 M groups, fp64 value) -
hash partition -> exchange over NVLink -> local final merge - STRONG scaling: the
1 B rows are split over the N ranks, N = 1 is the same table on one GPU (no
exchange).  It is the largest named configuration that fits one GPU.  One step =
one full pass (every row partitioned, exchanged and merged into dense group
columns).  The Exchange being replaced sits under every aggregate of
k_means.py:23-48.

Printed JSON (one line, rank 0):
  value      rows/s over all ranks, inputs resident in HBM, CUDA events, max over ranks
  roofline   HBM roofline of the step on SURVEY.md §8d's algorithmic bytes (18.5 B/row
             at N = 1, 50.5 B/row at N > 1), with the bytes the three-pass design really
             moves beside it (design_*), per-phase rooflines, and the NVLink roofline
  checks     computed after the timed region on the last step's output: row
             conservation, grand total, and a per-key diff (count exact, sum bit-exact
             on the integer-valued table, key ownership = hash partition) of >= 1 M
             sampled keys against oracle/etl_oracle.py - the only place the N > 1
             exchange can be checked against the oracle on real hardware
  e2e        the same metric through dist.shuffle_groupby_sum_host: pinned HOST
             key / value buffers copied in and the group columns copied out inside
             the timed region
  sub        configs 2, 3 and 5-shape (one GPU each; config 5 on every rank) and the
             drop-in load_csv (train_tf_ps.py:75-149, N = 1), each with its roofline,
             clocks, an oracle check and its own end-to-end / CPU figures
  cpu_baseline  oracle/etl_oracle.c's groupBy (pthreads, all host cores) on a bounded
             sample of the same table (rank 0, N = 1 only)
`--impl reference`: PySpark local[N] is probed for (import pyspark, java, baseline/_ref);
this image has neither, so the arm times the oracle port on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from tools.clocks import ClockSampler  # noqa: E402

METRIC = "ETL rows/sec"
UNIT = "rows/s"
FILL = [1738.2895095771808, 1724.1278848441589, 1762.6607185343803]  # the reference's means
NVLINK_GBS = 900.0   # per direction per GPU


def measured_peak_gbs():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(name: str):
    """Per-launch DRAM bytes from the committed ncu --set full captures (profiles/traffic.json), else None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            return json.load(fh).get(name)
    except Exception:
        return None


def config4_workload(rows_total: int, groups: int) -> dict:
    """The `config` both arms print (same keys, same values)."""
    return {"workload": "config4: groupBy(key).agg(count, sum) of int64 key / fp64 value rows - hash partition, "
                        "exchange, final merge (the Exchange under k_means.py:23-48)",
            "rows_total": rows_total, "groups": groups, "value_mode": "exact (integer-valued fp64: sums bit-exact)"}


def probe_pyspark() -> dict:
    """SURVEY.md §8c run-time upgrade: can the reference's own engine run here?"""
    found = {"pyspark": False, "java": shutil.which("java") is not None,
             "baseline_ref": os.path.isdir(os.path.join(ROOT, "baseline", "_ref"))}
    try:
        ref = os.path.join(ROOT, "baseline", "_ref")
        if found["baseline_ref"] and ref not in sys.path:
            sys.path.insert(0, ref)
        import pyspark  # noqa: F401

        found["pyspark"] = True
    except Exception:
        pass
    found["runnable"] = found["pyspark"] and found["java"]
    return found


def pyspark_groupby_pass(keys: np.ndarray, vals: np.ndarray, cores: int):
    """Only reached when probe_pyspark() finds a JVM and pyspark: the reference
    engine's own groupBy under local[N] on the sample -> (seconds, keys, count, sum)."""
    import pandas as pd
    from pyspark.sql import SparkSession
    from pyspark.sql import functions as F

    spark = (SparkSession.builder.master(f"local[{cores}]").appName("bench-reference")
             .config("spark.sql.execution.arrow.pyspark.enabled", "true").getOrCreate())
    df = spark.createDataFrame(pd.DataFrame({"k": keys, "v": vals})).cache()
    df.count()
    t0 = time.perf_counter()
    out = df.groupBy("k").agg(F.count("v").alias("c"), F.sum("v").alias("s")).toPandas()
    dt = time.perf_counter() - t0
    return dt, out["k"].to_numpy(), out["c"].to_numpy(), out["s"].to_numpy()


# ----------------------------------------------------------------------------
# reference arm / CPU baseline: the oracle port on the host cores
# ----------------------------------------------------------------------------

def cpu_groupby_timer(n: int, groups: int, seed: int):
    """-> (run_once() -> seconds, threads): oracle/etl_oracle.c's groupBy on the first n rows of the table."""
    from oracle import c_oracle as CO

    keys = CO.synth_column(seed, 0, 3, n, modulus=groups)
    vals = CO.synth_column(seed, 1, 0, n)
    scratch = (np.empty(n, np.int64), np.empty(n, np.int64), np.empty(n, np.float64))
    state = {}

    def run_once():
        t0 = time.perf_counter()
        k, c, s = CO.groupby_sum(keys, vals, scratch)
        dt = time.perf_counter() - t0
        state["groups"], state["rows"], state["sum"] = k.shape[0], int(c.sum()), float(s.sum())
        return dt

    return run_once, state, CO.threads(), (keys, vals)


def cpu_sample_size(rows_total: int, groups: int, budget_s: float, passes: int):
    """Largest power-of-ten-ish sample (same 10 rows per group) whose `passes` passes fit the budget, by a
    calibration pass on 10 M rows; capped by host memory (64 B per row of working set)."""
    from oracle import c_oracle as CO  # noqa: F401

    cal = 10_000_000
    run, _, _, _ = cpu_groupby_timer(cal, max(1, cal * groups // rows_total), 3)
    run()
    per_row = min(run(), run()) / cal
    n = int(budget_s / (passes * per_row * 1.3))     # 1.3: larger samples leave the caches
    try:
        avail = int(next(l for l in open("/proc/meminfo") if l.startswith("MemAvailable")).split()[1]) * 1024
        n = min(n, avail // 2 // 64)
    except Exception:
        pass
    n = max(1_000_000, min(n, rows_total, 200_000_000))
    return n // 1_000_000 * 1_000_000


def run_reference(args):
    """`--impl reference`: PySpark local[N] if the probe finds it, else the oracle port
    (oracle/etl_oracle.c, pthreads over every host core) of the same step on a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import pyspark_tf_gke_b200.synth as S

    cfg = S.CONFIG4
    rows_total, groups = args.rows or cfg["rows"], args.groups or cfg["groups"]
    probe = probe_pyspark()
    total_steps = args.steps + args.warmup
    n = cpu_sample_size(rows_total, groups, 100.0, total_steps)
    g = max(1, n * groups // rows_total)
    run, state, threads, (keys, vals) = cpu_groupby_timer(n, g, cfg["seed"])
    kind, note = "port", ("PySpark local[N] not runnable on this image (probe: pyspark %s, java %s, baseline/_ref %s); "
                          "oracle port oracle/etl_oracle.c on all host cores"
                          % (probe["pyspark"], probe["java"], probe["baseline_ref"]))
    if probe["runnable"]:
        try:
            def run():  # noqa: F811
                dt, k, c, s = pyspark_groupby_pass(keys, vals, threads)
                state["groups"], state["rows"], state["sum"] = k.shape[0], int(c.sum()), float(s.sum())
                return dt
            run()
            kind, note = "reference", f"PySpark local[{threads}] groupBy on the sample"
        except Exception as exc:  # fall back to the port, say why
            run = cpu_groupby_timer(n, g, cfg["seed"])[0]
            note = f"PySpark found but failed ({type(exc).__name__}: {exc}); oracle port instead"
    for _ in range(args.warmup):
        run()
    t0 = time.perf_counter()
    timed = 0.0
    for _ in range(args.steps):
        timed += run()
    dt = time.perf_counter() - t0
    assert state["rows"] == n
    value = n * args.steps / timed
    sample = f"first {n} rows / {g} groups of the config-4 table per step (same 10 rows per group)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * timed / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config4_workload(rows_total, groups),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample, "note": note,
                         "wall_s": dt},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "pyspark_probe": probe,
    }
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------

class Ctx:
    def __init__(self, args):
        import torch
        import torch.distributed as dist

        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run for --gpus > 1")
        torch.cuda.set_device(self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        import pyspark_tf_gke_b200 as pkg

        self.pkg = pkg
        self.eng = pkg.Engine(self.local)
        self.device = self.eng.device

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        if self.world == 1:
            return float(x)
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, x: int) -> int:
        if self.world == 1:
            return int(x)
        t = self.torch.tensor([x], dtype=self.torch.int64, device=self.device)
        self.dist.all_reduce(t)
        return int(t.item())

    def gather_var(self, t):
        """Variable-length 1-D device tensors of every rank -> list of host arrays (on every rank)."""
        torch, dist = self.torch, self.dist
        if self.world == 1:
            return [t.cpu().numpy()]
        sizes = torch.zeros(self.world, dtype=torch.int64, device=self.device)
        sizes[self.rank] = t.shape[0]
        dist.all_reduce(sizes)
        sizes = sizes.cpu().tolist()
        m = max(max(sizes), 1)
        padded = torch.zeros(m, dtype=t.dtype, device=self.device)
        padded[:t.shape[0]] = t
        out = torch.empty(self.world * m, dtype=t.dtype, device=self.device)
        dist.all_gather_into_tensor(out, padded)
        host = out.cpu().numpy().reshape(self.world, m)
        return [host[r, :sizes[r]].copy() for r in range(self.world)]


def timed_steps(ctx, step, steps: int, warmup: int):
    """W warm-ups, then K steps bracketed by barrier + synchronize, CUDA events per step, clocks sampled
    during the timed region -> (total_ms max over ranks, per-step ms of this rank, clocks).  Leaves the
    number of this library's kernels launched inside the timed region in ctx.launches."""
    torch = ctx.torch
    for _ in range(max(warmup, 3)):
        step()
    ctx.barrier()
    sampler = ClockSampler(ctx.local)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ctx.barrier()
    sampler.start()
    l0 = ctx.eng.launches
    ev[0].record()
    for i in range(steps):
        step()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ctx.launches = ctx.eng.launches - l0
    clocks = sampler.stop()
    total_ms = ctx.max_over_ranks(ev[0].elapsed_time(ev[-1]))
    per = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
    return total_ms, per, clocks


def pcie_peaks(ctx, nbytes: int = 1 << 30):
    """Measured pinned-memory H2D / D2H GB/s on this box (for e2e.pcie_frac)."""
    torch = ctx.torch
    h = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    d = torch.empty(nbytes, dtype=torch.uint8, device=ctx.device)
    out = {}
    for name, (dst, src) in {"h2d": (d, h), "d2h": (h, d)}.items():
        dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        best = 1e30
        for _ in range(3):
            t0 = time.perf_counter()
            dst.copy_(src, non_blocking=True)
            torch.cuda.synchronize()
            best = min(best, time.perf_counter() - t0)
        out[name] = nbytes / best / 1e9
    return out


# ------------------------------------------------------------------ config 4
def run_config4(ctx, args):
    torch = ctx.torch
    import pyspark_tf_gke_b200.dist as D
    import pyspark_tf_gke_b200.synth as S
    from oracle import etl_oracle as O     # the checker of `checks` below - never inside a timed region

    eng, world, rank = ctx.eng, ctx.world, ctx.rank
    cfg = S.CONFIG4
    rows_total, G = args.rows or cfg["rows"], args.groups or cfg["groups"]
    n = rows_total // world // 32 * 32
    rows_total = n * world
    row0 = rank * n
    key = eng.synth_column(cfg["seed"], 0, S.KIND_KEY64, n, row0=row0, modulus=G, with_valid=False).data
    val = eng.synth_column(cfg["seed"], 1, S.KIND_EXACT, n, row0=row0, with_valid=False).data
    cap = G // world + G // (world * 8) + 4096
    peer = D.BucketExchange(eng, n, groups_hint=G) if world > 1 else None
    if peer is not None:
        peer.timing = True               # events around the NVLink transfer of every step (first copy starts .. last group landed)
    state = {}

    def step():
        marks = {}

        def mark(name):
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            marks[name] = e
        mark("start")
        state["pg"] = None
        state["pg"] = D.shuffle_groupby_sum(eng, key, val, cap, groups_hint=G, peer=peer, mark=mark)
        state["marks"] = marks
        state.setdefault("all_marks", []).append(marks)
        if peer is not None:
            state.setdefault("windows", []).append(peer.window)

    total_ms, per_step, clocks = timed_steps(ctx, step, args.steps, args.warmup)
    launches = ctx.sum_over_ranks(ctx.launches)
    # phases of the timed steps (mean over steps, max over ranks)
    order = ["start", "init", "partition", "exchange", "level1", "level2", "merge"]
    phase_ms = {}
    for a, b in zip(order[:-1], order[1:]):
        vals_ms = [m[a].elapsed_time(m[b]) for m in state["all_marks"][-args.steps:] if a in m and b in m]
        if vals_ms:
            phase_ms[b] = ctx.max_over_ranks(statistics.mean(vals_ms))
    ms_per_step = total_ms / args.steps
    value = rows_total / (ms_per_step * 1e-3)
    transfer_ms = None
    if peer is not None:
        transfer_ms = ctx.max_over_ranks(statistics.mean(a.elapsed_time(b) for a, b in state["windows"][-args.steps:]))

    # ---- checks (not timed): the last step's output against the oracle
    res = eng.groupby_result(state["pg"], sort=False)
    gkeys, grows, gsum = res["keys"], res["rows"], res["cols"][0]["sum"]
    groups_found = ctx.sum_over_ranks(gkeys.shape[0])
    rows_found = ctx.sum_over_ranks(int(grows.sum().item()))
    col_sum, _ = eng.sum_count([ctx.pkg.Column(val)])
    tot = torch.stack([gsum.sum(), col_sum[0]])
    if world > 1:
        ctx.dist.all_reduce(tot)
    T = max(G // 64, min(G, 1 << 20))                    # sample: every key below T (>= 1 M keys of config 4)
    in_sel = key < T
    in_k, in_v = ctx.gather_var(key[in_sel]), ctx.gather_var(val[in_sel].view(torch.int64))
    out_sel = gkeys < T
    out_k, out_r = ctx.gather_var(gkeys[out_sel]), ctx.gather_var(grows[out_sel])
    out_s = ctx.gather_var(gsum[out_sel].view(torch.int64))
    checks = {"rows_conserved": rows_found == rows_total, "grand_total_exact": float(tot[0].item()) == float(tot[1].item()),
              "distinct_keys_found": groups_found, "spilled_records": int(res.get("spilled_records", 0))}
    if rank == 0:
        want = O.groupby_agg_fast(np.concatenate(in_k), [np.concatenate(in_v).view(np.float64)], [None])
        got_k = np.concatenate(out_k)
        o = np.argsort(got_k, kind="stable")
        got_k, got_r, got_s = got_k[o], np.concatenate(out_r)[o], np.concatenate(out_s)[o]
        same_keys = got_k.shape == want["keys"].shape and bool((got_k == want["keys"]).all())
        checks.update({
            "sample_keys": int(want["keys"].shape[0]), "sample_rows": int(sum(a.shape[0] for a in in_k)),
            "sample_keys_match": same_keys,
            "sample_counts_exact": same_keys and bool((got_r == want["rows"]).all()),
            "sample_sums_bit_exact": same_keys and got_s.tobytes() == want["cols"][0]["sum"].view(np.int64).tobytes(),
            "sample_keys_on_owner_rank": all(bool((O.partition_dest(out_k[r], world) == r).all()) for r in range(world)),
            "oracle": "oracle/etl_oracle.py groupby_agg_fast + partition_dest",
        })
        checks["all_ok"] = all(v for k, v in checks.items() if isinstance(v, bool))
    del res, gkeys, grows, gsum, in_sel, out_sel
    state.clear()

    # ---- end to end: pinned host buffers in, group columns out, through the public host API
    e2e = None
    if not args.no_e2e:
        try:
            h_key = torch.empty(n, dtype=torch.int64, pin_memory=True)
            h_val = torch.empty(n, dtype=torch.float64, pin_memory=True)
            h_key.copy_(key)
            h_val.copy_(val)
            h_out = (torch.empty(cap, dtype=torch.int64, pin_memory=True), torch.empty(cap, dtype=torch.int64, pin_memory=True),
                     torch.empty(cap, dtype=torch.float64, pin_memory=True))
            torch.cuda.synchronize()
            del key, val
            key = val = None
            e2e_steps = max(2, min(args.steps, 3))
            g_local = 0
            for _ in range(2):
                g_local = D.shuffle_groupby_sum_host(eng, h_key, h_val, cap, groups_hint=G, peer=peer, out=h_out)
            ctx.barrier()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                g_local = D.shuffle_groupby_sum_host(eng, h_key, h_val, cap, groups_hint=G, peer=peer, out=h_out)
            torch.cuda.synchronize()
            e2e_s = ctx.max_over_ranks(time.perf_counter() - t0)
            h2d, d2h = ctx.sum_over_ranks(n * 16), ctx.sum_over_ranks(g_local * 24)
            ok = ctx.sum_over_ranks(g_local) == groups_found and ctx.sum_over_ranks(int(h_out[1][:g_local].sum())) == rows_total
            pcie = pcie_peaks(ctx)
            floor_s = max(n * 16 / pcie["h2d"], g_local * 24 / pcie["d2h"]) / 1e9
            e2e = {"value": rows_total * e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                   "steps": e2e_steps, "ms_per_step": 1e3 * e2e_s / e2e_steps,
                   "api": "dist.shuffle_groupby_sum_host (pinned host key/value in, key/count/sum columns out)",
                   "output_matches_device_run": bool(ok), "pcie_gbs": pcie,
                   "pcie_floor_ms": 1e3 * floor_s, "pcie_frac": floor_s / (e2e_s / e2e_steps)}
            del h_key, h_val, h_out
        except Exception as exc:   # e.g. the host cannot pin 16 GB
            e2e = {"value": None, "unit": UNIT, "error": f"{type(exc).__name__}: {exc}"}
    slab_bytes_moved = 16.0 * sum(peer.slabs) * peer.buckets * (world - 1) if world > 1 else 0.0
    waves = peer.waves if world > 1 else 0
    del key, val, peer

    line = None
    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        alg_per_row = 18.5 if world == 1 else 50.5
        alg = alg_per_row * n                             # per GPU per step
        achieved = alg / (ms_per_step * 1e-3) / 1e9
        # what the three-pass design moves per GPU: partition levels read + write 16 B records,
        # the merge reads them once and writes 32 B per group
        g_local = groups_found / world
        design = {"level1": 32.0 * n, "level2": 32.0 * n, "merge": 16.0 * n + 32.0 * g_local}
        if world > 1:
            # N > 1 (dist.BucketExchange): "partition" = the local (rank, bucket) scatter - the first level, done by the
            # sender, wave by wave (the slabs of wave w already cross NVLink while wave w + 1 is being cut, so this
            # phase overlaps the transfer); "exchange" = what is left of the slab copies over NVLink (send buffer
            # read, receive buffer written) after the last wave has been cut, with the second level and the merge
            # of the bucket groups that have landed running behind them; what is left of those after the last group
            # has landed shows up as the (small) level2 / merge phases.  nvlink.transfer_ms is the whole transfer.
            design = {"partition": 32.0 * n, "exchange": 32.0 * n + 32.0 * n + 16.0 * n + 32.0 * g_local}
        phases = {}
        for name, ms in phase_ms.items():
            entry = {"ms": ms}
            if name in design and ms > 0:
                entry["design_bytes"] = design[name]
                entry["gbs"] = design[name] / (ms * 1e-3) / 1e9
                entry["frac"] = entry["gbs"] / peak
            phases[name] = entry
        design_total = sum(design.values())
        roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu_traffic("config4_step@1gpu") if world == 1 else None, "peak_source": peak_src,
                "kernel": "whole step: scatter_pipelined_kernel x2 (two partition levels) + groupby_merge_kernel",
                "algorithmic_bytes_per_row": alg_per_row, "algorithmic_bytes_per_step_per_gpu": alg,
                "design_bytes_per_row": design_total / n, "design_gbs": design_total / (ms_per_step * 1e-3) / 1e9,
                "design_frac": design_total / (ms_per_step * 1e-3) / 1e9 / peak, "phases": phases,
                "mean_step_ms": statistics.mean(per_step), "min_step_ms": min(per_step)}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": dict(config4_workload(rows_total, G), rows_per_gpu=n,
                           l2=f"no flush: {16 * n / 1e9:.1f} GB of input per GPU per step >> 126 MB L2",
                           parallelism=(f"rows split over {world} GPUs; each rank partitions its rows by (owner rank, bucket) in "
                                        f"{waves} waves, stores the slabs of a wave into the owners' symmetric-memory buffers "
                                        "over NVLink bucket group by bucket group while the next wave is being cut, and "
                                        "merges every group as soon as its last wave has landed")
                           if world > 1 else "one GPU, no exchange"),
            "roofline": roof, "checks": checks, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
        }
        if world > 1 and transfer_ms:
            sent = 16.0 * (n - n / world)
            gbs = sent / (transfer_ms * 1e-3) / 1e9
            line["nvlink"] = {"bound": "nvlink", "sent_bytes_per_gpu": sent, "transfer_ms": transfer_ms,
                              "exchange_ms": phase_ms.get("exchange"), "waves": waves,
                              "achieved": gbs, "peak": NVLINK_GBS, "unit": "GB/s", "frac": gbs / NVLINK_GBS,
                              "slab_room_bytes_per_gpu": slab_bytes_moved,
                              "note": "sent = payload records leaving the GPU (only the filled part of a slab moves); transfer_ms = "
                                      "first copy kernel of the first wave starts .. last bucket group of the last wave has landed "
                                      "on every rank (CUDA events, mean over the timed steps, max over ranks)"}
    return line


# ------------------------------------------------------------------ sub lines
def sub_config2(ctx, steps: int):
    """configs[1]: 10 M-row filter + 3 x impute/cast projection (k_means.py:27-28,49-51), one launch per step."""
    torch, eng = ctx.torch, ctx.eng
    import pyspark_tf_gke_b200.synth as S
    from oracle import c_oracle as CO, etl_oracle as O

    cfg = S.CONFIG2
    n = cfg["rows"]
    key = eng.synth_column(cfg["seed"], 0, S.KIND_KEY32, n, modulus=cfg["categories"], null_rate=cfg["key_null"])
    cols = [eng.synth_column(cfg["seed"], 1 + c, S.KIND_EXACT, n, null_rate=cfg["val_null"][c]) for c in range(3)]
    out_code = eng.empty(n, torch.int32)
    out_vals = [eng.empty(n, torch.float64) for _ in range(3)]
    out_rows = eng.empty(1, torch.int64)
    _, per, clocks = timed_steps(ctx, lambda: eng.project_filter_impute(key, cols, FILL, out_code, out_vals, out_rows), steps, 3)
    kept = int(out_rows.item())
    words = key.valid.shape[0]
    alg = n * 28 + 4 * words * 4 + kept * 28
    peak, _ = measured_peak_gbs()
    mean_ms = statistics.mean(per)
    # oracle check on the first 1 M rows (same generator on the host)
    m = 1_000_000
    code, kv = S.host_column(cfg["seed"], 0, S.KIND_KEY32, m, modulus=cfg["categories"], null_rate=cfg["key_null"])
    hv = [S.host_column(cfg["seed"], 1 + c, S.KIND_EXACT, m, null_rate=cfg["val_null"][c]) for c in range(3)]
    exp_code, exp_vals = O.project_filter_impute(code, kv, [v for v, _ in hv], [vv for _, vv in hv], FILL)
    k = exp_code.shape[0]
    ok = out_code[:k].cpu().numpy().tobytes() == exp_code.tobytes() and all(
        a[:k].cpu().numpy().tobytes() == b.tobytes() for a, b in zip(out_vals, exp_vals))
    # host-buffer entry (pinned), copies inside the timed region
    h_in = [torch.empty(n, dtype=torch.int32, pin_memory=True), torch.empty(words, dtype=torch.int32, pin_memory=True)]
    h_vals = [torch.empty(n, dtype=torch.float64, pin_memory=True) for _ in range(3)]
    h_vv = [torch.empty(words, dtype=torch.int32, pin_memory=True) for _ in range(3)]
    h_oc = torch.empty(n, dtype=torch.int32, pin_memory=True)
    h_ov = [torch.empty(n, dtype=torch.float64, pin_memory=True) for _ in range(3)]
    h_in[0].copy_(key.data)
    h_in[1].copy_(key.valid)
    for c in range(3):
        h_vals[c].copy_(cols[c].data)
        h_vv[c].copy_(cols[c].valid)
    torch.cuda.synchronize()
    np_args = (h_in[0].numpy(), h_in[1].numpy().view(np.uint32), [t.numpy() for t in h_vals],
               [t.numpy().view(np.uint32) for t in h_vv], FILL, h_oc.numpy(), [t.numpy() for t in h_ov])
    for _ in range(2):
        rows = eng.project_filter_impute_host(*np_args)
    t0 = time.perf_counter()
    for _ in range(5):
        eng.project_filter_impute_host(*np_args)
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / 5
    pcie = pcie_peaks(ctx, 1 << 28)
    h2d, d2h = n * 28 + 4 * words * 4, kept * 28
    floor_s = max(h2d / pcie["h2d"], d2h / pcie["d2h"]) / 1e9
    # CPU: the compiled oracle on the same table
    hk, hkv = h_in[0].numpy(), h_in[1].numpy().view(np.uint32)
    oc, ov = np.empty(n, np.int32), [np.empty(n, np.float64) for _ in range(3)]
    cpu = 1e30
    for _ in range(3):
        t0 = time.perf_counter()
        CO.project_filter_impute(hk, hkv, np_args[2], np_args[3], FILL, oc, ov)
        cpu = min(cpu, time.perf_counter() - t0)
    return {"workload": "config2: 10M-row filter + impute/cast projection (k_means.py:27-28,49-51)", "rows": n, "kept_rows": kept,
            "value": n / (mean_ms * 1e-3), "unit": UNIT, "ms_per_step": mean_ms, "steps": steps,
            "roofline": {"bound": "hbm", "achieved": alg / (mean_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg / (mean_ms * 1e-3) / 1e9 / peak, "kernel": "project_filter_impute_kernel",
                         "algorithmic_bytes_per_launch": alg, "traffic": ncu_traffic("project_filter_impute_kernel@config2")},
            "e2e": {"value": n / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "etl_project_filter_impute_host (pinned host buffers)", "rows_match": rows == kept,
                    "pcie_gbs": pcie, "pcie_floor_ms": 1e3 * floor_s, "pcie_frac": floor_s / e2e_s},
            "cpu_baseline": {"value": n / cpu, "unit": UNIT, "cores": CO.threads(), "kind": "port",
                             "sample": "the full 10M-row table, best of 3 passes of oracle/etl_oracle.c"},
            "checks": {"first_1M_rows_bit_exact_vs_oracle": bool(ok)}, "clocks": clocks}


def sub_config3(ctx, steps: int):
    """configs[2]: 100 M rows, 1 K groups, sum/count/mean of 3 fp64 columns with nulls, one GPU."""
    torch, eng = ctx.torch, ctx.eng
    import pyspark_tf_gke_b200.synth as S
    from oracle import etl_oracle as O

    cfg = S.CONFIG3
    n, G = cfg["rows"], cfg["groups"]
    key = eng.synth_column(cfg["seed"], 0, S.KIND_KEY32, n, modulus=G, with_valid=False)
    cols = [eng.synth_column(cfg["seed"], 1 + c, S.KIND_EXACT, n, null_rate=cfg["val_null"][c]) for c in range(3)]
    state = {}

    def step():
        gt = eng.groupby_new(4096, 3)
        eng.groupby_accumulate(gt, key.data, None, [c.data for c in cols], [c.valid for c in cols], groups_hint=G, key_range=G)
        state["gt"] = gt

    _, per, clocks = timed_steps(ctx, step, steps, 3)
    res = eng.groupby_result(state["gt"], sort=True)
    # oracle diff, group by group, on every row of 10 sampled keys (~1 M rows)
    idx = torch.nonzero(key.data < 10).flatten()
    hk = key.data[idx].cpu().numpy()
    hv = [c.data[idx].cpu().numpy() for c in cols]
    hvv = [(((c.valid[idx >> 5] >> (idx & 31).to(torch.int32)) & 1) != 0).cpu().numpy() for c in cols]
    want = O.groupby_agg_fast(hk, hv, hvv)
    ok = res["keys"][:10].tolist() == want["keys"].tolist() and res["rows"][:10].tolist() == want["rows"].tolist()
    for c in range(3):
        ok = ok and res["cols"][c]["count"][:10].tolist() == want["cols"][c]["count"].tolist()
        ok = ok and res["cols"][c]["sum"][:10].cpu().numpy().tobytes() == want["cols"][c]["sum"].tobytes()
    ok = ok and res["keys"].shape[0] == G and int(res["rows"].sum()) == n
    alg = n * 28 + 3 * ((n + 31) // 32) * 4
    peak, _ = measured_peak_gbs()
    mean_ms = statistics.mean(per)
    return {"workload": "config3: 100M-row groupBy on a 1K-group int32 key, 3 x sum/count/mean", "rows": n, "groups": G,
            "value": n / (mean_ms * 1e-3), "unit": UNIT, "ms_per_step": mean_ms, "steps": steps,
            "roofline": {"bound": "hbm", "achieved": alg / (mean_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg / (mean_ms * 1e-3) / 1e9 / peak, "kernel": "groupby_dense_kernel<3>",
                         "algorithmic_bytes_per_launch": alg, "traffic": ncu_traffic("groupby_dense_kernel@config3")},
            "checks": {"sampled_groups_exact_vs_oracle": bool(ok), "sample_rows": int(hk.shape[0])}, "clocks": clocks}


def sub_config5(ctx, steps: int, rows: int = 10_000_000):
    """configs[4] shape: CSV text -> decode -> null count, means, filter/impute, StringIndexer -> features
    f32 [N, 153]; one byte range of the text per rank (weak over ranks), text resident in HBM; the upload of
    the text from host memory is timed apart (e2e)."""
    torch, eng, pkg = ctx.torch, ctx.eng, ctx.pkg
    import gzip

    with gzip.open(os.path.join(ROOT, "tests", "golden", "health_workloads.csv.gz"), "rb") as fh:
        raw = fh.read()
    head, body = raw.split(b"\n", 1)
    reps = max(1, rows // 18155)
    raw, rows = head + b"\n" + body * reps, 18155 * reps
    pkg.table.upload_text(eng, raw[:1 << 20])       # creates the engine's pinned ring (one-time cost, not per file)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    text = pkg.table.upload_text(eng, raw)
    torch.cuda.synchronize()
    h2d_s = time.perf_counter() - t0
    wl = pkg.KMeansWorkload(eng, fit_kmeans=False)      # the ETL stage only (k_means.py:22-77)
    wl.logger.setLevel("ERROR")
    state = {}

    def step(device_text=text):
        table = pkg.read_csv(eng, raw, ["value", "lower_ci", "upper_ci"], ["measure_name"], device_text=device_text)
        model, _ = wl.k_means(table)
        state["F"], state["model"] = model.features(torch.float32), model

    total_ms, per, clocks = timed_steps(ctx, step, steps, 3)
    launches = ctx.launches // steps
    model, F = state["model"], state["F"]
    # known answers of the default input, unchanged by repeating the body
    ok = (len(model.labels) == 30 and model.null_key_count == 0 and model.rows_after_filter == ctx.world * rows and
          all(abs(a - b) <= 1e-9 * abs(b) for a, b in zip(model.means, FILL)) and tuple(F.shape) == (rows, 153))
    nbytes = len(raw)
    alg = nbytes + 2 * 28.5 * rows + 612 * rows
    peak, _ = measured_peak_gbs()
    ms = total_ms / steps
    # end to end: the CSV bytes start in HOST memory (upload inside the timed region), features end on the device
    # and the stage's scalars (counts, means, label table) on the host
    del text
    e2e_steps = 3
    step(device_text=None)
    ctx.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step(device_text=None)
    torch.cuda.synchronize()
    e2e_s = ctx.max_over_ranks(time.perf_counter() - t0) / e2e_steps
    out = {"workload": "config5 shape: CSV text -> decode -> aggregate/project/index -> features f32 [N,153]",
           "records_per_gpu": rows, "text_bytes_per_gpu": nbytes, "n_gpus": ctx.world,
           "value": ctx.world * rows / (ms * 1e-3), "unit": "records/s", "ms_per_step": ms, "steps": steps,
           "text_gb_per_s": ctx.world * nbytes / 1e9 / (ms * 1e-3), "launches_per_step": launches,
           "roofline": {"bound": "hbm", "achieved": alg / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": alg / (ms * 1e-3) / 1e9 / peak, "kernel": "whole stage (decode + project + features)",
                        "algorithmic_bytes_per_step_per_gpu": alg},
           "upload": {"seconds": h2d_s, "gb_per_s": nbytes / 1e9 / h2d_s,
                      "api": "table.upload_text (host bytes -> pinned ring -> HBM)"},
           "e2e": {"value": ctx.world * rows / e2e_s, "unit": "records/s", "ms_per_step": 1e3 * e2e_s, "steps": e2e_steps,
                   "h2d_bytes_per_step": ctx.world * nbytes, "d2h_bytes_per_step": 0,
                   "api": "read_csv(host bytes) -> KMeansWorkload.k_means -> features(float32) on the device"},
           "checks": {"known_answers_of_default_input": bool(ok), "means": model.means}, "clocks": clocks}
    if ctx.rank == 0 and ctx.world == 1:
        out["cpu_baseline"] = cpu_config5(raw, rows)
    return out


def sub_load_csv(ctx, steps: int, rows: int = 10_000_000):
    """a11 / a12 (train_tf_ps.py:75-149, the one row whose parity is pinned to the reference's own outputs): CSV
    text -> `X float32[N, 3]`, `y int32[N]`, sorted vocabulary through the drop-in `load_csv`; the reference does
    this in a single-threaded Python loop.  `value` with the text resident in HBM, `e2e` = `load_csv(host bytes)`
    -> NumPy arrays in host memory; checked bit for bit against the oracle's restatement on a 1 M-row sample."""
    torch, eng, pkg = ctx.torch, ctx.eng, ctx.pkg
    import gzip

    import pyspark_tf_gke_b200.train_input as TI
    from oracle import etl_oracle as O

    with gzip.open(os.path.join(ROOT, "tests", "golden", "health_workloads.csv.gz"), "rb") as fh:
        raw = fh.read()
    head, body = raw.split(b"\n", 1)
    reps = max(1, rows // 18155)
    sample_reps = min(reps, 55)
    sample = head + b"\n" + body * sample_reps
    raw, rows = head + b"\n" + body * reps, 18155 * reps
    TI._engine = eng                                   # load_csv's default engine = this process's
    text = pkg.table.upload_text(eng, raw)
    state = {}

    def step():
        state["out"] = TI.load_csv_device(raw, engine=eng, device_text=text)

    total_ms, per, clocks = timed_steps(ctx, step, steps, 3)
    launches = ctx.launches // steps
    X, y, vocab = state["out"]
    kept = int(X.shape[0])
    ms = total_ms / steps
    del text, X, y
    state.clear()
    TI.load_csv(raw)
    e2e_steps = 3
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        hX, hy, hvocab = TI.load_csv(raw)
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    # parity + the CPU arm on the same sample: the reference's algorithm as restated in oracle/etl_oracle.py (one core)
    sX, sy, svocab = TI.load_csv(sample)
    t0 = time.perf_counter()
    wX, wy, wvocab = O.load_csv(sample.decode("utf-8"))
    cpu_s = time.perf_counter() - t0
    ok = (sX.dtype == np.float32 and sy.dtype == np.int32 and sX.tobytes() == wX.tobytes() and sy.tobytes() == wy.tobytes() and
          svocab == wvocab and hvocab == wvocab and hX.shape == (kept, 3) and kept == wX.shape[0] * reps // sample_reps)
    nbytes = len(raw)
    alg = nbytes + 2 * 28.5 * rows + 16.0 * kept      # text read; decoded columns written and read again; X / y written
    peak, _ = measured_peak_gbs()
    return {"workload": "load_csv (train_tf_ps.py:75-149): CSV text -> X f32 [N,3], y i32 [N], vocab", "records": rows,
            "kept_rows": kept, "labels": len(vocab), "text_bytes": nbytes, "value": rows / (ms * 1e-3), "unit": "records/s",
            "ms_per_step": ms, "steps": steps, "launches_per_step": launches,
            "roofline": {"bound": "hbm", "achieved": alg / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg / (ms * 1e-3) / 1e9 / peak, "kernel": "whole function (decode + pack)",
                         "algorithmic_bytes_per_step": alg},
            "e2e": {"value": rows / e2e_s, "unit": "records/s", "ms_per_step": 1e3 * e2e_s, "steps": e2e_steps,
                    "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": 16 * kept,
                    "api": "train_input.load_csv(host bytes) -> NumPy X, y in host memory, vocab"},
            "cpu_baseline": {"value": 18155 * sample_reps / cpu_s, "unit": "records/s", "cores": 1, "kind": "port",
                             "sample": f"{18155 * sample_reps} records of the same text through oracle/etl_oracle.py load_csv "
                                       "(the reference's per-row Python loop, restated; golden-pinned to the reference)"},
            "checks": {"bit_identical_to_oracle_on_sample": bool(ok), "sample_records": 18155 * sample_reps},
            "clocks": clocks}


def cpu_config5(raw: bytes, rows: int, sample_reps: int = 55):
    """The same stage on the host cores, bounded sample (~1 M records): pyarrow.csv (multi-threaded) decode,
    pyarrow dictionary encode + the compiled oracle's projection (all cores), NumPy feature assembly."""
    import pyarrow as pa
    import pyarrow.compute as pc
    import pyarrow.csv as pacsv
    from oracle import c_oracle as CO, etl_oracle as O

    head, body = raw.split(b"\n", 1)
    block = body[:len(body) // max(1, rows // 18155)]
    text = head + b"\n" + block * sample_reps
    n = 18155 * sample_reps
    best = 1e30
    for _ in range(2):
        t0 = time.perf_counter()
        t = pacsv.read_csv(pa.BufferReader(text), convert_options=pacsv.ConvertOptions(
            include_columns=["measure_name", "value", "lower_ci", "upper_ci"], strings_can_be_null=True))
        key = pc.dictionary_encode(t["measure_name"]).combine_chunks()
        codes = key.indices.to_numpy(zero_copy_only=False).astype(np.int32)
        kv = O.pack_validity(~np.asarray(key.is_null()))
        vals, valid, means = [], [], []
        for c in ("value", "lower_ci", "upper_ci"):
            col = t[c].combine_chunks()
            v = col.to_numpy(zero_copy_only=False).astype(np.float64)
            ok = ~np.asarray(col.is_null())
            vals.append(np.nan_to_num(v))
            valid.append(O.pack_validity(ok))
            s, k = CO.sum_count(vals[-1], valid[-1], kv)
            means.append(s / k)
        kept, out_code, out_vals = CO.project_filter_impute(codes, kv, vals, valid, means)
        counts = np.bincount(out_code[:kept], minlength=len(key.dictionary))
        order = np.lexsort((np.asarray(key.dictionary.to_pylist(), dtype=object), -counts))
        remap = np.empty(len(order), np.int32)
        remap[order] = np.arange(len(order), dtype=np.int32)
        F = O.assemble_features_dense(remap[out_code[:kept]], len(order), [v[:kept] for v in out_vals], 5, dtype=np.float32)
        best = min(best, time.perf_counter() - t0)
    assert F.shape == (n, 153)
    return {"value": n / best, "unit": "records/s", "cores": CO.threads(), "kind": "port",
            "sample": f"{n} records ({len(text) / 1e6:.0f} MB of the same text): pyarrow.csv decode ({pa.cpu_count()} threads) + "
                      "oracle/etl_oracle.c sums and projection + NumPy feature assembly, best of 2"}


def run_ours(args):
    ctx = Ctx(args)
    torch = ctx.torch
    line = run_config4(ctx, args)
    torch.cuda.empty_cache()
    sub = {}
    if not args.no_sub:
        for name, fn in (("config2", sub_config2), ("config3", sub_config3), ("config5", sub_config5), ("load_csv", sub_load_csv)):
            if ctx.world > 1 and name != "config5":
                continue      # single-GPU configurations: reported by the N = 1 run
            try:
                sub[name] = fn(ctx, args.sub_steps)
            except Exception as exc:
                sub[name] = {"error": f"{type(exc).__name__}: {exc}"}
            torch.cuda.empty_cache()
    if ctx.rank == 0:
        line["sub"] = sub
        if ctx.world == 1 and not args.no_cpu_baseline:
            import pyspark_tf_gke_b200.synth as S

            rows_total, G = line["config"]["rows_total"], line["config"]["groups"]
            n = cpu_sample_size(rows_total, G, 20.0, 4)
            g = max(1, n * G // rows_total)
            run, state, threads, _ = cpu_groupby_timer(n, g, S.CONFIG4["seed"])
            run()
            best = min(run() for _ in range(3))
            line["cpu_baseline"] = {"value": n / best, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"first {n} rows / {g} groups of the config-4 table (same 10 rows per group), "
                                              "best of 3 passes of oracle/etl_oracle.c's groupBy"}
        print(json.dumps(line), flush=True)
    ctx.eng.close()
    if ctx.world > 1:
        ctx.dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=0, help="total rows (default: config 4's 1 B)")
    ap.add_argument("--groups", type=int, default=0, help="distinct keys (default: config 4's 100 M)")
    ap.add_argument("--sub-steps", type=int, default=10)
    ap.add_argument("--no-sub", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
