#!/usr/bin/env python
"""Per-kernel roofline measurements for every row of SURVEY.md §8 (one GPU).

    python tools/measure_kernels.py [--out profiles/r01_kernels.json] [--only name,...] [--full]

For each kernel: synthetic input of the BASELINE.json shape resident in HBM,
>= 3 warm-up launches, K timed launches bracketed by CUDA events on the launching
stream (inputs are larger than L2 everywhere), algorithmic bytes per launch
 / mean launch time = achieved GB/s, against the measured copy peak
of MEASURED_PEAKS.json.  A CPU figure for the same operation (NumPy or pyarrow,
bounded sample, thread count stated - never oracle/) is taken beside it.
These are kernel numbers; the headline line stays with bench.py."""
from __future__ import annotations

import argparse
import gzip
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

import pyspark_tf_gke_b200 as pkg  # noqa: E402
import pyspark_tf_gke_b200.synth as S  # noqa: E402


def peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


LAST_CLOCKS = None   # SM clock / throttle reasons sampled during the most recent timed() region; entry() records them


def timed(fn, steps=20, warmup=3):
    global LAST_CLOCKS
    from tools.clocks import ClockSampler

    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    sampler = ClockSampler(torch.cuda.current_device())
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    sampler.start()
    ev[0].record()
    for i in range(steps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    LAST_CLOCKS = sampler.stop()
    ts = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
    return sum(ts) / len(ts), min(ts)


def cpu_time(fn, reps=3):
    fn()
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
    return best


def entry(name, rows, alg_bytes, mean_ms, min_ms, cpu=None, note=""):
    p, src = peak()
    gbs = alg_bytes / (mean_ms * 1e-3) / 1e9
    e = {"kernel": name, "rows": rows, "algorithmic_bytes": alg_bytes, "mean_ms": mean_ms, "min_ms": min_ms,
         "rows_per_s": rows / (mean_ms * 1e-3), "achieved_gbs": gbs, "peak_gbs": p, "peak_source": src, "frac": gbs / p,
         "note": note, "clocks": LAST_CLOCKS}
    if cpu:
        e["cpu"] = cpu
    print(json.dumps(e), flush=True)
    return e


def table2(eng, n, seed=1, kind=S.KIND_EXACT):
    cfg = S.CONFIG2
    key = eng.synth_column(seed, 0, S.KIND_KEY32, n, modulus=cfg["categories"], null_rate=cfg["key_null"])
    cols = [eng.synth_column(seed, 1 + c, kind, n, null_rate=cfg["val_null"][c]) for c in range(3)]
    return key, cols


def m_null_count(eng, args):
    n = 100_000_000
    key, _ = table2(eng, n)
    out = eng.empty(1, torch.int64)
    mean, mn = timed(lambda: eng.null_count(key, out))
    host = key.valid.cpu().numpy().view(np.uint8)
    t = cpu_time(lambda: n - int(np.unpackbits(host[: n // 8]).sum()))
    return entry("null_count_kernel (k_means.py:23)", n, n // 8, mean, mn,
                 {"rows_per_s": n / t, "impl": "numpy unpackbits+sum", "cores": 1}, "1 bit per row; launch-latency bound")


def m_sum_count(eng, args):
    n = 100_000_000
    key, cols = table2(eng, n)
    s, c = eng.empty(3, torch.float64), eng.empty(3, torch.int64)
    mean, mn = timed(lambda: eng.sum_count(cols, row_valid=key.valid, out_sum=s, out_count=c))
    m = 10_000_000
    hv = [col.data[:m].cpu().numpy() for col in cols]
    hm = [np.unpackbits(col.valid[: m // 32].cpu().numpy().view(np.uint8), bitorder="little").astype(bool) for col in cols]
    t = cpu_time(lambda: [float(v[k].sum()) for v, k in zip(hv, hm)])
    return entry("column_sum_count_kernel x3 (k_means.py:45-48)", n, n * 24 + 4 * (n // 8), mean, mn,
                 {"rows_per_s": m / t, "impl": "numpy masked sum, 10M-row sample", "cores": 1})


def m_histogram(eng, args):
    n = 100_000_000
    out = []
    for bins in (30, 1000):
        codes = eng.synth_column(2, 0, S.KIND_KEY32, n, modulus=bins, null_rate=0.01)
        mean, mn = timed(lambda: eng.code_histogram(codes, bins))
        m = 10_000_000
        hc = codes.data[:m].cpu().numpy()
        t = cpu_time(lambda: np.bincount(hc, minlength=bins))
        out.append(entry(f"code_histogram_kernel bins={bins} (StringIndexer.fit, k_means.py:34,73)", n, n * 4 + n // 8, mean,
                         mn, {"rows_per_s": m / t, "impl": "numpy bincount, 10M-row sample", "cores": 1}))
    return out


def m_groupby3(eng, args):
    cfg = S.CONFIG3
    n = cfg["rows"]
    key = eng.synth_column(cfg["seed"], 0, S.KIND_KEY32, n, modulus=cfg["groups"], with_valid=False)
    cols = [eng.synth_column(cfg["seed"], 1 + c, S.KIND_EXACT, n, null_rate=cfg["val_null"][c]) for c in range(3)]
    gt = eng.groupby_new(4096, 3)

    def run():
        eng.groupby_accumulate(gt, key.data, None, [c.data for c in cols], [c.valid for c in cols], groups_hint=cfg["groups"],
                               key_range=int(os.environ.get("GB_KEY_RANGE", cfg["groups"])))

    mean, mn = timed(run, steps=10)
    m = 10_000_000
    import pyarrow as pa
    import pyarrow.compute  # noqa: F401

    tbl = pa.table({"k": key.data[:m].cpu().numpy(), **{f"v{c}": cols[c].data[:m].cpu().numpy() for c in range(3)}})
    t = cpu_time(lambda: tbl.group_by("k").aggregate([(f"v{c}", a) for c in range(3) for a in ("sum", "count")]))
    return entry("groupby_dense_kernel config3: 100M rows, 1K groups, 3x sum/count", n, n * 28 + 3 * (n // 8), mean, mn,
                 {"rows_per_s": m / t, "impl": "pyarrow Table.group_by, 10M-row sample", "cores": pa.cpu_count()})


def m_groupby4(eng, args):
    cfg = S.CONFIG4
    n = cfg["rows"] if args.full else cfg["rows"] // 4
    groups = cfg["groups"] if args.full else cfg["groups"] // 4
    key = eng.synth_column(cfg["seed"], 0, S.KIND_KEY64, n, modulus=groups, with_valid=False)
    val = eng.synth_column(cfg["seed"], 1, S.KIND_EXACT, n, with_valid=False)
    cap = 1 << int(np.ceil(np.log2(groups * 2)))
    gt = eng.groupby_new(cap, 1)
    torch.cuda.synchronize()

    def run():
        eng.groupby_accumulate(gt, key.data, None, [val.data], groups_hint=groups)

    mean, mn = timed(run, steps=3, warmup=1)
    return entry(f"groupby_global_kernel config4 (1 GPU): {n} rows, {groups} groups, sum/count", n, int(n * 18.5), mean, mn,
                 None, "random 32-byte table entries: sector traffic >> algorithmic bytes")


def m_partition(eng, args):
    n = 125_000_000
    key = eng.synth_column(3, 0, S.KIND_KEY64, n, modulus=10**8, with_valid=False)
    val = eng.synth_column(3, 1, S.KIND_EXACT, n, with_valid=False)
    mean, mn = timed(lambda: eng.hash_partition(key.data, val.data, 8), steps=10)
    return entry("hash_partition (count + scatter) P=8, 125M rows (config 4 map side)", n, n * (8 + 16 + 16), mean, mn, None,
                 "count pass reads keys (8 B), scatter pass reads 16 B and writes 16 B per row")


def m_merge(eng, args):
    """Config 4 on one GPU stage by stage, count-free slab path (csrc/slabs.cu + merge.cu), then the two
    count passes of the counted path for comparison."""
    from pyspark_tf_gke_b200 import _lib
    cfg = S.CONFIG4
    n = cfg["rows"] if args.full else cfg["rows"] // 4
    groups = cfg["groups"] if args.full else cfg["groups"] // 4
    key = eng.synth_column(cfg["seed"], 0, S.KIND_KEY64, n, modulus=groups, with_valid=False).data
    val = eng.synth_column(cfg["seed"], 1, S.KIND_EXACT, n, with_valid=False).data
    B, sub, rpk, cap1, cap2 = eng.slab_plan(n, groups)
    fine = B * sub
    res = []
    counters = torch.zeros(5, dtype=torch.int64, device=eng.device)
    cur1 = torch.zeros(B, dtype=torch.int64, device=eng.device)
    level1 = torch.empty((B * cap1, 2), dtype=torch.int64, device=eng.device)

    def l1():
        cur1.zero_()
        eng.slab_partition_rows(key, val, n, 1, B, [level1.data_ptr()], 0, cap1, cap1, cur1, counters[4:])

    mean, mn = timed(l1, steps=3, warmup=1)
    res.append(entry(f"level 1: scatter_pipelined_kernel<columns, slab> into {B} bucket slabs, {n} rows", n, n * 32, mean, mn, None,
                     "reads 16 B and writes 16 B per row; no count pass"))
    cur2 = torch.zeros(fine, dtype=torch.int64, device=eng.device)
    pieces = torch.empty((fine * cap2, 2), dtype=torch.int64, device=eng.device)

    def l2():
        cur2.zero_()
        eng.slab_partition_records(level1, B, cap1, cur1, None, cap1, B, sub, pieces, cap2, cur2, counters[4:])

    mean, mn = timed(l2, steps=3, warmup=1)
    res.append(entry(f"level 2: scatter_pipelined_kernel<records, slab> {B} x {sub} = {fine} piece slabs", n, n * 32, mean, mn, None,
                     "reads 16 B and writes 16 B per record; no count pass"))
    cap = groups + groups // 8
    pg = eng.new_partitioned_groups(cap, n // 8, fine, counters=counters)
    mean, mn = timed(lambda: eng.merge_slab_pieces(pieces, cur2, cap2, 0, fine, False, pg), steps=3, warmup=1)
    g, spilled, _, _, overflow = counters.cpu().tolist()
    # CPU side of config 4 on a sample that fits the host comfortably: same rows-per-group ratio (10)
    import pyarrow as pa
    m = 20_000_000
    sk = eng.synth_column(cfg["seed"], 0, S.KIND_KEY64, m, modulus=m // 10, with_valid=False).data.cpu().numpy()
    sv = eng.synth_column(cfg["seed"], 1, S.KIND_EXACT, m, with_valid=False).data.cpu().numpy()
    tbl = pa.table({"k": sk, "v": sv})
    t = cpu_time(lambda: tbl.group_by("k").aggregate([("v", "sum"), ("v", "count")]), reps=2)
    res.append(entry(f"groupby_merge_kernel: {fine} pieces, {g} groups, {spilled} spilled, {overflow} slab overflow", n, n * 16 + g * 32,
                     mean, mn,
                     {"rows_per_s": m / t, "impl": f"pyarrow Table.group_by, {m} rows / {m // 10} groups (whole groupBy, not just the merge)",
                      "cores": pa.cpu_count()},
                     "reads 16 B per record, writes 32 B per group; table updates stay in shared memory"))
    del level1, pieces, pg
    # the counted path's passes (exact sizes; what a slab overflow falls back to)
    mean, mn = timed(lambda: eng.hash_partition(key, val, 1, B), steps=3, warmup=1)
    res.append(entry(f"counted level 1: partition_count + scatter into {B} buckets, {n} rows", n, n * 40, mean, mn, None,
                     "count pass reads keys (8 B), scatter reads 16 B and writes 16 B per row"))
    return res


def health_like_text(target_rows):
    with gzip.open(os.path.join(ROOT, "tests", "golden", "health_workloads.csv.gz"), "rb") as fh:
        raw = fh.read()
    head, body = raw.split(b"\n", 1)
    reps = max(1, target_rows // 18155)
    return head + b"\n" + body * reps, 18155 * reps


def m_csv(eng, args):
    raw, rows = health_like_text(10_000_000)
    n = len(raw)
    text = pkg.table.upload_text(eng, raw)
    torch.cuda.synchronize()
    res = []
    mean, mn = timed(lambda: eng.csv_index_rows(text, n), steps=10)
    res.append(entry("csv_block_summary + span scan (row index, pass 1: reads the text, keeps 32 B of masks per 64 B)", rows,
                     n + n // 2, mean, mn))
    total = int(eng.csv_index_rows(text, n).item())
    assert total == rows + 1
    mean, mn = timed(lambda: eng.csv_row_offsets(text, n, total), steps=10)
    res.append(entry("csv_row_offsets_kernel (pass 2: reads the kept masks, writes row offsets)", rows, n // 4 + 8 * total,
                     mean, mn))
    row_start = eng.csv_row_offsets(text, n, total)
    cap = 1 << 16

    def parse(flags=4):   # 4 = ETL_CSV_USE_INDEX_MASKS: walk the masks the index pass kept (what read_csv does)
        d = (torch.zeros(cap, dtype=torch.int64, device=eng.device), torch.zeros(cap, dtype=torch.int64, device=eng.device))
        return eng.csv_parse_columns(text, n, row_start, 1, rows, [5, 6, 7, 2], [0, 0, 0, 1], [None, None, None, d], flags=flags)

    mean0, mn0 = timed(lambda: parse(0), steps=5)
    res.append(entry("csv_parse_rows_kernel<text scan> (4 columns; fallback when the index masks are gone)", rows,
                     n + rows * 28.5, mean0, mn0))
    mean, mn = timed(parse, steps=5)
    import io

    import pyarrow as pa
    import pyarrow.csv as pacsv

    sample = raw[: n // 10]
    sample = sample[: sample.rfind(b"\n") + 1]
    t = cpu_time(lambda: pacsv.read_csv(io.BytesIO(sample), convert_options=pacsv.ConvertOptions(
        include_columns=["measure_name", "value", "lower_ci", "upper_ci"])), reps=2)
    res.append(entry("csv_parse_rows_kernel (4 columns: 3 numeric + dictionary key)", rows, n + rows * 28.5, mean, mn,
                     {"rows_per_s": (rows / 10) / t, "bytes_per_s": len(sample) / t,
                      "impl": "pyarrow.csv.read_csv, 1/10 sample", "cores": pa.cpu_count()},
                     "instruction-issue bound (byte-level tokenising + numeral conversion), not HBM bound"))
    return res


def m_features(eng, args):
    n = 10_000_000
    idx = eng.synth_column(1, 0, S.KIND_KEY32, n, modulus=31, with_valid=False).data
    nums = [eng.synth_column(1, 1 + c, S.KIND_EXACT, n, with_valid=False).data for c in range(3)]
    res = []
    for dtype, es in ((torch.float64, 8), (torch.float32, 4)):
        out = torch.empty((n, 153), dtype=dtype, device=eng.device)
        mean, mn = timed(lambda: eng.assemble_dense(idx, nums, 30, 5, dtype=dtype, out=out), steps=5)
        res.append(entry(f"assemble_dense_kernel D=153 {str(dtype).split('.')[-1]} (k_means.py:38,64-77)", n,
                         n * (28 + 153 * es), mean, mn, None, "write bound"))
        del out
    mean, mn = timed(lambda: eng.assemble_sparse(idx, nums, 30, 5), steps=5)
    res.append(entry("assemble_sparse_kernel (ELL, 8 slots)", n, n * (28 + 8 * 12), mean, mn))
    lab = eng.synth_column(4, 0, S.KIND_KEY32, n, modulus=15, null_rate=0.083)
    cols = [eng.synth_column(4, 1 + c, S.KIND_EXACT, n, null_rate=0.074) for c in range(3)]
    X, y, r = eng.pack_tf(lab, cols)
    kept = int(r.item())
    mean, mn = timed(lambda: eng.pack_tf(lab, cols), steps=10)
    res.append(entry("tf_tile_count + tf_tile_scan + tf_pack (load_csv layout, train_tf_ps.py:120-149)", n,
                     n * 28 + 8 * (n // 8) + kept * 16, mean, mn, None,
                     "includes the torch allocation of X/y each call; bitmaps are read twice"))
    return res


def m_parquet(eng, args):
    import io

    import pyarrow as pa
    import pyarrow.parquet as pq

    n = 10_000_000
    rng = np.random.default_rng(0)
    names = np.array([f"measure {i}" for i in range(30)])
    cols = {"measure_name": pa.array(rng.choice(names, n))}
    for c, rate in zip(("value", "lower_ci", "upper_ci"), (0.071, 0.074, 0.074)):
        cols[c] = pa.array(np.floor(rng.random(n) * 100001), mask=rng.random(n) < rate)
    buf = io.BytesIO()
    import pyspark_tf_gke_b200.parquet as P

    res = []
    for codec in ("NONE", "snappy"):
        buf = io.BytesIO()
        pq.write_table(pa.table(cols), buf, compression=codec, use_dictionary=["measure_name"])
        raw = buf.getvalue()
        rows, plans, end = P._plan(raw, ["value", "lower_ci", "upper_ci", "measure_name"])
        data = pkg.table.upload_text(eng, raw, extra=max(0, end - P._padded_len(len(raw))))
        torch.cuda.synchronize()
        inflate = [j for plan in plans.values() for j in plan.inflate]
        jobs = [P.column_job(plan, name == "measure_name") for name, plan in plans.items()]

        def run():  # device part only: page plans are host metadata work, the file is already in HBM
            eng.parquet_decode_columns(data, rows, jobs, status=eng.parquet_inflate(data, inflate))

        mean, mn = timed(run, steps=5)
        t = cpu_time(lambda: pq.read_table(io.BytesIO(raw)), reps=2)
        res.append(entry(f"pq_snappy + pq_hybrid_decode + pq_gather, 4 columns (3 double + dictionary string), compression={codec}",
                         n, len(raw) + n * 28.5, mean, mn,
                         {"rows_per_s": n / t, "impl": "pyarrow.parquet.read_table", "cores": pa.cpu_count()},
                         f"{len(raw) / 1e6:.0f} MB file, {len(inflate)} compressed pages; includes per-column host glue"))
    return res


def m_kmeans(eng, args):
    n, L, repeats, k = 10_000_000, 30, 5, 25
    idx = eng.synth_column(1, 0, S.KIND_KEY32, n, modulus=31, with_valid=False).data
    nums = [eng.synth_column(1, 1 + c, S.KIND_EXACT, n, with_valid=False).data for c in range(3)]
    centers = torch.rand((k, repeats * L + 3), dtype=torch.float64, device=eng.device)
    cluster = torch.zeros(n, dtype=torch.int32, device=eng.device)
    res = []
    mean, mn = timed(lambda: eng.kmeans_assign(idx, nums, centers, L, repeats, cluster=cluster), steps=10)
    res.append(entry("kmeans_prepare + kmeans_assign k=25, D=153 (k_means.py:83-87, one Lloyd pass)", n, n * (28 + 12), mean, mn,
                     None, "reads 28 B and writes cluster + cost (12 B) per row; the dense [N,153] matrix is never built"))
    # `cluster` now holds the assignment to the random centres above (skewed, as real assignments are)
    mean, mn = timed(lambda: eng.kmeans_accumulate(idx, nums, cluster, k, L), steps=10)
    res.append(entry("kmeans_accumulate k=25 (per-cluster counts, label histogram, numeric sums)", n, n * 32, mean, mn))
    return res


def m_trainer(eng, args):
    from pyspark_tf_gke_b200 import trainer as T

    n, C, B = 12284, 15, 64   # the default input's training split (80 % of 15355 rows), batch 64
    rng = np.random.default_rng(0)
    X = torch.from_numpy(rng.random((n, 3)).astype(np.float32) * 1000).to(eng.device)
    y = torch.from_numpy(rng.integers(0, C, n).astype(np.int32)).to(eng.device)
    steps = n // B
    order = torch.from_numpy(T.shuffle_buffer_order(n, 3000, steps * B, rng)).to(eng.device)
    params = torch.from_numpy(T.glorot_uniform_params(C, rng)).to(eng.device)
    m, v = torch.zeros_like(params), torch.zeros_like(params)
    mean, mn = timed(lambda: eng.mlp_train_steps(X, y, order, B, C, params, m, v), steps=5)
    e = entry(f"mlp_train_kernel: one epoch = {steps} steps of batch {B} (train_tf_ps.py:328-343,664-672)", steps * B,
              steps * B * 16, mean, mn, None, "single CTA, launch-latency-bound problem: %.2f us per optimiser step" % (mean * 1e3 / steps))
    return e


ALL = {"trainer": m_trainer, "kmeans": m_kmeans, "parquet": m_parquet, "null_count": m_null_count, "sum_count": m_sum_count, "histogram": m_histogram, "groupby3": m_groupby3,
       "groupby4": m_groupby4, "merge": m_merge, "partition": m_partition, "csv": m_csv, "features": m_features}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--only", default="")
    ap.add_argument("--full", action="store_true", help="config 4 at the full 1B rows / 100M groups")
    args = ap.parse_args()
    eng = pkg.Engine(0)
    names = [s for s in args.only.split(",") if s] or list(ALL)
    results = []
    for name in names:
        r = ALL[name](eng, args)
        results.extend(r if isinstance(r, list) else [r])
        torch.cuda.empty_cache()
    if args.out:
        with open(args.out, "w") as fh:
            json.dump({"gpu": torch.cuda.get_device_name(0), "results": results}, fh, indent=1)


if __name__ == "__main__":
    main()
