#!/bin/bash
# ncu evidence for round 2 (one GPU; every program below is first run WITHOUT ncu and must exit 0):
#   1. launch list of the bench command (per-launch gpu__time_duration, cold cache, serialised)
#   2. --set full captures of the three config-4 kernels at full size, of groupby_dense (config 3) and of the
#      projection (config 2)
# Outputs land in gpurun_out/; the summaries (profiles/ncu_summary.py) are committed under profiles/.
set -x
out=gpurun_out
python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --sub-steps 2 > $out/r02_prof_bench.json 2> $out/r02_prof_bench.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/r02_bench_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --sub-steps 2 > $out/r02_ncu_bench.log 2>&1
python tools/measure_shuffle.py --steps 1 --warmup 1 > $out/r02_prof_shuffle.json 2> $out/r02_prof_shuffle.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:'scatter_pipelined_kernel|groupby_merge_kernel' -c 6 -f -o $out/r02_config4 \
    python tools/measure_shuffle.py --steps 1 --warmup 1 > $out/r02_ncu_config4.log 2>&1
python tools/measure_kernels.py --only groupby3 > $out/r02_prof_groupby3.json 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'groupby_dense_kernel' -c 2 -f -o $out/r02_groupby_dense \
    python tools/measure_kernels.py --only groupby3 > $out/r02_ncu_groupby3.log 2>&1
ls -la $out/*.ncu-rep
