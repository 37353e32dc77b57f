#!/bin/bash
# Knob sweep of config 4 on one GPU (tools/measure_shuffle.py): tile size / threads of the scatter kernel,
# slab vs counted path.  Appends JSON lines to gpurun_out/r02_sweep.jsonl.
out=gpurun_out/r02_sweep.jsonl
run() { echo "# $*" >> $out; env "$@" python tools/measure_shuffle.py --steps 3 --warmup 1 >> $out 2>> gpurun_out/r02_sweep.err; }
python tools/measure_shuffle.py --steps 3 --warmup 1 --counted >> $out 2>> gpurun_out/r02_sweep.err
run ETL_PART_TILE=2048
run ETL_PART_TILE=2048 ETL_PART_THREADS=512
run ETL_PART_TILE=2048 ETL_PART_THREADS=256
run ETL_PART_TILE=4096 ETL_PART_THREADS=512
run ETL_PART_TILE=4096 ETL_PART_THREADS=1024
run ETL_MERGE_SLOTS=2048 ETL_MERGE_GROUPS=768
