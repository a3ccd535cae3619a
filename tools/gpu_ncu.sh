#!/bin/bash
# ncu evidence of one bench step on the B200 box (after the plain command exited 0): launch list + `--set full` of the hot kernels.
#   tools/gpu_ncu.sh <tag>
set -u
TAG=$1
OUT=gpurun_out
mkdir -p $OUT
python bench.py --steps 5 --warmup 3 --no-cpu > $OUT/bench_${TAG}_nocpu.json 2> $OUT/bench_${TAG}_nocpu.err || { tail -30 $OUT/bench_${TAG}_nocpu.err; exit 2; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$TAG.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu > $OUT/ncu_launches_$TAG.log 2>&1 || { tail -5 $OUT/ncu_launches_$TAG.log; exit 3; }
ncu --set full --clock-control none --import-source on -k regex:'k_insert|k_filter_build|k_seed_scan|k_process_pairs|k_extend_tasks|k_commit_pairs|k_table_init' \
    --launch-skip 8 --launch-count 8 -f -o $OUT/full_$TAG python bench.py --steps 1 --warmup 1 --no-cpu > $OUT/ncu_full_$TAG.log 2>&1 || { tail -5 $OUT/ncu_full_$TAG.log; exit 4; }
ls -la $OUT/full_$TAG.ncu-rep
