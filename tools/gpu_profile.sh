#!/bin/bash
# Runs on the B200 box (through gpurun): GPU tests, a plain bench run, then -- only after those exited 0 -- the ncu launch
# list and one `--set full` capture of every hot kernel of one bench step.  Outputs land in gpurun_out/.
#   tools/gpu_profile.sh <tag> [tests]
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
if [ "${2:-}" = "tests" ]; then
  python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_$TAG.log 2>&1 || { tail -30 $OUT/pytest_gpu_$TAG.log; exit 1; }
  tail -2 $OUT/pytest_gpu_$TAG.log
fi
python bench.py --steps 5 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err || { tail -30 $OUT/bench_$TAG.err; exit 2; }
python - <<PY
import json
d = json.loads(open("$OUT/bench_$TAG.json").read().strip().splitlines()[-1])
print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["value"], "stages", {k: round(v["ms"], 2) for k, v in d["stages"].items()})
print("extend", d["stages"]["extend"]); print("roofline", d["roofline"]); print("cpu", d.get("cpu_baseline"))
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$TAG.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu > $OUT/ncu_launches_$TAG.log 2>&1 || { tail -5 $OUT/ncu_launches_$TAG.log; exit 3; }
ncu --set full --clock-control none --import-source on -k regex:'k_table_init|k_insert|k_multi|k_occupancy|k_seed_scan|k_replay|k_process_pairs|k_extend_tasks|k_commit_pairs' \
    --launch-skip 14 --launch-count 14 -f -o $OUT/full_$TAG python bench.py --steps 1 --warmup 1 --no-cpu > $OUT/ncu_full_$TAG.log 2>&1 || { tail -5 $OUT/ncu_full_$TAG.log; exit 4; }
ls -la $OUT/full_$TAG.ncu-rep
