#!/bin/bash
# one bench line per BASELINE.json shape (the driver's own bench line is c3):  tools/gpu_workloads.sh <tag> [shapes...]
TAG=$1; shift
mkdir -p gpurun_out
for w in "$@"; do
  steps=2; [ "$w" = "c2" ] && steps=1
  timeout 1500 python bench.py --workload $w --steps $steps --warmup 1 > gpurun_out/wl_${TAG}_$w.json 2> gpurun_out/wl_${TAG}_$w.err || { echo "$w FAILED"; tail -5 gpurun_out/wl_${TAG}_$w.err; continue; }
  python - <<PY
import json
d = json.loads(open("gpurun_out/wl_${TAG}_$w.json").read().strip().splitlines()[-1])
e = d["stages"]["extend"]
print("[$w] value %.1f Mbp/s  %.1f ms/step  e2e %.1f  ovl/s %.0f (%d per step) | build %.1f scan %.1f group %.1f ext %.1f (first %.1f tasks %.1f commit %.1f retry %.1f) | cpu %s" % (
  d["value"], d["ms_per_step"], d["e2e"]["value"], d["overlaps_per_sec"], d["overlaps_per_step"], d["stages"]["build"]["ms"], d["stages"]["seed_scan"]["ms"],
  d["stages"]["group_runs"]["ms"], e["ms"], e["ms_first_extensions"], e["ms_speculative_tasks"], e["ms_commit"], e.get("ms_retry", 0), d.get("cpu_baseline")))
PY
  grep "pair stage" gpurun_out/wl_${TAG}_$w.err | tail -2
done
