#!/bin/bash
# quick A/B on the B200 box: bench runs (no CPU leg) under different env settings;  tools/gpu_quick.sh <tag> "ENV=.. ENV=.." ...
TAG=$1; shift
mkdir -p gpurun_out
i=0
for envs in "$@"; do
  env $envs python bench.py --steps ${STEPS:-3} --warmup 3 --no-cpu > gpurun_out/q_${TAG}_$i.json 2> gpurun_out/q_${TAG}_$i.err || tail -20 gpurun_out/q_${TAG}_$i.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/q_${TAG}_$i.json").read().strip().splitlines()[-1])
e = d["stages"]["extend"]
print("[$envs] value %.1f ms %.2f e2e %.1f (%.2f ms) | build %.2f scan %.2f group %.2f ext %.2f (first %.2f tasks %.2f commit %.2f) ovl %d" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"],
      d["stages"]["build"]["ms"], d["stages"]["seed_scan"]["ms"], d["stages"]["group_runs"]["ms"], e["ms"], e["ms_first_extensions"], e["ms_speculative_tasks"], e["ms_commit"], d["overlaps_per_step"]))
print("   e2e stages", {k: round(v, 2) for k, v in d["e2e"].get("stages_ms", {}).items()})
PY
  i=$((i+1))
done
