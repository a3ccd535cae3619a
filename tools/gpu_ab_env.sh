#!/bin/bash
# A/B of one workload under env settings, same box;  tools/gpu_ab_env.sh <tag> <workload> <steps> "ENV=.." "ENV=.." ...
TAG=$1; W=$2; ST=$3; shift 3
mkdir -p gpurun_out
i=0
for envs in "$@"; do
  env $envs python bench.py --workload $W --steps $ST --warmup 2 --no-cpu > gpurun_out/ab_${TAG}_${W}_$i.json 2> gpurun_out/ab_${TAG}_${W}_$i.err || tail -5 gpurun_out/ab_${TAG}_${W}_$i.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/ab_${TAG}_${W}_$i.json").read().strip().splitlines()[-1]); e = d["stages"]["extend"]
print("$W [$envs] %.2f %s | %.2f ms/step | ext %.2f (first %.2f tasks %.2f commit %.2f retry %.2f) ovl %d" % (d["value"], d["unit"], d["ms_per_step"], e["ms"], e["ms_first_extensions"], e["ms_speculative_tasks"], e["ms_commit"], e["ms_retry"], d["overlaps_per_step"]))
PY
  i=$((i+1))
done
