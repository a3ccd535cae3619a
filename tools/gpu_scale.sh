#!/bin/bash
# bench lines of the BASELINE.json shapes at N GPUs of this box (device-timed, max over ranks; no CPU leg);  tools/gpu_scale.sh <tag> <N> [workloads...]
TAG=$1; N=$2; shift 2
WL=${@:-c3 c4 c5}
mkdir -p gpurun_out
for w in $WL; do
  out=gpurun_out/scale_${TAG}_${w}_n${N}
  if [ "$N" = 1 ]; then
    python bench.py --workload $w --steps ${STEPS:-8} --warmup 3 --no-cpu > $out.json 2> $out.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --workload $w --steps ${STEPS:-8} --warmup 3 --no-cpu > $out.json 2> $out.err
  fi
  rc=$?
  [ $rc != 0 ] && { echo "$w N=$N rc=$rc"; tail -5 $out.err; }
  python - <<PY
import json
try:
    d = json.loads(open("$out.json").read().strip().splitlines()[-1])
    e = d["stages"]["extend"]
    print("$w N=$N: %.1f %s | %.2f ms/step | e2e %.1f (%.2f ms) | ovl/step %d | build %.2f scan %.2f group %.2f ext %.2f (tasks %.2f) | clocks %s" % (d["value"], d["unit"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"],
          d["overlaps_per_step"], d["stages"]["build"]["ms"], d["stages"]["seed_scan"]["ms"], d["stages"]["group_runs"]["ms"], e["ms"], e["ms_speculative_tasks"], d["clocks"]))
except Exception as ex:
    print("$w N=$N: no line (%s)" % ex)
PY
done
