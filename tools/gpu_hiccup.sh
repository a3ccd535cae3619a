#!/bin/bash
# where do the random 5-100 ms gaps of a step come from?  host clock of the pair stage's phases (OIC_PHASE_CLOCKS) over many steps,
# with and without the per-step cudaMemGetInfo;  tools/gpu_hiccup.sh <tag> [steps]
TAG=${1:-hic}; STEPS=${2:-30}
mkdir -p gpurun_out
for rep in 0 1 2; do
for mode in always epoch; do
  E=""; [ $mode = always ] && E="OIC_MEMINFO_ALWAYS=1"
  env OIC_PHASE_CLOCKS=1 $E python bench.py --steps $STEPS --warmup 3 --no-cpu > gpurun_out/h_${TAG}_${mode}_$rep.json 2> gpurun_out/h_${TAG}_${mode}_$rep.err
  python - <<PY
import json, re
d = json.loads(open("gpurun_out/h_${TAG}_${mode}_$rep.json").read().strip().splitlines()[-1])
rows = [tuple(float(x) for x in re.findall(r"([0-9.]+) ms", l)) for l in open("gpurun_out/h_${TAG}_${mode}_$rep.err") if "pair stage host clock" in l]
rows = rows[3:]
import statistics as st
f = lambda i: "med %.2f max %.2f" % (st.median(r[i] for r in rows), max(r[i] for r in rows))
print("$mode rep $rep: %.2f ms/step, e2e %.2f | n=%d plan %s | kernels %s | copy %s" % (d["ms_per_step"], d["e2e"]["ms_per_step"], len(rows), f(0), f(1), f(2)))
PY
done; done
