#!/bin/bash
# 2-GPU box diagnostic: does a GPU run the same job slower when its neighbours are busy, and is NCCL involved?
#   A: rank 0's job alone on GPU 0 (GPU 1 idle)      B: rank 0's and rank 1's jobs as two INDEPENDENT processes at once (no NCCL)
#   C: the 2-rank run under torchrun (NCCL build)    nvidia-smi samples clocks / power / throttle reasons of every GPU at 10 Hz throughout
TAG=${1:-diag}; STEPS=${2:-12}
mkdir -p gpurun_out
nvidia-smi --query-gpu=timestamp,index,clocks.sm,clocks.mem,power.draw,power.limit,temperature.gpu,clocks_throttle_reasons.active --format=csv,noheader -lms 100 > gpurun_out/smi_${TAG}.csv &
SMI=$!
show() { python - "$@" <<'PY'
import json, sys
for f in sys.argv[2:]:
    d = json.loads(open(f).read().strip().splitlines()[-1]); e = d["stages"]["extend"]
    print("%s %s: %.2f ms/step | build %.2f scan %.2f ext %.2f (first %.2f tasks %.2f commit %.2f) cells %.3f G clocks %s" % (sys.argv[1], f.split("/")[-1], d["ms_per_step"], d["stages"]["build"]["ms"],
          d["stages"]["seed_scan"]["ms"], e["ms"], e["ms_first_extensions"], e["ms_speculative_tasks"], e["ms_commit"], d["work_per_step_rank0"]["ext_cells"] / 1e9, d["clocks"]))
PY
}
date +%T.%N; echo "== A: alone"
CUDA_VISIBLE_DEVICES=0 python bench.py --steps $STEPS --warmup 3 --no-cpu --as-rank 0/2 > gpurun_out/d_${TAG}_A0.json 2> gpurun_out/d_${TAG}_A0.err
show A gpurun_out/d_${TAG}_A0.json
CUDA_VISIBLE_DEVICES=1 python bench.py --steps $STEPS --warmup 3 --no-cpu --as-rank 1/2 > gpurun_out/d_${TAG}_A1.json 2> gpurun_out/d_${TAG}_A1.err
show A gpurun_out/d_${TAG}_A1.json
date +%T.%N; echo "== B: two independent processes at once"
CUDA_VISIBLE_DEVICES=0 python bench.py --steps $((STEPS*3)) --warmup 3 --no-cpu --as-rank 0/2 > gpurun_out/d_${TAG}_B0.json 2> gpurun_out/d_${TAG}_B0.err &
P0=$!
CUDA_VISIBLE_DEVICES=1 python bench.py --steps $((STEPS*3)) --warmup 3 --no-cpu --as-rank 1/2 > gpurun_out/d_${TAG}_B1.json 2> gpurun_out/d_${TAG}_B1.err &
P1=$!
wait $P0 $P1
show B gpurun_out/d_${TAG}_B0.json gpurun_out/d_${TAG}_B1.json
date +%T.%N; echo "== C: torchrun 2 ranks"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps $STEPS --warmup 3 --no-cpu > gpurun_out/d_${TAG}_C.json 2> gpurun_out/d_${TAG}_C.err
show C gpurun_out/d_${TAG}_C.json
date +%T.%N
kill $SMI
python - <<PY
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/smi_${TAG}.csv")) if len(r) >= 8]
by = collections.defaultdict(list)
for r in rows: by[r[1].strip()].append(r)
for g, rs in sorted(by.items()):
    busy = [r for r in rs if float(r[4].split()[0]) > 300]
    sm = sorted(int(r[2].split()[0]) for r in busy) or [0]
    pw = sorted(float(r[4].split()[0]) for r in busy) or [0]
    print("gpu", g, "samples", len(rs), "busy", len(busy), "sm MHz min/med/max", sm[0], sm[len(sm)//2], sm[-1], "power W med/max", pw[len(pw)//2], pw[-1], "limit", rs[0][5].strip(),
          "reasons", collections.Counter(r[7].strip() for r in busy).most_common(4))
PY
