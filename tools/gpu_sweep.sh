#!/bin/bash
# Full triangular sweep of a BASELINE.json shape through the DROP-IN EXECUTABLE on the B200 box: seeded reads -> FASTA ->
# the reference's gatekeeperCreate -> gkpStore -> overlapInCorePartition (ours, byte-identical job files) -> overlapInCore --sweep.
# Reports wall time, overlaps/s and query Mbp/s (sum over jobs of the bases of the -r range) as one JSON line in
# gpurun_out/sweep_<tag>.json.   tools/gpu_sweep.sh <tag> <genome_bp> <coverage> <hash_block_bp> [gpus] [extra overlapInCore args]
set -u
TAG=$1; GENOME=$2; COV=$3; BL=$4; GPUS=${5:-1}; shift 5 2>/dev/null || shift $#
EXTRA="$*"
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/gpurun_out
W=${SWEEP_DIR:-/tmp/oic_sweep_$TAG}
mkdir -p $OUT $W
cd $W
t0=$(date +%s.%N)
if [ ! -d c.gkpStore ]; then
  $ROOT/tools/gen_reads --genome-len $GENOME --coverage $COV --len-min ${LEN_MIN:-5000} --len-max ${LEN_MAX:-15000} --err ${READ_ERR:-0.02} \
      --seed ${SEED:-1003} --out reads.fasta || exit 1
  printf "name reads\npreset pacbio-corrected\n$W/reads.fasta\n" > in.gkp
  $ROOT/oracle/_ref/gatekeeperCreate -minlength 500 -o c.gkpStore in.gkp > gk.log 2>&1 || { tail gk.log; exit 1; }
  rm -f reads.fasta
fi
t1=$(date +%s.%N)
$ROOT/canu_b200/bin/overlapInCorePartition -g c.gkpStore -bl $BL -bs 0 -rs 2000000 -rl 0 -ol 500 -o ovl > part.log 2>&1 || { tail part.log; exit 1; }
NJOBS=$(wc -l < ovl.ovlopt)
rm -rf out; mkdir out
t2=$(date +%s.%N)
OIC_PHASE_CLOCKS=${OIC_PHASE_CLOCKS:-} $ROOT/canu_b200/bin/overlapInCore -t 16 -k 22 --hashbits 23 --hashload 0.75 --maxerate 0.06 --minlength 500 \
    --gpus $GPUS $EXTRA --sweep ovl --sweep-out out c.gkpStore > sweep.log 2>&1 || { tail -20 sweep.log; exit 2; }
t3=$(date +%s.%N)
python - <<PY
import glob, json, os, struct
import numpy as np
n_ovl = 0
for f in sorted(glob.glob("out/*/*.stats")):
    for line in open(f):
        if "Total overlaps produced" in line:
            n_ovl += int(line.split("=")[1])
ovb_bytes = sum(os.path.getsize(f) for f in glob.glob("out/*/*.ovb"))
n_ovb = len(glob.glob("out/*/*.ovb"))
# query bases per job = bases of the reads of its -r range (read lengths from the store's 16-byte read records, ST/gkStore.H)
recs = np.fromfile("c.gkpStore/reads", dtype=np.uint64).reshape(-1, 2)
lens = ((recs[:, 0] >> np.uint64(43)) & np.uint64((1 << 21) - 1)).astype(np.int64)
cum = np.concatenate([[0], np.cumsum(lens)])
qb = 0
for line in open("ovl.ovlopt"):
    t = line.split()
    a, b = [int(x) for x in t[t.index("-r") + 1].split("-")]
    b = min(b, len(lens) - 1)
    qb += int(cum[b + 1] - cum[a])
wall = $t3 - $t2
import re
stamps = [(float(m.group(2)), m.group(1)) for m in (re.search(r"^(Job|Bye|GPU \d+: context up|Loaded)[^\[]*\[([0-9.]+) s\]", l) for l in open("sweep.log")) if m]
first_job = min([t for t, k in stamps if k == "Job"] or [0.0])
bye = max([t for t, k in stamps if k == "Bye"] or [0.0])
d = {"tag": "$TAG", "genome_bp": $GENOME, "coverage": $COV, "hash_block_bp": $BL, "gpus": $GPUS, "jobs": $NJOBS, "ovb_files": n_ovb,
     "reads": int(len(lens) - 1), "read_bases": int(lens.sum()), "query_bases_scanned": qb, "overlaps": n_ovl, "ovb_bytes": ovb_bytes,
     "sweep_wall_s": wall, "overlaps_per_sec": n_ovl / wall, "query_mbp_per_sec": qb / 1e6 / wall,
     "cuda_startup_s": first_job, "jobs_phase_s": bye - first_job, "jobs_phase_query_mbp_per_sec": qb / 1e6 / max(bye - first_job, 1e-9),
     "store_build_s": $t1 - $t0, "extra_args": "$EXTRA",
     "note": "wall time of ONE overlapInCore --sweep process: store open + 2-bit reads to the GPU(s) + every job's hash build, search, "
             ".ovb (snappy) / .counts / .stats files"}
open("$OUT/sweep_$TAG.json", "w").write(json.dumps(d) + "\n")
print(json.dumps(d))
PY
grep -c "Hash block" sweep.log
grep -v "Hash block" sweep.log | grep "\[" | awk "NR<=6 || NR%5==0" ; cp sweep.log $OUT/sweep_$TAG.log; tail -3 sweep.log
