#!/bin/bash
# Frequent k-mer list of a synthetic store on the B200 box: the reference's meryl (as the pipeline runs it, Meryl.pm:434-450,634)
# against canu_b200/bin/frequentMers; outputs must be byte-identical.  tools/gpu_kmers.sh <tag> <genome_bp> <coverage> [K] [T]
set -u
TAG=$1; GENOME=$2; COV=$3; K=${4:-22}; T=${5:-20}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/gpurun_out; W=/tmp/oic_kmers_$TAG
mkdir -p $OUT $W; cd $W
$ROOT/tools/gen_reads --genome-len $GENOME --coverage $COV --len-min 5000 --len-max 15000 --err 0.02 --seed 1003 --out reads.fasta || exit 1
printf "name reads\npreset pacbio-corrected\n$W/reads.fasta\n" > in.gkp
$ROOT/oracle/_ref/gatekeeperCreate -minlength 500 -o c.gkpStore in.gkp > gk.log 2>&1 || { tail gk.log; exit 1; }
rm -f reads.fasta
NT=$(nproc)
t0=$(date +%s.%N)
$ROOT/oracle/_ref/meryl -B -C -L 2 -v -m $K -threads $NT -memory 65536 -s c.gkpStore -o ref.ms > meryl.log 2>&1 || { tail meryl.log; exit 1; }
t1=$(date +%s.%N)
$ROOT/oracle/_ref/meryl -Dt -n $T -s ref.ms > ref.fasta 2>> meryl.log
t2=$(date +%s.%N)
$ROOT/canu_b200/bin/frequentMers -m $K -n $T -s c.gkpStore -o gpu.fasta 2> gpu.log || { cat gpu.log; exit 2; }
t3=$(date +%s.%N)
cmp ref.fasta gpu.fasta && SAME=true || SAME=false
python - <<PY
import json, re
log = open("gpu.log").read()
m = re.search(r"(\d+) reads, (\d+) bases: (\d+) canonical .* \((\d+) pass\w*, ([0-9.]+) ms", log)
d = {"tag": "$TAG", "genome_bp": $GENOME, "coverage": $COV, "K": $K, "min_count": $T, "reads": int(m.group(1)), "bases": int(m.group(2)),
     "kmers_listed": int(m.group(3)), "gpu_passes": int(m.group(4)), "gpu_count_ms": float(m.group(5)),
     "gpu_tool_wall_s": $t3 - $t2, "meryl_build_s": $t1 - $t0, "meryl_dump_s": $t2 - $t1, "meryl_threads": $NT,
     "identical_output": "$SAME" == "true"}
d["gpu_count_Gbases_per_s"] = d["bases"] / 1e9 / (d["gpu_count_ms"] * 1e-3)
open("$OUT/kmers_$TAG.json", "w").write(json.dumps(d) + "\n")
print(json.dumps(d))
PY
