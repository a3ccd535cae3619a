#!/bin/bash
# SURVEY.md 8(f) N3 at full size: the overlap store of a whole sweep.  Runs the sweep (tools/gpu_sweep.sh), then builds the store from its
# .ovb files with canu_b200/bin/ovStoreBuild (GPU filter + mirror + sort) and with the reference's ovStoreBuild (oracle/_ref, host cores),
# times both, and compares every file of the two stores byte for byte.   tools/gpu_store.sh <tag> <genome_bp> <coverage> <hash_block_bp>
set -u
TAG=$1; GENOME=$2; COV=$3; BL=$4
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/gpurun_out
W=/tmp/oic_sweep_$TAG
bash $ROOT/tools/gpu_sweep.sh $TAG $GENOME $COV $BL 1 > $OUT/store_${TAG}_sweep.log 2>&1 || { tail -5 $OUT/store_${TAG}_sweep.log; exit 1; }
cd $W
ls out/*/*.ovb > ovb.list
t0=$(date +%s.%N)
$ROOT/canu_b200/bin/ovStoreBuild -O mine.ovlStore -G c.gkpStore $(cat ovb.list) > mine.log 2>&1 || { tail mine.log; exit 2; }
t1=$(date +%s.%N)
$ROOT/canu_b200/bin/ovStoreBuild -O mine2.ovlStore -G c.gkpStore $(cat ovb.list) > mine2.log 2>&1 || { tail mine2.log; exit 2; }
t2=$(date +%s.%N)
$ROOT/oracle/_ref/ovStoreBuild -O ref.ovlStore -G c.gkpStore -M 8 $(cat ovb.list) > ref.log 2>&1 || { tail ref.log; exit 3; }
t3=$(date +%s.%N)
same=true
for f in $(ls ref.ovlStore); do cmp -s ref.ovlStore/$f mine.ovlStore/$f || { same=false; echo "DIFFERENT: $f"; }; done
[ "$(ls ref.ovlStore | sort | tr '\n' ' ')" = "$(ls mine.ovlStore | sort | tr '\n' ' ')" ] || { same=false; echo "file lists differ"; }
python - <<PY
import json, os, re
sw = json.load(open("$OUT/sweep_$TAG.json"))
log = open("mine2.log").read()
m = re.findall(r"([0-9.]+) ms", log)
d = {"tag": "$TAG", "overlaps_in": sw["overlaps"], "ovb_files": sw["ovb_files"], "store_files": sorted(os.listdir("ref.ovlStore")),
     "store_bytes": sum(os.path.getsize(os.path.join("ref.ovlStore", f)) for f in os.listdir("ref.ovlStore")),
     "ours_wall_s_first": $t1 - $t0, "ours_wall_s": $t2 - $t1, "reference_wall_s": $t3 - $t2, "speedup_wall": ($t3 - $t2) / ($t2 - $t1),
     "byte_identical": "$same" == "true", "host_cores": os.cpu_count(),
     "ours_log_tail": log.strip().splitlines()[-12:], "reference": "oracle/_ref/ovStoreBuild -M 8 (the reference's own tool, same .ovb files)",
     "note": "wall time of the whole tool: read + decompress the .ovb files, filter + mirror + sort, write info / index / data / evalueLen / overlapScores"}
json.dump(d, open("$OUT/store_$TAG.json", "w"), indent=1)
print(json.dumps({k: d[k] for k in ("overlaps_in", "ours_wall_s_first", "ours_wall_s", "reference_wall_s", "speedup_wall", "byte_identical", "store_bytes")}))
print("\n".join(d["ours_log_tail"]))
PY
