#!/bin/bash
# one kernel-work iteration on the B200 box: parity first (CUDA path vs goldens/oracle + the full-size c3 job vs the reference binary), then
# quick bench lines under the given env settings;  tools/gpu_iter.sh <tag> "ENV=.." ...
TAG=$1; shift
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py "tests/test_gpu_fullsize_ref.py::test_c3_bench_job_records_equal_reference" -x -q > gpurun_out/pytest_$TAG.log 2>&1
echo "parity rc=$?"; tail -4 gpurun_out/pytest_$TAG.log
bash tools/gpu_quick.sh $TAG "$@"
