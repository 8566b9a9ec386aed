#!/bin/bash
# compute-sanitizer over the small parity tests.  Usage (under gpurun): bash tools/gpu_sanitize.sh <tag> <memcheck|racecheck|initcheck|synccheck>
TAG=${1:-r02}; TOOL=${2:-memcheck}
mkdir -p gpurun_out
TESTS="tests/test_gpu_semantics.py tests/test_gpu_hashlog.py tests/test_gpu_selection.py"
timeout 600 python -m pytest $TESTS -m gpu -x -q > gpurun_out/${TAG}_sanitize_plain.log 2>&1 || { tail -5 gpurun_out/${TAG}_sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool ${TOOL} --error-exitcode 9 --print-limit 20 \
  python -m pytest $TESTS -m gpu -x -q > gpurun_out/${TAG}_sanitize_${TOOL}.log 2>&1
echo "rc=$?" >> gpurun_out/${TAG}_sanitize_${TOOL}.log
tail -15 gpurun_out/${TAG}_sanitize_${TOOL}.log
