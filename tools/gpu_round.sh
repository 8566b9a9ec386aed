#!/bin/bash
# One GPU-box visit: parity tests, kernel timings of C1-C4, launch list + full ncu capture of the
# named case.  Usage (under gpurun): bash tools/gpu_round.sh <tag> <case> [kernel regex]
TAG=${1:-r01}; CASE=${2:-q1}; KREGEX=${3:-vgpu_fused}
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/${TAG}_pytest.log
cat gpurun_out/${TAG}_pytest.log | tail -5
(VGPU_FUSED=0 timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/${TAG}_pytest_interp.log
tail -3 gpurun_out/${TAG}_pytest_interp.log
timeout 600 python tools/kbench.py > gpurun_out/${TAG}_kbench.jsonl 2> gpurun_out/${TAG}_kbench.err
cat gpurun_out/${TAG}_kbench.jsonl; tail -3 gpurun_out/${TAG}_kbench.err
CMD="python tools/kbench.py --only ${CASE} --iters 1 --warmup 3"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG}_${CASE}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:${KREGEX} -s 3 -c 1 -f -o gpurun_out/${TAG}_${CASE} $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
tail -3 gpurun_out/${TAG}_ncu2.log
