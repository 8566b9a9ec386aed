#!/bin/bash
# One short GPU-box visit: the GPU parity tests (optionally a subset), kernel timings, and (optional) the ncu launch list
# of one kbench case.  Usage (under gpurun): bash tools/gpu_quick.sh <tag> "<pytest args>" "<kbench cases>" [ncu case]
TAG=${1:-r02}; PYT=${2:-tests}; CASES=${3:-orderkey,randkey}; NCU=${4:-}
mkdir -p gpurun_out
(timeout 1500 python -m pytest $PYT -m gpu -q 2>&1 | tail -60) > gpurun_out/${TAG}_pytest.log
tail -30 gpurun_out/${TAG}_pytest.log
timeout 600 python tools/kbench.py --only $CASES > gpurun_out/${TAG}_kbench.jsonl 2> gpurun_out/${TAG}_kbench.err
cat gpurun_out/${TAG}_kbench.jsonl; tail -5 gpurun_out/${TAG}_kbench.err
if [ -n "$NCU" ]; then
  CMD="python tools/kbench.py --only $NCU --iters 2 --warmup 1"
  $CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_${NCU}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
  python tools/launch_summary.py gpurun_out/${TAG}_${NCU}_launches.csv | tail -40
fi
