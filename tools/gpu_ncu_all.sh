#!/bin/bash
# One `ncu --set full` capture per single-GPU configuration's scan kernel (+ the HASH drain kernel).
# Usage under gpurun: bash tools/gpu_ncu_all.sh <tag>
TAG=${1:-r01}
mkdir -p gpurun_out
for CASE in q6 q6fp c4 orderkey; do
  CMD="python tools/kbench.py --only ${CASE} --iters 1 --warmup 3"
  $CMD > gpurun_out/${TAG}_plain_${CASE}.log 2>&1 || { echo "$CASE failed without ncu"; continue; }
  ncu --set full --clock-control none --import-source on -k regex:vgpu_fused -s 3 -c 1 -f -o gpurun_out/${TAG}_${CASE} $CMD > gpurun_out/${TAG}_ncu_${CASE}.log 2>&1
  tail -1 gpurun_out/${TAG}_ncu_${CASE}.log
done
CMD="python tools/kbench.py --only orderkey --iters 1 --warmup 3"
ncu --set full --clock-control none --import-source on -k regex:drain_overflow -s 4 -c 1 -f -o gpurun_out/${TAG}_drain $CMD > gpurun_out/${TAG}_ncu_drain.log 2>&1
tail -1 gpurun_out/${TAG}_ncu_drain.log
