#!/bin/bash
# One ncu --set full capture of a kbench case's kernel.  Usage (under gpurun): bash tools/gpu_ncu.sh <tag> <case> <kernel regex> [skip]
TAG=${1:-r02}; CASE=${2:-orderkey}; KREGEX=${3:-vgpu_fused}; SKIP=${4:-3}
mkdir -p gpurun_out
export VGPU_HASH_ENGINE=${VGPU_HASH_ENGINE:-log}
CMD="python tools/kbench.py --only ${CASE} --iters 1 --warmup 2 --engines ${VGPU_HASH_ENGINE}"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:${KREGEX} -s ${SKIP} -c 1 -f -o gpurun_out/${TAG}_${CASE} $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
tail -3 gpurun_out/${TAG}_ncu2.log; tail -3 gpurun_out/${TAG}_plain.log | cut -c1-300
