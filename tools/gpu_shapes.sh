#!/bin/bash
# kbench of one case under several launch shapes of the plan-specialised kernel.  Usage: bash tools/gpu_shapes.sh <tag> <case> "<shape> <shape> ..."
TAG=${1:-r02}; CASE=${2:-orderkey}; SHAPES=${3:-"512,2 384,2"}
mkdir -p gpurun_out
for s in $SHAPES; do
  echo "shape $s" >> gpurun_out/${TAG}_shapes.jsonl
  VGPU_FUSED_SHAPE=$s timeout 300 python tools/kbench.py --only $CASE --engines log >> gpurun_out/${TAG}_shapes.jsonl 2>> gpurun_out/${TAG}_shapes.err
done
cut -c1-420 gpurun_out/${TAG}_shapes.jsonl
