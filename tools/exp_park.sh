#!/bin/bash
for cfg in "0 0" "5 0" "0 1" ; do set -- $cfg
  echo "== VGPU_BUCKET_BITS=$1 VGPU_NO_PARKING=$2"
  VGPU_BUCKET_BITS=$1 VGPU_NO_PARKING=$2 timeout 300 python tools/kbench.py --only orderkey --iters 7 2>&1 | head -1 | cut -c1-330
done | tee gpurun_out/exp_park.log
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
VGPU_BUCKET_BITS=0 timeout 600 python -m pytest tests -m gpu -x -q -k "orderkey or nulls" 2>&1 | tail -3
