#!/bin/bash
# C3 (GROUP BY l_orderkey) kernel time for the table placements: plain hash (0) and 2^b-slot buckets.
mkdir -p gpurun_out
for b in ${BITS:-0 4 5 6}; do
  echo "== VGPU_BUCKET_BITS=$b"
  VGPU_BUCKET_BITS=$b timeout 300 python tools/kbench.py --only orderkey --iters 7 2>&1 | cut -c1-260
done | tee gpurun_out/exp_buckets.log
VGPU_BUCKET_BITS=5 timeout 600 python -m pytest tests -m gpu -x -q  2>&1 | tail -3
