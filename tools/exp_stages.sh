#!/bin/bash
# Where the C3 kernel's time goes: stage 1 = scan + run aggregation only, 2 = + home-slot loads, 0 = full.
for b in ${BITS:-0 5}; do for st in ${STAGES:-1 2 0}; do
  echo "== VGPU_BUCKET_BITS=$b VGPU_DEBUG_STAGE=$st"
  VGPU_CACHE_DIR=/tmp/vc_$st VGPU_BUCKET_BITS=$b VGPU_DEBUG_STAGE=$st timeout 300 python tools/kbench.py --only orderkey --iters 7 2>&1 | head -1 | cut -c1-200
done; done | tee gpurun_out/exp_stages.log
