#!/bin/bash
for b in 3 4 5 6 7; do
  echo "== VGPU_BUCKET_BITS=$b"
  VGPU_BUCKET_BITS=$b timeout 300 python tools/kbench.py --only orderkey --iters 7 2>&1 | head -1 | cut -c1-330
done | tee gpurun_out/exp_sweep.log
CMD="python tools/kbench.py --only orderkey --iters 1 --warmup 3"
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/sweep_launches.csv $CMD > /dev/null 2>&1
