#!/bin/bash
# The reference's own GPU hash-table build (velox/experimental/gpu/tests/HashTableTest.cu, built by oracle/Makefile into
# oracle/_ref/ref_gpu_hashtable) on this box, next to this repo's C3: 2^27 slots at 0.5 occupancy = 67 M distinct keys.
TAG=${1:-r02}
mkdir -p gpurun_out
for part in 0 1; do for tags in 0 1; do
  echo "partitioned=$part use_tags=$tags" >> gpurun_out/${TAG}_prior_art.log
  VFLAG_table_size=134217728 VFLAG_partitioned=$part VFLAG_use_tags=$tags timeout 120 oracle/_ref/ref_gpu_hashtable >> gpurun_out/${TAG}_prior_art.log 2>&1
done; done
cat gpurun_out/${TAG}_prior_art.log
python tools/kbench.py --only orderkey --engines log 2>/dev/null | cut -c1-330 | tee -a gpurun_out/${TAG}_prior_art.log
