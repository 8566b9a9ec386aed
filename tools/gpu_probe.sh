#!/bin/bash
# fused-scan probe on one GPU + ncu capture of the fused kernel.  Usage: bash tools/gpu_probe.sh <tag> [world]
TAG=${1:-r02}; W=${2:-8}
mkdir -p gpurun_out
python tools/fused_scan_probe.py $W 10 > gpurun_out/${TAG}_probe.json 2> gpurun_out/${TAG}_probe.err && cat gpurun_out/${TAG}_probe.json
tail -3 gpurun_out/${TAG}_probe.err
ncu --set full --clock-control none --import-source on -k regex:vgpu_fused -s 7 -c 1 -f -o gpurun_out/${TAG}_fusedsend python tools/fused_scan_probe.py $W 2 > gpurun_out/${TAG}_probe_ncu.log 2>&1
tail -3 gpurun_out/${TAG}_probe_ncu.log
