#!/bin/bash
# bench.py at N GPUs the way the driver launches it (gpurun --gpus N), plus the multi-GPU parity check.
N=${1:-2}; TAG=${2:-r02}; STEPS=${3:-20}
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
if [ "$N" -gt 1 ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 tests/multi_gpu_check.py > gpurun_out/${TAG}_mgpu_check_n$N.log 2>&1
  echo "check rc=$?"; grep -E "parity ok|Error|error|assert" gpurun_out/${TAG}_mgpu_check_n$N.log | head -12
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus $N --steps $STEPS --warmup 5 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
else
  timeout 900 python bench.py --gpus 1 --steps $STEPS --warmup 5 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
fi
echo "bench rc=$?"; tail -c 5000 gpurun_out/${TAG}_bench_n$N.json; tail -5 gpurun_out/${TAG}_bench_n$N.err | cut -c1-300
