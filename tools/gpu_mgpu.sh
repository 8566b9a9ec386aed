#!/bin/bash
# Multi-GPU visit (gpurun --gpus N): parity of the exchange over NCCL / CUDA IPC, then the C5 group-by leg timing.
N=${1:-2}; TAG=${2:-r02}
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 tests/multi_gpu_check.py > gpurun_out/${TAG}_mgpu_check_n$N.log 2>&1
echo "check rc=$?"; grep -E "parity ok|Error|error|assert|Traceback" gpurun_out/${TAG}_mgpu_check_n$N.log | head -20; tail -5 gpurun_out/${TAG}_mgpu_check_n$N.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 tools/mgpu_bench.py > gpurun_out/${TAG}_mgpu_bench_n$N.log 2>&1
echo "bench rc=$?"; grep -E "^\{" gpurun_out/${TAG}_mgpu_bench_n$N.log; tail -3 gpurun_out/${TAG}_mgpu_bench_n$N.log | cut -c1-300
