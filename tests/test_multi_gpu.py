"""Launches tests/multi_gpu_check.py on 2 GPUs (one process per GPU, NCCL) when the box has them."""
import os
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.gpu
def test_two_gpu_partition_alltoall_final():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(HERE, "multi_gpu_check.py")]
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert proc.returncode == 0, proc.stdout[-4000:]
    assert proc.stdout.count("multi-gpu parity ok") == 8, proc.stdout[-4000:]  # 5 + the 3 fused-exchange runs
