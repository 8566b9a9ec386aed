"""The C++ operator shim (velox-1_b200/host: Velox's Operator API over the C-ABI) through its own test
binary, tests/host/host_tests (built by __graft_entry__.build()).  `cpu` needs no device: expression
parser, lowering, plan node types, registration seams, vector decoding.  `gpu` runs PlanBuilder plans
through the registered GPU operators and compares with oracle/liboracle.so and plain C++ answers."""
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(HERE, "host", "host_tests")


def run(mode):
    if not os.path.exists(BIN):
        subprocess.run(["make"], cwd=os.path.join(HERE, "host"), check=True)
    proc = subprocess.run([BIN, mode], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert proc.returncode == 0, proc.stdout[-4000:]
    assert "0 failed" in proc.stdout
    return proc.stdout


def test_host_logic_without_device():
    out = run("cpu")
    assert "registration" in out and "lowering" in out


def test_gpu_operator_fails_loudly_without_device():
    """No CPU fallback: on a box without a GPU a plan that needs a GPU operator must raise."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a box without a GPU")
    proc = subprocess.run([BIN, "gpu"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert proc.returncode != 0
    assert "vgpu_init" in proc.stdout or "CUDA" in proc.stdout


@pytest.mark.gpu
def test_gpu_operators_against_oracle():
    out = run("gpu")
    for name in ("FilterProject", "Q6", "Q1", "high cardinality", "dictionary VARCHAR", "edge cases"):
        assert name in out
