import importlib
import os
import sys

import pytest

# a plan-specialised kernel that fails to compile is a test failure, not a silent fall-back to the interpreter
os.environ.setdefault("VGPU_STRICT", "1")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def pkg():
    return importlib.import_module("velox-1_b200")


@pytest.fixture(scope="session")
def engine(pkg):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    eng = importlib.import_module("velox-1_b200.engine")
    pkg.abi.check(pkg.abi.lib().vgpu_init(0))
    return eng
