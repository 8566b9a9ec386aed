"""libvgpu.so loads without a GPU and exports every entry point include/vgpu.h declares; the
compute entry points fail loudly (no CPU fallback) when there is no device."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared():
    text = open(os.path.join(ROOT, "include", "vgpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\bint\s+(vgpu_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported():
    names = declared()
    assert len(names) >= 40
    lib = C.CDLL(os.path.join(ROOT, "velox-1_b200", "libvgpu.so"))
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.vgpu_abi_version() == 1


def test_host_library_loads():
    lib = C.CDLL(os.path.join(ROOT, "velox-1_b200", "libvelox_b200_host.so"))
    assert lib is not None


def test_no_cpu_fallback_without_device(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a box without a GPU")
    rc = pkg.abi.lib().vgpu_init(0)
    assert rc != 0  # VGPU_ERR_CUDA: nothing else takes over
    with pytest.raises(Exception):
        pkg.abi.check(rc)
