"""B200-native FilterProject -> HashAggregation (the Velox hot path), Python host side.

The product is libvgpu.so (csrc/, C-ABI in include/vgpu.h) and the C++ operator shim (host/);
this package is the thin Python host used by the tests and bench.py: ctypes prototypes (abi),
typed expressions and their lowering (expr), and device-buffer helpers built on torch (engine).
There is no CPU implementation in here: every compute call goes through the C-ABI.
"""
from . import abi, expr  # noqa: F401

__all__ = ["abi", "expr", "engine"]


def __getattr__(name):
    if name == "engine":
        import importlib
        return importlib.import_module(__name__ + ".engine")
    raise AttributeError(name)
