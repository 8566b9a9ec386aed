"""CPU checks of the short-string -> integer mapping VGPU_VARCHAR keys use (include/vgpu.h vgpu_number_of_string /
vgpu_string_of_number; the device's sv_number in csrc/fused_prelude.cuh computes the same value from a StringView):
the oracle's restatement of VectorHasher::stringAsNumber (velox/exec/VectorHasher.h:324-329), the Python host helpers
and the C header helpers (compiled here with gcc) agree, round-trip, and keep strings of different lengths apart."""
import ctypes
import os
import subprocess
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CASES = [b"", b"a", b"\x00", b"\x00\x00", b"ab", b"ba", b"A", b"N", b"R", b"AIR", b"REG AIR", b"\xff" * 7, b"a\x00b", b"1234567"]


def test_oracle_known_answers():
    import np_oracle as orc
    # loadPartialWord is little-endian; each length has its own range (marker bit at 8 * size)
    assert orc.string_as_number(b"") == 0
    assert orc.string_as_number(b"a") == 0x100 + 0x61
    assert orc.string_as_number(b"ab") == 0x10000 + 0x6261
    assert orc.string_as_number(b"\x00") == 0x100 and orc.string_as_number(b"\x00\x00") == 0x10000
    assert orc.string_as_number(b"\xff" * 7) == (1 << 56) + (1 << 56) - 1
    assert orc.string_as_number(b"12345678") is None
    nums = [orc.string_as_number(s) for s in CASES]
    assert len(set(nums)) == len(CASES)


def test_python_helpers_match_the_oracle(pkg):
    import importlib
    import np_oracle as orc
    engine = importlib.import_module("velox-1_b200.engine")
    for s in CASES:
        n = orc.string_as_number(s)
        assert engine.number_of_string(s) == n
        assert engine.string_of_number(n) == s
    views = engine.string_views(CASES + [b"a string that is far too long"])
    raw = views.view(np.uint8).reshape(-1, 16)
    for i, s in enumerate(CASES):
        assert int(raw[i, 0:4].view(np.uint32)[0]) == len(s) and bytes(raw[i, 4:4 + len(s)]) == s and not raw[i, 4 + len(s):].any()
    assert int(raw[-1, 0:4].view(np.uint32)[0]) == 29 and bytes(raw[-1, 4:8]) == b"a st"


def test_c_header_helpers_match_the_oracle():
    import np_oracle as orc
    src = r'''
#include "vgpu.h"
long long num(const char* d, unsigned n) { return vgpu_number_of_string(d, n); }
int str(long long v, char* out) { return vgpu_string_of_number(v, out); }
'''
    with tempfile.TemporaryDirectory() as tmp:
        c = os.path.join(tmp, "h.c")
        so = os.path.join(tmp, "h.so")
        open(c, "w").write(src)
        subprocess.run(["gcc", "-shared", "-fPIC", "-O1", "-I", os.path.join(ROOT, "include"), "-o", so, c], check=True)
        lib = ctypes.CDLL(so)
        lib.num.restype = ctypes.c_longlong
        lib.num.argtypes = [ctypes.c_char_p, ctypes.c_uint]
        lib.str.argtypes = [ctypes.c_longlong, ctypes.c_char_p]
        for s in CASES:
            n = lib.num(s, len(s))
            assert n == orc.string_as_number(s)
            buf = ctypes.create_string_buffer(8)
            assert lib.str(n, buf) == len(s) and buf.raw[:len(s)] == s
        assert lib.num(b"12345678", 8) == -1
        buf = ctypes.create_string_buffer(8)
        assert lib.str(5, buf) == -1 and lib.str(1 << 60, buf) == -1
