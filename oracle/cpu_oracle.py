"""TEST INFRASTRUCTURE — ctypes wrapper of oracle/liboracle.so (velox_cpu.cpp), the C++ CPU
restatement of the reference path that bench.py times as the CPU baseline.  Only tests/,
smoke() and bench.py's cpu_baseline / --impl reference legs import this."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "liboracle.so")
SUFFIX = {0: "b", 3: "i32", 4: "i64", 6: "f64"}
NP = {0: np.uint8, 3: np.int32, 4: np.int64, 6: np.float64}
KIND = {"sum_f64": 0, "sum_i64": 1, "count": 2, "min": 3, "max": 4}


class _Col(C.Structure):
    _fields_ = [("type", C.c_int32), ("pad", C.c_int32), ("values", C.c_void_p), ("nulls", C.c_void_p)]


class _Acc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("input", C.c_int32)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            raise FileNotFoundError(f"{SO} missing: run `make -C oracle liboracle.so`")
        _lib = C.CDLL(SO)
        _lib.vcpu_run.restype = C.c_int64
        _lib.vcpu_run.argtypes = [C.c_char_p, C.POINTER(C.c_char_p), C.c_int32, C.POINTER(_Col), C.c_int32, C.c_int64,
                                  C.POINTER(C.c_int32), C.c_int32, C.POINTER(_Acc), C.c_int32, C.c_int32, C.c_int32,
                                  C.c_int32, C.c_void_p, C.c_int64, C.POINTER(C.c_double)]
        _lib.vcpu_error.restype = C.c_char_p
    return _lib


def to_sexpr(e):
    """Expression tree (duck-typed like np_oracle) -> the s-expression text velox_cpu.cpp parses."""
    t = SUFFIX[e.type]
    if hasattr(e, "index"):
        return f"c{e.index}:{t}"
    if not hasattr(e, "fn"):
        if e.value is None:
            return f"null:{t}"
        if e.type == 6:
            return f"{float(e.value)!r}:{t}"
        if e.type == 0:
            return f"{'true' if e.value else 'false'}:{t}"
        return f"{int(e.value)}:{t}"
    return "(" + " ".join([e.fn] + [to_sexpr(a) for a in e.args]) + ")"


class OracleError(Exception):
    pass


def run(filter_expr, outputs, cols, n, key_outputs, accs, ignore_null_keys=False, threads=1, batch_rows=4096,
        max_groups=None):
    """cols: [(type, values ndarray, null mask (True = NULL) or None)]; accs: [(kind name, output index)].
    Returns ({key tuple: [(value, non-null count)]}, seconds)."""
    keep = []
    carr = (_Col * max(1, len(cols)))()
    for i, (t, v, nl) in enumerate(cols):
        v = np.ascontiguousarray(v, dtype=NP[t])
        keep.append(v)
        carr[i].type = t
        carr[i].values = v.ctypes.data
        if nl is not None:
            nb = np.ascontiguousarray(nl, dtype=np.uint8)
            keep.append(nb)
            carr[i].nulls = nb.ctypes.data
    outs = (C.c_char_p * max(1, len(outputs)))(*[to_sexpr(o).encode() for o in outputs])
    out_types = [o.type for o in outputs]
    keys = (C.c_int32 * max(1, len(key_outputs)))(*key_outputs)
    aarr = (_Acc * max(1, len(accs)))(*[_Acc(KIND[k], i) for k, i in accs])
    nk, na = len(key_outputs), len(accs)
    words = 2 * nk + 2 * na
    cap = max_groups if max_groups is not None else max(1024, n)
    buf = np.zeros((cap, max(words, 1)), dtype=np.int64)
    secs = C.c_double()
    got = lib().vcpu_run(to_sexpr(filter_expr).encode() if filter_expr is not None else b"", outs, len(outputs), carr,
                         len(cols), n, keys, nk, aarr, na, 1 if ignore_null_keys else 0, threads, batch_rows,
                         buf.ctypes.data, cap, C.byref(secs))
    if got < 0:
        raise OracleError(lib().vcpu_error().decode())
    res = {}
    for r in buf[: min(got, cap)]:
        key = tuple(None if r[nk + k] else int(r[k]) for k in range(nk))
        vals = []
        for a, (kind, idx) in enumerate(accs):
            raw, nn = r[2 * nk + 2 * a], int(r[2 * nk + 2 * a + 1])
            t = out_types[idx] if idx >= 0 else 4
            fp = kind == "sum_f64" or (kind in ("min", "max") and t == 6)
            vals.append((float(raw.view(np.float64)) if fp else int(raw), nn))
        res[key] = vals
    return res, secs.value
