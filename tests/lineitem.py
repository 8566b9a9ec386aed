"""Test helper: lineitem columns from the reference's own dbgen (oracle/_ref/libdbgen_ref.so, built
by oracle/Makefile from the sources under /root/reference; the prebuilt .so travels to the GPU
box).  TEST INFRASTRUCTURE — the product never imports this."""
import ctypes as C
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libdbgen_ref.so")

COLUMNS = [
    ("l_orderkey", np.int64), ("l_partkey", np.int64), ("l_suppkey", np.int64), ("l_linenumber", np.int32),
    ("l_quantity", np.float64), ("l_extendedprice", np.float64), ("l_discount", np.float64), ("l_tax", np.float64),
    ("l_returnflag", np.int32), ("l_linestatus", np.int32), ("l_shipdate", np.int32), ("l_commitdate", np.int32),
    ("l_receiptdate", np.int32), ("l_shipinstruct", np.int32), ("l_shipmode", np.int32),
]
DICTS = {"l_returnflag": ["A", "N", "R"], "l_linestatus": ["F", "O"],
         "l_shipinstruct": ["COLLECT COD", "DELIVER IN PERSON", "NONE", "TAKE BACK RETURN"],
         "l_shipmode": ["AIR", "FOB", "MAIL", "RAIL", "REG AIR", "SHIP", "TRUCK"]}


class _Cols(C.Structure):
    _fields_ = [(name, C.c_void_p) for name, _ in COLUMNS]


_lib = None


def ref_lib():
    global _lib
    if _lib is None:
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(f"{REF_SO} missing: run `make -C oracle ref` where /root/reference exists")
        _lib = C.CDLL(REF_SO)
        _lib.ref_order_count.restype = C.c_int64
        _lib.ref_order_count.argtypes = [C.c_double]
        _lib.ref_gen_lineitem.restype = C.c_int64
        _lib.ref_gen_lineitem.argtypes = [C.c_double, C.c_int64, C.c_int64, C.POINTER(_Cols)]
    return _lib


def gen_range(sf, order_offset, max_orders, wanted=None):
    """Lineitems of orders [order_offset, order_offset + max_orders): dict name -> ndarray."""
    lib = ref_lib()
    cap = 7 * max_orders
    arrays = {}
    cols = _Cols()
    for name, dt in COLUMNS:
        if wanted is None or name in wanted:
            arrays[name] = np.empty(cap, dtype=dt)
            setattr(cols, name, arrays[name].ctypes.data)
    n = lib.ref_gen_lineitem(sf, order_offset, max_orders, C.byref(cols))
    return {k: v[:n] for k, v in arrays.items()}


def lineitem(sf, wanted=None, threads=None, cache=True):
    """Whole lineitem table at scale factor sf (cached under /tmp)."""
    key = "all" if wanted is None else "-".join(sorted(wanted))
    path = f"/tmp/vgpu_lineitem_sf{sf}_{abs(hash(key)) % 10**8}.npz"
    if cache and os.path.exists(path):
        with np.load(path) as z:
            return {k: z[k] for k in z.files}
    lib = ref_lib()
    orders = lib.ref_order_count(sf)
    gen_range(sf, 0, 1, wanted)  # one serial call: dbgen initialises its static tables lazily
    threads = threads or min(32, os.cpu_count() or 1)
    chunk = max(1, (orders + threads * 4 - 1) // (threads * 4))
    offsets = list(range(0, orders, chunk))
    with ThreadPoolExecutor(threads) as pool:
        parts = list(pool.map(lambda o: gen_range(sf, o, min(chunk, orders - o), wanted), offsets))
    out = {k: np.concatenate([p[k] for p in parts]) for k in parts[0]}
    if cache:
        try:
            np.savez(path, **out)
        except OSError:
            pass
    return out
