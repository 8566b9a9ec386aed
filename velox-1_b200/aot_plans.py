"""Pipeline shapes whose plan-specialised kernels build() compiles ahead of time with nvcc
(-gencode arch=compute_100a,code=sm_100a) into velox-1_b200/aot/<hash>.cubin.

Every other shape is compiled at first use with NVRTC from the same generator (csrc/fused.cu) and
the same prelude (csrc/fused_prelude.cuh); shapes the generator declines run on the micro-op
interpreter (csrc/pipeline.cu).  The list below is the single-GPU configurations of BASELINE.json
 plus the smoke test's shape, so the GPU box needs no compiler for
them.  Each entry is described exactly the way a caller binds it: expressions, column encodings /
nullability, aggregation mode, keys and accumulators.

Run as a script (or through __graft_entry__.build()) on a box WITHOUT a GPU: the library's dry-run
mode generates the kernel text, nvcc compiles it.
"""
import ctypes as C
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

from . import abi, expr as ex

HERE = os.path.dirname(os.path.abspath(__file__))
AOT_DIR = os.path.join(HERE, "aot")
PRELUDE = os.path.join(HERE, "csrc", "fused_prelude.cuh")
I32, I64, D = abi.INTEGER, abi.BIGINT, abi.DOUBLE

DATE_1994_01_01, DATE_1995_01_01, DATE_1998_09_02 = 8766, 9131, 10471


def _fake_column(type, encoding=abi.FLAT, nulls=False, base_nulls=False):
    """A column descriptor with plausible (aligned, never dereferenced) addresses."""
    c = abi.Column()
    c.type, c.encoding = type, encoding
    c.values = 0x10000
    c.indices = 0x20000 if encoding == abi.DICTIONARY else None
    c.nulls = 0x30000 if nulls else None
    c.base_nulls = 0x40000 if base_nulls else None
    c.base_length = 8 if encoding == abi.DICTIONARY else 0
    return c


def q1():
    rf, ls = ex.col(0, I32), ex.col(1, I32)
    qty, ep, disc, tax, sd = ex.col(2, D), ex.col(3, D), ex.col(4, D), ex.col(5, D), ex.col(6, I32)
    f = ex.call("lte", sd, ex.lit(DATE_1998_09_02, I32))
    disc_price = ex.call("multiply", ep, ex.call("minus", ex.lit(1.0, D), disc))
    charge = ex.call("multiply", disc_price, ex.call("plus", ex.lit(1.0, D), tax))
    outs = [rf, ls, qty, ep, disc_price, charge, disc]
    accs = [(abi.ACC_SUM_F64, 2), (abi.ACC_SUM_F64, 3), (abi.ACC_SUM_F64, 4), (abi.ACC_SUM_F64, 5),
            (abi.ACC_SUM_F64, 2), (abi.ACC_COUNT, 2), (abi.ACC_SUM_F64, 3), (abi.ACC_COUNT, 3),
            (abi.ACC_SUM_F64, 6), (abi.ACC_COUNT, 6), (abi.ACC_COUNT, -1)]
    keys = [(0, 0, 0, 2), (1, 0, 0, 1)]
    cols = [_fake_column(t) for t in (I32, I32, D, D, D, D, I32)]
    return dict(name="C2 Q1", filter=f, outputs=outs, cols=cols, mode=abi.AGG_ARRAY, keys=keys, accs=accs)


def q6_exprs():
    sd, ep, qty, disc = ex.col(0, I32), ex.col(1, D), ex.col(2, D), ex.col(3, D)
    f = ex.call("and",
                ex.call("gte", sd, ex.lit(DATE_1994_01_01, I32)),
                ex.call("lt", sd, ex.lit(DATE_1995_01_01, I32)),
                ex.call("between", disc, ex.lit(0.05, D), ex.lit(0.07, D)),
                ex.call("lt", qty, ex.lit(24.0, D)))
    return f, [ex.call("multiply", ep, disc)]


def q6_global():
    f, outs = q6_exprs()
    cols = [_fake_column(t) for t in (I32, D, D, D)]
    return dict(name="C1 Q6 global sum", filter=f, outputs=outs, cols=cols, mode=abi.AGG_GLOBAL, keys=[],
                accs=[(abi.ACC_SUM_F64, 0), (abi.ACC_COUNT, -1)])


def q6_filter_project(nulls_out):
    f, outs = q6_exprs()
    cols = [_fake_column(t) for t in (I32, D, D, D)]
    return dict(name="C1 Q6 FilterProject", filter=f, outputs=outs, cols=cols, mode=None, out_nulls=nulls_out)


def orderkey(with_count):
    outs = [ex.col(0, I64), ex.col(1, D)]
    cols = [_fake_column(I64), _fake_column(D)]
    accs = [(abi.ACC_SUM_F64, 1)] + ([(abi.ACC_COUNT, -1)] if with_count else [])
    return dict(name="C3 GROUP BY l_orderkey", filter=None, outputs=outs, cols=cols, mode=abi.AGG_HASH,
                keys=[(0, 64, 0, 0)], accs=accs)


def c4():
    cols = [_fake_column(I32, abi.DICTIONARY, nulls=True), _fake_column(I32, abi.DICTIONARY, nulls=True),
            _fake_column(D), _fake_column(I32), _fake_column(D, nulls=True)]
    outs = [ex.col(0, I32), ex.col(1, I32), ex.col(2, D), ex.col(3, I32), ex.col(4, D)]
    keys = [(0, 0, 0, 6), (1, 0, 0, 3)]
    accs = [(abi.ACC_MIN, 2), (abi.ACC_MAX, 2), (abi.ACC_MIN, 3), (abi.ACC_MAX, 3), (abi.ACC_COUNT, 4),
            (abi.ACC_COUNT, -1)]
    return dict(name="C4 dictionary keys with nulls", filter=None, outputs=outs, cols=cols, mode=abi.AGG_ARRAY,
                keys=keys, accs=accs)


def case_when():
    """TPC-H Q12 / Q14 shape: sums of CASE WHEN expressions grouped by a small key (special forms SELECT / COALESCE)."""
    kk, p, e, dd = ex.col(0, I32), ex.col(1, I32), ex.col(2, D), ex.col(3, D)
    rev = ex.call("multiply", e, ex.call("minus", ex.lit(1.0, D), dd))
    high = ex.call("if", ex.call("or", ex.call("eq", p, ex.lit(0, I32)), ex.call("eq", p, ex.lit(1, I32))),
                   ex.lit(1, I64), ex.lit(0, I64))
    promo = ex.call("if", ex.call("lt", p, ex.lit(2, I32)), rev, ex.lit(0.0, D))
    outs = [kk, high, promo, rev, ex.call("coalesce", ex.call("if", ex.call("gt", dd, ex.lit(0.05, D)), e), ex.lit(0.0, D))]
    keys = [(0, 0, 0, 6)]
    accs = [(abi.ACC_SUM_I64, 1), (abi.ACC_SUM_F64, 2), (abi.ACC_SUM_F64, 3), (abi.ACC_SUM_F64, 4), (abi.ACC_COUNT, -1)]
    cols = [_fake_column(t) for t in (I32, I32, D, D)]
    return dict(name="N3 sum(case when)", filter=None, outputs=outs, cols=cols, mode=abi.AGG_ARRAY, keys=keys, accs=accs)


def orderkey_streaming(world):
    """C5: the C3 scan with the exchange fused in (entries leave for the owners' windows), `world` ranks."""
    plan = orderkey(False)
    plan["name"] = f"C5 GROUP BY l_orderkey, exchange fused in, {world} ranks"
    plan["dry_send"] = f"{world},1"
    return plan


def plans():
    return [q1(), q6_global(), q6_filter_project(False), q6_filter_project(True), orderkey(False), orderkey(True), c4(),
            case_when(), orderkey_streaming(2), orderkey_streaming(4), orderkey_streaming(8)]


def generate(plan):
    """Dry-run the plan through the C-ABI; returns (source body, hash) or None when declined."""
    from . import engine  # Program / Aggregation wrappers (no device use in dry-run mode)
    lib = abi.lib()
    prog = engine.Program(plan["filter"], plan["outputs"], n_columns=len(plan["cols"]))
    arr = (abi.Column * len(plan["cols"]))(*plan["cols"])
    n = 1 << 20
    if plan["mode"] is None:
        outs = (abi.OutColumn * len(plan["outputs"]))()
        for i in range(len(plan["outputs"])):
            outs[i].values = 0x50000
            outs[i].nulls = 0x60000 if plan.get("out_nulls") else None
        abi.check(lib.vgpu_filter_project(prog.handle, arr, len(plan["cols"]), n, None, 0, 0x70000, 0x80000, 0x90000,
                                          outs, None))
    else:
        agg = engine.Aggregation(plan["mode"], plan["keys"], plan["accs"])
        if plan.get("dry_send"):
            os.environ["VGPU_DRY_SEND"] = plan["dry_send"]
        try:
            abi.check(lib.vgpu_agg_add_input(agg.handle, prog.handle, arr, len(plan["cols"]), n, None, 0, None))
        finally:
            os.environ.pop("VGPU_DRY_SEND", None)
    note = abi.last_kernel_note()
    if note != "dry run":
        return None, note
    return abi.last_kernel_source(), note


def build_all(verbose=True):
    """Generates and compiles every listed plan; returns the list of cubin paths."""
    lib = abi.lib()
    abi.check(lib.vgpu_debug_set_dry_run(1))
    os.makedirs(AOT_DIR, exist_ok=True)
    prelude = open(PRELUDE).read()
    jobs = []
    try:
        for plan in plans():
            got, note = generate(plan)
            if got is None:
                raise RuntimeError(f"{plan['name']}: generator declined the shape ({note})")
            body, h = got
            jobs.append((plan["name"], body, h))
    finally:
        abi.check(lib.vgpu_debug_set_dry_run(0))

    def compile_one(job):
        name, body, h = job
        cu = os.path.join(AOT_DIR, f"{h:016x}.cu")
        cubin = os.path.join(AOT_DIR, f"{h:016x}.cubin")
        if os.path.exists(cubin) and os.path.exists(cu) and open(cu).read() == prelude + "\n" + body:
            return cubin
        with open(cu, "w") as f:
            f.write(prelude + "\n" + body)
        cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
               "-cubin", "-o", cubin, cu]
        proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if proc.returncode != 0:
            raise RuntimeError(f"{name}: nvcc failed\n{proc.stdout[-6000:]}")
        return cubin

    with ThreadPoolExecutor(4) as pool:
        out = list(pool.map(compile_one, jobs))
    keep = {f"{h:016x}" for _, _, h in jobs}
    for fn in os.listdir(AOT_DIR):  # drop cubins of shapes no longer listed / of older generators
        if fn.split(".")[0] not in keep:
            os.remove(os.path.join(AOT_DIR, fn))
    if verbose:
        for (name, _, h), path in zip(jobs, out):
            print(f"aot: {name:40s} -> aot/{h:016x}.cubin")
    return out


if __name__ == "__main__":
    build_all()
