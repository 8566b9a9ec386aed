"""CPU tests of the oracles' special forms (if / CASE WHEN, coalesce) and of left-to-right default-null argument
evaluation: hand-computed KATs, numpy oracle == C++ oracle (which narrows selectivity vectors the way SwitchExpr.cpp /
CoalesceExpr.cpp / Expr.cpp do), both == the reference's vendored DuckDB on the SQL spelling, and the lowering of the
new ops."""
import numpy as np
import pytest

import cpu_oracle as cpu
import duck_ref
import helpers as H
import np_oracle as orc
from helpers import abi, ex, D, I32, I64

NULL = None


def col64(vals):
    v = np.array([0 if x is None else x for x in vals], dtype=np.int64)
    nn = np.array([x is not None for x in vals])
    return v, nn


def as_list(v, nl):
    return [None if n else x.item() for x, n in zip(v, nl)]


def test_if_kat():
    a, a_nn = col64([10, 20, NULL, 40, 50])
    b, b_nn = col64([2, 0, 5, NULL, 0])
    x, y = ex.col(0, I64), ex.col(1, I64)
    cols = [(a, a_nn), (b, b_nn)]
    nz = ex.call("neq", y, ex.lit(0, I64))
    v, nl, er = orc.eval_expr(ex.call("if", nz, ex.call("divide", x, y), ex.lit(-1, I64)), cols, 5)
    assert not er.any()
    assert as_list(v, nl) == [5, -1, None, -1, -1]          # row 3: NULL condition takes ELSE
    v, nl, er = orc.eval_expr(ex.call("if", nz, ex.call("divide", x, y)), cols, 5)
    assert not er.any() and as_list(v, nl) == [5, None, None, None, None]
    v, nl, er = orc.eval_expr(ex.call("divide", x, y), cols, 5)
    assert er.tolist() == [False, True, False, False, True]  # outside the if the same division raises


def test_coalesce_kat():
    a, a_nn = col64([1, NULL, NULL, 4])
    b, b_nn = col64([9, 8, NULL, 0])
    x, y = ex.col(0, I64), ex.col(1, I64)
    cols = [(a, a_nn), (b, b_nn)]
    v, nl, er = orc.eval_expr(ex.call("coalesce", x, y), cols, 4)
    assert as_list(v, nl) == [1, 8, None, 4]
    v, nl, er = orc.eval_expr(ex.call("coalesce", x, y, ex.lit(7, I64)), cols, 4)
    assert as_list(v, nl) == [1, 8, 7, 4]
    # 100 / b raises on row 3 (b == 0), but a is not NULL there, so that argument is never evaluated
    v, nl, er = orc.eval_expr(ex.call("coalesce", x, ex.call("divide", ex.lit(100, I64), y)), cols, 4)
    assert not er.any() and as_list(v, nl) == [1, 12, None, 4]


def test_null_argument_hides_later_argument_error_kat():
    a, a_nn = col64([NULL, 2, NULL, 4])
    b, b_nn = col64([0, 1, 0, 0])
    x, y = ex.col(0, I64), ex.col(1, I64)
    cols = [(a, a_nn), (b, b_nn)]
    v, nl, er = orc.eval_expr(ex.call("plus", x, ex.call("divide", ex.lit(10, I64), y)), cols, 4)
    assert er.tolist() == [False, False, False, True] and nl.tolist() == [True, False, True, False]
    v, nl, er = orc.eval_expr(ex.call("plus", ex.call("divide", ex.lit(10, I64), y), x), cols, 4)
    assert er.tolist() == [True, False, True, True]


def _random_inputs(n=20_000, seed=1):
    rng = np.random.default_rng(seed)
    a = rng.integers(-50, 50, n).astype(np.int64)
    b = rng.integers(-3, 4, n).astype(np.int64)
    d = np.round(rng.normal(0, 10, n), 2)
    return n, a, b, d, rng.random(n) > 0.2, rng.random(n) > 0.2, rng.random(n) > 0.3


def _exprs():
    a, b, d = ex.col(0, I64), ex.col(1, I64), ex.col(2, D)
    nz = ex.call("neq", b, ex.lit(0, I64))
    case = ex.switch([(ex.call("lt", a, ex.lit(-10, I64)), ex.lit(0, I64)),
                      (ex.call("lt", a, ex.lit(10, I64)), ex.call("plus", a, ex.lit(100, I64))),
                      (ex.call("eq", b, ex.lit(0, I64)), b)], default=ex.call("multiply", a, b))
    return [ex.call("if", nz, ex.call("divide", a, b), ex.lit(-1, I64)),
            case,
            ex.call("coalesce", a, b, ex.lit(7, I64)),
            ex.call("if", ex.call("gt", d, ex.lit(0.0, D)), d, ex.call("negate", d)),
            ex.call("coalesce", d, ex.lit(0.0, D)),
            ex.call("plus", a, ex.call("if", nz, ex.call("divide", ex.lit(100, I64), b)))]


def test_cpp_oracle_matches_numpy_on_special_forms():
    n, a, b, d, a_nn, b_nn, d_nn = _random_inputs()
    outs = _exprs()
    filt = ex.call("gte", ex.call("coalesce", ex.col(2, D), ex.lit(-1.0, D)), ex.lit(-1.0, D))
    accs = [("sum_i64", 0), ("sum_i64", 1), ("sum_i64", 2), ("sum_f64", 3), ("sum_f64", 4), ("sum_i64", 5), ("count", 5), ("count", -1)]
    sel, o = orc.filter_project(filt, outs, [(a, a_nn), (b, b_nn), (d, d_nn)], n)
    want = orc.aggregate([], [(k, None if i < 0 else o[i]) for k, i in accs], sel, global_agg=True)[()]
    cols = [(4, a, ~a_nn), (4, b, ~b_nn), (6, d, ~d_nn)]  # cpu.run takes NULL masks
    got, _ = cpu.run(filt, outs, cols, n, [], accs, threads=2, batch_rows=1000)
    assert len(got) == 1
    for (gv, gc), (wv, wc), (kind, _) in zip(got[()], want, accs):
        assert gc == wc
        if kind == "sum_f64":
            assert abs(gv - wv) <= 1e-9 * max(1.0, abs(wv))
        else:
            assert gv == wv


@pytest.mark.skipif(not duck_ref.available(), reason="oracle/_ref/libduckdb_ref.so not built")
def test_special_forms_match_reference_duckdb():
    """The SQL spelling of the same expressions in the reference's own DuckDB (integer division written with a
    guarded divisor: DuckDB's `/` on integers does not raise, so only guarded forms are comparable)."""
    n, a, b, d, a_nn, b_nn, d_nn = _random_inputs(n=8000, seed=4)
    outs = _exprs()
    sel, o = orc.filter_project(None, outs, [(a, a_nn), (b, b_nn), (d, d_nn)], n)
    assert sel.all()
    DK = duck_ref
    with DK.Session(threads=1) as s:
        s.create_table("t", [("id", DK.INTEGER, np.arange(n, dtype=np.int32)), ("a", DK.BIGINT, a, None, a_nn),
                             ("b", DK.BIGINT, b, None, b_nn), ("d", DK.DOUBLE, d, None, d_nn)])
        # Presto integer division truncates toward zero; DuckDB 0.4's `/` on BIGINT does too
        vals, valid, _ = s.query(
            "select case when b <> 0 then a / b else -1 end, "
            "case when a < -10 then 0 when a < 10 then a + 100 when b = 0 then b else a * b end, "
            "coalesce(a, b, 7), case when d > 0.0 then d else -d end, coalesce(d, 0.0), "
            "a + case when b <> 0 then 100 / b end from t order by id")
    for k in range(len(outs)):
        wv, wn = o[k]
        assert np.array_equal(valid[:, k], ~wn), f"output {k} nulls"
        live = ~wn
        assert np.array_equal(vals[live, k], np.asarray(wv)[live].astype(np.float64)), f"output {k}"


def test_lowering_of_special_forms_validates():
    """SELECT / COALESCE ops pass the C-ABI program validation (no GPU needed); a BOOLEAN-valued if is rejected."""
    pkg = H.pkg
    outs = _exprs()
    low = ex.lower(None, outs)
    kinds = [op.opcode for op in low.ops]
    assert abi.OP_SELECT in kinds and abi.OP_COALESCE in kinds
    desc, keep = ex.to_desc(low, 3)
    prog = abi.C.c_void_p() if hasattr(abi, "C") else None
    import ctypes as C
    handle = C.c_void_p()
    rc = abi.lib().vgpu_program_create(C.byref(desc), C.byref(handle))
    assert rc == 0, abi.lib().vgpu_last_error()
    abi.lib().vgpu_program_destroy(handle)
    with pytest.raises(AssertionError):
        ex.call("if", ex.call("is_null", ex.col(0, I64)), ex.call("is_null", ex.col(1, I64)), ex.call("is_null", ex.col(2, D)))
