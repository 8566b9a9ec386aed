"""-m gpu tests of expression / selection semantics at the edges the reference tests cover
(ConjunctTest.cpp 3x3 truth tables, ComparisonsTest.cpp between KAT, FilterProjectTest.cpp
dictionary / constant / all-pass / all-fail, SumTest.cpp overflow, ragged and empty batches)."""
import numpy as np
import pytest
import torch

import helpers as H
import np_oracle as orc
from helpers import abi, ex, D, I32, I64, B

pytestmark = pytest.mark.gpu
NPT = orc.NP


def check(engine, filt, outs, spec, n, in_rows=None):
    """spec: list of (type, values, not_null or None).  Runs GPU + oracle, compares everything."""
    cols = [engine.flat_column(t, v, nn) for t, v, nn in spec]
    prog = engine.Program(filt, outs, n_columns=len(spec))
    rows_t = None
    ospec = [(v, nn) for _, v, nn in spec]
    total = n
    if in_rows is not None:
        rows_t = torch.from_numpy(in_rows.astype(np.int32)).cuda()
        ospec = [(np.asarray(v)[in_rows], None if nn is None else np.asarray(nn)[in_rows]) for v, nn in ospec]
        total = len(in_rows)
    res = engine.filter_project(prog, cols, n, in_rows=rows_t)
    sel, want = orc.filter_project(filt, outs, ospec, total)
    got_sel = engine.unpack_bits(res.sel_bits.cpu().numpy().view(np.uint64), total)
    assert np.array_equal(got_sel, sel)
    assert res.num_selected == int(sel.sum())
    assert np.array_equal(res.sel_idx[: res.num_selected].cpu().numpy(), np.nonzero(sel)[0])
    for k, o in enumerate(outs):
        wv, wn = want[k]
        nn_bits = engine.unpack_bits(res.nulls[k].cpu().numpy().view(np.uint64), total)
        assert np.array_equal(~nn_bits[sel], wn[sel]), f"output {k} nulls"
        live = sel & ~wn
        if o.type == B:
            gv = engine.unpack_bits(res.values[k].cpu().numpy().view(np.uint64), total)
        else:
            gv = res.values[k].cpu().numpy()[:total]
        a, b = gv[live], np.asarray(wv)[live]
        if a.dtype.kind == "f":
            assert np.array_equal(a.view(f"u{a.itemsize}"), b.astype(a.dtype).view(f"u{a.itemsize}")), f"output {k}"
        else:
            assert np.array_equal(a, b), f"output {k}"
    return res


@pytest.mark.parametrize("n", [1, 31, 64, 1023, 1024, 1025, 2049, 5000, 148 * 1024 + 3])
def test_ragged_sizes(engine, n):
    rng = np.random.default_rng(n)
    a = rng.integers(-100, 100, n).astype(np.int32)
    x = rng.normal(size=n)
    nn = rng.random(n) > 0.2
    f = ex.call("and", ex.call("gt", ex.col(0, I32), ex.lit(-20, I32)), ex.call("lt", ex.col(1, D), ex.lit(0.7, D)))
    outs = [ex.call("multiply", ex.col(1, D), ex.call("plus", ex.col(1, D), ex.lit(2.5, D))), ex.col(0, I32)]
    check(engine, f, outs, [(I32, a, None), (D, x, nn)], n)


def test_empty_batch(engine):
    prog = engine.Program(ex.call("gt", ex.col(0, I32), ex.lit(0, I32)), [ex.col(0, I32)])
    res = engine.filter_project(prog, [engine.flat_column(I32, np.zeros(0, np.int32))], 0)
    assert res.num_selected == 0


def test_conjunct_truth_tables(engine):
    # all 3x3 combinations of TRUE / FALSE / NULL (velox/expression/tests/ConjunctTest.cpp:30-110)
    vals = [(True, True), (False, True), (False, False)]  # (value, not_null)
    a = np.array([vals[i][0] for i in range(3) for _ in range(3)])
    an = np.array([vals[i][1] for i in range(3) for _ in range(3)])
    b = np.array([vals[j][0] for _ in range(3) for j in range(3)])
    bn = np.array([vals[j][1] for _ in range(3) for j in range(3)])
    spec = [(B, a, an), (B, b, bn)]
    ca, cb = ex.col(0, B), ex.col(1, B)
    for fn in ("and", "or"):
        e = ex.call(fn, ca, cb)
        res = check(engine, e, [e, ex.call("not", e), ex.call("is_null", e)], spec, 9)
    # SQL: T and N = N, F and N = F, T or N = T, F or N = N
    e = ex.call("and", ca, cb)
    v, nl, _ = orc.eval_expr(e, [(a, an), (b, bn)], 9)
    assert list(nl) == [False, False, True, False, False, False, True, False, True]
    assert list(v & ~nl) == [True] + [False] * 8


def test_between_kat(engine):
    # velox/functions/prestosql/tests/ComparisonsTest.cpp:57-68
    x = np.array([0, 1, 4, 5, 10, -1, -4, -5, -10, 0], dtype=np.int32)
    nn = np.array([True] * 9 + [False])
    e = ex.call("between", ex.col(0, I32), ex.lit(-1, I32), ex.lit(5, I32))
    res = check(engine, e, [e], [(I32, x, nn)], 10)
    sel = engine.unpack_bits(res.sel_bits.cpu().numpy().view(np.uint64), 10)
    assert list(sel) == [True, True, True, True, False, True, False, False, False, False]
    # null bounds make the result NULL, unlike AND of two comparisons
    lo = np.array([0, 5, 0], dtype=np.int64)
    hi_nn = np.array([True, False, True])
    e2 = ex.call("between", ex.col(0, I64), ex.col(1, I64), ex.col(2, I64))
    check(engine, None, [e2, ex.call("is_null", e2)], [(I64, np.array([1, 1, 9]), None), (I64, lo, None), (I64, np.array([3, 3, 3]), hi_nn)], 3)


def test_comparisons_all_types_and_nan(engine):
    rng = np.random.default_rng(3)
    n = 4099
    for t in (abi.TINYINT, abi.SMALLINT, I32, I64, abi.REAL, D):
        if t in (abi.REAL, D):
            a = rng.normal(size=n).astype(NPT[t]); b = rng.normal(size=n).astype(NPT[t])
            a[::17] = np.nan; b[::23] = np.nan; b[::5] = a[::5]
        else:
            info = np.iinfo(NPT[t])
            a = rng.integers(info.min, info.max, n, dtype=NPT[t], endpoint=True)
            b = rng.integers(info.min, info.max, n, dtype=NPT[t], endpoint=True)
            b[::5] = a[::5]
        nn = rng.random(n) > 0.1
        outs = [ex.call(fn, ex.col(0, t), ex.col(1, t)) for fn in ("eq", "neq", "lt", "lte", "gt", "gte")]
        check(engine, None, outs, [(t, a, nn), (t, b, None)], n)


def test_double_arithmetic_bit_exact(engine):
    rng = np.random.default_rng(5)
    n = 10007
    a = rng.normal(scale=1e3, size=n); b = rng.normal(size=n); b[::13] = 0.0
    x, y = ex.col(0, D), ex.col(1, D)
    outs = [ex.call("plus", x, y), ex.call("minus", x, y), ex.call("multiply", x, y), ex.call("divide", x, y),
            ex.call("mod", x, y), ex.call("negate", x),
            ex.call("multiply", ex.call("multiply", x, ex.call("minus", ex.lit(1.0, D), y)), ex.call("plus", ex.lit(1.0, D), y))]
    check(engine, None, outs, [(D, a, None), (D, b, rng.random(n) > 0.05)], n)


def test_real_arithmetic_rounds_like_float(engine):
    rng = np.random.default_rng(6)
    n = 5003
    a = rng.normal(scale=100, size=n).astype(np.float32); b = rng.normal(size=n).astype(np.float32)
    x, y = ex.col(0, abi.REAL), ex.col(1, abi.REAL)
    outs = [ex.call("plus", x, y), ex.call("multiply", x, y), ex.call("divide", x, y), ex.call("minus", x, y)]
    check(engine, None, outs, [(abi.REAL, a, None), (abi.REAL, b, None)], n)


def test_checked_integer_arithmetic(engine):
    a = np.array([1, 2**62, -2**63, 7, 100, -7], dtype=np.int64)
    b = np.array([2, 2**62, -1, 0, 3, 2], dtype=np.int64)
    x, y = ex.col(0, I64), ex.col(1, I64)
    ok = np.array([True, False, False, False, True, True])
    # rows that would raise are filtered out first: no error (ConjunctExpr/deferred errors)
    f = ex.col(2, B)
    outs = [ex.call("plus", x, y), ex.call("multiply", x, y), ex.call("divide", x, y), ex.call("mod", x, y), ex.call("minus", x, y)]
    check(engine, f, outs, [(I64, a, None), (I64, b, None), (B, ok, None)], 6)
    for fn, rows, msg in (("plus", [1], "overflow"), ("divide", [3], "division by zero"), ("divide", [2], "overflow"),
                          ("multiply", [1], "overflow")):
        sel = np.zeros(6, dtype=bool); sel[rows] = True
        prog = engine.Program(ex.col(2, B), [ex.call(fn, x, y)])
        cols = [engine.flat_column(I64, a), engine.flat_column(I64, b), engine.flat_column(B, sel)]
        with pytest.raises(abi.VgpuError) as ei:
            engine.filter_project(prog, cols, 6)
        assert ei.value.status == abi.ERR_USER and msg in ei.value.message
    # int32 arithmetic overflows at int32 bounds (checkedPlus<int32_t>)
    prog = engine.Program(None, [ex.call("plus", ex.col(0, I32), ex.col(0, I32))])
    with pytest.raises(abi.VgpuError):
        engine.filter_project(prog, [engine.flat_column(I32, np.array([2**30, 2**30 + 5], np.int32))], 2)


def test_error_suppressed_by_conjunct(engine):
    # (b <> 0) AND (a / b > 1): the division error on b == 0 is dropped because the other conjunct
    # decides the row (velox/expression/ConjunctExpr.cpp:67-93); OR symmetric.
    a = np.array([10, 10, 10, 3], dtype=np.int64); b = np.array([0, 2, 5, 0], dtype=np.int64)
    x, y = ex.col(0, I64), ex.col(1, I64)
    div = ex.call("gt", ex.call("divide", x, y), ex.lit(1, I64))
    e = ex.call("and", ex.call("neq", y, ex.lit(0, I64)), div)
    res = check(engine, e, [x], [(I64, a, None), (I64, b, None)], 4)
    assert res.num_selected == 2
    e2 = ex.call("and", div, ex.call("neq", y, ex.lit(0, I64)))  # order must not matter
    check(engine, e2, [x], [(I64, a, None), (I64, b, None)], 4)
    with pytest.raises(abi.VgpuError):
        engine.filter_project(engine.Program(div, [x]), [engine.flat_column(I64, a), engine.flat_column(I64, b)], 4)


def test_casts(engine):
    d = np.array([0.5, 1.5, -0.5, -2.5, 2.4999, 1e9, -1e9, 0.49999999999999994, 123456.5], dtype=np.float64)
    outs = [ex.cast(ex.col(0, D), I64), ex.cast(ex.col(0, D), I32), ex.cast(ex.col(1, I32), D), ex.cast(ex.col(1, I32), I64),
            ex.cast(ex.col(0, D), abi.REAL), ex.cast(ex.col(1, I32), abi.SMALLINT)]
    i = np.array([1, -1, 0, 32767, -32768, 5, 6, 7, 8], dtype=np.int32)
    check(engine, None, outs, [(D, d, None), (I32, i, None)], 9)
    for bad in (np.nan, 3e9, -3e9):
        prog = engine.Program(None, [ex.cast(ex.col(0, D), I32)])
        with pytest.raises(abi.VgpuError) as ei:
            engine.filter_project(prog, [engine.flat_column(D, np.array([1.0, bad]))], 2)
        assert ei.value.status == abi.ERR_USER
    with pytest.raises(abi.VgpuError):
        engine.filter_project(engine.Program(None, [ex.cast(ex.col(0, I32), abi.SMALLINT)]),
                              [engine.flat_column(I32, np.array([40000], np.int32))], 1)


def test_dictionary_constant_and_in_rows(engine):
    rng = np.random.default_rng(11)
    n = 7001
    base = rng.normal(size=50); base_nn = rng.random(50) > 0.2
    idx = rng.integers(0, 50, n).astype(np.int32); outer_nn = rng.random(n) > 0.1
    flat = rng.integers(0, 10, n).astype(np.int32)
    cols = [engine.dictionary_column(D, base, idx, outer_nn, base_nn), engine.flat_column(I32, flat),
            engine.constant_column(D, 2.0, n), engine.constant_column(D, None, n)]
    x, k, c2, cn = ex.col(0, D), ex.col(1, I32), ex.col(2, D), ex.col(3, D)
    f = ex.call("gt", ex.call("multiply", x, c2), ex.lit(-1.0, D))
    outs = [ex.call("plus", x, c2), ex.call("plus", x, cn), k]
    prog = engine.Program(f, outs)
    ov, onn = orc.decode(base, idx, outer_nn, base_nn)
    ospec = [(ov, onn), (flat, None), (np.full(n, 2.0), None), (np.zeros(n), np.zeros(n, bool))]
    for in_rows in (None, np.sort(rng.choice(n, 3000, replace=False))):
        total = n if in_rows is None else len(in_rows)
        rt = None if in_rows is None else torch.from_numpy(in_rows.astype(np.int32)).cuda()
        res = engine.filter_project(prog, cols, n, in_rows=rt)
        osp = ospec if in_rows is None else [(v[in_rows], None if m is None else m[in_rows]) for v, m in ospec]
        sel, want = orc.filter_project(f, outs, osp, total)
        assert np.array_equal(engine.unpack_bits(res.sel_bits.cpu().numpy().view(np.uint64), total), sel)
        for kk in range(3):
            nn_bits = engine.unpack_bits(res.nulls[kk].cpu().numpy().view(np.uint64), total)
            assert np.array_equal(~nn_bits[sel], want[kk][1][sel])
            live = sel & ~want[kk][1]
            assert np.array_equal(res.values[kk].cpu().numpy()[:total][live], want[kk][0][live])


def test_all_pass_all_fail(engine):
    n = 3000
    a = np.arange(n, dtype=np.int32)
    for lo, expect in ((-1, n), (n, 0)):
        prog = engine.Program(ex.call("gt", ex.col(0, I32), ex.lit(lo, I32)), [ex.col(0, I32)])
        res = engine.filter_project(prog, [engine.flat_column(I32, a)], n)
        assert res.num_selected == expect


def _switch_inputs(n=4099, seed=3):
    rng = np.random.default_rng(seed)
    a = rng.integers(-50, 50, n).astype(np.int64)
    b = rng.integers(-3, 4, n).astype(np.int64)          # many zeros: a / b raises where b == 0
    d = np.round(rng.normal(0, 10, n), 2)
    a_nn, b_nn, d_nn = rng.random(n) > 0.2, rng.random(n) > 0.2, rng.random(n) > 0.3
    return n, [(I64, a, a_nn), (I64, b, b_nn), (D, d, d_nn)]


def test_if_takes_branch_per_row_and_drops_untaken_errors(engine):
    """if(b <> 0, a / b, -1): the division error on b == 0 is in the branch those rows do not take
    (SwitchExpr.cpp:64-160); a NULL condition takes ELSE; no ELSE gives NULL."""
    n, spec = _switch_inputs()
    a, b, d = ex.col(0, I64), ex.col(1, I64), ex.col(2, D)
    nz = ex.call("neq", b, ex.lit(0, I64))
    outs = [ex.call("if", nz, ex.call("divide", a, b), ex.lit(-1, I64)),
            ex.call("if", nz, ex.call("divide", a, b)),                                  # no ELSE
            ex.call("if", ex.call("gt", d, ex.lit(0.0, D)), d, ex.call("negate", d)),      # abs()
            ex.call("if", ex.call("is_null", a), ex.lit(0, I64), a),
            ex.call("if", ex.call("lt", a, b), ex.lit(1.5, D), ex.lit(-2.5, D))]           # two literals
    res = check(engine, None, outs, spec, n)
    # the same division outside the if raises
    with pytest.raises(abi.VgpuError):
        cols = [engine.flat_column(t, v, nn) for t, v, nn in spec]
        engine.filter_project(engine.Program(None, [ex.call("divide", a, b)], n_columns=3), cols, n)
    # an error in the condition itself is kept
    with pytest.raises(abi.VgpuError):
        bad = ex.call("if", ex.call("gt", ex.call("divide", a, b), ex.lit(0, I64)), a, b)
        engine.filter_project(engine.Program(None, [bad], n_columns=3), cols, n)
    # ... unless a conjunct decides the row first
    guarded = ex.call("if", ex.call("and", nz, ex.call("gt", ex.call("divide", a, b), ex.lit(0, I64))), a, b)
    check(engine, None, [guarded], spec, n)
    assert res.num_selected == n


def test_case_when_chain_and_coalesce(engine):
    n, spec = _switch_inputs(seed=11)
    a, b, d = ex.col(0, I64), ex.col(1, I64), ex.col(2, D)
    case = ex.switch([(ex.call("lt", a, ex.lit(-10, I64)), ex.lit(0, I64)),
                      (ex.call("lt", a, ex.lit(10, I64)), ex.call("plus", a, ex.lit(100, I64))),
                      (ex.call("eq", b, ex.lit(0, I64)), b)], default=ex.call("multiply", a, b))
    case_no_else = ex.switch([(ex.call("gt", d, ex.lit(5.0, D)), d), (ex.call("lt", d, ex.lit(-5.0, D)), ex.lit(0.0, D))])
    outs = [case, case_no_else,
            ex.call("coalesce", a, b), ex.call("coalesce", a, b, ex.lit(7, I64)),
            ex.call("coalesce", d, ex.lit(0.0, D)),
            # b's error (division by zero) only counts where a is NULL
            ex.call("coalesce", a, ex.call("if", ex.call("neq", b, ex.lit(0, I64)), ex.call("divide", ex.lit(100, I64), b), ex.lit(0, I64)))]
    filt = ex.call("gte", ex.call("coalesce", d, ex.lit(-1.0, D)), ex.lit(-1.0, D))
    check(engine, filt, outs, spec, n)


def test_null_argument_hides_later_argument_error(engine):
    """Default-null calls evaluate arguments left to right and drop a row at its first NULL argument, so an
    error in a LATER argument does not surface on it (Expr.cpp:371-441 evalArgsDefaultNulls)."""
    a = np.array([1, 2, 3, 4], dtype=np.int64); a_nn = np.array([False, True, False, True])
    b = np.array([0, 1, 0, 2], dtype=np.int64)
    x, y = ex.col(0, I64), ex.col(1, I64)
    spec = [(I64, a, a_nn), (I64, b, None)]
    e = ex.call("plus", x, ex.call("divide", ex.lit(10, I64), y))   # rows 0 and 2: x NULL, 10 / 0
    res = check(engine, None, [e, ex.call("lt", x, ex.call("divide", ex.lit(10, I64), y))], spec, 4)
    assert res.num_selected == 4
    # with the error in the FIRST argument the row raises although the second is NULL
    with pytest.raises(abi.VgpuError):
        cols = [engine.flat_column(t, v, nn) for t, v, nn in spec]
        engine.filter_project(engine.Program(None, [ex.call("plus", ex.call("divide", ex.lit(10, I64), y), x)], n_columns=2), cols, 4)


def test_sum_of_case_when_aggregation(engine):
    """TPC-H Q12 / Q14 shape: sum(case when ... then x else 0 end) grouped by a small key."""
    rng = np.random.default_rng(5)
    n = 200_003
    k = rng.integers(0, 7, n).astype(np.int32)
    pri = rng.integers(0, 5, n).astype(np.int32)
    ep = np.round(rng.uniform(900, 100000, n), 2)
    disc = rng.integers(0, 11, n) / 100.0
    import importlib
    plan = importlib.import_module("velox-1_b200.aot_plans").case_when()   # the shape build() prebuilds a kernel for
    outs, keys, accs = plan["outputs"], plan["keys"], plan["accs"]
    cols = [engine.flat_column(I32, k), engine.flat_column(I32, pri), engine.flat_column(D, ep), engine.flat_column(D, disc)]
    prog = engine.Program(None, outs)
    agg = engine.Aggregation(abi.AGG_ARRAY, keys, accs)
    agg.add_input(prog, cols, n)
    got = H.decode_rows(agg.extract(), keys, accs, [o.type for o in outs], abi.AGG_ARRAY)
    want = H.run_oracle_agg(None, outs, [(k, None), (pri, None), (ep, None), (disc, None)], n, keys, accs)
    H.assert_groups_equal(got, want, accs)
