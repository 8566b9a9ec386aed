"""CPU tests that pin the oracles: the reference's own goldens / known-answer tests for this path
(SURVEY.md section 8c), and the C++ restatement (oracle/velox_cpu.cpp) against the numpy one."""
import numpy as np
import pytest

import cpu_oracle as cpu
import helpers as H
import lineitem as L
import np_oracle as orc
from helpers import abi, ex, D, I32, I64, B


def test_dbgen_goldens():
    # velox/connectors/tpch/tests/TpchConnectorTest.cpp:178-192; velox/tpch/gen/tests/TpchGenTest.cpp:311-351,394-398
    d = L.lineitem(0.01, cache=False)
    assert len(d["l_orderkey"]) == 60175
    first = L.gen_range(1.0, 0, 100)
    assert first["l_orderkey"][0] == 1 and first["l_shipdate"][0] == 9568       # 1996-03-13
    assert first["l_orderkey"][-1] == 388 and first["l_shipdate"][-1] == 8393   # 1992-12-24
    nxt = L.gen_range(1.0, 100, 100)
    assert nxt["l_orderkey"][0] == 389 and nxt["l_shipdate"][0] == 8868         # 1994-04-13
    assert nxt["l_orderkey"][-1] == 800 and nxt["l_shipdate"][-1] == 10430      # 1998-07-23
    assert (first["l_orderkey"] == 1).sum() == 6 and (first["l_orderkey"] == 2).sum() == 1
    # reproducible across offsets (TpchGenTest.cpp:365-410)
    again = L.gen_range(1.0, 50, 100)
    assert np.array_equal(again["l_extendedprice"][: 10], L.gen_range(1.0, 50, 10)["l_extendedprice"][:10])


def test_between_kat():
    # velox/functions/prestosql/tests/ComparisonsTest.cpp:57-68
    x = np.array([0, 1, 4, 5, 10, -1, -4, -5, -10], dtype=np.int32)
    e = ex.call("between", ex.col(0, I32), ex.lit(-1, I32), ex.lit(5, I32))
    v, nl, er = orc.eval_expr(e, [(x, None)], len(x))
    assert list(v) == [True, True, True, True, False, True, False, False, False]


def test_conjunct_truth_table():
    # velox/expression/tests/ConjunctTest.cpp:30-110 (results as DuckDB / SQL define them)
    T, F, N = (True, True), (False, True), (False, False)
    want_and = {(T, T): T, (T, F): F, (T, N): N, (F, T): F, (F, F): F, (F, N): F, (N, T): N, (N, F): F, (N, N): N}
    want_or = {(T, T): T, (T, F): T, (T, N): T, (F, T): T, (F, F): F, (F, N): N, (N, T): T, (N, F): N, (N, N): N}
    for fn, table in (("and", want_and), ("or", want_or)):
        for (a, b), w in table.items():
            cols = [(np.array([a[0]]), np.array([a[1]])), (np.array([b[0]]), np.array([b[1]]))]
            v, nl, _ = orc.eval_expr(ex.call(fn, ex.col(0, B), ex.col(1, B)), cols, 1)
            assert bool(nl[0]) == (not w[1]), (fn, a, b)
            if w[1]:
                assert bool(v[0]) == w[0], (fn, a, b)


def test_sum_overflow_kat():
    # velox/functions/prestosql/aggregates/tests/SumTest.cpp:610-620: integer sums throw, doubles do not
    big = np.array([2**62, 2**62], dtype=np.int64)
    with pytest.raises(orc.UserError):
        orc.aggregate([], [("sum_i64", (big, np.zeros(2, bool)))], np.ones(2, bool), global_agg=True)
    r = orc.aggregate([], [("sum_f64", (np.array([1.7e308, 1.7e308]), np.zeros(2, bool)))], np.ones(2, bool), global_agg=True)
    assert np.isinf(r[()][0][0])
    with pytest.raises(cpu.OracleError):
        cpu.run(None, [ex.col(0, I64)], [(I64, big, None)], 2, [], [("sum_i64", 0)])


def test_wave_group_sums_kat():
    # velox/experimental/wave/exec/tests/AggregationTest.cpp:58-93: keys i % 3, values i, i in [0, 10)
    i = np.arange(10, dtype=np.int64)
    keys = [(i % 3, np.zeros(10, bool))]
    r = orc.aggregate(keys, [("sum_i64", (i, np.zeros(10, bool))), ("count", None), ("sum_f64", (i, np.zeros(10, bool)))], np.ones(10, bool))
    assert {k[0]: v[0][0] for k, v in r.items()} == {0: 18, 1: 12, 2: 15}
    assert {k[0]: v[1][0] for k, v in r.items()} == {0: 4, 1: 3, 2: 3}
    c, _ = cpu.run(None, [ex.col(0, I64), ex.col(1, I64)], [(I64, i % 3, None), (I64, i, None)], 10, [0],
                   [("sum_i64", 1), ("count", -1), ("sum_f64", 1)])
    assert {k[0]: v[0][0] for k, v in c.items()} == {0: 18, 1: 12, 2: 15}


def test_empty_and_all_null_groups():
    # SumTest.cpp:509-538: empty grouped input -> no rows; global over empty input -> one NULL row
    e = np.zeros(0)
    assert orc.aggregate([(e, e.astype(bool))], [("sum_f64", (e, e.astype(bool)))], e.astype(bool)) == {}
    g = orc.aggregate([], [("sum_f64", (e, e.astype(bool))), ("count", None)], e.astype(bool), global_agg=True)
    assert g[()][0][1] == 0 and g[()][1][0] == 0
    c, _ = cpu.run(None, [ex.col(0, D)], [(D, e, None)], 0, [], [("sum_f64", 0), ("count", -1)])
    assert c[()][0][1] == 0 and c[()][1][0] == 0


@pytest.fixture(scope="module")
def sf01():
    return L.lineitem(0.1)


def np_names(accs):
    return [(H.KIND_NAME[k], i) for k, i in accs]


def test_cpp_oracle_matches_numpy_on_q6_q1(sf01):
    n = len(sf01["l_shipdate"])
    f, outs = H.q6_exprs()
    accs = [(abi.ACC_SUM_F64, 0), (abi.ACC_COUNT, -1)]
    want = H.run_oracle_agg(f, outs, [(sf01[c], None) for c in H.Q6_COLS], n, [], accs)
    for threads in (1, 4):
        got, _ = cpu.run(f, outs, [(t, sf01[c], None) for c, t in zip(H.Q6_COLS, [I32, D, D, D])], n, [], np_names(accs), threads=threads)
        H.assert_groups_equal(got, want, accs)
    f, outs = H.q1_exprs()
    want = H.run_oracle_agg(f, outs, [(sf01[c], None) for c in H.Q1_COLS], n, H.Q1_KEYS, H.Q1_ACCS)
    cols = [(t, sf01[c], None) for c, t in zip(H.Q1_COLS, [I32, I32, D, D, D, D, I32])]
    for threads in (1, 3):
        got, _ = cpu.run(f, outs, cols, n, [0, 1], np_names(H.Q1_ACCS), threads=threads)
        H.assert_groups_equal(got, want, H.Q1_ACCS)
    assert len(want) == 4


def test_cpp_oracle_matches_numpy_high_cardinality_and_nulls(sf01):
    n = 200_001
    ok, ep, mode = sf01["l_orderkey"][:n], sf01["l_extendedprice"][:n], sf01["l_shipmode"][:n]
    knull = (np.arange(n) % 11) == 0
    vnull = (np.arange(n) % 7) == 0
    outs = [ex.col(0, I64), ex.col(1, D), ex.col(2, I32)]
    accs = [(abi.ACC_SUM_F64, 1), (abi.ACC_MIN, 1), (abi.ACC_MAX, 1), (abi.ACC_COUNT, 1), (abi.ACC_COUNT, -1)]
    for key_out, kcol in ((0, (ok, knull)), (2, (mode, knull))):
        ocols = [(ok, ~knull if key_out == 0 else None), (ep, ~vnull), (mode, ~knull if key_out == 2 else None)]
        want = H.run_oracle_agg(None, outs, ocols, n, [(key_out, 0, 0, 0)], accs)
        ccols = [(I64, ok, knull if key_out == 0 else None), (D, ep, vnull), (I32, mode, knull if key_out == 2 else None)]
        for threads in (1, 4):
            got, _ = cpu.run(None, outs, ccols, n, [key_out], np_names(accs), threads=threads)
            H.assert_groups_equal(got, want, accs)
        got, _ = cpu.run(None, outs, ccols, n, [key_out], np_names(accs), ignore_null_keys=True)
        want2 = H.run_oracle_agg(None, outs, ocols, n, [(key_out, 0, 0, 0)], accs, ignore_null_keys=True)
        H.assert_groups_equal(got, want2, accs)
        assert (None,) in want and (None,) not in want2


def test_sf1_anchor_counts():
    # BASELINE.md section 2 anchors (computed with the reference's dbgen): Q6 row count and sum at SF1
    d = L.lineitem(1.0, wanted=H.Q6_COLS)
    n = len(d["l_shipdate"])
    assert n == 6_001_215  # velox/tpch/gen/TpchGen.cpp:33-34
    f, outs = H.q6_exprs()
    got, secs = cpu.run(f, outs, [(t, d[c], None) for c, t in zip(H.Q6_COLS, [I32, D, D, D])], n, [],
                        [("sum_f64", 0), ("count", -1)], threads=4)
    assert got[()][1][0] == 248_078
    assert abs(got[()][0][0] - 568264314.8435) < 1e-3
