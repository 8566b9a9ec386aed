"""Pins the oracle to the reference's OWN test oracle: the vendored DuckDB (velox/external/duckdb) that
velox/exec/tests/utils/QueryAssertions.cpp (DuckDbQueryRunner, assertQuery) checks every operator answer against.
oracle/_ref/libduckdb_ref.so is compiled by oracle/Makefile from the sources under /root/reference; the same SQL the
reference's tests run (TpchQueryBuilder.cpp's Q1 / Q6 plans are asserted against the TPC-H SQL text) is run here and
compared with np_oracle + the C++ oracle on the same rows.  CPU only."""
import numpy as np
import pytest

import duck_ref
import helpers as H
import lineitem
import np_oracle as orc

pytestmark = pytest.mark.skipif(not duck_ref.available(), reason="oracle/_ref/libduckdb_ref.so not built")

RTOL = 1e-12  # double sums over <= 30k rows per group: summation order differs, values agree far below this


@pytest.fixture(scope="module")
def table():
    return lineitem.lineitem(0.01)


@pytest.fixture(scope="module")
def duck(table):
    t = table
    s = duck_ref.Session(threads=2)
    D = duck_ref
    s.create_table("lineitem", [
        ("l_orderkey", D.BIGINT, t["l_orderkey"]), ("l_partkey", D.BIGINT, t["l_partkey"]),
        ("l_quantity", D.DOUBLE, t["l_quantity"]), ("l_extendedprice", D.DOUBLE, t["l_extendedprice"]),
        ("l_discount", D.DOUBLE, t["l_discount"]), ("l_tax", D.DOUBLE, t["l_tax"]),
        ("l_returnflag", D.VARCHAR_CODES, t["l_returnflag"], lineitem.DICTS["l_returnflag"]),
        ("l_linestatus", D.VARCHAR_CODES, t["l_linestatus"], lineitem.DICTS["l_linestatus"]),
        ("l_shipdate", D.DATE, t["l_shipdate"]),
        ("l_shipinstruct", D.VARCHAR_CODES, t["l_shipinstruct"], lineitem.DICTS["l_shipinstruct"]),
        ("l_shipmode", D.VARCHAR_CODES, t["l_shipmode"], lineitem.DICTS["l_shipmode"])])
    yield s
    s.close()


def input_cols(t, names):
    """Operator input columns: (values, not-null mask or None)."""
    return [(np.asarray(t[c]), None) for c in names]


def oracle_cols(t, names):
    """Evaluated columns as orc.aggregate takes them: (values, NULL mask)."""
    n = len(t[names[0]])
    return [(np.asarray(t[c]), np.zeros(n, dtype=bool)) for c in names]


def test_q1_sql_matches_oracle(duck, table):
    """TPC-H Q1 text as the reference's TpchQueryBuilder test asserts it (velox/dwio/parquet/tests/ParquetTpchTest.cpp
    assertQuery(tpchPlan, duckDbSql); plan shape TpchQueryBuilder.cpp:189-244)."""
    vals, valid, strs = duck.query(
        "select l_returnflag, l_linestatus, sum(l_quantity), sum(l_extendedprice), "
        "sum(l_extendedprice * (1 - l_discount)), sum(l_extendedprice * (1 - l_discount) * (1 + l_tax)), "
        "avg(l_quantity), avg(l_extendedprice), avg(l_discount), count(*) "
        "from lineitem where l_shipdate <= date '1998-09-02' group by l_returnflag, l_linestatus "
        "order by l_returnflag, l_linestatus", strings=(0, 1))
    f, outs = H.q1_exprs()
    n = len(table["l_orderkey"])
    res = H.run_oracle_agg(f, outs, input_cols(table, H.Q1_COLS), n, H.Q1_KEYS, H.Q1_ACCS)
    assert len(res) == len(vals) == 4
    assert valid[:, 2:].all()
    for r in range(len(vals)):
        key = (lineitem.DICTS["l_returnflag"].index(strs[0][r]), lineitem.DICTS["l_linestatus"].index(strs[1][r]))
        a = res[key]
        sums = [a[0][0], a[1][0], a[2][0], a[3][0]]
        avgs = [a[4][0] / a[5][0], a[6][0] / a[7][0], a[8][0] / a[9][0]]
        np.testing.assert_allclose(vals[r, 2:6], sums, rtol=RTOL)
        np.testing.assert_allclose(vals[r, 6:9], avgs, rtol=RTOL)
        assert int(vals[r, 9]) == a[10][0]


def test_q6_sql_matches_oracle(duck, table):
    """TPC-H Q6 (TpchQueryBuilder.cpp:461-486)."""
    vals, valid, _ = duck.query(
        "select sum(l_extendedprice * l_discount), count(*) from lineitem where l_shipdate >= date '1994-01-01' "
        "and l_shipdate < date '1995-01-01' and l_discount between 0.05 and 0.07 and l_quantity < 24.0")
    f, outs = H.q6_exprs()
    n = len(table["l_orderkey"])
    res = H.run_oracle_agg(f, outs, input_cols(table, H.Q6_COLS), n, [], [(H.abi.ACC_SUM_F64, 0), (H.abi.ACC_COUNT, -1)])
    (s, _), (c, _) = res[()]
    assert int(vals[0, 1]) == c
    np.testing.assert_allclose(vals[0, 0], s, rtol=RTOL)


def test_orderkey_groupby_matches_oracle(duck, table):
    """High-cardinality BIGINT key (BASELINE config C3 shape): every group, exact count and sum within RTOL."""
    vals, valid, _ = duck.query("select l_orderkey, sum(l_extendedprice), count(*), min(l_extendedprice), "
                                "max(l_partkey) from lineitem group by l_orderkey order by l_orderkey")
    n = len(table["l_orderkey"])
    ok = oracle_cols(table, ["l_orderkey", "l_extendedprice", "l_partkey"])
    sel = np.ones(n, dtype=bool)
    res = orc.aggregate([ok[0]], [("sum_f64", ok[1]), ("count", None), ("min", ok[1]), ("max", ok[2])], sel)
    assert len(res) == len(vals)
    keys = sorted(res)
    want = np.array([[k[0], res[k][0][0], res[k][1][0], res[k][2][0], res[k][3][0]] for k in keys], dtype=np.float64)
    assert np.array_equal(vals[:, 0], want[:, 0])
    assert np.array_equal(vals[:, 2], want[:, 2])
    np.testing.assert_allclose(vals[:, 1], want[:, 1], rtol=RTOL)
    assert np.array_equal(vals[:, 3], want[:, 3])  # min/max are selections, bit-exact
    assert np.array_equal(vals[:, 4], want[:, 4])


def test_two_varchar_keys_min_max(duck, table):
    """Two dictionary-encoded VARCHAR keys, min/max of DOUBLE and DATE (config C4 shape)."""
    vals, valid, strs = duck.query("select l_shipmode, l_shipinstruct, count(*), min(l_extendedprice), "
                                   "max(l_extendedprice), min(l_shipdate), max(l_shipdate) from lineitem "
                                   "where l_quantity < 0.3 group by 1, 2 order by 1, 2", strings=(0, 1))
    n = len(table["l_orderkey"])
    c = oracle_cols(table, ["l_shipmode", "l_shipinstruct", "l_extendedprice", "l_shipdate"])
    sel = table["l_quantity"] < 0.3
    res = orc.aggregate([c[0], c[1]], [("count", None), ("min", c[2]), ("max", c[2]), ("min", c[3]), ("max", c[3])], sel)
    assert len(res) == len(vals) == 28
    for r in range(len(vals)):
        key = (lineitem.DICTS["l_shipmode"].index(strs[0][r]), lineitem.DICTS["l_shipinstruct"].index(strs[1][r]))
        assert [a[0] for a in res[key]] == [int(vals[r, 2]), vals[r, 3], vals[r, 4], int(vals[r, 5]), int(vals[r, 6])]


@pytest.fixture(scope="module")
def nulls_session():
    """A small table with NULLs in keys and values (the shape of AggregationTest.cpp's makeVectors + nullEvery data)."""
    rng = np.random.default_rng(7)
    n = 5000
    k = rng.integers(0, 13, n).astype(np.int32)
    k_nn = rng.random(n) > 0.1
    a = rng.integers(-1000, 1000, n).astype(np.int64)
    a_nn = rng.random(n) > 0.2
    b = np.round(rng.normal(0, 100, n), 3)
    b_nn = rng.random(n) > 0.3
    c = rng.integers(0, 5, n).astype(np.int32)
    c_nn = rng.random(n) > 0.15
    # group 12 only ever sees NULL values: its sum/min/max/avg must be NULL, count 0
    a_nn[k == 12] = False
    b_nn[k == 12] = False
    s = duck_ref.Session(threads=1)
    D = duck_ref
    s.create_table("t", [("k", D.INTEGER, k, None, k_nn), ("a", D.BIGINT, a, None, a_nn),
                         ("b", D.DOUBLE, b, None, b_nn), ("c", D.INTEGER, c, None, c_nn)])
    data = dict(k=(k, ~k_nn), a=(a, ~a_nn), b=(b, ~b_nn), c=(c, ~c_nn), n=n)
    yield s, data
    s.close()


def test_null_keys_and_null_inputs(nulls_session):
    """NULL is a group of its own; aggregates skip NULL inputs; all-NULL input gives NULL (count 0)
    — AggregationTest.cpp (singleBigintKey / aggregateOfNulls) asserts the same SQL against DuckDB."""
    s, d = nulls_session
    vals, valid, _ = s.query("select k, sum(a), count(a), min(b), max(b), sum(b), count(*), avg(a) from t "
                             "group by k order by k nulls first")
    sel = np.ones(d["n"], dtype=bool)
    res = orc.aggregate([d["k"]], [("sum_i64", d["a"]), ("count", d["a"]), ("min", d["b"]), ("max", d["b"]),
                                   ("sum_f64", d["b"]), ("count", None)], sel)
    assert len(res) == len(vals) == 14
    assert not valid[0, 0]  # the NULL group sorts first
    for r in range(len(vals)):
        key = (None,) if not valid[r, 0] else (int(vals[r, 0]),)
        acc = res[key]
        for col, (v, cnt), exact in [(1, acc[0], True), (3, acc[2], True), (4, acc[3], True), (5, acc[4], False)]:
            assert bool(valid[r, col]) == (cnt > 0), (key, col)
            if cnt:
                if exact:
                    assert vals[r, col] == v
                else:
                    np.testing.assert_allclose(vals[r, col], v, rtol=1e-11, atol=1e-9)
        assert int(vals[r, 2]) == acc[1][0]
        assert int(vals[r, 6]) == acc[5][0]
        if acc[0][1]:
            np.testing.assert_allclose(vals[r, 7], acc[0][0] / acc[0][1], rtol=1e-13)
        else:
            assert not valid[r, 7]
    key12 = [r for r in range(len(vals)) if valid[r, 0] and int(vals[r, 0]) == 12][0]
    assert not valid[key12, 1] and not valid[key12, 3] and int(vals[key12, 2]) == 0


def test_three_valued_filter_and_null_propagation(nulls_session):
    """Kleene AND / OR with NULL operands, NOT, BETWEEN and arithmetic NULL propagation
    (velox/expression/ConjunctExpr.cpp; ExprTest.cpp asserts these against DuckDB the same way)."""
    s, d = nulls_session
    vals, valid, _ = s.query("select count(*), sum(a + c), count(a + c), sum(b * 2.0) from t "
                             "where (a > 100 or c < 2) and not (k = 3) and b between -150.0 and 150.0")
    ex = H.ex
    k, a, b, c = ex.col(0, H.I32), ex.col(1, H.I64), ex.col(2, H.D), ex.col(3, H.I32)
    f = ex.call("and",
                ex.call("or", ex.call("gt", a, ex.lit(100, H.I64)), ex.call("lt", c, ex.lit(2, H.I32))),
                ex.call("not", ex.call("eq", k, ex.lit(3, H.I32))),
                ex.call("between", b, ex.lit(-150.0, H.D), ex.lit(150.0, H.D)))
    outs = [ex.call("plus", a, ex.cast(c, H.I64)),
            ex.call("multiply", b, ex.lit(2.0, H.D))]
    cols = [(d[x][0], ~d[x][1]) for x in "kabc"]
    sel, o = orc.filter_project(f, outs, cols, d["n"])
    res = orc.aggregate([], [("count", None), ("sum_i64", o[0]), ("count", o[0]), ("sum_f64", o[1])], sel, global_agg=True)[()]
    assert int(vals[0, 0]) == res[0][0]
    assert int(vals[0, 1]) == res[1][0]
    assert int(vals[0, 2]) == res[2][0]
    np.testing.assert_allclose(vals[0, 3], res[3][0], rtol=1e-12)
    assert 0 < res[0][0] < d["n"]


def test_ignore_null_keys_equals_is_not_null_filter(nulls_session):
    """ignoreNullKeys (partial aggregation below a join, OperatorUtils.cpp:141-157) == WHERE k IS NOT NULL."""
    s, d = nulls_session
    vals, valid, _ = s.query("select k, sum(a), count(*) from t where k is not null group by k order by k")
    sel = np.ones(d["n"], dtype=bool)
    res = orc.aggregate([d["k"]], [("sum_i64", d["a"]), ("count", None)], sel, ignore_null_keys=True)
    assert len(res) == len(vals) == 13
    for r in range(len(vals)):
        acc = res[(int(vals[r, 0]),)]
        if acc[0][1]:
            assert int(vals[r, 1]) == acc[0][0]
        assert int(vals[r, 2]) == acc[1][0]
