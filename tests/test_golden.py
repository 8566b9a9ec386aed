"""Committed golden vectors (tests/golden/lineitem_golden.json, made by tests/golden/make_golden.py from the
reference's own dbgen): exact integer answers for lineitem SF0.01.  CPU tests pin the oracle to them; the
GPU tests pin the CUDA path to them without going through the oracle."""
import json
import os

import numpy as np
import pytest

import helpers as H
import lineitem
import np_oracle as orc

HERE = os.path.dirname(os.path.abspath(__file__))
abi, ex = H.abi, H.ex
FLAGS, STATUS = lineitem.DICTS["l_returnflag"], lineitem.DICTS["l_linestatus"]


@pytest.fixture(scope="module")
def golden():
    with open(os.path.join(HERE, "golden", "lineitem_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="module")
def sf001():
    if not os.path.exists(lineitem.REF_SO):
        pytest.skip("oracle/_ref/libdbgen_ref.so not built")
    return lineitem.lineitem(0.01, cache=False)


def test_reference_goldens_in_fixture(golden):
    """The reference's own expectations (TpchGenTest.cpp:311-351,394-398; TpchConnectorTest.cpp:178-192)."""
    f = golden["first_200_orders_sf1"]
    keys = np.array(f["l_orderkey"])
    assert golden["sf0.01"]["rows"] == 60175
    assert keys[0] == 1 and f["l_shipdate"][0] == 9568            # 1996-03-13
    assert (keys == 1).sum() == 6 and (keys == 2).sum() == 1
    first100 = keys <= 388
    assert keys[first100][-1] == 388 and np.array(f["l_shipdate"])[first100][-1] == 8393   # 1992-12-24
    assert keys[~first100][0] == 389 and np.array(f["l_shipdate"])[~first100][0] == 8868   # 1994-04-13
    assert keys[-1] == 800 and f["l_shipdate"][-1] == 10430                                 # 1998-07-23


def test_dbgen_reproduces_fixture(golden):
    if not os.path.exists(lineitem.REF_SO):
        pytest.skip("oracle/_ref/libdbgen_ref.so not built")
    live = lineitem.gen_range(1.0, 0, 200)
    f = golden["first_200_orders_sf1"]
    for name, v in live.items():
        got = v if v.dtype.kind in "iu" else np.rint(v * 100.0).astype(np.int64)
        assert np.array_equal(got, np.array(f[name])), name


def q1_check(got, golden):
    g = golden["sf0.01"]["q1"]
    assert len(got) == len(g)
    for (rf, ls), vals in got.items():
        w = g[f"{FLAGS[rf]},{STATUS[ls]}"]
        assert vals[10][0] == w["count"]                                   # count(*): exact
        assert vals[5][0] == w["count"]                                    # avg's count
        for idx, key, scale in ((0, "sum_qty_raw", 1e-2), (1, "sum_eprice_cents", 1e-2), (2, "sum_disc_price_cents_x100", 1e-4),
                                (3, "sum_charge_cents_x10000", 1e-6), (8, "sum_disc_pct", 1e-2)):
            want = w[key] * scale                                          # Velox units (quantity is raw*0.01: reference quirk)
            assert abs(vals[idx][0] - want) <= 1e-9 * abs(want), (rf, ls, key)


def test_oracle_q1_q6_match_golden(golden, sf001):
    n = len(sf001["l_orderkey"])
    f, outs = H.q1_exprs()
    got = H.run_oracle_agg(f, outs, [(sf001[c], None) for c in H.Q1_COLS], n, H.Q1_KEYS, H.Q1_ACCS)
    q1_check(got, golden)
    f6, o6 = H.q6_exprs()
    sel, vals = orc.filter_project(f6, o6, [(sf001[c], None) for c in H.Q6_COLS], n)
    assert int(sel.sum()) == golden["sf0.01"]["q6"]["rows"]
    want = golden["sf0.01"]["q6"]["sum_eprice_disc_cents_pct"] * 1e-4
    assert abs(float(vals[0][0][sel].sum()) - want) <= 1e-9 * want


@pytest.mark.gpu
def test_gpu_q1_q6_match_golden(engine, golden, sf001):
    n = len(sf001["l_orderkey"])
    f, outs = H.q1_exprs()
    types = [abi.INTEGER, abi.INTEGER, abi.DOUBLE, abi.DOUBLE, abi.DOUBLE, abi.DOUBLE, abi.INTEGER]
    cols = [engine.flat_column(t, sf001[c]) for c, t in zip(H.Q1_COLS, types)]
    agg = engine.Aggregation(abi.AGG_ARRAY, H.Q1_KEYS, H.Q1_ACCS)
    agg.add_input(engine.Program(f, outs), cols, n)
    q1_check(H.decode_rows(agg.extract(), H.Q1_KEYS, H.Q1_ACCS, [o.type for o in outs], abi.AGG_ARRAY), golden)
    f6, o6 = H.q6_exprs()
    cols6 = [engine.flat_column(t, sf001[c]) for c, t in zip(H.Q6_COLS, [abi.INTEGER, abi.DOUBLE, abi.DOUBLE, abi.DOUBLE])]
    g = engine.Aggregation(abi.AGG_GLOBAL, [], [(abi.ACC_SUM_F64, 0), (abi.ACC_COUNT, -1)])
    g.add_input(engine.Program(f6, o6), cols6, n)
    (vals,) = H.decode_rows(g.extract(), [], [(abi.ACC_SUM_F64, 0), (abi.ACC_COUNT, -1)], [abi.DOUBLE], abi.AGG_GLOBAL).values()
    assert vals[1][0] == golden["sf0.01"]["q6"]["rows"]
    want = golden["sf0.01"]["q6"]["sum_eprice_disc_cents_pct"] * 1e-4
    assert abs(vals[0][0] - want) <= 1e-9 * want


@pytest.mark.gpu
def test_gpu_groupbys_match_golden(engine, golden, sf001):
    n = len(sf001["l_orderkey"])
    I32, I64, D = abi.INTEGER, abi.BIGINT, abi.DOUBLE
    # GROUP BY l_orderkey: group count exact, per-group sums folded into an exact checksum of cents
    accs = [(abi.ACC_SUM_F64, 1), (abi.ACC_COUNT, -1)]
    agg = engine.Aggregation(abi.AGG_HASH, [(0, 64, 0, 0)], accs)
    agg.add_input(engine.Program(None, [ex.col(0, I64), ex.col(1, D)]),
                  [engine.flat_column(I64, sf001["l_orderkey"]), engine.flat_column(D, sf001["l_extendedprice"])], n)
    rows = agg.extract()
    w = golden["sf0.01"]["orderkey_groupby"]
    assert len(rows) == w["groups"]
    cents = np.rint(rows[:, 3].view(np.float64) * 100.0).astype(np.int64)
    assert int(cents.sum()) == w["sum_cents"]
    assert int((rows[:, 0].astype(object) * cents).sum() % (1 << 61)) == w["checksum_key_times_cents"]
    assert int(rows[:, 4].sum()) == n
    # GROUP BY l_shipmode, l_shipinstruct: min / max / count, exact
    keys = [(0, 0, 0, 6), (1, 0, 0, 3)]
    accs = [(abi.ACC_MIN, 2), (abi.ACC_MAX, 2), (abi.ACC_MIN, 3), (abi.ACC_MAX, 3), (abi.ACC_COUNT, -1)]
    outs = [ex.col(0, I32), ex.col(1, I32), ex.col(2, D), ex.col(3, I32)]
    cols = [engine.flat_column(I32, sf001["l_shipmode"]), engine.flat_column(I32, sf001["l_shipinstruct"]),
            engine.flat_column(D, sf001["l_extendedprice"]), engine.flat_column(I32, sf001["l_shipdate"])]
    agg = engine.Aggregation(abi.AGG_ARRAY, keys, accs)
    agg.add_input(engine.Program(None, outs), cols, n)
    got = H.decode_rows(agg.extract(), keys, accs, [o.type for o in outs], abi.AGG_ARRAY)
    want = golden["sf0.01"]["shipmode_shipinstruct"]
    assert len(got) == len(want) == 28
    for (mo, ins), v in got.items():
        g = want[f"{lineitem.DICTS['l_shipmode'][mo]}|{lineitem.DICTS['l_shipinstruct'][ins]}"]
        assert round(v[0][0] * 100) == g["min_eprice_cents"] and round(v[1][0] * 100) == g["max_eprice_cents"]
        assert v[2][0] == g["min_shipdate"] and v[3][0] == g["max_shipdate"] and v[4][0] == g["count"]
