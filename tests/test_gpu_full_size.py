"""BASELINE.json's full sizes (lineitem SF10, 59,986,052 rows, synthetic columns generated in HBM) through
size-independent properties: batch-split invariance, count conservation, agreement with independent torch
reductions.  The oracle is too slow at this size; exact answers at small sizes are in the other GPU tests."""
import importlib

import numpy as np
import pytest
import torch

import helpers as H

pytestmark = pytest.mark.gpu
abi, ex = H.abi, H.ex
I32, I64, D = abi.INTEGER, abi.BIGINT, abi.DOUBLE
SF = 10.0


@pytest.fixture(scope="module")
def sf10(engine):
    tpch = importlib.import_module("velox-1_b200.tpch_synth")
    names = ["l_orderkey", "l_returnflag", "l_linestatus", "l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_shipdate",
             "l_shipmode", "l_shipinstruct"]
    return tpch.lineitem(SF, names), tpch.row_count(SF)


def q1_groups(engine, data, lo, hi, agg=None):
    f, outs = H.q1_exprs()
    types = [I32, I32, D, D, D, D, I32]
    cols = [engine.DeviceColumn(t, abi.FLAT, data[c][lo:hi], length=hi - lo) for c, t in zip(H.Q1_COLS, types)]
    agg = agg or engine.Aggregation(abi.AGG_ARRAY, H.Q1_KEYS, H.Q1_ACCS)
    agg.add_input(engine.Program(f, outs), cols, hi - lo)
    return agg, outs


def test_q1_sf10_split_invariance_and_counts(engine, sf10):
    data, n = sf10
    whole, outs = q1_groups(engine, data, 0, n)
    cut = 23_456_789  # not a multiple of any tile size
    parts, _ = q1_groups(engine, data, 0, cut)
    q1_groups(engine, data, cut, n, parts)
    a = H.decode_rows(whole.extract(), H.Q1_KEYS, H.Q1_ACCS, [o.type for o in outs], abi.AGG_ARRAY)
    b = H.decode_rows(parts.extract(), H.Q1_KEYS, H.Q1_ACCS, [o.type for o in outs], abi.AGG_ARRAY)
    H.assert_groups_equal(a, b, H.Q1_ACCS)                      # counts exact, FP sums 1e-9
    passing = data["l_shipdate"] <= H.DATE_1998_09_02
    assert sum(v[-1][0] for v in a.values()) == int(passing.sum())
    for (rf, ls), v in a.items():                               # every group against independent torch reductions
        m = passing & (data["l_returnflag"] == rf) & (data["l_linestatus"] == ls)
        assert v[-1][0] == int(m.sum())
        want = float(data["l_extendedprice"][m].sum())
        assert abs(v[1][0] - want) <= 1e-9 * want


def test_q6_sf10_selection_count_and_sum(engine, sf10):
    data, n = sf10
    f, outs = H.q6_exprs()
    cols = [engine.DeviceColumn(t, abi.FLAT, data[c], length=n) for c, t in zip(H.Q6_COLS, [I32, D, D, D])]
    res = engine.filter_project(engine.Program(f, outs), cols, n)
    m = (data["l_shipdate"] >= H.DATE_1994_01_01) & (data["l_shipdate"] < H.DATE_1995_01_01) & (data["l_discount"] >= 0.05) & \
        (data["l_discount"] <= 0.07) & (data["l_quantity"] < 24.0)
    assert res.num_selected == int(m.sum())
    idx = res.sel_idx[: res.num_selected].long()
    assert torch.equal(idx, torch.nonzero(m).reshape(-1))       # ascending row ids, bit-exact selection
    assert torch.equal(res.values[0][idx], (data["l_extendedprice"] * data["l_discount"])[idx])  # IEEE multiply: bit-exact


def test_orderkey_sf10_conservation(engine, sf10):
    data, n = sf10
    groups = int(torch.unique_consecutive(data["l_orderkey"]).numel())
    accs = [(abi.ACC_SUM_F64, 1), (abi.ACC_COUNT, -1)]
    agg = engine.Aggregation(abi.AGG_HASH, [(0, 64, 0, 0)], accs, expected_groups=groups)
    cols = [engine.DeviceColumn(I64, abi.FLAT, data["l_orderkey"], length=n), engine.DeviceColumn(D, abi.FLAT, data["l_extendedprice"], length=n)]
    agg.add_input(engine.Program(None, [ex.col(0, I64), ex.col(1, D)]), cols, n)
    assert agg.num_groups() == groups
    rows = agg.extract()
    assert int(rows[:, 4].sum()) == n                           # every row counted once
    assert len(np.unique(rows[:, 0])) == groups                 # no group twice
    total = float(data["l_extendedprice"].sum())
    assert abs(float(rows[:, 3].view(np.float64).sum()) - total) <= 1e-9 * total
    # spot-check 1000 groups against a direct reduction
    pick = rows[:: max(1, len(rows) // 1000)][:1000]
    keys = torch.from_numpy(pick[:, 0].copy()).cuda()
    lo = torch.searchsorted(data["l_orderkey"], keys)
    hi = torch.searchsorted(data["l_orderkey"], keys, right=True)
    assert torch.equal((hi - lo).cpu(), torch.from_numpy(pick[:, 4].copy()))
    csum = torch.cumsum(data["l_extendedprice"], 0)
    for i in range(0, 1000, 97):
        want = float(data["l_extendedprice"][int(lo[i]):int(hi[i])].sum())
        assert abs(float(pick[i, 3:4].view(np.float64)[0]) - want) <= 1e-9 * max(want, 1.0)
    del csum
