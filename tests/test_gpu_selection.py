"""-m gpu: row selections (the batch's SelectivityVector as an index list, velox/vector/SelectivityVector.h) and sliced
columns whose base address is only element-aligned stay on the plan-specialised kernels (gather loads / per-row loads)
and give the oracle's result, for the ARRAY, GLOBAL and HASH sinks."""
import numpy as np
import pytest
import torch

import helpers as H
from helpers import abi, ex, D, I32, I64

pytestmark = pytest.mark.gpu


def fused():
    return abi.last_kernel_note().startswith("built from")


def q1_data(rng, n):
    return [rng.integers(0, 3, n).astype(np.int32), rng.integers(0, 2, n).astype(np.int32), rng.integers(1, 51, n) * 0.01,
            rng.integers(90000, 10500000, n) * 0.01, rng.integers(0, 11, n) * 0.01, rng.integers(0, 9, n) * 0.01,
            rng.integers(8036, 10592, n).astype(np.int32)]


@pytest.mark.parametrize("frac", [0.5, 0.03])
def test_q1_with_row_selection(engine, frac):
    rng = np.random.default_rng(5)
    n = 400_003
    data = q1_data(rng, n)
    rows = np.sort(rng.choice(n, int(n * frac), replace=False)).astype(np.int32)
    f, outs = H.q1_exprs()
    types = [I32, I32, D, D, D, D, I32]
    cols = [engine.flat_column(t, v) for t, v in zip(types, data)]
    prog = engine.Program(f, outs)
    agg = engine.Aggregation(abi.AGG_ARRAY, H.Q1_KEYS, H.Q1_ACCS)
    agg.add_input(prog, cols, n, in_rows=torch.from_numpy(rows).cuda())
    assert fused(), abi.last_kernel_note()
    got = H.decode_rows(agg.extract(), H.Q1_KEYS, H.Q1_ACCS, [o.type for o in outs], abi.AGG_ARRAY)
    want = H.run_oracle_agg(f, outs, [(v[rows], None) for v in data], len(rows), H.Q1_KEYS, H.Q1_ACCS)
    H.assert_groups_equal(got, want, H.Q1_ACCS)


@pytest.mark.parametrize("offset", [1, 3, 5])
def test_sliced_columns_keep_the_fused_kernel(engine, offset):
    # a slice starting `offset` rows into the buffers: 8-byte columns are 8-byte aligned, 4-byte columns 4-byte aligned
    rng = np.random.default_rng(7)
    n = 300_000
    data = q1_data(rng, n + offset)
    f, outs = H.q1_exprs()
    types = [I32, I32, D, D, D, D, I32]
    full = [engine.flat_column(t, v) for t, v in zip(types, data)]
    cols = [engine.DeviceColumn(c.type, abi.FLAT, c.values[offset:], length=n) for c in full]
    prog = engine.Program(f, outs)
    agg = engine.Aggregation(abi.AGG_ARRAY, H.Q1_KEYS, H.Q1_ACCS)
    agg.add_input(prog, cols, n)
    assert fused(), abi.last_kernel_note()
    got = H.decode_rows(agg.extract(), H.Q1_KEYS, H.Q1_ACCS, [o.type for o in outs], abi.AGG_ARRAY)
    want = H.run_oracle_agg(f, outs, [(v[offset:], None) for v in data], n, H.Q1_KEYS, H.Q1_ACCS)
    H.assert_groups_equal(got, want, H.Q1_ACCS)


def test_hash_groupby_sliced_and_selected(engine):
    rng = np.random.default_rng(9)
    n = 200_001
    k = np.repeat(np.arange(1, n // 3 + 2, dtype=np.int64) * 5, 3)[: n + 1]
    v = rng.normal(size=n + 1)
    vnn = rng.random(n + 1) > 0.1
    accs = [(abi.ACC_SUM_F64, 1), (abi.ACC_COUNT, -1), (abi.ACC_MIN, 1), (abi.ACC_MAX, 1)]
    keys = [(0, 64, 0, 0)]
    outs = [ex.col(0, I64), ex.col(1, D)]
    prog = engine.Program(None, outs)
    kc, vc = engine.flat_column(I64, k), engine.flat_column(D, v)
    # sliced by one row (values only; the null bitmap of a sliced vector is rebuilt by the host, so none here)
    cols = [engine.DeviceColumn(I64, abi.FLAT, kc.values[1:], length=n), engine.DeviceColumn(D, abi.FLAT, vc.values[1:], length=n)]
    agg = engine.Aggregation(abi.AGG_HASH, keys, accs)
    agg.add_input(prog, cols, n)
    assert fused(), abi.last_kernel_note()
    assert agg.info()["engine"] == 1
    got = H.decode_rows(agg.extract(), keys, accs, [I64, D], abi.AGG_HASH)
    H.assert_groups_equal(got, H.run_oracle_agg(None, outs, [(k[1:], None), (v[1:], None)], n, keys, accs), accs)
    # a row selection over nullable input
    rows = np.sort(rng.choice(n, n // 4, replace=False)).astype(np.int32)
    cols = [engine.flat_column(I64, k[:n]), engine.flat_column(D, v[:n], vnn[:n])]
    agg = engine.Aggregation(abi.AGG_HASH, keys, accs)
    agg.add_input(prog, cols, n, in_rows=torch.from_numpy(rows).cuda())
    assert fused(), abi.last_kernel_note()
    got = H.decode_rows(agg.extract(), keys, accs, [I64, D], abi.AGG_HASH)
    H.assert_groups_equal(got, H.run_oracle_agg(None, outs, [(k[:n][rows], None), (v[:n][rows], vnn[:n][rows])], len(rows), keys, accs), accs)


def test_q6_global_with_selection_and_dictionary(engine):
    rng = np.random.default_rng(11)
    n = 250_000
    sd = rng.integers(8036, 10592, n).astype(np.int32)
    ep = rng.integers(90000, 10500000, n) * 0.01
    qty = rng.integers(1, 51, n) * 0.01
    base = np.arange(11) * 0.01
    didx = rng.integers(0, 11, n).astype(np.int32)
    dnn = rng.random(n) > 0.05
    f, outs = H.q6_exprs()
    cols = [engine.flat_column(I32, sd), engine.flat_column(D, ep), engine.flat_column(D, qty), engine.dictionary_column(D, base, didx, dnn)]
    rows = np.sort(rng.choice(n, n // 2, replace=False)).astype(np.int32)
    accs = [(abi.ACC_SUM_F64, 0), (abi.ACC_COUNT, -1)]
    prog = engine.Program(f, outs)
    agg = engine.Aggregation(abi.AGG_GLOBAL, [], accs)
    agg.add_input(prog, cols, n, in_rows=torch.from_numpy(rows).cuda())
    assert fused(), abi.last_kernel_note()
    got = H.decode_rows(agg.extract(), [], accs, [D], abi.AGG_GLOBAL)
    want = H.run_oracle_agg(f, outs, [(sd[rows], None), (ep[rows], None), (qty[rows], None), (base[didx][rows], dnn[rows])], len(rows), [], accs)
    H.assert_groups_equal(got, want, accs)
