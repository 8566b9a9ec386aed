"""-m gpu tests of the HASH aggregation engines (csrc/hashagg.cu log engine, csrc/agg.cu table engine) against the
numpy oracle: sorted and unsorted keys, runs that cross block / batch edges, NULL keys and the key that equals the
table's empty marker, filters, nullable accumulator inputs, engine migration, and the accumulator edge KATs of
velox/functions/prestosql/aggregates/tests (NaN in min / max, +-0.0, +-inf, sum(bigint) at the overflow boundary).
Find-or-insert by key equality with NULL as a key value: velox/exec/GroupingSet.cpp:240-307, HashTable.cpp:326-342."""
import os

import numpy as np
import pytest

import helpers as H
from helpers import abi, ex, D, I32, I64

pytestmark = pytest.mark.gpu

GB_ACCS = [(abi.ACC_SUM_F64, 1), (abi.ACC_COUNT, -1), (abi.ACC_MIN, 1), (abi.ACC_MAX, 1)]
RAW_KEY = [(0, 64, 0, 0)]


def run_hash(engine, filt, outs, spec, n, keys, accs, batches=1, ignore=False, hint=0):
    """HASH aggregation over `batches` equal slices; returns (decoded groups, engine info)."""
    prog = engine.Program(filt, outs, n_columns=len(spec))
    agg = engine.Aggregation(abi.AGG_HASH, keys, accs, ignore_null_keys=ignore, expected_groups=hint)
    step = ((n + batches - 1) // batches + 63) // 64 * 64 if batches > 1 else n
    for lo in range(0, max(n, 1), max(step, 1)):
        hi = min(n, lo + step)
        cols = [engine.flat_column(t, np.asarray(v)[lo:hi], None if nn is None else np.asarray(nn)[lo:hi]) for t, v, nn in spec]
        agg.add_input(prog, cols, hi - lo)
    agg.finalize()
    info = agg.info()
    if os.environ.get("VGPU_HASH_ENGINE", "log") != "table" and os.environ.get("VGPU_FUSED", "1") != "0":
        assert info["engine"] == 1, ("the log engine was expected", info, abi.last_kernel_note())
    got = H.decode_rows(agg.extract(), keys, accs, [o.type for o in outs], abi.AGG_HASH, nullable=not ignore)
    return got, info


def oracle(filt, outs, spec, n, keys, accs, ignore=False):
    return H.run_oracle_agg(filt, outs, [(v, nn) for _, v, nn in spec], n, keys, accs, ignore_null_keys=ignore)


def sorted_runs(rng, n_groups, max_run=7, start=1, stride=3):
    keys = start + stride * np.arange(n_groups, dtype=np.int64)
    return np.repeat(keys, rng.integers(1, max_run + 1, n_groups))


@pytest.mark.parametrize("batches", [1, 2, 5])
def test_sorted_runs_take_the_sorted_path(engine, batches):
    rng = np.random.default_rng(7)
    k = sorted_runs(rng, 40_000)
    n = len(k)
    v = rng.integers(1, 10_000, n) * 0.01
    spec = [(I64, k, None), (D, v, None)]
    outs = [ex.col(0, I64), ex.col(1, D)]
    got, info = run_hash(engine, None, outs, spec, n, RAW_KEY, GB_ACCS, batches=batches)
    H.assert_groups_equal(got, oracle(None, outs, spec, n, RAW_KEY, GB_ACCS), GB_ACCS)
    if info["engine"] == 1:
        # sorted input: only runs touching a 64-row half-block edge go through the deduplication
        assert info["sorted"] == 1 and info["front_entries"] > 0
        assert info["front_entries"] + info["deduped_groups"] == 40_000
        assert info["deduped_entries"] < n // 8


def test_one_long_run_spanning_blocks_and_batches(engine):
    # a single key over many blocks and two batches, then distinct keys: every entry of the long run is an edge entry
    k = np.concatenate([np.full(5000, 42, dtype=np.int64), 1000 + np.arange(3000, dtype=np.int64)])
    v = np.ones(len(k))
    spec = [(I64, k, None), (D, v, None)]
    outs = [ex.col(0, I64), ex.col(1, D)]
    for batches in (1, 2, 3):
        got, info = run_hash(engine, None, outs, spec, len(k), RAW_KEY, GB_ACCS, batches=batches)
        assert got[(42,)][0] == (5000.0, 5000) and got[(42,)][1][0] == 5000
        assert len(got) == 3001


@pytest.mark.parametrize("batches", [1, 3])
def test_unsorted_keys_with_duplicates(engine, batches):
    rng = np.random.default_rng(11)
    n = 300_011
    k = rng.integers(-50_000, 50_000, n).astype(np.int64)
    v = rng.normal(size=n)
    spec = [(I64, k, None), (D, v, None)]
    outs = [ex.col(0, I64), ex.col(1, D)]
    got, info = run_hash(engine, None, outs, spec, n, RAW_KEY, GB_ACCS, batches=batches)
    H.assert_groups_equal(got, oracle(None, outs, spec, n, RAW_KEY, GB_ACCS), GB_ACCS)
    if info["engine"] == 1:
        assert info["sorted"] == 0 and info["deduped_groups"] == len(got) - 1  # key -1 lives in its dedicated row


def test_sorted_then_unsorted_batch(engine):
    # the first batch is sorted, the second repeats its keys: the sorted flag must fall and every entry be deduplicated
    rng = np.random.default_rng(13)
    a = sorted_runs(rng, 9000)
    k = np.concatenate([a, a[::-1]])
    n = len(k)
    v = rng.integers(0, 100, n).astype(np.float64)
    spec = [(I64, k, None), (D, v, None)]
    outs = [ex.col(0, I64), ex.col(1, D)]
    prog = engine.Program(None, outs)
    agg = engine.Aggregation(abi.AGG_HASH, RAW_KEY, GB_ACCS)
    half = len(a)
    agg.add_input(prog, [engine.flat_column(I64, k[:half]), engine.flat_column(D, v[:half])], half)
    agg.add_input(prog, [engine.flat_column(I64, k[half:]), engine.flat_column(D, v[half:])], n - half)
    got = H.decode_rows(agg.extract(), RAW_KEY, GB_ACCS, [o.type for o in outs], abi.AGG_HASH)
    H.assert_groups_equal(got, oracle(None, outs, spec, n, RAW_KEY, GB_ACCS), GB_ACCS)
    assert len(got) == 9000


def test_equal_keys_across_batches_sorted(engine):
    # batch 2 starts with the key batch 1 ended with (the run is split by the batch edge) and the stream stays sorted
    k1 = np.repeat(np.arange(1, 2001, dtype=np.int64), 3)
    k2 = np.repeat(np.arange(2000, 4001, dtype=np.int64), 2)
    k = np.concatenate([k1, k2])
    v = np.arange(len(k), dtype=np.float64)
    spec = [(I64, k, None), (D, v, None)]
    outs = [ex.col(0, I64), ex.col(1, D)]
    prog = engine.Program(None, outs)
    agg = engine.Aggregation(abi.AGG_HASH, RAW_KEY, GB_ACCS)
    agg.add_input(prog, [engine.flat_column(I64, k1), engine.flat_column(D, v[:len(k1)])], len(k1))
    agg.add_input(prog, [engine.flat_column(I64, k2), engine.flat_column(D, v[len(k1):])], len(k2))
    agg.finalize()
    info = agg.info()
    got = H.decode_rows(agg.extract(), RAW_KEY, GB_ACCS, [o.type for o in outs], abi.AGG_HASH)
    H.assert_groups_equal(got, oracle(None, outs, spec, len(k), RAW_KEY, GB_ACCS), GB_ACCS)
    assert got[(2000,)][1][0] == 5
    if info["engine"] == 1:
        assert info["sorted"] == 1


def test_null_keys_and_empty_marker_key(engine):
    # NULL is a key value of its own; -1 (all ones) is the hash tables' empty marker and must still be a normal group
    rng = np.random.default_rng(17)
    n = 50_003
    k = rng.integers(-3, 40, n).astype(np.int64)
    knn = rng.random(n) > 0.05
    v = rng.normal(size=n)
    vnn = rng.random(n) > 0.3
    spec = [(I64, k, knn), (D, v, vnn)]
    outs = [ex.col(0, I64), ex.col(1, D)]
    for ignore in (False, True):
        for batches in (1, 2):
            got, _ = run_hash(engine, None, outs, spec, n, RAW_KEY, GB_ACCS, batches=batches, ignore=ignore)
            want = oracle(None, outs, spec, n, RAW_KEY, GB_ACCS, ignore=ignore)
            assert ((-1,) in got) and (((None,) in got) == (not ignore))
            H.assert_groups_equal(got, want, GB_ACCS)


def test_filter_and_computed_key(engine):
    rng = np.random.default_rng(19)
    n = 120_001
    a = rng.integers(0, 5000, n).astype(np.int64)
    b = rng.integers(0, 7, n).astype(np.int64)
    v = rng.normal(size=n)
    spec = [(I64, a, None), (I64, b, None), (D, v, None)]
    f = ex.call("gt", ex.col(2, D), ex.lit(-0.5, D))
    outs = [ex.call("plus", ex.call("multiply", ex.col(0, I64), ex.lit(10, I64)), ex.col(1, I64)), ex.col(2, D)]
    got, info = run_hash(engine, f, outs, spec, n, RAW_KEY, GB_ACCS, batches=2)
    H.assert_groups_equal(got, oracle(f, outs, spec, n, RAW_KEY, GB_ACCS), GB_ACCS)
    if info["engine"] == 1:
        assert info["sorted"] == 0  # no probe with a filter / computed key: everything is deduplicated


def test_nullable_input_appears_in_a_later_batch(engine):
    # batch 1: the sum's input has no null bitmap (lean entry layout); batch 2 brings NULLs -> the layout has to widen
    rng = np.random.default_rng(23)
    n = 40_000
    k = rng.integers(0, 3000, n).astype(np.int64)
    v = rng.normal(size=n)
    vnn = np.ones(n, dtype=bool)
    vnn[n // 2:] = rng.random(n - n // 2) > 0.5
    accs = [(abi.ACC_SUM_F64, 1), (abi.ACC_MIN, 1)]
    outs = [ex.col(0, I64), ex.col(1, D)]
    prog = engine.Program(None, outs)
    agg = engine.Aggregation(abi.AGG_HASH, RAW_KEY, accs)
    h = n // 2
    agg.add_input(prog, [engine.flat_column(I64, k[:h]), engine.flat_column(D, v[:h])], h)
    agg.add_input(prog, [engine.flat_column(I64, k[h:]), engine.flat_column(D, v[h:], vnn[h:])], n - h)
    got = H.decode_rows(agg.extract(), RAW_KEY, accs, [o.type for o in outs], abi.AGG_HASH)
    H.assert_groups_equal(got, oracle(None, outs, [(I64, k, None), (D, v, vnn)], n, RAW_KEY, accs), accs)


def test_merge_rows_after_logged_batches(engine):
    # packed rows folded into an aggregation that already logged input (log -> table migration)
    rng = np.random.default_rng(29)
    n = 30_000
    k = rng.integers(0, 2000, n).astype(np.int64)
    v = rng.integers(0, 50, n).astype(np.float64)
    outs = [ex.col(0, I64), ex.col(1, D)]
    prog = engine.Program(None, outs)
    a = engine.Aggregation(abi.AGG_HASH, RAW_KEY, GB_ACCS)
    b = engine.Aggregation(abi.AGG_HASH, RAW_KEY, GB_ACCS)
    h = n // 2
    a.add_input(prog, [engine.flat_column(I64, k[:h]), engine.flat_column(D, v[:h])], h)
    b.add_input(prog, [engine.flat_column(I64, k[h:]), engine.flat_column(D, v[h:])], n - h)
    import torch
    rows = torch.from_numpy(b.extract()).cuda()
    a.merge_rows(rows)
    got = H.decode_rows(a.extract(), RAW_KEY, GB_ACCS, [o.type for o in outs], abi.AGG_HASH)
    H.assert_groups_equal(got, oracle(None, outs, [(I64, k, None), (D, v, None)], n, RAW_KEY, GB_ACCS), GB_ACCS)


def test_reset_reuses_the_log(engine):
    rng = np.random.default_rng(31)
    outs = [ex.col(0, I64), ex.col(1, D)]
    prog = engine.Program(None, outs)
    agg = engine.Aggregation(abi.AGG_HASH, RAW_KEY, GB_ACCS)
    for rep in range(3):
        k = sorted_runs(rng, 5000 + 1000 * rep, start=rep)
        v = rng.normal(size=len(k))
        agg.reset()
        agg.add_input(prog, [engine.flat_column(I64, k), engine.flat_column(D, v)], len(k))
        got = H.decode_rows(agg.extract(), RAW_KEY, GB_ACCS, [o.type for o in outs], abi.AGG_HASH)
        H.assert_groups_equal(got, oracle(None, outs, [(I64, k, None), (D, v, None)], len(k), RAW_KEY, GB_ACCS), GB_ACCS)


def test_both_engines_agree_on_many_cells(engine):
    # wide entries (sum_i64, counts, min / max of ints and doubles) through the log engine and the table engine
    rng = np.random.default_rng(37)
    n = 200_003
    k = rng.integers(0, 20_000, n).astype(np.int64)
    x = rng.integers(-10**9, 10**9, n).astype(np.int64)
    xnn = rng.random(n) > 0.2
    y = rng.normal(size=n) * 1e6
    ynn = rng.random(n) > 0.2
    spec = [(I64, k, None), (I64, x, xnn), (D, y, ynn)]
    outs = [ex.col(0, I64), ex.col(1, I64), ex.col(2, D)]
    accs = [(abi.ACC_SUM_I64, 1), (abi.ACC_COUNT, 1), (abi.ACC_MIN, 1), (abi.ACC_MAX, 1),
            (abi.ACC_SUM_F64, 2), (abi.ACC_MIN, 2), (abi.ACC_MAX, 2), (abi.ACC_COUNT, 2), (abi.ACC_COUNT, -1)]
    want = oracle(None, outs, spec, n, RAW_KEY, accs)
    saved = os.environ.get("VGPU_HASH_ENGINE")
    try:
        for eng in ("log", "table"):
            os.environ["VGPU_HASH_ENGINE"] = eng
            got, info = run_hash(engine, None, outs, spec, n, RAW_KEY, accs, batches=2)
            assert info["engine"] == (1 if eng == "log" else 2)
            H.assert_groups_equal(got, want, accs)
    finally:
        if saved is None:
            os.environ.pop("VGPU_HASH_ENGINE", None)
        else:
            os.environ["VGPU_HASH_ENGINE"] = saved


# ---- accumulator edge KATs -------------------------------------------------------------------------------------
def bits(x):
    return np.float64(x).view(np.uint64)


@pytest.mark.parametrize("mode", [abi.AGG_ARRAY, abi.AGG_HASH])
def test_min_max_nan_zero_inf(engine, mode):
    """MinMaxAggregates.cpp:116-199 + SimpleNumericAggregate.h:118-138: any non-null input clears the group's null flag,
    NaN never replaces the running value (comparisons with NaN are false) -> a NaN-only group is lowest() / max(),
    not NULL; -0.0 == +0.0 (which zero survives depends on the input order in the reference, so zeros compare by ==)."""
    nan, inf = float("nan"), float("inf")
    groups = {
        0: [nan, nan, nan],            # NaN only
        1: [nan, 3.0, -2.0, nan],      # NaN first
        2: [5.0, nan],                 # NaN last
        3: [-0.0, 0.0, -0.0],          # signed zeros
        4: [inf, 1.0, -inf],           # infinities
        5: [-inf],
        6: [1e308, -1e308, 1.7976931348623157e308],
    }
    k = np.array([g for g, vs in groups.items() for _ in vs], dtype=np.int32)
    v = np.array([x for vs in groups.values() for x in vs], dtype=np.float64)
    nn = np.ones(len(k), dtype=bool)
    # group 7: one NULL input only -> min / max are NULL
    k = np.append(k, np.int32(7)); v = np.append(v, 0.0); nn = np.append(nn, False)
    rng = np.random.default_rng(41)
    perm = rng.permutation(len(k))
    k, v, nn = k[perm], v[perm], nn[perm]
    accs = [(abi.ACC_MIN, 1), (abi.ACC_MAX, 1), (abi.ACC_COUNT, 1), (abi.ACC_COUNT, -1)]
    outs = [ex.col(0, I32), ex.col(1, D)]
    keys = [(0, 0, 0, 7)] if mode == abi.AGG_ARRAY else [(0, 5, 0, 7)]
    prog = engine.Program(None, outs)
    agg = engine.Aggregation(mode, keys, accs)
    agg.add_input(prog, [engine.flat_column(I32, k), engine.flat_column(D, v, nn)], len(k))
    got = H.decode_rows(agg.extract(), keys, accs, [o.type for o in outs], mode)
    want = oracle(None, outs, [(I32, k, None), (D, v, nn)], len(k), keys, accs)
    dmax, dlow = np.finfo(np.float64).max, np.finfo(np.float64).min
    assert want[(0,)][0] == (dmax, 3) and want[(0,)][1] == (dlow, 3)      # the oracle restates the reference here
    assert set(got) == set(want)
    for g in want:
        for a in range(2):
            (gv, gn), (wv, wn) = got[g][a], want[g][a]
            assert (gn > 0) == (wn > 0), (g, a, gn, wn)
            if wn:
                assert gv == wv, (g, a, gv, wv)
                if g[0] != 3:
                    assert bits(gv) == bits(wv), (g, a, gv, wv)
        assert got[g][2][0] == want[g][2][0] and got[g][3][0] == want[g][3][0]
    assert got[(7,)][0][1] == 0 and got[(7,)][1][1] == 0                  # NULL: no non-null input


@pytest.mark.parametrize("mode", [abi.AGG_ARRAY, abi.AGG_HASH])
def test_sum_bigint_overflow_boundary(engine, mode):
    """SumAggregateBase.h:150-158 checks every add in input order.  The device folds partial sums in an unspecified
    order, so the contract is: a group whose TRUE sum leaves int64 always raises; a group all of whose
    prefix sums in every order stay inside int64 never raises; [MAX, -5, 5]-style inputs whose true sum fits but some
    orders overflow MAY raise (the reference raises for the orders that overflow, too)."""
    imax, imin = np.iinfo(np.int64).max, np.iinfo(np.int64).min
    outs = [ex.col(0, I32), ex.col(1, I64)]
    keys = [(0, 0, 0, 3)] if mode == abi.AGG_ARRAY else [(0, 4, 0, 3)]
    accs = [(abi.ACC_SUM_I64, 1)]
    prog = engine.Program(None, outs)

    def run(k, x):
        agg = engine.Aggregation(mode, keys, accs)
        agg.add_input(prog, [engine.flat_column(I32, np.asarray(k, dtype=np.int32)), engine.flat_column(I64, np.asarray(x, dtype=np.int64))], len(k))
        rows = agg.extract()
        return H.decode_rows(rows, keys, accs, [o.type for o in outs], mode)

    # exactly at the boundary, same-sign addends: no order can overflow
    got = run([0, 0, 0, 1, 1], [imax - 10, 4, 6, imin + 7, -7])
    assert got[(0,)][0][0] == imax and got[(1,)][0][0] == imin and got[(0,)][0][1] > 0
    # true sum overflows: always an error, in both directions
    for k, x in (([0, 0], [imax, 1]), ([2, 2, 2], [imin, -1, 0]), ([1, 1, 1, 1], [imax // 2, imax // 2, 1, 1])):
        with pytest.raises(abi.VgpuError) as ei:
            run(k, x)
        assert ei.value.status == abi.ERR_USER and "overflow" in ei.value.message
    # mixed signs whose true sum fits: either the exact sum or the reference's own order-dependent error
    for x in ([imax, -5, 5], [imax, 5, -5]):
        try:
            got = run([3, 3, 3], x)
            assert got[(3,)][0][0] == imax
        except abi.VgpuError as e:
            assert e.status == abi.ERR_USER and "overflow" in e.message


def test_int32_max_key_next_to_null_in_a_two_key_hash(engine):
    """A 33-bit INTEGER key field: INT32_MAX must not wrap onto the NULL id."""
    i32 = np.iinfo(np.int32)
    a = np.array([i32.max, i32.max, 0, i32.min, 5, 0], dtype=np.int32)
    ann = np.array([True, True, False, True, True, False])
    b = np.array([1, 1, 1, 2, 2, 1], dtype=np.int32)
    v = np.arange(6, dtype=np.float64) + 1
    keys = [(0, 33, i32.min, i32.max), (1, 4, 0, 7)]
    accs = [(abi.ACC_SUM_F64, 2), (abi.ACC_COUNT, -1)]
    outs = [ex.col(0, I32), ex.col(1, I32), ex.col(2, D)]
    spec = [(I32, a, ann), (I32, b, None), (D, v, None)]
    got, _ = run_hash(engine, None, outs, spec, 6, keys, accs)
    want = oracle(None, outs, spec, 6, keys, accs)
    assert (i32.max, 1) in got and (None, 1) in got and len(got) == 4
    H.assert_groups_equal(got, want, accs)
    # a declared domain that does not fit its field is rejected
    with pytest.raises(abi.VgpuError):
        run_hash(engine, None, outs, spec, 6, [(0, 32, i32.min, i32.max), (1, 4, 0, 7)], accs)


@pytest.mark.parametrize("cardinality,batches", [(4, 1), (50, 3), (1000, 1)])
def test_unsorted_keys_of_low_cardinality_take_the_combining_pass(engine, cardinality, batches):
    # many more entries than keys, nothing sorted: finalize folds every chunk in shared memory before partitioning
    # (fin_combine_kernel) and a partition's CTA folds runs of one key in registers; NULL keys and the marker key included
    rng = np.random.default_rng(1000 + cardinality)
    n = 400_003
    k = rng.integers(-1, cardinality - 1, n).astype(np.int64)
    knn = rng.random(n) > 0.02
    v = rng.integers(-1000, 1000, n) * 0.125
    vnn = rng.random(n) > 0.1
    spec = [(I64, k, knn), (D, v, vnn)]
    outs = [ex.col(0, I64), ex.col(1, D)]
    got, info = run_hash(engine, None, outs, spec, n, RAW_KEY, GB_ACCS, batches=batches)
    H.assert_groups_equal(got, oracle(None, outs, spec, n, RAW_KEY, GB_ACCS), GB_ACCS)
    assert len(got) == cardinality + 1  # + the NULL key
    if info["engine"] == 1 and cardinality <= 50:
        assert info["deduped_entries"] < n // 4, info  # the count after the combining pass (one entry per key and chunk; ranges end in partial chunks)
