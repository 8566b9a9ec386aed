"""-m gpu: the window exchange (csrc/exchange.cu) with all ranks emulated on ONE GPU (vgpu_debug_exchange_local), so the
partition + store kernel, the dedicated-group slots, the per-source regions, the count matrix indexing, the ownership
function and the receiver-side deduplication are covered on a 1-GPU box.  Shape of the reference's
partialAggregation -> PartitionedOutput -> Exchange -> finalAggregation (velox/exec/tests/MultiFragmentTest.cpp:180-212).
The real multi-process run over NCCL / CUDA IPC is tests/multi_gpu_check.py (needs >= 2 GPUs)."""
import numpy as np
import pytest

import helpers as H
from helpers import abi, ex, D, I32, I64

pytestmark = pytest.mark.gpu
GB_ACCS = [(abi.ACC_SUM_F64, 1), (abi.ACC_COUNT, -1), (abi.ACC_MIN, 1), (abi.ACC_MAX, 1)]
RAW_KEY = [(0, 64, 0, 0)]


def exchange(engine, world, mode, filt, outs, types, shards, keys, accs, fused=False, batches=1, steps=1):
    """shards[r] = list of (values, not_null) per column for rank r.  Returns the per-rank decoded final groups.
    fused: the partial aggregations stream their entries into the owners' windows from inside the scan
    (vgpu_agg_attach_exchange); batches: add_input calls per rank; steps: reset + scan + exchange rounds."""
    prog = engine.Program(filt, outs, n_columns=len(types))
    group = engine.LocalExchangeGroup(world) if fused else None
    partials = [engine.Aggregation(mode, keys, accs) for _ in range(world)]
    finals = [engine.Aggregation(mode, keys, accs) for _ in range(world)]
    for _ in range(steps):
        for r in range(world):
            p, f = partials[r], finals[r]
            p.reset(); f.reset()
            if fused:
                group.attach(r, p)
            n = len(shards[r][0][0])
            step = ((n + batches - 1) // batches + 63) // 64 * 64 if batches > 1 else max(n, 1)
            for lo in range(0, max(n, 1), step):
                hi = min(n, lo + step)
                cols = [engine.flat_column(t, np.asarray(v)[lo:hi], None if nn is None else np.asarray(nn)[lo:hi]) for t, (v, nn) in zip(types, shards[r])]
                p.add_input(prog, cols, hi - lo)
            empty = [engine.flat_column(t, np.asarray(v)[:0]) for t, (v, nn) in zip(types, shards[r])]
            f.add_input(prog, empty, 0)
        if fused:
            group.shuffle(partials, finals)
        else:
            engine.exchange_local(partials, finals)
        engine.check_errors()
    got = [H.decode_rows(f.extract(), keys, accs, [o.type for o in outs], mode) for f in finals]
    if fused:
        group.close()
    return got


def union_of(per_rank):
    u = {}
    for g in per_rank:
        assert not (u.keys() & g.keys()), "a group is final on two ranks"
        u.update(g)
    return u


@pytest.mark.parametrize("fused", [False, True])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_hash_exchange_sorted_shards(engine, world, fused):
    # every rank scans its own order range (disjoint, sorted keys): the shape of config C5
    rng = np.random.default_rng(100 + world)
    shards, allk, allv = [], [], []
    for r in range(world):
        k = np.repeat(1 + 4 * (r * 20_000 + np.arange(20_000, dtype=np.int64)), rng.integers(1, 8, 20_000))
        v = rng.integers(1, 100_000, len(k)) * 0.01
        shards.append([(k, None), (v, None)])
        allk.append(k); allv.append(v)
    outs = [ex.col(0, I64), ex.col(1, D)]
    got = exchange(engine, world, abi.AGG_HASH, None, outs, [I64, D], shards, RAW_KEY, GB_ACCS, fused=fused,
                   batches=3 if fused and world == 3 else 1, steps=2 if world == 2 else 1)
    k, v = np.concatenate(allk), np.concatenate(allv)
    want = H.run_oracle_agg(None, outs, [(k, None), (v, None)], len(k), RAW_KEY, GB_ACCS)
    H.assert_groups_equal(union_of(got), want, GB_ACCS)
    if world > 1:
        sizes = [len(g) for g in got]
        assert min(sizes) > 0.8 * len(want) / world, sizes  # hash partitioning spreads the groups


@pytest.mark.parametrize("fused", [False, True])
@pytest.mark.parametrize("world", [2, 5])
def test_hash_exchange_shared_keys_nulls_and_marker(engine, world, fused):
    # the same keys on every rank, NULL keys and the key -1 (the tables' empty marker), nullable input
    rng = np.random.default_rng(200 + world)
    shards, cols = [], [[], [], [], []]
    for r in range(world):
        n = 30_000 + 17 * r
        k = rng.integers(-2, 3000, n).astype(np.int64)
        knn = rng.random(n) > 0.03
        v = rng.normal(size=n)
        vnn = rng.random(n) > 0.25
        shards.append([(k, knn), (v, vnn)])
        for c, x in zip(cols, (k, knn, v, vnn)):
            c.append(x)
    outs = [ex.col(0, I64), ex.col(1, D)]
    got = exchange(engine, world, abi.AGG_HASH, None, outs, [I64, D], shards, RAW_KEY, GB_ACCS, fused=fused, batches=2 if world == 5 else 1)
    k, knn, v, vnn = [np.concatenate(c) for c in cols]
    want = H.run_oracle_agg(None, outs, [(k, knn), (v, vnn)], len(k), RAW_KEY, GB_ACCS)
    u = union_of(got)
    assert (None,) in u and (-1,) in u
    H.assert_groups_equal(u, want, GB_ACCS)


@pytest.mark.parametrize("world", [2, 8])
def test_dense_exchange_q1_and_global(engine, world):
    rng = np.random.default_rng(300 + world)
    f, outs = H.q1_exprs()
    types = [I32, I32, D, D, D, D, I32]
    shards, whole = [], None
    for r in range(world):
        n = 40_000 + r
        d = [rng.integers(0, 3, n).astype(np.int32), rng.integers(0, 2, n).astype(np.int32), rng.integers(1, 51, n) * 0.01,
             rng.integers(90000, 10500000, n) * 0.01, rng.integers(0, 11, n) * 0.01, rng.integers(0, 9, n) * 0.01,
             rng.integers(8036, 10592, n).astype(np.int32)]
        shards.append([(x, None) for x in d])
        whole = d if whole is None else [np.concatenate([a, b]) for a, b in zip(whole, d)]
    got = exchange(engine, world, abi.AGG_ARRAY, f, outs, types, shards, H.Q1_KEYS, H.Q1_ACCS)
    want = H.run_oracle_agg(f, outs, [(x, None) for x in whole], len(whole[0]), H.Q1_KEYS, H.Q1_ACCS)
    H.assert_groups_equal(union_of(got), want, H.Q1_ACCS)
    # global aggregation: one row over all ranks, on the rank that owns group 0
    f6, outs6 = H.q6_exprs()
    accs6 = [(abi.ACC_SUM_F64, 0), (abi.ACC_COUNT, -1)]
    sh6 = [[(s[6][0], None), (s[3][0], None), (s[2][0], None), (s[4][0], None)] for s in shards]
    got6 = exchange(engine, world, abi.AGG_GLOBAL, f6, outs6, [I32, D, D, D], sh6, [], accs6)
    assert sum(len(g) for g in got6) == 1
    want6 = H.run_oracle_agg(f6, outs6, [(whole[6], None), (whole[3], None), (whole[2], None), (whole[4], None)], len(whole[0]), [], accs6)
    H.assert_groups_equal(union_of(got6), want6, accs6)


def test_fused_exchange_really_streams_and_falls_back(engine):
    # attached: nothing is logged locally (the partial's own groups are only the dedicated ones); a window that is too
    # small makes the scan log locally and the exchange partition afterwards, with the same result
    rng = np.random.default_rng(77)
    k = np.sort(rng.integers(0, 50_000, 200_000)).astype(np.int64)
    v = rng.normal(size=len(k))
    outs = [ex.col(0, I64), ex.col(1, D)]
    prog = engine.Program(None, outs, n_columns=2)
    want = H.run_oracle_agg(None, outs, [(k, None), (v, None)], len(k), RAW_KEY, GB_ACCS)
    for window, streams in ((256 << 20, True), (1 << 20, False)):
        group = engine.LocalExchangeGroup(2, window)
        partials = [engine.Aggregation(abi.AGG_HASH, RAW_KEY, GB_ACCS) for _ in range(2)]
        finals = [engine.Aggregation(abi.AGG_HASH, RAW_KEY, GB_ACCS) for _ in range(2)]
        half = len(k) // 2
        for r, (lo, hi) in enumerate(((0, half), (half, len(k)))):
            group.attach(r, partials[r])
            partials[r].add_input(prog, [engine.flat_column(I64, k[lo:hi]), engine.flat_column(D, v[lo:hi])], hi - lo)
            finals[r].add_input(prog, [engine.flat_column(I64, k[:0]), engine.flat_column(D, v[:0])], 0)
        if streams:
            group.shuffle(partials, finals)
            engine.check_errors()
            got = [H.decode_rows(f.extract(), RAW_KEY, GB_ACCS, [I64, D], abi.AGG_HASH) for f in finals]
            H.assert_groups_equal(union_of(got), want, GB_ACCS)
            partials[0].finalize()
            assert partials[0].num_groups() == 0  # nothing stayed behind
        else:
            # too small for the per-CTA sub-regions: the entries were logged locally instead
            partials[0].finalize()
            assert partials[0].num_groups() > 1000
        group.close()


@pytest.mark.parametrize("world", [4, 8])
def test_fused_exchange_narrow_entries_many_ranks(engine, world):
    # one accumulator = two-word entries: the shape of config C5, whose streaming kernel keeps per-owner buffers for up
    # to 8 ranks; unsorted keys shared between the ranks plus sorted private ranges, two batches, two rounds
    rng = np.random.default_rng(400 + world)
    accs = [(abi.ACC_SUM_F64, 1)]
    shards, allk, allv = [], [], []
    for r in range(world):
        own = np.repeat(10_000_000 + 3 * (r * 30_000 + np.arange(30_000, dtype=np.int64)), rng.integers(1, 6, 30_000))
        shared = rng.integers(-5, 4000, 20_000).astype(np.int64)
        k = np.concatenate([own, shared])
        v = rng.integers(1, 1000, len(k)) * 0.5
        shards.append([(k, None), (v, None)])
        allk.append(k); allv.append(v)
    outs = [ex.col(0, I64), ex.col(1, D)]
    got = exchange(engine, world, abi.AGG_HASH, None, outs, [I64, D], shards, RAW_KEY, accs, fused=True, batches=2, steps=2)
    k, v = np.concatenate(allk), np.concatenate(allv)
    want = H.run_oracle_agg(None, outs, [(k, None), (v, None)], len(k), RAW_KEY, accs)
    H.assert_groups_equal(union_of(got), want, accs)
