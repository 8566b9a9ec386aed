"""-m gpu tests of VGPU_VARCHAR columns (include/vgpu.h): flat and dictionary-encoded StringView arrays read on the
device as the integer the reference maps short string keys to (velox/exec/VectorHasher.h:324-329 stringAsNumber,
:605-625 valueId's range mode), used as grouping keys of the HASH and ARRAY aggregations.  The oracle groups by the
numbers computed by its own restatement of stringAsNumber."""
import numpy as np
import pytest

import helpers as H
from helpers import abi, ex, D, I64

pytestmark = pytest.mark.gpu

ACCS = [(abi.ACC_SUM_F64, 1), (abi.ACC_COUNT, -1), (abi.ACC_MIN, 1), (abi.ACC_MAX, 1)]
RAW_KEY = [(0, 64, 0, 0)]
V = abi.VARCHAR


def random_strings(rng, n, distinct, max_len=7, alphabet=b"ab\x00\xffXYZ09 "):
    pool = [b""] + [bytes(rng.choice(list(alphabet), size=int(rng.integers(1, max_len + 1))).tolist()) for _ in range(distinct - 1)]
    return [pool[i] for i in rng.integers(0, len(pool), n)], pool


def numbers(strings):
    import np_oracle as orc
    return orc.strings_as_numbers(strings)


def run(engine, mode, keys, key_col, values, n, filt=None, in_rows=None, ignore=False):
    outs = [ex.col(0, I64), ex.col(1, D)]
    prog = engine.Program(filt, outs, n_columns=2)
    agg = engine.Aggregation(mode, keys, ACCS, ignore_null_keys=ignore)
    import torch
    sel = None if in_rows is None else torch.from_numpy(np.asarray(in_rows, dtype=np.int32)).cuda()
    agg.add_input(prog, [key_col, engine.flat_column(D, values)], n, in_rows=sel)
    agg.finalize()
    agg.check_errors()
    return H.decode_rows(agg.extract(), keys, ACCS, [I64, D], mode, nullable=not ignore)


def want(strings, nn, values, keys, filt=None, ignore=False, rows=None):
    num = numbers([s if len(s) <= 7 else b"" for s in strings])
    cols = [(num, nn), (values, None)]
    n = len(strings)
    if rows is not None:
        cols = [(num[rows], None if nn is None else nn[rows]), (values[rows], None)]
        n = len(rows)
    return H.run_oracle_agg(filt, [ex.col(0, I64), ex.col(1, D)], cols, n, keys, ACCS, ignore_null_keys=ignore)


def test_number_helpers_round_trip(engine):
    import np_oracle as orc
    for s in [b"", b"a", b"\x00", b"\x00\x00", b"ab", b"\xff" * 7, b"N", b"AIR", b"REG AIR"]:
        assert engine.number_of_string(s) == orc.string_as_number(s)
        assert engine.string_of_number(orc.string_as_number(s)) == s
    assert orc.string_as_number(b"a") == 0x161 and orc.string_as_number(b"") == 0 and orc.string_as_number(b"12345678") is None


@pytest.mark.parametrize("n,distinct", [(1, 1), (777, 5), (200_003, 3000)])
def test_flat_varchar_key_hash(engine, n, distinct):
    rng = np.random.default_rng(n)
    strings, _ = random_strings(rng, n, distinct)
    nn = rng.random(n) > 0.1
    vals = rng.integers(-1000, 1000, n) * 0.25
    col = engine.flat_column(V, engine.string_views(strings), nn)
    got = run(engine, abi.AGG_HASH, RAW_KEY, col, vals, n)
    H.assert_groups_equal(got, want(strings, nn, vals, RAW_KEY), ACCS)
    # group keys come back as numbers that name their strings
    seen = {engine.string_of_number(k[0]) for k in got if k[0] is not None}
    assert seen == {s for s, ok in zip(strings, nn) if ok}


@pytest.mark.parametrize("base_len", [3, 32, 33, 5000])
def test_dictionary_varchar_key(engine, base_len):
    rng = np.random.default_rng(base_len)
    n = 50_021
    _, pool = random_strings(rng, 0, base_len + 1)
    base = pool[:base_len]
    idx = rng.integers(0, base_len, n).astype(np.int32)
    nn = rng.random(n) > 0.05
    base_nn = rng.random(base_len) > 0.1
    vals = rng.normal(size=n)
    col = engine.dictionary_column(V, engine.string_views(base), idx, nn, base_nn)
    got = run(engine, abi.AGG_HASH, RAW_KEY, col, vals, n)
    strings = [base[i] for i in idx]
    H.assert_groups_equal(got, want(strings, nn & base_nn[idx], vals, RAW_KEY), ACCS)


def test_single_character_flags_in_array_mode(engine):
    # Q1's l_returnflag as a flat VARCHAR column: numbers 256 + 'A' .. 256 + 'R' declared as the key's range
    rng = np.random.default_rng(3)
    n = 100_000
    strings = [bytes([c]) for c in rng.choice(list(b"ANR"), n)]
    vals = rng.integers(1, 50, n).astype(np.float64)
    keys = [(0, 0, 256 + ord("A"), 256 + ord("R"))]
    col = engine.flat_column(V, engine.string_views(strings))
    got = run(engine, abi.AGG_ARRAY, keys, col, vals, n)
    H.assert_groups_equal(got, want(strings, None, vals, keys), ACCS)
    assert {engine.string_of_number(k[0]) for k in got} == {b"A", b"N", b"R"}


def test_selection_and_filter(engine):
    rng = np.random.default_rng(5)
    n = 30_000
    strings, _ = random_strings(rng, n, 200)
    vals = rng.normal(size=n)
    rows = np.sort(rng.choice(n, n // 3, replace=False)).astype(np.int32)
    filt = ex.call("gt", ex.col(1, D), ex.lit(0.0, D))
    col = engine.flat_column(V, engine.string_views(strings))
    got = run(engine, abi.AGG_HASH, RAW_KEY, col, vals, n, filt=filt, in_rows=rows)
    H.assert_groups_equal(got, want(strings, None, vals, RAW_KEY, filt=filt, rows=rows), ACCS)


def test_long_strings_only_matter_where_a_passing_row_reads_them(engine):
    rng = np.random.default_rng(9)
    n = 10_000
    strings, _ = random_strings(rng, n, 50)
    vals = np.abs(rng.normal(size=n)) + 1.0
    nn = np.ones(n, dtype=bool)
    # a long string under a NULL and one in a row the filter drops: neither is read
    strings[17] = b"far too long for a number"; nn[17] = False
    strings[4242] = b"12345678"; vals[4242] = -1.0
    filt = ex.call("gt", ex.col(1, D), ex.lit(0.0, D))
    col = engine.flat_column(V, engine.string_views(strings), nn)
    got = run(engine, abi.AGG_HASH, RAW_KEY, col, vals, n, filt=filt)
    H.assert_groups_equal(got, want(strings, nn, vals, RAW_KEY, filt=filt), ACCS)
    # the same long string in a passing row: reported, never silently grouped
    vals[4242] = 1.0
    with pytest.raises(abi.VgpuError) as e:
        run(engine, abi.AGG_HASH, RAW_KEY, engine.flat_column(V, engine.string_views(strings), nn), vals, n, filt=filt)
    assert e.value.status == abi.ERR_UNSUPPORTED and "7 bytes" in str(e.value)
