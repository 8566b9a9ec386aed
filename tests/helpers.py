"""Shared helpers of the parity tests: query shapes of BASELINE.json's configs, running them through
the C-ABI (GPU) and through the numpy oracle, and comparing the two."""
import importlib

import numpy as np

pkg = importlib.import_module("velox-1_b200")
abi, ex = pkg.abi, pkg.expr

D, I32, I64, B = abi.DOUBLE, abi.INTEGER, abi.BIGINT, abi.BOOLEAN
KIND_NAME = {abi.ACC_SUM_F64: "sum_f64", abi.ACC_SUM_I64: "sum_i64", abi.ACC_COUNT: "count",
             abi.ACC_MIN: "min", abi.ACC_MAX: "max"}

DATE_1994_01_01, DATE_1995_01_01, DATE_1998_09_02 = 8766, 9131, 10471


def q6_exprs():
    """Config C1 (TpchQueryBuilder.cpp:461-486 shape): columns shipdate, extendedprice, quantity, discount."""
    sd, ep, qty, disc = ex.col(0, I32), ex.col(1, D), ex.col(2, D), ex.col(3, D)
    f = ex.call("and",
                ex.call("gte", sd, ex.lit(DATE_1994_01_01, I32)),
                ex.call("lt", sd, ex.lit(DATE_1995_01_01, I32)),
                ex.call("between", disc, ex.lit(0.05, D), ex.lit(0.07, D)),
                ex.call("lt", qty, ex.lit(24.0, D)))
    return f, [ex.call("multiply", ep, disc)]


Q6_COLS = ["l_shipdate", "l_extendedprice", "l_quantity", "l_discount"]


def q1_exprs():
    """Config C2 (TpchQueryBuilder.cpp:189-244): columns returnflag, linestatus, quantity,
    extendedprice, discount, tax, shipdate.  Outputs: 2 keys, then qty, eprice, disc_price, charge, disc."""
    rf, ls = ex.col(0, I32), ex.col(1, I32)
    qty, ep, disc, tax, sd = ex.col(2, D), ex.col(3, D), ex.col(4, D), ex.col(5, D), ex.col(6, I32)
    f = ex.call("lte", sd, ex.lit(DATE_1998_09_02, I32))
    disc_price = ex.call("multiply", ep, ex.call("minus", ex.lit(1.0, D), disc))
    charge = ex.call("multiply", disc_price, ex.call("plus", ex.lit(1.0, D), tax))
    return f, [rf, ls, qty, ep, disc_price, charge, disc]


Q1_COLS = ["l_returnflag", "l_linestatus", "l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_shipdate"]
# sum(qty) sum(eprice) sum(disc_price) sum(charge) avg(qty) avg(eprice) avg(disc) count(*):
# avg = SUM_F64 + COUNT over the same input (AverageAggregateBase.h:65-69)
Q1_ACCS = [(abi.ACC_SUM_F64, 2), (abi.ACC_SUM_F64, 3), (abi.ACC_SUM_F64, 4), (abi.ACC_SUM_F64, 5),
           (abi.ACC_SUM_F64, 2), (abi.ACC_COUNT, 2), (abi.ACC_SUM_F64, 3), (abi.ACC_COUNT, 3),
           (abi.ACC_SUM_F64, 6), (abi.ACC_COUNT, 6), (abi.ACC_COUNT, -1)]
Q1_KEYS = [(0, 0, 0, 2), (1, 0, 0, 1)]  # (output, bits, min, max): dictionary ids of 3 / 2 strings


def oracle_cols(cols):
    """[(values, not_null mask or None)] -> decoded oracle columns."""
    return [(v, nn) for v, nn in cols]


def run_oracle_agg(filter_expr, outputs, cols, n, keys, accs, ignore_null_keys=False):
    """Oracle result {key tuple: [(value, nn)]} for accumulators given as (kind, output index)."""
    import np_oracle as orc  # imported here so that bench.py / tools can use the query shapes without loading the oracle
    sel, outs = orc.filter_project(filter_expr, outputs, cols, n)
    key_cols = [outs[k[0]] for k in keys]
    acc_in = [(KIND_NAME[kind], None if idx < 0 else outs[idx]) for kind, idx in accs]
    return orc.aggregate(key_cols, acc_in, sel, ignore_null_keys=ignore_null_keys, global_agg=not keys)


def decode_rows(rows, keys, accs, out_types, mode, nullable=True):
    """Packed group rows from vgpu_agg_extract -> {key tuple: [(value, nn)]}.
    ARRAY/GLOBAL keys are mixed-radix ids (id 0 = NULL when nullable); HASH keys are bit fields or
    the raw value."""
    res = {}
    na = len(accs)
    for r in rows:
        key, flags = int(r[0]), int(r[1])
        kt = []
        if mode == abi.AGG_GLOBAL:
            pass
        elif mode == abi.AGG_ARRAY:
            g = key
            for (_, _, lo, hi) in keys:
                card = hi - lo + 1 + (1 if nullable else 0)
                d = g % card
                g //= card
                if nullable:
                    kt.append(None if d == 0 else d - 1 + lo)
                else:
                    kt.append(d + lo)
        else:
            shift = 0
            for (_, bits, lo, hi) in keys:
                if bits == 64:
                    kt.append(None if flags & 1 else int(np.int64(np.uint64(key & (2**64 - 1)))))
                else:
                    d = (key >> shift) & ((1 << bits) - 1)
                    shift += bits
                    if nullable:
                        kt.append(None if d == 0 else d - 1 + lo)
                    else:
                        kt.append(d + lo)
        vals = []
        for a, (kind, idx) in enumerate(accs):
            raw = np.int64(r[3 + a])
            nn = int(r[3 + na + a])
            t = out_types[idx] if idx >= 0 else I64
            if kind == abi.ACC_SUM_F64 or (kind in (abi.ACC_MIN, abi.ACC_MAX) and t in (abi.REAL, abi.DOUBLE)):
                v = float(raw.view(np.float64))
            else:
                v = int(raw)
            vals.append((v, nn))
        res[tuple(kt)] = vals
    return res


def assert_groups_equal(got, want, accs, rtol=1e-9):
    """Bit-exact for counts / integer sums / min / max / group sets; rtol for FP sums (north_star)."""
    assert set(got.keys()) == set(want.keys()), (sorted(got.keys(), key=str)[:5], sorted(want.keys(), key=str)[:5])
    for k, w in want.items():
        g = got[k]
        for a, (kind, _) in enumerate(accs):
            gv, gn = g[a]
            wv, wn = w[a]
            if kind == abi.ACC_COUNT:
                assert gv == wv, (k, a, gv, wv)
                continue
            # the ABI's "non-null input count" of SUM / MIN / MAX only says NULL (0) or not (> 0): the log engine
            # does not count rows unless a COUNT accumulator asks for it
            assert (gn > 0) == (wn > 0), (k, a, "non-null count", gn, wn)
            if wn == 0:
                continue  # NULL result: value is unspecified
            if kind == abi.ACC_SUM_F64:
                assert abs(gv - wv) <= rtol * max(abs(wv), 1e-300), (k, a, gv, wv)
            else:
                assert gv == wv, (k, a, gv, wv)
