"""TEST INFRASTRUCTURE — numpy restatement of the reference's FilterProject -> HashAggregation
semantics.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this;
the product (velox-1_b200/) never does.

Each function cites the reference code it follows (paths under the reference tree).  Pinned
against the reference's own known-answer tests in tests/test_oracle.py (between KAT, the 3x3
AND/OR truth table, sum overflow, Wave's hand-computed group sums) and against lineitem data
produced by the reference's own dbgen (oracle/_ref).

Values are numpy arrays; a column is (values, not_null) with not_null a bool array or None.
Expression nodes are duck-typed: .type (vgpu_type numbering == velox TypeKind), Col.index,
Lit.value (None = NULL), Call.fn/.args.
"""
import numpy as np

BOOLEAN, TINYINT, SMALLINT, INTEGER, BIGINT, REAL, DOUBLE = range(7)
NP = {BOOLEAN: np.bool_, TINYINT: np.int8, SMALLINT: np.int16, INTEGER: np.int32, BIGINT: np.int64,
      REAL: np.float32, DOUBLE: np.float64}
INT_TYPES = (TINYINT, SMALLINT, INTEGER, BIGINT)


class UserError(Exception):
    """VeloxUserError analogue (velox/common/base/Exceptions.h)."""


def decode(values, indices=None, not_null=None, base_not_null=None):
    """DecodedVector (velox/vector/DecodedVector.cpp:60-92): flatten one dictionary wrap.
    valueAt(i) = base[indices[i]]; row i is null when the wrapper says so or the base value is."""
    if indices is None:
        return values, not_null
    n = len(indices)
    nn = np.ones(n, dtype=bool) if not_null is None else np.asarray(not_null, dtype=bool).copy()
    safe = np.where(nn, indices, 0)
    out = np.asarray(values)[safe]
    if base_not_null is not None:
        nn &= np.asarray(base_not_null, dtype=bool)[safe]
    return out, (None if nn.all() else nn)


def string_as_number(s):
    """VectorHasher::stringAsNumber (velox/exec/VectorHasher.h:324-329) over bits::loadPartialWord
    (velox/common/base/BitUtil.h: the first `size` bytes as a little-endian word): a string of at most 7 bytes as an
    int64; every length has its own range, so "a" != "a\\0".  Longer strings have no number (kUnmappable upstream)."""
    if len(s) > 7:
        return None
    word = 0
    for i, b in enumerate(bytes(s)):
        word |= b << (8 * i)
    return word if len(s) == 0 else word + (1 << (len(s) * 8))


def strings_as_numbers(strings):
    """A VARCHAR column of short strings as the BIGINT column the device reads it as."""
    return np.array([string_as_number(s) for s in strings], dtype=np.int64)


# ---- expression evaluation -------------------------------------------------------------------
def _limits(t):
    i = np.iinfo(NP[t])
    return int(i.min), int(i.max)


def _checked(fn, t, a, b, live):
    """Checked integer arithmetic (velox/common/base/CheckedArithmetic.h:27-95).  Returns
    (result, error mask); only `live` rows can raise."""
    lo, hi = _limits(t)
    A = a.astype(object)
    B = b.astype(object) if b is not None else None
    err = np.zeros(len(a), dtype=bool)
    if fn == "plus":
        r = A + B
    elif fn == "minus":
        r = A - B
    elif fn == "multiply":
        r = A * B
    elif fn == "negate":
        r = -A
    elif fn in ("divide", "mod"):
        zero = (b == 0)
        err |= zero
        Bs = np.where(zero, 1, b).astype(object)
        sign = lambda x: 1 if x >= 0 else -1
        if fn == "divide":  # C++ truncating division
            r = np.array([(abs(int(x)) // abs(int(y))) * sign(x) * sign(y) for x, y in zip(A, Bs)], dtype=object)
        else:  # checkedModulus: b == -1 -> 0; result takes the dividend's sign
            r = np.array([0 if y == -1 else (abs(int(x)) % abs(int(y))) * sign(x) for x, y in zip(A, Bs)], dtype=object)
    else:
        raise ValueError(fn)
    r = np.asarray(r, dtype=object)
    over = np.array([(x < lo or x > hi) for x in r], dtype=bool) if len(r) else np.zeros(0, dtype=bool)
    err |= over
    err &= live
    out = np.array([int(x) if not e else 0 for x, e in zip(r, err | over)], dtype=NP[t]) if len(r) else np.zeros(0, NP[t])
    return out, err


def _fast_int(fn, t, a, b, live):
    """int64 fast path with explicit overflow tests; falls back to _checked when anything trips."""
    if t not in INT_TYPES:
        return None
    with np.errstate(all="ignore"):
        A = a.astype(np.int64)
        B = b.astype(np.int64) if b is not None else None
        lo, hi = _limits(t)
        if fn == "plus":
            r = A + B
            bad = ((A ^ r) & (B ^ r)) < 0
        elif fn == "minus":
            r = A - B
            bad = ((A ^ B) & (A ^ r)) < 0
        else:
            return None
        bad |= (r < lo) | (r > hi)
        if (bad & live).any():
            return None
        return r.astype(NP[t]), np.zeros(len(a), dtype=bool)


def eval_expr(e, cols, n):
    """Returns (values, null mask, error mask) for every row (selection is applied by callers).
    Follows velox/expression/Expr.cpp:371-441 (default-null), ConjunctExpr.cpp:24-174 (AND/OR),
    velox/functions/prestosql/{Arithmetic.h:45-86,Comparisons.h:43-185}, type/Conversions.h."""
    t = e.type
    if hasattr(e, "index"):
        v, nn = cols[e.index]
        null = np.zeros(n, dtype=bool) if nn is None else ~np.asarray(nn, dtype=bool)
        return np.asarray(v), null, np.zeros(n, dtype=bool)
    if not hasattr(e, "fn"):
        if e.value is None:
            return np.zeros(n, dtype=NP[t]), np.ones(n, dtype=bool), np.zeros(n, dtype=bool)
        return np.full(n, e.value, dtype=NP[t]), np.zeros(n, dtype=bool), np.zeros(n, dtype=bool)
    fn = e.fn
    if fn in ("and", "or"):
        parts = [eval_expr(a, cols, n) for a in e.args]
        decided = np.zeros(n, dtype=bool)
        any_err = np.zeros(n, dtype=bool)
        any_null = np.zeros(n, dtype=bool)
        for v, nl, er in parts:
            ok = ~nl & ~er
            decided |= ok & (~v if fn == "and" else v)
            any_err |= er
            any_null |= nl
        err = ~decided & any_err
        null = ~decided & ~err & any_null
        val = ~decided & ~err & ~null if fn == "and" else decided
        return val, null, err
    if fn == "if":
        # SwitchExpr.cpp:64-160: THEN is evaluated on the rows whose condition is TRUE, ELSE on the rest
        # (FALSE or NULL), so an error in the branch a row does not take never surfaces.
        (cv, cn, ce), (av, an, ae), (bv, bn, be) = [eval_expr(a, cols, n) for a in e.args]
        take = cv & ~cn & ~ce
        err = ce | (take & ae) | (~take & ~ce & be)
        null = np.where(take, an, bn) & ~err
        return np.where(take, av, bv).astype(NP[t]), null, err
    if fn == "coalesce":
        # CoalesceExpr.cpp:39-81: argument k is evaluated only where arguments 0..k-1 were NULL.
        v, null, err = eval_expr(e.args[0], cols, n)
        v = np.asarray(v).astype(NP[t])
        for a in e.args[1:]:
            bv, bn, be = eval_expr(a, cols, n)
            need = null & ~err
            v = np.where(need, bv, v).astype(NP[t])
            err = err | (need & be)
            null = np.where(need, bn, null) & ~err
        return v, null, err
    args = [eval_expr(a, cols, n) for a in e.args]
    if fn == "is_null":
        return args[0][1].copy(), np.zeros(n, dtype=bool), args[0][2].copy()
    # Expr::evalArgsDefaultNulls (Expr.cpp:371-441): arguments are evaluated left to right and a row whose
    # argument k is NULL (without an error) is deselected before argument k+1 is evaluated, so an error in a
    # later argument does not surface on it.
    null = np.zeros(n, dtype=bool)
    err = np.zeros(n, dtype=bool)
    for _, nl, er in args:
        err |= er & ~null
        null |= nl & ~err
    live = ~null
    a = args[0][0]
    b = args[1][0] if len(args) > 1 else None
    at = e.args[0].type
    with np.errstate(all="ignore"):
        if fn == "not":
            return ~a & live, null, err
        if fn in ("eq", "neq", "lt", "lte", "gt", "gte"):
            op = {"eq": np.equal, "neq": np.not_equal, "lt": np.less, "lte": np.less_equal,
                  "gt": np.greater, "gte": np.greater_equal}[fn]
            return op(a, b) & live, null, err
        if fn == "between":
            c = args[2][0]
            return (a >= b) & (a <= c) & live, null, err
        if fn == "cast":
            if at in (REAL, DOUBLE) and t in INT_TYPES:
                lo, hi = _limits(t)
                x = a.astype(np.float64)
                bad = np.isnan(x)
                r = np.trunc(x)  # std::round: half away from zero
                r = r + np.sign(x) * (np.abs(x - r) >= 0.5)
                bad |= ~(r >= -9223372036854775808.0) | ~(r < 9223372036854775808.0)
                ri = np.where(bad, 0, r).astype(np.int64)
                bad |= (ri < lo) | (ri > hi)
                return np.where(bad, 0, ri).astype(NP[t]), null, err | (bad & live)
            if at in INT_TYPES and t in INT_TYPES:
                lo, hi = _limits(t)
                x = a.astype(np.int64)
                bad = (x < lo) | (x > hi)
                return np.where(bad, 0, x).astype(NP[t]), null, err | (bad & live)
            return a.astype(NP[t]), null, err
        if t in (REAL, DOUBLE):
            if fn == "plus":
                r = a + b
            elif fn == "minus":
                r = a - b
            elif fn == "multiply":
                r = a * b
            elif fn == "divide":
                r = a / b
            elif fn == "negate":
                r = -a
            elif fn == "mod":
                r = np.where(b == 0, np.nan, np.fmod(a, np.where(b == 0, 1, b)))
            else:
                raise ValueError(fn)
            return r.astype(NP[t]), null, err
        fast = _fast_int(fn, t, a, b, live)
        if fast is not None:
            return fast[0], null, err | fast[1]
        r, bad = _checked(fn, t, a, b, live)
        return r, null, err | bad


def filter_project(filter_expr, outputs, cols, n):
    """FilterProject::getOutput (velox/exec/FilterProject.cpp:125-191) + processFilterResults
    (velox/exec/OperatorUtils.cpp:187-210): a row passes when the filter is TRUE and not NULL.
    Returns (selected bool mask, [(values, null mask)] per output); raises UserError when a row that
    is still alive carries an evaluation error."""
    if filter_expr is None:
        sel = np.ones(n, dtype=bool)
    else:
        v, nl, er = eval_expr(filter_expr, cols, n)
        if er.any():
            raise UserError("filter evaluation error")
        sel = v & ~nl
    outs = []
    for o in outputs:
        v, nl, er = eval_expr(o, cols, n)
        if (er & sel).any():
            raise UserError("projection evaluation error")
        outs.append((v, nl))
    return sel, outs


# ---- aggregation -----------------------------------------------------------------------------
def aggregate(keys, accs, sel, ignore_null_keys=False, global_agg=False):
    """HashAggregation over one or more batches already concatenated
    (velox/exec/GroupingSet.cpp:240-307 grouped, :513-559 global; NULL is a key value unless
    ignore_null_keys, velox/exec/OperatorUtils.cpp:141-157).

    keys: [(values, null mask)]; accs: [(kind, (values, null mask) or None)] with kind in
    sum_f64 | sum_i64 | count | min | max.  Returns {key tuple (None = NULL): [(value, non-null count)]}.
    Accumulator semantics: velox/functions/lib/aggregates/SumAggregateBase.h:25-170 (nulls ignored,
    checked int64 add), CountAggregate.cpp:27-149, MinMaxAggregates.cpp:116-280 (init lowest()/max(),
    update `result < value` / `result > value`, so NaN never replaces).  A value whose non-null count
    is 0 is NULL (velox/exec/Aggregate.h:305-337); count is never NULL."""
    sel = np.asarray(sel, dtype=bool).copy()
    n = len(sel)
    if ignore_null_keys:
        for _, nl in keys:
            sel &= ~nl
    rows = np.nonzero(sel)[0]
    if global_agg or not keys:
        gid = np.zeros(len(rows), dtype=np.int64)
        uniq = [()]
        ngroups = 1
    else:
        cols = []
        for v, nl in keys:
            vv = np.asarray(v)[rows]
            nn = np.asarray(nl)[rows]
            cols.append(np.where(nn, 0, vv).astype(np.int64))
            cols.append(nn.astype(np.int64))
        if len(rows):
            stacked = np.stack(cols, axis=1)
            u, gid = np.unique(stacked, axis=0, return_inverse=True)
            gid = gid.reshape(-1)
            uniq = [tuple(None if u[g, 2 * k + 1] else int(u[g, 2 * k]) for k in range(len(keys))) for g in range(len(u))]
            ngroups = len(u)
        else:
            uniq, ngroups, gid = [], 0, np.zeros(0, dtype=np.int64)
    out = {k: [] for k in uniq}
    for kind, src in accs:
        if src is None:
            v = None
            live = np.ones(len(rows), dtype=bool)
        else:
            v = np.asarray(src[0])[rows]
            live = ~np.asarray(src[1])[rows]
        cnt = np.zeros(ngroups, dtype=np.int64)
        np.add.at(cnt, gid[live], 1)
        if kind == "count":
            res = cnt
        elif kind == "sum_f64":
            res = np.zeros(ngroups, dtype=np.float64)
            np.add.at(res, gid[live], v[live].astype(np.float64))  # sequential, input order
        elif kind == "sum_i64":
            res = np.zeros(ngroups, dtype=np.int64)
            x = v[live].astype(np.int64)
            approx = np.zeros(ngroups, dtype=np.float64)
            np.add.at(approx, gid[live], np.abs(x.astype(np.float64)))
            if (approx >= 2.0**62).any():  # exact check, in row order like checkedPlus
                acc = [0] * ngroups
                for g, val in zip(gid[live], x):
                    acc[g] += int(val)
                    if acc[g] > 2**63 - 1 or acc[g] < -2**63:
                        raise UserError("integer overflow")
                res = np.array(acc, dtype=np.int64)
            else:
                np.add.at(res, gid[live], x)
        elif kind in ("min", "max"):
            fp = np.issubdtype(v.dtype, np.floating)
            if fp:
                init = np.finfo(np.float64).max if kind == "min" else np.finfo(np.float64).min
                res = np.full(ngroups, init, dtype=np.float64)
                x = v[live].astype(np.float64)
                ok = ~np.isnan(x)
                (np.minimum if kind == "min" else np.maximum).at(res, gid[live][ok], x[ok])
            else:
                init = np.iinfo(np.int64).max if kind == "min" else np.iinfo(np.int64).min
                res = np.full(ngroups, init, dtype=np.int64)
                (np.minimum if kind == "min" else np.maximum).at(res, gid[live], v[live].astype(np.int64))
        else:
            raise ValueError(kind)
        for g, k in enumerate(uniq):
            out[k].append((res[g].item(), int(cnt[g])))
    return out
