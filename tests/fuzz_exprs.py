"""Seeded random expression trees over the device's op set (the reference runs an expression fuzzer against its own
evaluator in CI, velox/expression/tests/ExpressionFuzzer.cpp; this is the same idea against our oracle).  Test helper."""
import numpy as np

from helpers import ex, D, I32, I64, B

N_ROWS = 3001


def inputs(seed):
    rng = np.random.default_rng(seed)
    n = N_ROWS
    spec = [(I64, rng.integers(-40, 40, n).astype(np.int64), rng.random(n) > 0.15),
            (I64, rng.integers(-3, 4, n).astype(np.int64), rng.random(n) > 0.15),
            (D, np.round(rng.normal(0, 20, n), 1), rng.random(n) > 0.2),
            (D, rng.integers(-2, 3, n).astype(np.float64), None),
            (I32, rng.integers(0, 9, n).astype(np.int32), rng.random(n) > 0.1),
            (B, rng.random(n) > 0.5, rng.random(n) > 0.25)]
    return n, spec


class Gen:
    def __init__(self, seed):
        self.rng = np.random.default_rng(1000 + seed)

    def pick(self, xs):
        return xs[int(self.rng.integers(0, len(xs)))]

    def value(self, t, depth):
        r = self.rng.random()
        if depth <= 0 or r < 0.25:
            if t == I64:
                return self.pick([ex.col(0, I64), ex.col(1, I64), ex.lit(int(self.rng.integers(-5, 6)), I64)])
            if t == I32:
                return self.pick([ex.col(4, I32), ex.lit(int(self.rng.integers(0, 9)), I32)])
            return self.pick([ex.col(2, D), ex.col(3, D), ex.lit(float(self.rng.integers(-4, 5)) / 2, D)])
        if r < 0.55:
            fns = ["plus", "minus", "multiply"] + (["divide"] if t == D or self.rng.random() < 0.3 else []) + (["mod"] if t != D and self.rng.random() < 0.3 else [])
            a, b = self.value(t, depth - 1), self.value(t, depth - 1)
            if not hasattr(a, "fn") and not hasattr(a, "index") and not hasattr(b, "fn") and not hasattr(b, "index"):
                b = ex.col({I64: 0, I32: 4, D: 2}[t], t)  # two literals would need folding
            return ex.call(self.pick(fns), a, b)
        if r < 0.72:
            return ex.call("if", self.boolean(depth - 1), self.value(t, depth - 1), *( [self.value(t, depth - 1)] if self.rng.random() < 0.7 else []))
        if r < 0.84:
            a = self.value(t, depth - 1)
            if not hasattr(a, "fn") and not hasattr(a, "index"):
                a = ex.col({I64: 1, I32: 4, D: 2}[t], t)
            return ex.call("coalesce", a, self.value(t, depth - 1))
        if r < 0.92:
            a = self.value(t, depth - 1)
            return ex.call("negate", a) if (hasattr(a, "fn") or hasattr(a, "index")) else a
        src = self.pick([x for x in (I64, I32, D) if x != t])
        a = self.value(src, depth - 1)
        return ex.cast(a if (hasattr(a, "fn") or hasattr(a, "index")) else ex.col({I64: 0, I32: 4, D: 2}[src], src), t)

    def boolean(self, depth):
        r = self.rng.random()
        if depth <= 0 or r < 0.15:
            return ex.col(5, B)
        if r < 0.5:
            t = self.pick([I64, D, I32])
            a = self.value(t, depth - 1)
            if not hasattr(a, "fn") and not hasattr(a, "index"):
                a = ex.col({I64: 0, I32: 4, D: 2}[t], t)
            return ex.call(self.pick(["eq", "neq", "lt", "lte", "gt", "gte"]), a, self.value(t, depth - 1))
        if r < 0.6:
            t = self.pick([I64, D])
            lo = int(self.rng.integers(-10, 0))
            return ex.call("between", ex.col({I64: 0, D: 2}[t], t), ex.lit(lo if t == I64 else float(lo), t), ex.lit(5 if t == I64 else 5.0, t))
        if r < 0.8:
            return ex.call(self.pick(["and", "or"]), self.boolean(depth - 1), self.boolean(depth - 1), *([self.boolean(depth - 1)] if self.rng.random() < 0.3 else []))
        if r < 0.9:
            return ex.call("not", self.boolean(depth - 1))
        return ex.call("is_null", self.value(self.pick([I64, D]), depth - 1))


def program(seed):
    """(filter or None, outputs) of bounded size."""
    g = Gen(seed)
    filt = g.boolean(2) if seed % 3 else None
    outs = [g.value(I64, 3), g.value(D, 3)]
    if seed % 2:
        outs.append(g.value(I32, 2))
    return filt, outs
