"""Typed expression trees (the ITypedExpr subset of the path) and their lowering to a vgpu program.

This is the Python-side mirror of the C++ host compiler (velox-1_b200/host/ExprLowering.cpp):
tests and bench.py build programs with it and hand them to the C-ABI.  Function names follow the
reference's registry (velox/functions/prestosql/registration/*.cpp, name mapping of
velox/duckdb/conversion/DuckParser.cpp:53-70): plus minus multiply divide mod negate, eq neq lt lte
gt gte between, and or not is_null, cast.
"""
import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

from . import abi

ARITH = {"plus": abi.OP_ADD, "minus": abi.OP_SUB, "multiply": abi.OP_MUL, "divide": abi.OP_DIV,
         "mod": abi.OP_MOD}
CMP = {"eq": abi.OP_EQ, "neq": abi.OP_NE, "lt": abi.OP_LT, "lte": abi.OP_LE, "gt": abi.OP_GT,
       "gte": abi.OP_GE}
FP_TYPES = (abi.REAL, abi.DOUBLE)


@dataclass(frozen=True)
class Expr:
    type: int


@dataclass(frozen=True)
class Col(Expr):
    index: int = 0


@dataclass(frozen=True)
class Lit(Expr):
    value: object = None  # None = typed NULL


@dataclass(frozen=True)
class Call(Expr):
    fn: str = ""
    args: Tuple[Expr, ...] = ()


def col(index, type):
    return Col(type, index)


def lit(value, type):
    return Lit(type, value)


def null(type):
    return Lit(type, None)


def call(fn, *args):
    """Builds a call, inferring the result type the way the reference's signatures do."""
    if fn in ARITH:
        a, b = args
        assert a.type == b.type, f"{fn}: operand types differ ({a.type} vs {b.type}); insert a cast"
        return Call(a.type, fn, (a, b))
    if fn == "negate":
        return Call(args[0].type, fn, tuple(args))
    if fn in CMP:
        a, b = args
        assert a.type == b.type, f"{fn}: operand types differ"
        return Call(abi.BOOLEAN, fn, (a, b))
    if fn == "between":
        assert args[0].type == args[1].type == args[2].type
        return Call(abi.BOOLEAN, fn, tuple(args))
    if fn in ("and", "or"):
        flat = []
        for x in args:  # ExprCompiler.cpp:131-158 flattens nested AND/OR of the same kind
            assert x.type == abi.BOOLEAN
            if isinstance(x, Call) and x.fn == fn:
                flat.extend(x.args)
            else:
                flat.append(x)
        return Call(abi.BOOLEAN, fn, tuple(flat))
    if fn == "not":
        assert args[0].type == abi.BOOLEAN
        return Call(abi.BOOLEAN, fn, tuple(args))
    if fn == "is_null":
        return Call(abi.BOOLEAN, fn, tuple(args))
    if fn == "if":
        # if(condition, then[, else]) — SwitchExpr.cpp:64-160; no ELSE = NULL for the remaining rows
        cond, then = args[0], args[1]
        other = args[2] if len(args) > 2 else null(then.type)
        assert cond.type == abi.BOOLEAN and then.type == other.type, "if: condition must be boolean, branches of one type"
        assert then.type != abi.BOOLEAN, "boolean-valued if is not supported on the device"
        return Call(then.type, fn, (cond, then, other))
    if fn == "coalesce":
        assert len(args) >= 2 and all(a.type == args[0].type for a in args) and args[0].type != abi.BOOLEAN
        return Call(args[0].type, fn, tuple(args))
    raise ValueError(f"unknown function {fn}")


def switch(cases, default=None):
    """CASE WHEN c1 THEN v1 WHEN c2 THEN v2 ... [ELSE default] END as nested if (SwitchExpr.cpp evaluates the
    conditions in order on the rows no earlier condition took — exactly what the nesting does)."""
    out = default if default is not None else null(cases[0][1].type)
    for cond, value in reversed(cases):
        out = call("if", cond, value, out)
    return out


def in_list(x, values):
    """x IN (v1, v2, ...) over literals: OR of eq, the form the device evaluates
    (velox/functions/prestosql/InPredicate.cpp; same three-valued result)."""
    terms = [call("eq", x, lit(v, x.type)) for v in values]
    return terms[0] if len(terms) == 1 else call("or", *terms)


def cast(x, to_type):
    return x if x.type == to_type else Call(to_type, "cast", (x,))


# ---- lowering ----------------------------------------------------------------------------------
@dataclass
class Lowered:
    ops: List[abi.Op] = field(default_factory=list)
    filter: Optional[abi.Operand] = None
    outputs: List[abi.Operand] = field(default_factory=list)
    n_columns: int = 0


class _Lowerer:
    def __init__(self):
        self.ops: List[abi.Op] = []
        self.free_v = list(range(12))[::-1]
        self.free_b = list(range(12))[::-1]
        self.uses = {}     # expr -> remaining uses (common sub-expression elimination)
        self.where = {}    # expr -> (is_bool, slot) of a materialised shared value
        self.pinned = set()   # slots of shared values with uses left
        self.keep = set()     # slots read by the sink (filter / outputs): never recycled
        self.max_col = -1

    def count(self, e):
        self.uses[e] = self.uses.get(e, 0) + 1
        if self.uses[e] == 1 and isinstance(e, Call):
            for a in e.args:
                self.count(a)

    @staticmethod
    def opnd(kind, type, index=0):
        return abi.Operand(kind, type, index)

    def alloc(self, is_bool):
        pool = self.free_b if is_bool else self.free_v
        if not pool:
            raise ValueError("expression needs more than 12 scratch slots")
        return pool.pop()

    def release(self, operand):
        key = (operand.type == abi.BOOLEAN, operand.index)
        if operand.kind == abi.OPND_SLOT and key not in self.pinned and key not in self.keep:
            (self.free_b if operand.type == abi.BOOLEAN else self.free_v).append(operand.index)

    def leaf(self, e):
        if isinstance(e, Col):
            self.max_col = max(self.max_col, e.index)
            return self.opnd(abi.OPND_COL, e.type, e.index), None
        if e.value is None:
            return self.opnd(abi.OPND_NULL, e.type), None
        return self.opnd(abi.OPND_IMM, e.type), e.value

    def gen(self, e, want_slot=False):
        """Emits code for e; returns (operand, imm).  operand.kind is ACC unless the value had to be
        (or already was) stored in a slot, or e is a leaf."""
        if not isinstance(e, Call):
            return self.leaf(e)
        if e in self.where:  # shared value computed earlier
            is_bool, slot = self.where[e]
            self.uses[e] -= 1
            o = self.opnd(abi.OPND_SLOT, e.type, slot)
            if self.uses[e] == 0:
                self.pinned.discard((is_bool, slot))
                del self.where[e]
            return o, None
        shared = self.uses.get(e, 1) > 1
        is_bool = e.type == abi.BOOLEAN
        if e.fn in ("and", "or"):
            res = self._gen_conjunct(e)
        else:
            res = self._gen_call(e)
        if shared or want_slot:
            if res.kind != abi.OPND_SLOT:
                slot = self.alloc(is_bool)
                self.ops[-1].dst_slot = slot
                res = self.opnd(abi.OPND_SLOT, e.type, slot)
            if shared:
                self.uses[e] -= 1
                self.where[e] = (is_bool, res.index)
                self.pinned.add((is_bool, res.index))
        return res, None

    def _emit(self, opcode, type, a, b=None, c=None, imm=None, imm2=None):
        op = abi.Op()
        op.opcode, op.type, op.dst_slot = opcode, type, -1
        none = self.opnd(abi.OPND_NONE, 0)
        op.a, op.b, op.c = a, b or none, c or none
        imm_type = next((o.type for o in (a, b) if o is not None and o.kind == abi.OPND_IMM), type)
        for value, fld, ty in ((imm, "imm", imm_type), (imm2, "imm2", c.type if c is not None else type)):
            if value is None:
                continue
            if ty in FP_TYPES:
                getattr(op, fld).d = float(value)
            else:
                getattr(op, fld).i = int(value)
        self.ops.append(op)
        is_bool = abi.OP_EQ <= opcode <= abi.OP_IS_NULL
        return self.opnd(abi.OPND_ACC, abi.BOOLEAN if is_bool else type)

    def _gen_args(self, args):
        """Evaluates call arguments so that at most the last non-leaf one stays in ACC."""
        calls = [i for i, a in enumerate(args) if isinstance(a, Call) and a not in self.where]
        out = [None] * len(args)
        for i, a in enumerate(args):
            keep_acc = bool(calls) and i == calls[-1]
            out[i] = self.gen(a, want_slot=isinstance(a, Call) and not keep_acc)
        return out

    def _gen_call(self, e):
        args = self._gen_args(e.args)
        ops = [o for o, _ in args]
        imms = [v for _, v in args]
        fn = e.fn
        if fn in ARITH or fn in CMP:
            if imms[0] is not None and imms[1] is not None:
                raise ValueError(f"{fn} of two literals: fold constants before lowering")
            t = e.args[0].type
            res = self._emit((ARITH.get(fn) or CMP.get(fn)), t, ops[0], ops[1],
                             imm=imms[0] if imms[0] is not None else imms[1])
        elif fn == "between":
            if imms[0] is not None:
                raise ValueError("between: literal value operand; fold first")
            t = e.args[0].type
            res = self._emit(abi.OP_BETWEEN, t, ops[0], ops[1], ops[2], imm=imms[1], imm2=imms[2])
        elif fn == "negate":
            res = self._emit(abi.OP_NEG, e.type, ops[0], imm=imms[0])
        elif fn == "cast":
            res = self._emit(abi.OP_CAST, e.type, ops[0], imm=imms[0])
        elif fn == "not":
            res = self._emit(abi.OP_NOT, abi.BOOLEAN, ops[0], imm=imms[0])
        elif fn == "is_null":
            res = self._emit(abi.OP_IS_NULL, abi.BOOLEAN, ops[0], imm=imms[0])
        elif fn == "if":
            if imms[0] is not None or ops[0].kind == abi.OPND_NULL:
                raise ValueError("if: constant condition; fold before lowering")
            res = self._emit(abi.OP_SELECT, e.type, ops[0], ops[1], ops[2], imm=imms[1], imm2=imms[2])
        elif fn == "coalesce":
            res = None
            # coalesce(a, b, c) = coalesce(a, coalesce(b, c)); arguments were evaluated left to right above
            # and stay in slots / leaves, so fold from the right
            acc_o, acc_imm = ops[-1], imms[-1]
            for k in range(len(ops) - 2, -1, -1):
                if imms[k] is not None:
                    raise ValueError("coalesce: literal before the last argument; fold before lowering")
                res = self._emit(abi.OP_COALESCE, e.type, ops[k], acc_o, imm=acc_imm)
                acc_o, acc_imm = res, None
        else:
            raise ValueError(fn)
        for o in ops:
            self.release(o)
        return res

    def _gen_conjunct(self, e):
        opcode = abi.OP_AND if e.fn == "and" else abi.OP_OR
        first, imm = self.gen(e.args[0], want_slot=isinstance(e.args[0], Call))
        running = None
        for x in e.args[1:]:
            o, v = self.gen(x)
            self._emit(opcode, abi.BOOLEAN, running or first, o, imm=v if v is not None else imm)
            imm = None
            if running is None:
                self.release(first)
                running = self.opnd(abi.OPND_SLOT, abi.BOOLEAN, self.alloc(True))
            self.release(o) if o.kind == abi.OPND_SLOT and o.index != running.index else None
            self.ops[-1].dst_slot = running.index
        return running or first


def lower(filter: Optional[Expr], outputs: List[Expr]) -> Lowered:
    """Lowers a filter and output expressions into one op list (shared sub-expressions once)."""
    lw = _Lowerer()
    if filter is not None:
        assert filter.type == abi.BOOLEAN
        lw.count(filter)
    for o in outputs:
        lw.count(o)
    res = Lowered()
    if filter is not None:
        f, imm = lw.gen(filter, want_slot=isinstance(filter, Call))
        if f.kind == abi.OPND_IMM:  # constant filter: materialise through NOT NOT
            raise ValueError("constant filter: fold it before lowering")
        lw.keep.add((True, f.index)) if f.kind == abi.OPND_SLOT else None
        res.filter = f
    for o in outputs:
        operand, imm = lw.gen(o, want_slot=isinstance(o, Call))
        if operand.kind == abi.OPND_IMM:
            # a literal output: MOV it into a slot so the sink can read it
            slot = lw.alloc(o.type == abi.BOOLEAN)
            if o.type == abi.BOOLEAN:
                raise ValueError("boolean literal outputs are not supported")
            lw._emit(abi.OP_MOV, o.type, operand, imm=imm)
            lw.ops[-1].dst_slot = slot
            operand = lw.opnd(abi.OPND_SLOT, o.type, slot)
        if operand.kind == abi.OPND_SLOT:
            lw.keep.add((operand.type == abi.BOOLEAN, operand.index))
        res.outputs.append(operand)
    res.ops = lw.ops
    res.n_columns = lw.max_col + 1
    return res


def to_desc(lowered: Lowered, n_columns: Optional[int] = None):
    """Returns (ProgramDesc, keepalive) for vgpu_program_create."""
    ops = (abi.Op * max(1, len(lowered.ops)))(*lowered.ops)
    outs = (abi.Operand * max(1, len(lowered.outputs)))(*lowered.outputs)
    d = abi.ProgramDesc()
    d.ops = C.cast(ops, C.POINTER(abi.Op))
    d.n_ops = len(lowered.ops)
    d.filter = lowered.filter if lowered.filter is not None else abi.Operand(abi.OPND_NONE, 0, 0)
    d.outputs = C.cast(outs, C.POINTER(abi.Operand))
    d.n_outputs = len(lowered.outputs)
    d.n_columns = lowered.n_columns if n_columns is None else n_columns
    return d, (ops, outs)
