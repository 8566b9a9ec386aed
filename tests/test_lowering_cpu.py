"""Host logic that needs no device: expression lowering (velox-1_b200/expr.py), program validation through
the C-ABI, and the kernel generator in dry-run mode for every prebuilt plan shape."""
import ctypes as C
import importlib

import pytest

import helpers as H

abi, ex = H.abi, H.ex
I32, I64, D, B = abi.INTEGER, abi.BIGINT, abi.DOUBLE, abi.BOOLEAN


def opcodes(lowered):
    return [op.opcode for op in lowered.ops]


def test_q1_lowering_shares_subexpressions():
    f, outs = H.q1_exprs()
    lw = ex.lower(f, outs)
    ops = opcodes(lw)
    assert ops.count(abi.OP_SUB) == 1 and ops.count(abi.OP_MUL) == 2 and ops.count(abi.OP_ADD) == 1   # (1 - disc) and disc_price once
    assert ops.count(abi.OP_LE) == 1 and lw.filter.type == B
    assert [o.kind for o in lw.outputs[:4]] == [abi.OPND_COL] * 4                                    # keys, qty, eprice: plain columns
    assert lw.outputs[4].kind == abi.OPND_SLOT and lw.outputs[5].kind == abi.OPND_SLOT
    assert lw.n_columns == 7


def test_q6_lowering_flattens_and():
    f, outs = H.q6_exprs()
    lw = ex.lower(f, outs)
    ops = opcodes(lw)
    assert ops.count(abi.OP_AND) == 3 and ops.count(abi.OP_BETWEEN) == 1 and ops.count(abi.OP_MUL) == 1
    between = [op for op in lw.ops if op.opcode == abi.OP_BETWEEN][0]
    assert between.imm.d == 0.05 and between.imm2.d == 0.07


def test_in_list_is_or_of_eq():
    e = ex.in_list(ex.col(0, I32), [3, 5, 8])
    assert e.fn == "or" and [a.fn for a in e.args] == ["eq"] * 3
    lw = ex.lower(e, [])
    assert opcodes(lw).count(abi.OP_EQ) == 3 and opcodes(lw).count(abi.OP_OR) == 2
    assert ex.in_list(ex.col(0, D), [1.5]).fn == "eq"


def test_type_checks_at_build_time():
    with pytest.raises(AssertionError):
        ex.call("plus", ex.col(0, I32), ex.col(1, D))         # no implicit casts (Velox signatures are exact)
    with pytest.raises(AssertionError):
        ex.call("and", ex.col(0, I32), ex.col(1, B))
    with pytest.raises(ValueError):
        ex.lower(None, [ex.call("plus", ex.lit(1, I64), ex.lit(2, I64))])   # constants are folded by the planner


def test_program_validation_through_the_abi():
    """vgpu_program_create rejects malformed programs with VGPU_ERR_INVALID (VELOX_CHECK analogue) — no device needed."""
    lib = abi.lib()
    abi.check(lib.vgpu_debug_set_dry_run(1))
    try:
        lw = ex.lower(None, [ex.call("plus", ex.col(0, D), ex.col(1, D))])
        lw.ops[0].a.index = 9                                  # column 9 of a 2-column program
        desc, keep = ex.to_desc(lw, 2)
        handle = C.c_void_p()
        rc = lib.vgpu_program_create(C.byref(desc), C.byref(handle))
        assert rc == abi.ERR_INVALID
        buf = C.create_string_buffer(512)
        lib.vgpu_last_error(buf, 512)
        assert buf.value
    finally:
        abi.check(lib.vgpu_debug_set_dry_run(0))


def test_generator_accepts_every_prebuilt_plan_shape():
    ap = importlib.import_module("velox-1_b200.aot_plans")
    lib = abi.lib()
    abi.check(lib.vgpu_debug_set_dry_run(1))
    try:
        seen = {}
        for plan in ap.plans():
            got, note = ap.generate(plan)
            assert got is not None, (plan["name"], note)
            body, h = got
            assert "vgpu_fused" in body and "__launch_bounds__" in body
            again, _ = ap.generate(plan)
            assert again[1] == h, "the kernel hash of a shape must be stable"
            seen[h] = plan["name"]
        assert len(seen) >= len(ap.plans()) - 1                # count(*) rides on the row counter: the two l_orderkey plans share a kernel
    finally:
        abi.check(lib.vgpu_debug_set_dry_run(0))
