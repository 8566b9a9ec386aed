// Plan-specialised kernels: source generation.
//
// For one pipeline shape — the ABI program, the encodings / nullability of the bound columns, the
// sink (FilterProject, ARRAY/GLOBAL group-by, HASH group-by) with its keys and accumulator cells —
// gen_source() writes a straight-line CUDA kernel over the helpers of fused_prelude.cuh.  A warp
// owns 128-row blocks; a thread owns rows 2*lane, 2*lane+1, 64+2*lane, 64+2*lane+1 of the block
// (every column load is one fully coalesced 16- or 8-byte access per thread), evaluates the
// expression tree in registers and folds the rows into
//   * register accumulators (few groups: one predicated add per group and cell),
//   * shared-memory tables with atomics (medium cardinality), or
//   * the HBM open-addressing table (HASH sink).
// fused_rt.cu compiles the text (prebuilt cubin from build(), else NVRTC) and launches it; the
// micro-op interpreter of pipeline.cu runs every shape this generator declines.
//
// The scalar semantics come from the same functions as the interpreter's (arith, cast_value,
// compare, and3/or3), called with compile-time opcodes so they fold to single instructions.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>

#include "fused.h"

namespace vg {

namespace {

struct Gen {
  const PipelineParams& p;
  const Program& prog;
  std::string s;
  FusedPlan& plan;
  bool ok = true;
  std::string why;

  Gen(const PipelineParams& pp, FusedPlan& pl) : p(pp), prog(*pp.prog), plan(pl) {}

  void out(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    s += buf;
  }
  void decline(const char* reason) { if (ok) { ok = false; why = reason; } }

  static std::string hex(uint64_t v) {
    char b[32];
    snprintf(b, sizeof(b), "0x%llxull", (unsigned long long)v);
    return b;
  }

  // Row layout of a warp's 128-row block.  pair: thread rows 2l, 2l+1, 64+2l, 64+2l+1 (two 16-byte loads per 8-byte
  // column).  quad (LOG sink): thread rows 4l .. 4l+3 (one 256-bit load per 8-byte column), so that runs of equal keys
  // fold inside the thread first and one segmented scan covers the whole block.
  bool quad = false;
  std::string row_of_j() const { return quad ? "(r0 + j)" : "((j < 2 ? r0 : r1) + (j & 1))"; }
  bool col_used[kMaxCols] = {false};
  void mark(const vgpu_operand& o) { if (o.kind == VGPU_OPND_COL && o.index < kMaxCols) col_used[o.index] = true; }

  // ---- operands ---------------------------------------------------------------------------
  std::string val(const vgpu_operand& o, uint64_t imm) {
    switch (o.kind) {
      case VGPU_OPND_ACC: return "acc";
      case VGPU_OPND_SLOT: return "s" + std::to_string(o.index);
      case VGPU_OPND_COL: return "c" + std::to_string(o.index);
      case VGPU_OPND_IMM: return "V{" + hex(imm) + ", false, false}";
      default: return "V{0ull, true, false}";
    }
  }
  std::string bval(const vgpu_operand& o, uint64_t imm) {
    switch (o.kind) {
      case VGPU_OPND_ACC: return "bacc";
      case VGPU_OPND_SLOT: return "bs" + std::to_string(o.index);
      case VGPU_OPND_COL: return "cb" + std::to_string(o.index);
      case VGPU_OPND_IMM: return imm ? "B{true, false, false}" : "B{false, false, false}";
      default: return "B{false, true, false}";
    }
  }

  // ---- column loads -------------------------------------------------------------------------
  // Alignment a column's walked array (values, or indices of a dictionary) needs for the block-wide vector loads.
  static int walk_log2(const DevCol& d) {
    const bool dict = d.encoding == VGPU_DICTIONARY;
    return dict ? 2 : (d.type == VGPU_BOOLEAN ? -1 : (d.type == VGPU_TINYINT ? 0 : d.type == VGPU_SMALLINT ? 1 : (d.type == VGPU_INTEGER || d.type == VGPU_REAL) ? 2 : d.type == VGPU_VARCHAR ? 4 : 3));
  }
  bool vector_loads_ok(const DevCol& d) const {
    if (p.in_rows) return false;
    const void* walk = d.encoding == VGPU_DICTIONARY ? (const void*)d.indices : d.values;
    const int lg = walk_log2(d);
    const uintptr_t need = (lg == 3 || lg == 4) ? 16 : lg == 2 ? (quad ? 16 : 8) : 1;
    return (reinterpret_cast<uintptr_t>(walk) & (need - 1)) == 0;
  }

  // full: whole 128-row block (vector loads where the column's alignment allows); else per-row guarded loads.
  // With a row selection (in_rows, the batch's SelectivityVector as an index list) every column is gathered
  // through RR[j] = in_rows[row]; a sliced column whose base is not vector-aligned is read row by row.
  void gen_loads(bool full) {
    if (p.in_rows) {
      out("    i64 RR[4];\n    _Pragma(\"unroll\") for (int j = 0; j < 4; ++j) { const i64 r = %s; RR[j] = r < n ? (i64)__ldg(p.in_rows + r) : 0; }\n", row_of_j().c_str());
    }
    for (int c = 0; c < p.n_cols; ++c) {
      if (!col_used[c]) continue;
      const DevCol& d = p.cols[c];
      if (d.encoding == VGPU_CONSTANT) continue;
      const bool dict = d.encoding == VGPU_DICTIONARY;
      const int lg = walk_log2(d);
      const char* src = dict ? "indices" : "values";
      const bool vec = full && vector_loads_ok(d);
      if (d.nulls) {
        if (p.in_rows) {
          out("    u32 N%d = 0u;\n    _Pragma(\"unroll\") for (int j = 0; j < 4; ++j) if (%s < n && bit1(p.nulls[%d], RR[j])) N%d |= 1u << j;\n", c, row_of_j().c_str(), c, c);
        } else if (quad) {
          if (full) out("    const u32 N%d = bits4(p.nulls[%d], r0);\n", c, c);
          else out("    const u32 N%d = r0 < n ? bits4(p.nulls[%d], r0) : 0u;\n", c, c);
        } else if (full) out("    const u32 N%d = bits2(p.nulls[%d], r0) | (bits2(p.nulls[%d], r1) << 2);\n", c, c, c);
        else out("    const u32 N%d = (r0 < n ? bits2(p.nulls[%d], r0) : 0u) | ((r1 < n ? bits2(p.nulls[%d], r1) : 0u) << 2);\n", c, c, c);
      }
      if (lg == -1) { // bit-packed BOOLEAN
        if (p.in_rows) {
          out("    u32 Q%d = 0u;\n    _Pragma(\"unroll\") for (int j = 0; j < 4; ++j) if (%s < n && bit1(reinterpret_cast<const u64*>(p.values[%d]), RR[j])) Q%d |= 1u << j;\n", c, row_of_j().c_str(), c, c);
        } else if (quad) {
          if (full) out("    const u32 Q%d = bits4(reinterpret_cast<const u64*>(p.values[%d]), r0);\n", c, c);
          else out("    const u32 Q%d = r0 < n ? bits4(reinterpret_cast<const u64*>(p.values[%d]), r0) : 0u;\n", c, c);
        } else if (full) out("    const u32 Q%d = bits2(reinterpret_cast<const u64*>(p.values[%d]), r0) | (bits2(reinterpret_cast<const u64*>(p.values[%d]), r1) << 2);\n", c, c, c);
        else out("    const u32 Q%d = (r0 < n ? bits2(reinterpret_cast<const u64*>(p.values[%d]), r0) : 0u) | ((r1 < n ? bits2(reinterpret_cast<const u64*>(p.values[%d]), r1) : 0u) << 2);\n", c, c, c);
        continue;
      }
      if (lg == 4) { // flat VARCHAR: one 16-byte StringView per row -> its number (sv_number)
        out("    u64 Q%d[4];\n", c);
        out("    _Pragma(\"unroll\") for (int j = 0; j < 4; ++j) {\n      const i64 r = %s;\n", row_of_j().c_str());
        out("      Q%d[j] = r < n ? sv_number(ld16(p.values[%d], %s * 16)) : 0ull;\n    }\n", c, c, p.in_rows ? "RR[j]" : "r");
        continue;
      }
      const bool real = !dict && d.type == VGPU_REAL;
      if (vec && quad && lg == 3) {
        const void* walk = dict ? (const void*)d.indices : d.values;
        if ((reinterpret_cast<uintptr_t>(walk) & 31u) == 0) {
          out("    const u64x4 qq%d = ld32(p.%s[%d], r0 * 8);\n", c, src, c);
          out("    const u64 Q%d[4] = {qq%d.x, qq%d.y, qq%d.z, qq%d.w};\n", c, c, c, c, c);
        } else { // 16-byte aligned only
          out("    const uint4 qa%d = ld16(p.%s[%d], r0 * 8), qb%d = ld16(p.%s[%d], r0 * 8 + 16);\n", c, src, c, c, src, c);
          out("    const u64 Q%d[4] = {pack64(qa%d.x, qa%d.y), pack64(qa%d.z, qa%d.w), pack64(qb%d.x, qb%d.y), pack64(qb%d.z, qb%d.w)};\n", c, c, c, c, c, c, c, c, c);
        }
      } else if (vec && quad && lg == 2) {
        out("    const uint4 qa%d = ld16(p.%s[%d], r0 * 4);\n", c, src, c);
        if (real) out("    const u64 Q%d[4] = {d2u((double)__uint_as_float(qa%d.x)), d2u((double)__uint_as_float(qa%d.y)), d2u((double)__uint_as_float(qa%d.z)), d2u((double)__uint_as_float(qa%d.w))};\n", c, c, c, c, c);
        else out("    const u64 Q%d[4] = {sext32(qa%d.x), sext32(qa%d.y), sext32(qa%d.z), sext32(qa%d.w)};\n", c, c, c, c, c);
      } else if (vec && !quad && lg == 3) {
        out("    const uint4 qa%d = ld16(p.%s[%d], r0 * 8), qb%d = ld16(p.%s[%d], r1 * 8);\n", c, src, c, c, src, c);
        out("    const u64 Q%d[4] = {pack64(qa%d.x, qa%d.y), pack64(qa%d.z, qa%d.w), pack64(qb%d.x, qb%d.y), pack64(qb%d.z, qb%d.w)};\n", c, c, c, c, c, c, c, c, c);
      } else if (vec && !quad && lg == 2) {
        out("    const uint2 qa%d = ld8(p.%s[%d], r0 * 4), qb%d = ld8(p.%s[%d], r1 * 4);\n", c, src, c, c, src, c);
        if (real) out("    const u64 Q%d[4] = {d2u((double)__uint_as_float(qa%d.x)), d2u((double)__uint_as_float(qa%d.y)), d2u((double)__uint_as_float(qb%d.x)), d2u((double)__uint_as_float(qb%d.y))};\n", c, c, c, c, c);
        else out("    const u64 Q%d[4] = {sext32(qa%d.x), sext32(qa%d.y), sext32(qb%d.x), sext32(qb%d.y)};\n", c, c, c, c, c);
      } else {
        const char* ty = lg == 0 ? "signed char" : lg == 1 ? "short" : lg == 2 ? (real ? "float" : "int") : "u64";
        const char* at = p.in_rows ? "RR[j]" : "r";
        out("    u64 Q%d[4];\n", c);
        out("    _Pragma(\"unroll\") for (int j = 0; j < 4; ++j) {\n");
        out("      const i64 r = %s;\n", row_of_j().c_str());
        if (real) out("      Q%d[j] = r < n ? d2u((double)reinterpret_cast<const float*>(p.%s[%d])[%s]) : 0ull;\n", c, src, c, at);
        else if (lg == 3) out("      Q%d[j] = r < n ? reinterpret_cast<const u64*>(p.%s[%d])[%s] : 0ull;\n", c, src, c, at);
        else out("      Q%d[j] = r < n ? (u64)(i64)reinterpret_cast<const %s*>(p.%s[%d])[%s] : 0ull;\n", c, ty, src, c, at);
        out("    }\n");
      }
    }
  }

  bool small_dict(int c) const { return p.cols[c].small_dict != 0 && fused_small_dict_enabled(); }

  // Kernel prologue: lane L keeps base value L of every small dictionary (and all lanes its null bits).
  void gen_small_dict_registers() {
    for (int c = 0; c < p.n_cols; ++c) {
      if (!col_used[c] || p.cols[c].encoding != VGPU_DICTIONARY || !small_dict(c)) continue;
      const DevCol& d = p.cols[c];
      const char* ty = d.type == VGPU_TINYINT ? "signed char" : d.type == VGPU_SMALLINT ? "short" : d.type == VGPU_INTEGER ? "int" : d.type == VGPU_REAL ? "float" : "u64";
      out("  u64 DB%d = 0ull;\n  if (lane < p.base_len[%d]) {\n", c, c);
      if (d.type == VGPU_VARCHAR) out("    DB%d = sv_number(ld16(p.values[%d], lane * 16));\n", c, c);
      else if (d.type == VGPU_REAL) out("    DB%d = d2u((double)__ldg(reinterpret_cast<const float*>(p.values[%d]) + lane));\n", c, c);
      else if (d.type == VGPU_BIGINT || d.type == VGPU_DOUBLE) out("    DB%d = __ldg(reinterpret_cast<const u64*>(p.values[%d]) + lane);\n", c, c);
      else out("    DB%d = (u64)(i64)__ldg(reinterpret_cast<const %s*>(p.values[%d]) + lane);\n", c, ty, c);
      out("  }\n");
      if (d.base_nulls) out("  const u32 DBN%d = (u32)__ldg(p.base_nulls[%d]);\n", c, c);
    }
  }

  // VARCHAR: bit 63 of the number marks a string longer than 7 bytes; it becomes the value's error latch, raised only
  // when a passing row consumes it (a NULL row's bytes are never looked at).
  void gen_varchar_fixup(int c) {
    if (p.cols[c].type != VGPU_VARCHAR) return;
    out("      c%d_.e = valid && !c%d_.n && (c%d_.v >> 63) != 0ull; c%d_.v &= 0x7fffffffffffffffull; if (c%d_.e) ec |= VGF_ERR_STRKEY;\n", c, c, c, c, c);
  }

  // Per-row decoded column values (inside the row loop).
  void gen_row_columns() {
    for (int c = 0; c < p.n_cols; ++c) {
      if (!col_used[c]) continue;
      const DevCol& d = p.cols[c];
      const bool is_bool = d.type == VGPU_BOOLEAN;
      if (d.encoding == VGPU_CONSTANT) {
        if (is_bool) out("      const B cb%d = {%s, %s, false};\n", c, d.constant ? "true" : "false", d.const_null ? "true" : "false");
        else out("      const V c%d = {%s, %s, false};\n", c, hex((uint64_t)d.constant).c_str(), d.const_null ? "true" : "false");
        continue;
      }
      std::string nn = d.nulls ? "!((N" + std::to_string(c) + " >> j) & 1u)" : "false";
      if (d.encoding == VGPU_DICTIONARY && small_dict(c)) {
        // base value from the lane that holds it: no second, dependent memory round trip after the index load
        // (every lane executes the shuffle; rows that are invalid or NULL just drop what they read)
        out("      V c%d_; {\n", c);
        out("        bool isn = %s;\n", nn.c_str());
        out("        const u32 ix = (u32)Q%d[j] & 31u;\n", c);
        if (d.type == VGPU_BIGINT || d.type == VGPU_DOUBLE || d.type == VGPU_REAL || d.type == VGPU_VARCHAR) out("        const u64 got = __shfl_sync(0xffffffffu, DB%d, (int)ix);\n", c);
        else out("        const u64 got = (u64)(i64)__shfl_sync(0xffffffffu, (int)DB%d, (int)ix);\n", c); // one 32-bit shuffle
        if (d.base_nulls) out("        isn = isn || !((DBN%d >> ix) & 1u);\n", c);
        out("        isn = isn || !valid;\n");
        out("        c%d_ = V{isn ? 0ull : got, isn, false};\n      }\n", c);
        gen_varchar_fixup(c);
        out("      const V c%d = c%d_;\n", c, c);
        continue;
      }
      if (d.encoding == VGPU_DICTIONARY) {
        out("      V c%d_; bool cv%d_ = false; {\n", c, c);
        out("        bool isn = %s; u64 raw = 0;\n", nn.c_str());
        out("        if (valid && !isn) {\n          const i64 ix = (i64)Q%d[j];\n", c);
        if (is_bool) out("          cv%d_ = bit1(reinterpret_cast<const u64*>(p.values[%d]), ix);\n", c, c);
        else {
          const char* ty = d.type == VGPU_TINYINT ? "signed char" : d.type == VGPU_SMALLINT ? "short" : d.type == VGPU_INTEGER ? "int" : d.type == VGPU_REAL ? "float" : "u64";
          if (d.type == VGPU_VARCHAR) out("          raw = sv_number(ld16(p.values[%d], ix * 16));\n", c);
          else if (d.type == VGPU_REAL) out("          raw = d2u((double)__ldg(reinterpret_cast<const float*>(p.values[%d]) + ix));\n", c);
          else if (d.type == VGPU_BIGINT || d.type == VGPU_DOUBLE) out("          raw = __ldg(reinterpret_cast<const u64*>(p.values[%d]) + ix);\n", c);
          else out("          raw = (u64)(i64)__ldg(reinterpret_cast<const %s*>(p.values[%d]) + ix);\n", ty, c);
        }
        if (d.base_nulls) out("          if (!bit1(p.base_nulls[%d], ix)) isn = true;\n", c);
        out("        }\n        c%d_ = V{raw, isn, false};\n      }\n", c);
        if (is_bool) out("      const B cb%d = {cv%d_ && !c%d_.n, c%d_.n, false};\n", c, c, c, c);
        else { gen_varchar_fixup(c); out("      const V c%d = c%d_;\n", c, c); }
        continue;
      }
      if (is_bool) out("      const B cb%d = {(bool)((Q%d >> j) & 1u), %s, false};\n", c, c, nn.c_str());
      else if (d.type == VGPU_VARCHAR) {
        out("      V c%d_ = {Q%d[j], %s, false};\n", c, c, nn.c_str());
        gen_varchar_fixup(c);
        out("      const V c%d = c%d_;\n", c, c);
      } else out("      const V c%d = {Q%d[j], %s, false};\n", c, c, nn.c_str());
    }
  }

  // ---- the ABI ops, one row --------------------------------------------------------------------
  void gen_ops() {
    for (int k = 0; k < p.n_slots; ++k) out("      V s%d = {0ull, false, false};\n", k);
    for (int k = 0; k < p.n_bslots; ++k) out("      B bs%d = {false, false, false};\n", k);
    out("      V acc = {0ull, false, false}; B bacc = {false, false, false};\n");
    for (const vgpu_op& op : prog.ops) {
      const int oc = op.opcode;
      const uint64_t imm = (uint64_t)op.imm.i, imm2 = (uint64_t)op.imm2.i;
      out("      {\n");
      if ((oc >= VGPU_OP_ADD && oc <= VGPU_OP_MOD)) {
        out("        const V a = %s;\n", val(op.a, imm).c_str());
        if (oc == VGPU_OP_NEG) out("        const V b = {0ull, false, false};\n");
        else out("        const V b = %s;\n", val(op.b, imm).c_str());
        out("        V r; r.n = a.n || b.n; r.e = a.e || (!a.n && b.e);\n");
        out("        const u32 er = arith(%d, %d, a.v, b.v, r.v);\n", oc, op.type);
        out("        if (er && !r.n && valid) { r.e = true; ec |= er; }\n");
        out("        acc = r;\n");
        if (op.dst_slot >= 0) out("        s%d = r;\n", op.dst_slot);
      } else if (oc == VGPU_OP_CAST) {
        out("        const V a = %s;\n", val(op.a, imm).c_str());
        out("        V r; r.n = a.n; r.e = a.e;\n");
        out("        const u32 er = cast_value(%d, %d, a.v, r.v);\n", op.a.type, op.type);
        out("        if (er && !r.n && valid) { r.e = true; ec |= er; }\n");
        out("        acc = r;\n");
        if (op.dst_slot >= 0) out("        s%d = r;\n", op.dst_slot);
      } else if (oc == VGPU_OP_MOV) {
        out("        acc = %s;\n", val(op.a, imm).c_str());
        if (op.dst_slot >= 0) out("        s%d = acc;\n", op.dst_slot);
      } else if (oc == VGPU_OP_SELECT) {
        out("        const B a = %s; const V b = %s, c = %s;\n", bval(op.a, imm).c_str(), val(op.b, imm).c_str(), val(op.c, imm2).c_str());
        out("        V r = (a.v && !a.n) ? b : c; if (a.e) r = V{0ull, false, true};\n");
        out("        acc = r;\n");
        if (op.dst_slot >= 0) out("        s%d = r;\n", op.dst_slot);
      } else if (oc == VGPU_OP_COALESCE) {
        out("        const V a = %s, b = %s;\n", val(op.a, imm).c_str(), val(op.b, imm).c_str());
        out("        const V r = (a.e || !a.n) ? a : b;\n");
        out("        acc = r;\n");
        if (op.dst_slot >= 0) out("        s%d = r;\n", op.dst_slot);
      } else if (oc >= VGPU_OP_EQ && oc <= VGPU_OP_BETWEEN) {
        if (op.type == VGPU_BOOLEAN) {
          out("        const B a = %s, b = %s;\n", bval(op.a, imm).c_str(), bval(op.b, imm).c_str());
          out("        B r; r.n = a.n || b.n; r.e = a.e || (!a.n && b.e); r.v = (a.v %s b.v) && !r.n;\n", oc == VGPU_OP_EQ ? "==" : "!=");
        } else if (oc == VGPU_OP_BETWEEN) {
          out("        const V a = %s, b = %s, c = %s;\n", val(op.a, imm).c_str(), val(op.b, imm).c_str(), val(op.c, imm2).c_str());
          out("        B r; r.n = a.n || b.n || c.n; r.e = a.e || (!a.n && (b.e || (!b.n && c.e)));\n");
          out("        r.v = compare(%d, %d, a.v, b.v) && compare(%d, %d, a.v, c.v) && !r.n;\n", VGPU_OP_GE, op.type, VGPU_OP_LE, op.type);
        } else {
          out("        const V a = %s, b = %s;\n", val(op.a, imm).c_str(), val(op.b, imm).c_str());
          out("        B r; r.n = a.n || b.n; r.e = a.e || (!a.n && b.e); r.v = compare(%d, %d, a.v, b.v) && !r.n;\n", oc, op.type);
        }
        out("        bacc = r;\n");
        if (op.dst_slot >= 0) out("        bs%d = r;\n", op.dst_slot);
      } else {
        if (oc == VGPU_OP_IS_NULL) {
          if (op.a.type == VGPU_BOOLEAN) out("        const B a = %s; B r = {a.n, false, a.e};\n", bval(op.a, imm).c_str());
          else out("        const V a = %s; B r = {a.n, false, a.e};\n", val(op.a, imm).c_str());
        } else if (oc == VGPU_OP_NOT) {
          out("        const B a = %s; B r = {!a.v, a.n, a.e};\n", bval(op.a, imm).c_str());
        } else {
          out("        const B r = %s(%s, %s);\n", oc == VGPU_OP_AND ? "and3" : "or3", bval(op.a, imm).c_str(), bval(op.b, imm).c_str());
        }
        out("        bacc = r;\n");
        if (op.dst_slot >= 0) out("        bs%d = r;\n", op.dst_slot);
      }
      out("      }\n");
    }
  }

  void gen_filter() {
    out("      bool pass = valid, raise = false;\n");
    if (prog.filter.kind != VGPU_OPND_NONE) {
      out("      { const B f = %s; raise = valid && f.e; pass = valid && f.v && !f.n && !f.e; }\n", bval(prog.filter, 1).c_str());
    }
  }

  std::string out_val(int idx) {
    const vgpu_operand& o = p.outputs[idx];
    if (o.kind == VGPU_OPND_IMM || o.kind == VGPU_OPND_ACC) decline("output operand kind");
    if (o.type == VGPU_BOOLEAN) return "";
    return val(o, 0);
  }

  // Group number (ARRAY) or normalized key (HASH) of the row.
  void gen_keys(bool hash) {
    if (hash) out("      u64 key = 0; bool knull = false;\n"); else out("      u32 g = 0;\n");
    for (int k = 0; k < p.n_keys; ++k) {
      const DevKey& ks = p.keys[k];
      const vgpu_operand& o = p.outputs[ks.output];
      if (o.type == VGPU_BOOLEAN) { decline("boolean grouping key"); return; }
      out("      { const V k = %s; raise = raise || (pass && k.e);\n", out_val(ks.output).c_str());
      if (p.ignore_null_keys) out("        pass = pass && !k.n;\n");
      if (hash && ks.bits == 64) {
        out("        key = k.v; knull = k.n; }\n");
        continue;
      }
      out("        const u64 u = k.v - %s;\n", hex((uint64_t)ks.min).c_str());
      out("        const bool bad = k.n ? %s : u > %s;\n", ks.nullable ? "false" : "true", hex((uint64_t)(ks.max - ks.min)).c_str());
      // 64-bit ids: a 33-bit INTEGER field must not wrap into the NULL id (id 0)
      out("        const u64 id = k.n ? 0ull : u + %uull;\n", ks.nullable ? 1u : 0u);
      if (hash) out("        key += id << %d;\n", (int)ks.shift);
      else out("        g += (u32)id * %uu;\n", (unsigned)ks.mult);
      out("        if (bad && pass) { ec |= VGF_ERR_KEYRANGE; raise = true; pass = false; } }\n");
    }
  }

  // Value and liveness of cell `c` for the row: declares x<c> (u64) and l<c> (bool).
  void gen_cell_input(int c) {
    const DevCell& cell = p.cells[c];
    if (cell.input < 0) {
      out("      const u64 x%d = 1ull; const bool l%d = pass;\n", c, c);
      return;
    }
    const vgpu_operand& o = p.outputs[cell.input];
    if (o.type == VGPU_BOOLEAN) {
      out("      const B xb%d = %s; raise = raise || (pass && xb%d.e);\n", c, bval(o, 0).c_str(), c);
      out("      const u64 x%d = %s; const bool l%d = pass && !xb%d.n;\n", c, cell.kind == kCellCount ? "1ull" : ("(u64)xb" + std::to_string(c) + ".v").c_str(), c, c);
      return;
    }
    out("      const V xv%d = %s; raise = raise || (pass && xv%d.e);\n", c, out_val(cell.input).c_str(), c);
    std::string x = "xv" + std::to_string(c) + ".v";
    const bool fp_in = o.type == VGPU_REAL || o.type == VGPU_DOUBLE;
    if (cell.kind == kCellCount) x = "1ull";
    else if (cell.kind == kCellSumF64 && !fp_in) x = "d2u((double)(i64)" + x + ")";
    std::string live = "pass && !xv" + std::to_string(c) + ".n";
    if (cell.kind == kCellMinF64 || cell.kind == kCellMaxF64) live += " && (u2d(" + x + ") == u2d(" + x + "))";
    out("      const u64 x%d = %s; const bool l%d = %s;\n", c, x.c_str(), c, live.c_str());
  }

  void gen_raise() {
    out("      if (raise) atomicOr(p.err_word, ec ? ec : VGF_ERR_OVERFLOW);\n");
  }

  // ---- whole kernels ----------------------------------------------------------------------------
  void gen_block_header(bool full) {
    if (quad) out("    const i64 r0 = blk * 128 + 4 * lane;\n");
    else out("    const i64 r0 = blk * 128 + 2 * lane, r1 = r0 + 64;\n");
    gen_loads(full);
  }
  void gen_row_begin(bool full) {
    out("    _Pragma(\"unroll\") for (int j = 0; j < 4; ++j) {\n");
    if (full) out("      const bool valid = true;\n");
    else out("      const bool valid = %s < n;\n", row_of_j().c_str());
    out("      u32 ec = 0;\n");
    gen_row_columns();
    gen_ops();
    gen_filter();
  }

  int live_cells[kMaxCells];
  int n_live = 0;

  // How the ARRAY / GLOBAL sink keeps its partial groups.
  enum ArrayStrategy {
    AS_REGS,    // registers: one predicated update per (group, cell)
    AS_THREAD,  // per-thread tables in shared memory: plain read-modify-write, no conflicts
    AS_ATOMIC   // per-warp (or per-CTA) tables in shared memory: atomics
  };
  int cell_boff[kMaxCells]; // AS_THREAD: byte offset of the cell's block [G][NT]

  void gen_array_body(bool full, ArrayStrategy st, int G) {
    // Thread-private tables turn every row into dependent shared-memory read-modify-writes; the warps
    // left to cover HBM latency are too few, so the warp's NEXT block is pulled into L2 one iteration
    // early (Q1 SF10: 0.49 -> 0.39 ms).  Register / atomic strategies measured 1-5 % slower with it.
    static const bool prefetch_all = getenv("VGPU_PREFETCH_ALL") && getenv("VGPU_PREFETCH_ALL")[0] == '1'; // tuning aid
    if (full && (st == AS_THREAD || prefetch_all)) gen_input_prefetch();
    gen_block_header(full);
    gen_row_begin(full);
    gen_keys(false);
    for (int i = 0; i < n_live; ++i) {
      const int c = live_cells[i];
      const int kind = p.cells[c].kind;
      gen_cell_input(c);
      if (st == AS_REGS) {
        if (kind == kCellCount) out("      _Pragma(\"unroll\") for (int gg = 0; gg < %d; ++gg) if (l%d && g == (u32)gg) A%d[gg] += 1u;\n", G, c, c);
        else out("      _Pragma(\"unroll\") for (int gg = 0; gg < %d; ++gg) if (l%d && g == (u32)gg) A%d[gg] = cell_comb<%d>(A%d[gg], x%d, accerr);\n", G, c, c, kind, c, c);
      } else if (st == AS_THREAD) {
        if (kind == kCellCount) out("      if (l%d) { u32* a = reinterpret_cast<u32*>(tabc + %d) + g * NT + threadIdx.x; *a = *a + 1u; }\n", c, cell_boff[i]);
        else out("      if (l%d) { u64* a = reinterpret_cast<u64*>(tabc + %d) + g * NT + threadIdx.x; *a = cell_comb<%d>(*a, x%d, accerr); }\n", c, cell_boff[i], kind, c);
      } else {
        if (kind == kCellCount) out("      if (l%d) atomicAdd(reinterpret_cast<u32*>(mytab + %d * %d + g), 1u);\n", c, i, G);
        else out("      if (l%d) smem_update<%d>(mytab + %d * %d + g, x%d, accerr);\n", c, kind, i, G, c);
      }
    }
    gen_raise();
    out("    }\n");
  }

  // Next block's input lines -> L2 (no registers held): the warp's following iteration then starts
  // from L2 latency instead of HBM latency.  One 128-byte line per lane per column.
  void gen_input_prefetch() {
    if (p.in_rows) return;
    static const int dist = getenv("VGPU_PREFETCH_DIST") ? std::max(1, atoi(getenv("VGPU_PREFETCH_DIST"))) : 1; // blocks ahead (tuning aid)
    const std::string ahead = p.mode == kSinkLog ? "(blk + " + std::to_string(dist) + ")" : "(blk + " + std::to_string(dist) + " * nwarps)";
    out("    if (%s < n_full) {\n", ahead.c_str());
    for (int c = 0; c < p.n_cols; ++c) {
      if (!col_used[c]) continue;
      const DevCol& d = p.cols[c];
      if (d.encoding == VGPU_CONSTANT) continue;
      const bool dict = d.encoding == VGPU_DICTIONARY;
      const int w = dict ? 4 : (d.type == VGPU_BIGINT || d.type == VGPU_DOUBLE) ? 8 : (d.type == VGPU_INTEGER || d.type == VGPU_REAL) ? 4 : 0;
      if (!w) continue;
      out("      if (lane < %d) prefetch_l2(reinterpret_cast<const char*>(p.%s[%d]) + %s * %d + lane * 128);\n",
          w, dict ? "indices" : "values", c, ahead.c_str(), 128 * w);
    }
    out("    }\n");
  }

  // Rows of the block -> KEY / KN / PASS and the per-cell X / L arrays (HASH and LOG sinks).
  void gen_keyed_rows(bool full) {
    if (full) gen_input_prefetch();
    gen_block_header(full);
    out("    u64 KEY[4]; bool KN[4], PASS[4];\n");
    for (int i = 0; i < n_live; ++i) out("    u64 X%d[4]; bool L%d[4];\n", live_cells[i], live_cells[i]);
    gen_row_begin(full);
    gen_keys(true);
    for (int i = 0; i < n_live; ++i) {
      const int c = live_cells[i];
      gen_cell_input(c);
      out("      X%d[j] = x%d; L%d[j] = l%d;\n", c, c, c, c);
    }
    out("      KEY[j] = key; KN[j] = knull; PASS[j] = pass;\n");
    gen_raise();
    out("    }\n");
    if (p.mode == kSinkLog && p.sorted_probe) {
      // the batch's last row is the "row before" of the next batch's first one (one warp of the grid gets here)
      out("    if (blk == n_blocks - 1) {\n      _Pragma(\"unroll\") for (int j = 0; j < 4; ++j) if (%s == n - 1) { p.sort_prev_out[0] = KEY[j]; p.sort_prev_out[1] = 1ull; }\n    }\n", row_of_j().c_str());
    }
  }

  // Run aggregation before any group state is touched.  Equal keys tend to arrive together (sorted or
  // clustered inputs such as l_orderkey), so per 64-row half block: (1) a thread folds its two
  // rows when their keys match, (2) a segmented scan over the lanes carries a run's partial
  // from lane to lane, (3) only the last lane of a run owns the run's segment.  Any two rows with
  // equal keys may be folded, so this is exact for every input; it just saves nothing when
  // neighbours differ.  Leaves up to four segments per thread: SK / SN (key), SA (active), SV<c> (cells)
  // and, with_home, SH (home slot of the key in the HBM table).
  void gen_run_aggregation(bool with_home) {
    out("    u64 SK[4]; bool SN[4], SA[4];\n");
    if (with_home) out("    u64 SH[4];\n");
    for (int i = 0; i < n_live; ++i) out("    u64 SV%d[4];\n", live_cells[i]);
    out("    _Pragma(\"unroll\") for (int pr = 0; pr < 2; ++pr) {\n");
    out("      const int j0 = 2 * pr, j1 = j0 + 1;\n");
    out("      const bool p0 = PASS[j0], p1 = PASS[j1];\n");
    out("      const bool merged = p0 && p1 && KEY[j0] == KEY[j1] && KN[j0] == KN[j1];\n");
    out("      // tail segment = the lane's last passing row (plus the first when merged); head = a distinct first row\n");
    out("      const bool hasT = p0 || p1, hasH = p0 && p1 && !merged;\n");
    out("      const u64 tk = p1 ? KEY[j1] : KEY[j0]; const bool tn = p1 ? KN[j1] : KN[j0];\n");
    out("      const u64 hk = KEY[j0]; const bool hn = KN[j0];\n");
    for (int i = 0; i < n_live; ++i) {
      const int c = live_cells[i];
      const int kind = p.cells[c].kind;
      out("      const u64 v0_%d = L%d[j0] ? X%d[j0] : cell_identity<%d>(), v1_%d = L%d[j1] ? X%d[j1] : cell_identity<%d>();\n", c, c, c, kind, c, c, c, kind);
      out("      u64 h%d = v0_%d;\n", c, c);
      out("      u64 t%d = p1 ? (merged ? cell_comb<%d>(v0_%d, v1_%d, accerr) : v1_%d) : v0_%d;\n", c, kind, c, c, c, c);
    }
    out("      const u64 fk = hasH ? hk : tk; const bool fn = hasH ? hn : tn;\n");
    out("      const u64 ptk = __shfl_up_sync(0xffffffffu, tk, 1);\n");
    out("      const u32 pfl = __shfl_up_sync(0xffffffffu, (u32)hasT | ((u32)tn << 1), 1);\n");
    out("      const bool link = lane > 0 && hasT && (pfl & 1u) && ptk == fk && (bool)((pfl >> 1) & 1u) == fn;\n");
    out("      bool f = !(link && !hasH); // the carry flows through this lane's tail only when it has no separate head\n");
    out("      _Pragma(\"unroll\") for (int d = 1; d < 32; d <<= 1) {\n");
    for (int i = 0; i < n_live; ++i) out("        const u64 u%d = __shfl_up_sync(0xffffffffu, t%d, d);\n", live_cells[i], live_cells[i]);
    out("        const bool uf = __shfl_up_sync(0xffffffffu, (u32)f, d) != 0u;\n");
    out("        if (lane >= d && !f) {\n");
    for (int i = 0; i < n_live; ++i) out("          t%d = cell_comb<%d>(u%d, t%d, accerr);\n", live_cells[i], p.cells[live_cells[i]].kind, live_cells[i], live_cells[i]);
    out("          f = uf;\n        }\n      }\n");
    for (int i = 0; i < n_live; ++i) {
      const int c = live_cells[i];
      out("      { const u64 c%d_ = __shfl_up_sync(0xffffffffu, t%d, 1); if (hasH && link) h%d = cell_comb<%d>(c%d_, h%d, accerr); }\n", c, c, c, p.cells[c].kind, c, c);
    }
    out("      const bool next_link = __shfl_down_sync(0xffffffffu, (u32)link, 1) != 0u && lane < 31;\n");
    out("      SK[j0] = hk; SN[j0] = hn; SA[j0] = hasH;\n");
    if (with_home) out("      SH[j0] = HOME[j0]; SH[j1] = p1 ? HOME[j1] : HOME[j0];\n");
    out("      SK[j1] = tk; SN[j1] = tn; SA[j1] = hasT && !next_link;\n");
    for (int i = 0; i < n_live; ++i) out("      SV%d[j0] = h%d; SV%d[j1] = t%d;\n", live_cells[i], live_cells[i], live_cells[i], live_cells[i]);
    out("    }\n");
  }

  void gen_hash_body(bool full) {
    gen_keyed_rows(full);
    // home slots of the rows' keys -> L2 while the run aggregation below computes
    // (computed once per row and handed to the probe below: the hash is ~25 instructions)
    out("    u64 HOME[4];\n");
    out("    _Pragma(\"unroll\") for (int j = 0; j < 4; ++j) {\n");
    out("      HOME[j] = home_slot(p, KEY[j]);\n");
    out("      if (PASS[j] && !KN[j] && (j == 0 || KEY[j] != KEY[j - 1])) prefetch_l2(p.table + HOME[j] * %d);\n", p.row_words);
    out("    }\n");
    gen_run_aggregation(true);
    // the (up to) four segments of the thread probe the table together: their first reads and
    // their claims overlap instead of forming four dependent round trips to HBM
    {
      out("    i64 SL[4]; bool PK[4];\n    hash_probe4<%d>(p, SK, SN, SA, SH, SL, PK, inserted, accerr);\n", p.row_words);
      // park what the first claim did not settle (or, buffer full, walk the table here)
      out("    if (p.ovf != nullptr) {\n      i64 OE[4]; ovf_reserve(p, PK, OE);\n");
      out("      _Pragma(\"unroll\") for (int sgm = 0; sgm < 4; ++sgm) {\n        if (!PK[sgm]) continue;\n");
      out("        if (OE[sgm] < (i64)p.ovf_cap) {\n          u64* e = p.ovf + OE[sgm] * %d;\n          e[0] = SK[sgm];\n", p.row_words);
      for (int c = 0; c < p.n_cells && 1 + c < p.row_words; ++c) {
        bool live = false;
        for (int i = 0; i < n_live; ++i) live = live || live_cells[i] == c;
        if (live) out("          e[%d] = SV%d[sgm];\n", 1 + c, c);
        else out("          e[%d] = cell_identity<%d>();\n", 1 + c, p.cells[c].kind);
      }
      out("        } else {\n          SL[sgm] = hash_walk<%d>(p, SK[sgm], home_slot(p, SK[sgm]), inserted, accerr);\n        }\n      }\n    }\n", p.row_words);
    }
    out("    _Pragma(\"unroll\") for (int sgm = 0; sgm < 4; ++sgm) {\n      if (SA[sgm] && SL[sgm] >= 0) {\n");
    for (int i = 0; i < n_live; ++i) {
      const int c = live_cells[i];
      out("        if (SV%d[sgm] != cell_identity<%d>()) hbm_update<%d>(p.table + SL[sgm] * %d + %d, SV%d[sgm], accerr);\n", c, p.cells[c].kind, p.cells[c].kind, p.row_words, 1 + c, c);
    }
    out("      }\n    }\n");
  }

  // LOG sink: the run-aggregated segments are appended to the CTA's region of the log.  A segment whose run cannot
  // continue in a neighbouring half block goes to the front of the region, one that touches a half-block edge to
  // the back; when the whole input turns out to be sorted by key, the front entries are complete groups and only
  // the back entries need deduplication (hashagg.cu).  NULL keys and the key equal to the empty marker are rare and
  // go to two dedicated rows with atomics.
  // ---- LOG sink: run aggregation + entry log ------------------------------------------------------------------
  // A warp walks a CONTIGUOUS range of 128-row blocks.  Inside a block, runs of equal keys are folded in the thread
  // (4 consecutive rows), then across lanes (one segmented scan); the block's LAST run is not logged but stays open and
  // enters the next block as "the row before lane 0's first row", so a run is logged once, where it ends, however
  // many blocks it covers.  What is left to deduplicate for a clustered input are the runs at the two ends of each
  // warp's range.  Plans whose keys and accumulator inputs are plain flat columns (owned_runs_ok) settle most of
  // those in the kernel as well: the warp reads on past its range, up to 31 rows, and finishes its open run; the
  // next warp applies the same test to its own first rows (they continue the row before them and the run is shorter
  // than 32 rows) and drops them.  Both sides evaluate the rule on the same rows, so it holds for unsorted inputs too
  // (where every entry is deduplicated anyway).  Longer runs and the ends of a batch stay "edge" (back) entries.
  bool owned_runs_ok() const {
    static const bool off = getenv("VGPU_OWNED_RUNS") && getenv("VGPU_OWNED_RUNS")[0] == '0'; // tuning aid
    if (off || !p.sorted_probe) return false;
    for (int i = 0; i < n_live; ++i) {
      const DevCell& cell = p.cells[live_cells[i]];
      if (cell.input < 0) continue;
      const vgpu_operand& o = p.outputs[cell.input];
      if (o.kind != VGPU_OPND_COL || o.type == VGPU_BOOLEAN) return false;
      const DevCol& d = p.cols[o.index];
      if (d.encoding != VGPU_FLAT || d.type == VGPU_BOOLEAN || d.type == VGPU_VARCHAR) return false;
    }
    return true;
  }
  std::string row_load(int c, const char* r) const {
    const DevCol& d = p.cols[c];
    char buf[256];
    if (d.type == VGPU_REAL) snprintf(buf, sizeof buf, "d2u((double)__ldg(reinterpret_cast<const float*>(p.values[%d]) + %s))", c, r);
    else if (d.type == VGPU_BIGINT || d.type == VGPU_DOUBLE) snprintf(buf, sizeof buf, "__ldg(reinterpret_cast<const u64*>(p.values[%d]) + %s)", c, r);
    else snprintf(buf, sizeof buf, "(u64)(i64)__ldg(reinterpret_cast<const %s*>(p.values[%d]) + %s)", d.type == VGPU_TINYINT ? "signed char" : d.type == VGPU_SMALLINT ? "short" : "int", c, r);
    return buf;
  }
  int log_kind(int c) const { return p.cells[c].kind; }

  // Kernel prologue: the warp's block range and its open run.
  void gen_log_prologue() {
    out("  const i64 per = (n_blocks + nwarps - 1) / nwarps;\n");
    out("  const i64 blk_begin = wid * per < n_blocks ? wid * per : n_blocks;\n");
    out("  const i64 blk_end = blk_begin + per < n_blocks ? blk_begin + per : n_blocks;\n");
    out("  u64 ck = 0ull; bool cn = false, chave = false, cedge = false; // the open run: key, is-null, exists, is an edge run\n");
    for (int i = 0; i < n_live; ++i) out("  u64 cv%d = cell_identity<%d>();\n", live_cells[i], log_kind(live_cells[i]));
  }

  // Stores one entry (key SKx, values SVx<c>) at offset OFFx of the CTA's region.
  // (send mode: into the warp's staging area instead, with the entry's kind beside it)
  bool send_mode() const { return p.mode == kSinkLog && p.send_world > 0; }
  static constexpr int kSendStage = 192; // fused_prelude.cuh kSendRing: 5 entries per lane per block + up to 31 pending
  // per warp: the staging area (entries, kinds), the per-owner fill counts, the per-owner buffers (fused_prelude.cuh kSendBuf)
  int send_warp_words() const { return kSendStage * p.log_entry_words + 4 + p.send_world * 48 * p.log_entry_words; }
  std::string send_targs() const { return std::to_string(p.log_entry_words) + ", " + std::to_string(p.send_world) + ", " + (p.send_hash ? "true" : "false"); }
  // Entry (key, values SV<c><vsuffix>) of kind `kind` in send mode: front entries into the warp's ring at `off`, back
  // entries straight to their owner.
  void gen_send_store(const char* kind, const char* off, const char* key, const char* vsuffix) {
    const int EW = p.log_entry_words;
    std::string vals;
    for (int i = 0; i < p.log_n_cells; ++i) {
      const int c = p.log_cells[i];
      vals += ", SV" + std::to_string(p.cells[c].skip ? 0 : c) + vsuffix;
    }
    for (int w = 1 + p.log_n_cells; w < EW; ++w) vals += ", 0ull";
    out("      if (%s == 1u) {\n        u32 at = sbeg + spend + %s; if (at >= %du) at -= %du;\n        u64* e = wl + (u64)at * %d;\n", kind, off, kSendStage, kSendStage, EW);
    if (EW == 2) {
      const int c = p.log_cells[0];
      out("        *reinterpret_cast<ulonglong2*>(e) = make_ulonglong2(%s, SV%d%s);\n", key, p.cells[c].skip ? 0 : c, vsuffix);
    } else {
      out("        const u64 ent[%d] = {%s%s};\n        _Pragma(\"unroll\") for (int w = 0; w < %d; ++w) e[w] = ent[w];\n", EW, key, vals.c_str(), EW);
    }
    out("      } else if (%s == 2u) {\n        const u64 ent[%d] = {%s%s};\n", kind, EW, key, vals.c_str());
    out("        send_back<%s>(scap, sfront, sback, sbase, ent, accerr);\n      }\n", send_targs().c_str());
  }
  // After the stores: whole steps of 32 entries leave the ring.
  void gen_send_steps(const char* added) {
    out("    __syncwarp();\n    spend += %s;\n", added);
    out("    while (spend >= 32u) {\n      send_step<%s>(scap, sfront, sback, sbase, wl, wbuf, wcnt, sbeg, 32u, accerr);\n", send_targs().c_str());
    out("      sbeg += 32u; if (sbeg >= %du) sbeg -= %du;\n      spend -= 32u;\n    }\n", kSendStage, kSendStage);
  }
  void gen_log_store(const char* cond, const char* off, const char* key, const char* vsuffix, const char* kind) {
    const int EW = p.log_entry_words;
    if (send_mode()) { gen_send_store(kind, off, key, vsuffix); return; }
    out("      if (%s) {\n        u64* e = region + (u64)%s * %d;\n", cond, off, EW);
    if (EW == 2) {
      const int c = p.log_cells[0];
      out("        *reinterpret_cast<ulonglong2*>(e) = make_ulonglong2(%s, SV%d%s);\n", key, p.cells[c].skip ? 0 : c, vsuffix);
    } else {
      out("        e[0] = %s;\n", key);
      for (int i = 0; i < p.log_n_cells; ++i) {
        const int c = p.log_cells[i];
        out("        e[%d] = SV%d%s;\n", 1 + i, p.cells[c].skip ? 0 : c, vsuffix);
      }
      for (int w = 1 + p.log_n_cells; w < EW; ++w) out("        e[%d] = 0ull;\n", w);
    }
    out("      }\n");
  }
  // The two dedicated groups (NULL key, the key equal to the empty marker) are updated in place.
  void gen_log_special(const char* cond, const char* isnull, const char* vsuffix) {
    out("      if (%s) {\n", cond);
    out("        u64* sp = p.log_special + (%s ? 0 : %d);\n        sp[0] = 1ull;\n", isnull, 1 + p.log_n_cells);
    for (int i = 0; i < p.log_n_cells; ++i) {
      const int c = p.log_cells[i];
      out("        hbm_update<%d>(sp + %d, SV%d%s, accerr);\n", log_kind(c), 1 + i, p.cells[c].skip ? 0 : c, vsuffix);
    }
    out("      }\n");
  }

  void gen_log_body(bool full) {
    gen_keyed_rows(full);
    const bool own = owned_runs_ok();
    out("    const bool first_iter = blk == blk_begin;\n");
    if (p.sorted_probe) {
      // adjacent rows must be in non-decreasing (unsigned) key order: inside the thread, lane to lane, and against the
      // row before the block (the open run's key; for the warp's first block the row itself, or the previous batch's last)
      out("    u64 before = ck; bool have_before = chave;\n");
      out("    if (first_iter && lane == 0) {\n");
      out("      if (blk > 0) { before = __ldg(reinterpret_cast<const u64*>(p.values[%d]) + (blk * 128 - 1)); have_before = true; }\n", p.sorted_key_col);
      out("      else { before = p.sort_prev_in[0]; have_before = p.sort_prev_in[1] != 0ull; }\n");
      out("    }\n    {\n");
      if (full) out("      const bool ok0 = true, ok1 = true, ok2 = true, ok3 = true;\n");
      else out("      const bool ok0 = r0 < n, ok1 = r0 + 1 < n, ok2 = r0 + 2 < n, ok3 = r0 + 3 < n;\n");
      out("      bool bad = (ok1 && KEY[0] > KEY[1]) || (ok2 && KEY[1] > KEY[2]) || (ok3 && KEY[2] > KEY[3]);\n");
      out("      const u64 pk = __shfl_up_sync(0xffffffffu, KEY[3], 1);\n");
      out("      bad = bad || (ok0 && (lane > 0 ? pk > KEY[0] : (have_before && before > KEY[0])));\n");
      out("      viol = viol || bad;\n    }\n");
    }
    // Thread-local fold.  rf<c> is the running fold of the run that is open at row j; the run that ENDS before row j
    // leaves its key and value in slot j - 1 (slot "4" for the run handed over by the previous block, which can only end
    // in lane 0).  Slot 3 is the thread's tail run: it may continue in the next lane.
    out("    u64 SK[4]; bool SN[4], SA[4];\n    u64 SK4 = 0ull; bool SN4 = false, SA4 = false;\n");
    for (int i = 0; i < n_live; ++i) out("    u64 SV%d[4], SV%d_4 = 0ull;\n", live_cells[i], live_cells[i]);
    out("    const bool cin = chave && lane == 0;\n");
    out("    u64 lastk = ck, firstk = 0ull; bool lastn = cn, firstn = false, have = cin;\n");
    for (int i = 0; i < n_live; ++i) out("    u64 rf%d = cin ? cv%d : cell_identity<%d>();\n", live_cells[i], live_cells[i], log_kind(live_cells[i]));
    out("    _Pragma(\"unroll\") for (int j = 0; j < 4; ++j) {\n");
    out("      const bool start = PASS[j] && !(have && lastk == KEY[j] && lastn == KN[j]);\n");
    out("      if (j == 0) { SK4 = lastk; SN4 = lastn; SA4 = start && have;\n");
    for (int i = 0; i < n_live; ++i) out("        SV%d_4 = rf%d;\n", live_cells[i], live_cells[i]);
    out("      } else { SK[j - 1] = lastk; SN[j - 1] = lastn; SA[j - 1] = start && have;\n");
    for (int i = 0; i < n_live; ++i) out("        SV%d[j - 1] = rf%d;\n", live_cells[i], live_cells[i]);
    out("      }\n");
    out("      if (PASS[j] && !have) { firstk = KEY[j]; firstn = KN[j]; }\n");
    for (int i = 0; i < n_live; ++i) {
      const int c = live_cells[i];
      out("      { const u64 v = L%d[j] ? X%d[j] : cell_identity<%d>(); rf%d = start ? v : cell_comb<%d>(rf%d, v, accerr); }\n", c, c, log_kind(c), c, log_kind(c), c);
    }
    out("      if (PASS[j]) { lastk = KEY[j]; lastn = KN[j]; have = true; }\n");
    out("    }\n");
    out("    SK[3] = lastk; SN[3] = lastn;\n");
    out("    const bool hasT = have, single = have && !(SA4 || SA[0] || SA[1] || SA[2]);\n");
    out("    const u64 ptk = __shfl_up_sync(0xffffffffu, lastk, 1);\n");
    out("    const u32 pfl = __shfl_up_sync(0xffffffffu, (u32)hasT | ((u32)lastn << 1), 1);\n");
    out("    const bool link = lane > 0 && hasT && (pfl & 1u) && ptk == firstk && (bool)((pfl >> 1) & 1u) == firstn;\n");
    out("    bool f = !(link && single); // the carry flows through this lane's tail only when the tail is its only run\n");
    for (int i = 0; i < n_live; ++i) out("    u64 t%d = rf%d;\n", live_cells[i], live_cells[i]);
    // (runs are short in practice: the scan stops as soon as no lane waits for a carry from further up)
    out("    _Pragma(\"unroll\") for (int d = 1; d < 32; d <<= 1) {\n");
    out("      if (__all_sync(0xffffffffu, f)) break;\n");
    for (int i = 0; i < n_live; ++i) out("      const u64 u%d = __shfl_up_sync(0xffffffffu, t%d, d);\n", live_cells[i], live_cells[i]);
    out("      const bool uf = __shfl_up_sync(0xffffffffu, (u32)f, d) != 0u;\n");
    out("      const bool take = lane >= d && !f;\n");
    for (int i = 0; i < n_live; ++i) out("      t%d = take ? cell_comb<%d>(u%d, t%d, accerr) : t%d;\n", live_cells[i], log_kind(live_cells[i]), live_cells[i], live_cells[i], live_cells[i]);
    out("      f = take ? uf : f;\n    }\n");
    // the head of a lane with several runs takes the previous lane's (scanned) tail
    for (int i = 0; i < n_live; ++i) {
      const int c = live_cells[i];
      out("    { const u64 carry = __shfl_up_sync(0xffffffffu, t%d, 1);\n", c);
      out("      if (link && !single) { if (SA[0]) SV%d[0] = cell_comb<%d>(carry, SV%d[0], accerr); else if (SA[1]) SV%d[1] = cell_comb<%d>(carry, SV%d[1], accerr); else SV%d[2] = cell_comb<%d>(carry, SV%d[2], accerr); } }\n",
          c, log_kind(c), c, c, log_kind(c), c, c, log_kind(c), c);
      out("    SV%d[3] = t%d;\n", c, c);
    }
    out("    const bool next_link = __shfl_down_sync(0xffffffffu, (u32)link, 1) != 0u && lane < 31;\n");
    out("    SA[3] = hasT && !next_link;\n");
    // the block's last run stays open
    out("    const u32 tm = __ballot_sync(0xffffffffu, hasT);\n");
    out("    const int LT = 31 - __clz((int)tm); // tm == 0: no rows passed and nothing was handed over\n");
    out("    bool nhave = tm != 0u;\n");
    out("    const u64 nk = __shfl_sync(0xffffffffu, SK[3], LT & 31);\n");
    out("    const bool nn = __shfl_sync(0xffffffffu, (u32)SN[3], LT & 31) != 0u;\n");
    for (int i = 0; i < n_live; ++i) out("    const u64 nv%d = __shfl_sync(0xffffffffu, SV%d[3], LT & 31);\n", live_cells[i], live_cells[i]);
    out("    if (lane == LT) SA[3] = false;\n");
    // an edge run that is still open stays an edge run (same key: for a clustered input that is the same run; for any
    // other input the entry is deduplicated anyway)
    out("    bool nedge = cedge && nhave && nk == ck && nn == cn;\n");
    out("    bool EDGE[4] = {false, false, false, false};\n");
    // the warp's first block: the run of its first row may continue the previous warp's (or batch's) last run
    out("    if (first_iter) {\n");
    out("      const u32 pm = __ballot_sync(0xffffffffu, PASS[0] || PASS[1] || PASS[2] || PASS[3]);\n");
    out("      const int Lf = __ffs((int)pm) - 1;\n");
    out("      const u64 KF = __shfl_sync(0xffffffffu, firstk, Lf & 31);\n");
    out("      const bool KFn = __shfl_sync(0xffffffffu, (u32)firstn, Lf & 31) != 0u;\n");
    if (own) {
      // position of the head run (the run of the block's first row): lanes 0 .. Lh-1 are that run and nothing else
      if (full) out("      const u32 LV = 32u; const u32 nvalid = 128u;\n");
      else out("      const u32 nvalid = (u32)(n - blk * 128); const u32 LV = (nvalid + 3u) >> 2;\n");
      out("      const u32 chain = __ballot_sync(0xffffffffu, single && (link || lane == 0));\n");
      out("      const u32 Lh = min(LV, (u32)__ffs((int)~chain) - 1u);\n");
      out("      const bool whole = Lh >= LV;\n");
      out("      const bool cont = __shfl_sync(0xffffffffu, (u32)(have_before && before == KEY[0]), 0) != 0u;\n");
      out("      const bool link_h = __shfl_sync(0xffffffffu, (u32)link, (int)(Lh & 31u)) != 0u && Lh < 32u;\n");
      out("      const u32 in_lane = __shfl_sync(0xffffffffu, SA[0] ? 1u : (SA[1] ? 2u : 3u), (int)(Lh & 31u));\n");
      out("      const u32 hlen = whole ? nvalid : 4u * Lh + ((Lh == 0u || link_h) ? in_lane : 0u);\n");
      out("      int hs = -1;\n");
      out("      if (!whole) {\n");
      out("        if (Lh > 0u && !link_h) { if (lane == (int)Lh - 1) hs = 3; }\n");
      out("        else if (lane == (int)Lh) hs = SA[0] ? 0 : (SA[1] ? 1 : 2);\n");
      out("      }\n");
      out("      if (cont && blk > 0 && hlen < 32u) {\n");
      out("        // the previous warp read these rows and folded them into its last run\n");
      out("        _Pragma(\"unroll\") for (int sgm = 0; sgm < 4; ++sgm) if (hs == sgm) SA[sgm] = false;\n");
      out("        if (whole) nhave = false;\n");
      out("      } else if (cont) {\n");
      out("        _Pragma(\"unroll\") for (int sgm = 0; sgm < 4; ++sgm) EDGE[sgm] = SK[sgm] == KF && SN[sgm] == KFn;\n");
      out("        nedge = nedge || (nk == KF && nn == KFn);\n");
      out("      }\n");
    } else {
      out("      if (pm != 0u) {\n");
      out("        _Pragma(\"unroll\") for (int sgm = 0; sgm < 4; ++sgm) EDGE[sgm] = SK[sgm] == KF && SN[sgm] == KFn;\n");
      out("        nedge = nedge || (nk == KF && nn == KFn);\n");
      out("      }\n");
    }
    out("    }\n");
    out("    u32 KIND[5];\n");
    out("    _Pragma(\"unroll\") for (int sgm = 0; sgm < 4; ++sgm) {\n");
    out("      const bool special = SN[sgm] || SK[sgm] == VGF_EMPTY_KEY;\n");
    out("      KIND[sgm] = !SA[sgm] || special ? 0u : (EDGE[sgm] ? 2u : 1u);\n");
    {
      // (the special-group update of slot sgm)
      out("      if (SA[sgm] && special) {\n");
      out("        u64* sp = p.log_special + (SN[sgm] ? 0 : %d);\n        sp[0] = 1ull;\n", 1 + p.log_n_cells);
      for (int i = 0; i < p.log_n_cells; ++i) {
        const int c = p.log_cells[i];
        out("        hbm_update<%d>(sp + %d, SV%d[sgm], accerr);\n", log_kind(c), 1 + i, p.cells[c].skip ? 0 : c);
      }
      out("      }\n");
    }
    out("    }\n");
    out("    {\n      const bool special = SN4 || SK4 == VGF_EMPTY_KEY;\n");
    out("      KIND[4] = !SA4 || special ? 0u : (cedge ? 2u : 1u);\n");
    gen_log_special("SA4 && special", "SN4", "_4");
    out("    }\n");
    if (send_mode()) out("    u32 OFF[5]; const u32 sadd = warp_compact_front<5>(KIND, OFF);\n");
    else out("    u32 OFF[5]; log_reserve<5>(logcur, (u32)p.log_region_cap, KIND, OFF);\n");
    out("    _Pragma(\"unroll\") for (int sgm = 0; sgm < 4; ++sgm) {\n      if (KIND[sgm] == 0u) continue;\n");
    if (send_mode()) {
      gen_send_store("KIND[sgm]", "OFF[sgm]", "SK[sgm]", "[sgm]");
    } else {
      const int EW = p.log_entry_words;
      out("      u64* e = region + (u64)OFF[sgm] * %d;\n", EW);
      if (EW == 2) {
        const int c = p.log_cells[0];
        out("      *reinterpret_cast<ulonglong2*>(e) = make_ulonglong2(SK[sgm], SV%d[sgm]);\n", p.cells[c].skip ? 0 : c);
      } else {
        out("      e[0] = SK[sgm];\n");
        for (int i = 0; i < p.log_n_cells; ++i) {
          const int c = p.log_cells[i];
          out("      e[%d] = SV%d[sgm];\n", 1 + i, p.cells[c].skip ? 0 : c);
        }
        for (int w = 1 + p.log_n_cells; w < EW; ++w) out("      e[%d] = 0ull;\n", w);
      }
    }
    out("    }\n");
    gen_log_store("KIND[4] != 0u", "OFF[4]", "SK4", "_4", "KIND[4]");
    if (send_mode()) gen_send_steps("sadd");
    out("    ck = nk; cn = nn; chave = nhave; cedge = nedge;\n");
    for (int i = 0; i < n_live; ++i) out("    cv%d = nv%d;\n", live_cells[i], live_cells[i]);
  }

  // After the warp's last block: its open run is logged.
  void gen_log_epilogue() {
    const bool own = owned_runs_ok();
    out("  if (chave) { // (warp-uniform)\n");
    out("    bool edge = true;\n");
    if (own) {
      // read on: the rows after the range that carry the open run's key (at most 31) belong to it; the next warp drops them
      out("    {\n");
      out("      const i64 r0e = blk_end * 128;\n");
      out("      const i64 avail = n - r0e;\n");
      out("      if (avail > 0) {\n");
      out("        const i64 r = r0e + lane;\n");
      out("        const bool in = lane < avail;\n");
      out("        const u64 k = in ? __ldg(reinterpret_cast<const u64*>(p.values[%d]) + r) : 0ull;\n", p.sorted_key_col);
      out("        const u32 eq = __ballot_sync(0xffffffffu, in && k == ck);\n");
      out("        const u32 take = min(32u, (u32)__ffs((int)~eq) - 1u);\n");
      out("        if (take < 32u) {\n");
      for (int i = 0; i < n_live; ++i) out("          u64 ex%d = cell_identity<%d>();\n", live_cells[i], log_kind(live_cells[i]));
      out("          if ((u32)lane < take) {\n");
      out("            const bool pass = true; bool raise = false;\n");
      bool declared[kMaxCols] = {false};
      for (int i = 0; i < n_live; ++i) {
        const DevCell& cell = p.cells[live_cells[i]];
        if (cell.input < 0) continue;
        const int c = p.outputs[cell.input].index;
        if (declared[c]) continue;
        declared[c] = true;
        if (p.cols[c].nulls) out("            const V c%d = {%s, !bit1(p.nulls[%d], r), false};\n", c, row_load(c, "r").c_str(), c);
        else out("            const V c%d = {%s, false, false};\n", c, row_load(c, "r").c_str());
      }
      for (int i = 0; i < n_live; ++i) {
        const int c = live_cells[i];
        gen_cell_input(c);
        out("            if (l%d) ex%d = x%d;\n", c, c, c);
      }
      out("            (void)raise;\n          }\n");
      out("          for (u32 d = 1; d < take; d <<= 1) {\n");
      for (int i = 0; i < n_live; ++i) {
        const int c = live_cells[i];
        out("            ex%d = cell_comb<%d>(ex%d, __shfl_down_sync(0xffffffffu, ex%d, d), accerr);\n", c, log_kind(c), c, c);
      }
      out("          }\n");
      for (int i = 0; i < n_live; ++i) {
        const int c = live_cells[i];
        out("          if (take > 0u) cv%d = cell_comb<%d>(cv%d, __shfl_sync(0xffffffffu, ex%d, 0), accerr);\n", c, log_kind(c), c, c);
      }
      out("          edge = cedge || (i64)take == avail; // complete, unless it began before the range or the batch ends here\n");
      out("        }\n      }\n    }\n");
    }
    out("    const bool special = cn || ck == VGF_EMPTY_KEY;\n");
    out("    u64 SK4 = ck; bool SN4 = cn;\n");
    for (int i = 0; i < n_live; ++i) out("    u64 SV%d_4 = cv%d;\n", live_cells[i], live_cells[i]);
    gen_log_special("lane == 0 && special", "SN4", "_4");
    out("    u32 K1[1] = {lane == 0 && !special ? (edge ? 2u : 1u) : 0u}, O1[1];\n");
    if (send_mode()) out("    const u32 sadd = warp_compact_front<1>(K1, O1);\n");
    else out("    log_reserve<1>(logcur, (u32)p.log_region_cap, K1, O1);\n");
    gen_log_store("K1[0] != 0u", "O1[0]", "SK4", "_4", "K1[0]");
    if (send_mode()) gen_send_steps("sadd");
    out("    (void)SN4;\n  }\n");
  }

  void gen_fp_body(bool full) {
    gen_block_header(full);
    out("    bool PASS[4];\n");
    for (int k = 0; k < p.n_outputs; ++k) {
      if (!p.outs[k].values) continue;
      out("    u64 O%d[4]; bool ON%d[4];\n", k, k);
    }
    gen_row_begin(full);
    for (int k = 0; k < p.n_outputs; ++k) {
      if (!p.outs[k].values) continue;
      const vgpu_operand& o = p.outputs[k];
      if (o.kind == VGPU_OPND_IMM || o.kind == VGPU_OPND_ACC) decline("output operand kind");
      if (o.type == VGPU_BOOLEAN) out("      { const B o = %s; raise = raise || (pass && o.e); O%d[j] = o.v; ON%d[j] = o.n; }\n", bval(o, 0).c_str(), k, k);
      else out("      { const V o = %s; raise = raise || (pass && o.e); O%d[j] = o.v; ON%d[j] = o.n; }\n", val(o, 0).c_str(), k, k);
    }
    out("      PASS[j] = pass;\n");
    gen_raise();
    out("    }\n");
    if (full) out("    const int nw = 4;\n");
    else out("    const int nw = (int)((n - blk * 128 + 31) >> 5);\n");
    out("    store_block_bits(p.sel_words, blk * 4, nw, PASS);\n");
    out("    { int cnt = (int)PASS[0] + (int)PASS[1] + (int)PASS[2] + (int)PASS[3];\n");
    out("      _Pragma(\"unroll\") for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);\n");
    out("      if (lane == 0 && cnt) atomicAdd(&p.tile_counts[blk >> 3], cnt); }\n");
    for (int k = 0; k < p.n_outputs; ++k) {
      if (!p.outs[k].values) continue;
      const vgpu_operand& o = p.outputs[k];
      const int t = o.type;
      if (t == VGPU_BOOLEAN) {
        out("    { bool m[4] = {O%d[0] != 0, O%d[1] != 0, O%d[2] != 0, O%d[3] != 0}; store_block_bits(reinterpret_cast<u32*>(p.out_values[%d]), blk * 4, nw, m); }\n", k, k, k, k, k);
      } else {
        if (t == VGPU_REAL) out("    _Pragma(\"unroll\") for (int j = 0; j < 4; ++j) O%d[j] = __float_as_uint((float)u2d(O%d[j]));\n", k, k);
        const int lg = t == VGPU_TINYINT ? 0 : t == VGPU_SMALLINT ? 1 : (t == VGPU_INTEGER || t == VGPU_REAL) ? 2 : 3;
        if (full && lg == 3) {
          out("    { uint4* d = reinterpret_cast<uint4*>(p.out_values[%d]);\n", k);
          out("      d[r0 >> 1] = make_uint4((u32)O%d[0], (u32)(O%d[0] >> 32), (u32)O%d[1], (u32)(O%d[1] >> 32));\n", k, k, k, k);
          out("      d[r1 >> 1] = make_uint4((u32)O%d[2], (u32)(O%d[2] >> 32), (u32)O%d[3], (u32)(O%d[3] >> 32)); }\n", k, k, k, k);
        } else if (full && lg == 2) {
          out("    { uint2* d = reinterpret_cast<uint2*>(p.out_values[%d]);\n", k);
          out("      d[r0 >> 1] = make_uint2((u32)O%d[0], (u32)O%d[1]); d[r1 >> 1] = make_uint2((u32)O%d[2], (u32)O%d[3]); }\n", k, k, k, k);
        } else {
          const char* ty = lg == 0 ? "unsigned char" : lg == 1 ? "unsigned short" : lg == 2 ? "u32" : "u64";
          out("    _Pragma(\"unroll\") for (int j = 0; j < 4; ++j) { const i64 r = (j < 2 ? r0 : r1) + (j & 1); if (r < n) reinterpret_cast<%s*>(p.out_values[%d])[r] = (%s)O%d[j]; }\n", ty, k, ty, k);
        }
      }
      if (p.outs[k].nulls) {
        out("    { bool m[4]; _Pragma(\"unroll\") for (int j = 0; j < 4; ++j) m[j] = !ON%d[j] && (((j < 2 ? r0 : r1) + (j & 1)) < n);\n", k);
        out("      store_block_bits(reinterpret_cast<u32*>(p.out_nulls[%d]), blk * 4, nw, m); }\n", k);
      }
    }
  }

  void generate() {
    // (VGPU_QUAD_ARRAY=1: tuning aid — the 4-consecutive-rows layout for the ARRAY / GLOBAL sinks as well)
    static const bool quad_array = getenv("VGPU_QUAD_ARRAY") && getenv("VGPU_QUAD_ARRAY")[0] == '1';
    quad = p.mode == kSinkLog || (quad_array && p.mode == kSinkArray);
    for (const auto& op : prog.ops) { mark(op.a); mark(op.b); mark(op.c); }
    for (const auto& o : prog.outputs) mark(o);
    mark(prog.filter);
    for (int c = 0; c < p.n_cols; ++c) {
      if (!col_used[c] || p.cols[c].encoding == VGPU_CONSTANT) continue;
      const DevCol& d = p.cols[c];
      const void* walk = d.encoding == VGPU_DICTIONARY ? (const void*)d.indices : d.values;
      const int lg = walk_log2(d);
      // element alignment is all a column needs (a sliced vector's base is rarely 16-byte aligned); bitmaps are read as 32-bit words
      if (lg > 0 && (reinterpret_cast<uintptr_t>(walk) & ((1u << lg) - 1u)) != 0) { decline("column not aligned to its element size"); return; }
      if (lg < 0 && (reinterpret_cast<uintptr_t>(walk) & 7u) != 0) { decline("unaligned bitmap"); return; }
      if (d.type == VGPU_VARCHAR && (reinterpret_cast<uintptr_t>(d.values) & 15u) != 0) { decline("VARCHAR values not 16-byte aligned"); return; }
      if (d.nulls && (reinterpret_cast<uintptr_t>(d.nulls) & 7u) != 0) { decline("unaligned nulls"); return; }
    }
    if (p.in_rows && p.mode == kSinkFilterProject) { decline("row selection with the FilterProject sink"); return; }
    n_live = 0;
    if (p.mode == kSinkLog) {
      // only what an entry stores: its own cells, plus the rows cell when a stored count cell aliases it in this launch
      bool need[kMaxCells] = {false};
      for (int i = 0; i < p.log_n_cells; ++i) need[p.cells[p.log_cells[i]].skip ? 0 : p.log_cells[i]] = true;
      for (int c = 0; c < p.n_cells; ++c) if (need[c]) live_cells[n_live++] = c;
      if (p.log_n_cells < 1 || p.log_entry_words < 1 + p.log_n_cells) { decline("log entry layout"); return; }
    } else {
      for (int c = 0; c < p.n_cells; ++c) if (!p.cells[c].skip) live_cells[n_live++] = c;
    }

    const int G = p.n_groups;
    int copies = 1, split = 1;
    // Registers the body needs: the 4 rows of every staged column, addressing, and (register
    // strategy) the accumulators.  Pick the widest CTA shape whose per-thread budget covers it.
    int need = 24;
    for (int c = 0; c < p.n_cols; ++c) {
      if (!col_used[c] || p.cols[c].encoding == VGPU_CONSTANT) continue;
      const DevCol& d = p.cols[c];
      need += (d.encoding == VGPU_DICTIONARY || d.type == VGPU_INTEGER || d.type == VGPU_REAL) ? 4 : (d.type == VGPU_BIGINT || d.type == VGPU_DOUBLE) ? 8 : 2;
      if (d.encoding == VGPU_DICTIONARY) need += 4;
    }
    if (p.mode == kSinkHash) need += 8 + 3 * n_live * 4;
    if (p.mode == kSinkLog) need += 12 + 4 * n_live * 4; // + the open run and its slot (64 registers spill; measured: 384 x 2 beats 512 x 2 with spills)
    int acc_regs = 0, group_bytes = 0; // per thread: accumulator registers / bytes of one thread table
    for (int i = 0; i < n_live; ++i) {
      const bool cnt = p.cells[live_cells[i]].kind == kCellCount;
      acc_regs += (cnt ? 1 : 2) * G;
      group_bytes += (cnt ? 4 : 8) * G;
    }
    static const int shapes[][2] = {{512, 2}, {384, 2}, {512, 1}, {384, 1}, {256, 1}};
    auto pick = [&](int regs_needed, int smem_per_thread, int& nt, int& ctas) {
      for (const auto& sh : shapes) {
        const int budget = std::min(255, (65536 / (sh[0] * sh[1])) & ~7);
        const int64_t smem = (int64_t)smem_per_thread * sh[0] + 1024;
        if (budget >= regs_needed && smem * sh[1] <= 226 * 1024) { nt = sh[0]; ctas = sh[1]; return true; }
      }
      return false;
    };
    ArrayStrategy st = AS_REGS;
    int NT = 256, CTAS = 1;
    if (p.mode != kSinkArray) {
      pick(need, 0, NT, CTAS);
    } else if (acc_regs <= 24) {
      pick(need + acc_regs, 0, NT, CTAS);
    } else {
      int nt_r = 0, ctas_r = 0, nt_t = 0, ctas_t = 0;
      const bool can_regs = G * n_live <= 48 && pick(need + acc_regs, 0, nt_r, ctas_r);
      const bool can_thread = pick(need + 8, group_bytes, nt_t, ctas_t);
      // thread tables win when they keep more warps resident than the register file would
      if (can_thread && (!can_regs || nt_t * ctas_t > nt_r * ctas_r)) { st = AS_THREAD; NT = nt_t; CTAS = ctas_t; }
      else if (can_regs) { st = AS_REGS; NT = nt_r; CTAS = ctas_r; }
      else {
        // shared tables with atomics: one wide CTA per SM gives the most warp-private copies
        st = AS_ATOMIC;
        if (need <= 64) { NT = 1024; CTAS = 1; } else pick(need + 8, 0, NT, CTAS);
      }
    }
    if (const char* e = getenv("VGPU_FUSED_SHAPE")) { // tuning aid: "threads,ctas_per_sm"
      int a = 0, b = 0;
      if (sscanf(e, "%d,%d", &a, &b) == 2 && a >= 64 && a <= 1024 && a % 32 == 0 && b >= 1 && b <= 8 && st != AS_THREAD) { NT = a; CTAS = b; }
    }
    plan.threads = NT;
    plan.smem_bytes = 0;
    if (send_mode()) {
      if (p.send_world > 8) { decline("fused exchange: more than 8 ranks"); return; }
      plan.smem_bytes = (NT / 32) * send_warp_words() * 8;
      if ((plan.smem_bytes + 1280) * CTAS > 227 * 1024) { decline("fused exchange: staging area exceeds shared memory"); return; }
    }
    out("#define NT %d\n", NT);
    out("using namespace vgf;\n");
    if (p.mode == kSinkArray && st == AS_ATOMIC) {
      const int64_t one = (int64_t)G * n_live * 8;
      const int64_t limit = 200 * 1024 / CTAS;
      if (one > limit) { decline("group table exceeds shared memory"); return; }
      copies = one * (NT / 32) <= limit ? NT / 32 : 1;
      // VGPU_ATOMIC_SPLIT=2|4 (tuning aid): several copies per warp, lanes taking them in turn, so that fewer lanes of a
      // warp meet in one cell or bank.  Measured on C4: 0.294 -> 0.325 ms with 2 copies (the larger carve-out costs more
      // than the conflicts did), so the default stays one copy per warp.
      split = 1;
      if (copies > 1) {
        int want = getenv("VGPU_ATOMIC_SPLIT") ? atoi(getenv("VGPU_ATOMIC_SPLIT")) : 1;
        for (split = 4; split > 1; split >>= 1) if (split <= want && one * copies * split <= limit) break;
        copies *= split;
      }
      plan.smem_bytes = (int)(one * copies);
    }
    if (p.mode == kSinkArray && st == AS_THREAD) {
      int off = 0;
      for (int i = 0; i < n_live; ++i) {
        cell_boff[i] = off;
        off += (p.cells[live_cells[i]].kind == kCellCount ? 4 : 8) * G * NT;
        off = (off + 15) & ~15;
      }
      plan.smem_bytes = off;
    }
    out("extern \"C\" __global__ void __launch_bounds__(NT, %d) vgpu_fused(const FusedParams p) {\n", CTAS);
    out("  const int lane = threadIdx.x & 31;\n  const i64 n = p.n_rows;\n");
    out("  const i64 n_blocks = (n + 127) >> 7, n_full = n >> 7;\n");
    out("  const i64 wid = (i64)blockIdx.x * (NT / 32) + (threadIdx.x >> 5), nwarps = (i64)gridDim.x * (NT / 32);\n");
    out("  u32 accerr = 0;\n");
    gen_small_dict_registers();
    if (p.mode == kSinkHash) out("  u32 inserted = 0;\n");
    if (p.mode == kSinkLog) {
      if (send_mode()) {
        const int W = p.send_world, words = send_warp_words(), EW = p.log_entry_words;
        out("  extern __shared__ __align__(16) u64 sendstage[];\n");
        out("  __shared__ u32 sfront[%d], sback[%d];\n  __shared__ u64* sbase[%d];\n", W, W, W);
        out("  const u32 scap = (u32)p.send_sub_cap;\n");
        out("  if (threadIdx.x < %d) {\n    const unsigned long long c = p.send_counts[blockIdx.x * %d + threadIdx.x];\n", W, W);
        out("    sfront[threadIdx.x] = (u32)c; sback[threadIdx.x] = (u32)(c >> 32);\n");
        out("    sbase[threadIdx.x] = p.send_base[threadIdx.x] + (u64)blockIdx.x * p.send_sub_cap * %d;\n  }\n  __syncthreads();\n", EW);
        out("  u64* const wl = sendstage + (threadIdx.x >> 5) * %d;\n", words);
        out("  u32* const wcnt = reinterpret_cast<u32*>(wl + %d);\n", kSendStage * EW);
        out("  u64* const wbuf = wl + %d;\n", kSendStage * EW + 4);
        out("  if ((threadIdx.x & 31) < 8) wcnt[threadIdx.x & 31] = 0u;\n  __syncwarp();\n");
        out("  u32 sbeg = 0u, spend = 0u; // the warp's ring: first pending entry, pending entries\n");
      } else {
        out("  __shared__ u32 logcur[2];\n  if (threadIdx.x == 0) { logcur[0] = 0u; logcur[1] = 0u; }\n  __syncthreads();\n");
        out("  u64* const region = p.log + (u64)blockIdx.x * p.log_region_cap * %d;\n", p.log_entry_words);
      }
      if (p.sorted_probe) out("  bool viol = false;\n");
    }
    if (p.mode == kSinkArray) {
      if (st == AS_REGS) {
        for (int i = 0; i < n_live; ++i) {
          const int c = live_cells[i];
          if (p.cells[c].kind == kCellCount) out("  u32 A%d[%d];\n  _Pragma(\"unroll\") for (int gg = 0; gg < %d; ++gg) A%d[gg] = 0u;\n", c, G, G, c);
          else out("  u64 A%d[%d];\n  _Pragma(\"unroll\") for (int gg = 0; gg < %d; ++gg) A%d[gg] = cell_identity<%d>();\n", c, G, G, c, p.cells[c].kind);
        }
      } else if (st == AS_THREAD) {
        out("  extern __shared__ __align__(16) char tabc[];\n");
        for (int i = 0; i < n_live; ++i) {
          const int kind = p.cells[live_cells[i]].kind;
          if (kind == kCellCount) out("  for (int gg = 0; gg < %d; ++gg) (reinterpret_cast<u32*>(tabc + %d))[gg * NT + threadIdx.x] = 0u;\n", G, cell_boff[i]);
          else out("  for (int gg = 0; gg < %d; ++gg) (reinterpret_cast<u64*>(tabc + %d))[gg * NT + threadIdx.x] = cell_identity<%d>();\n", G, cell_boff[i], kind);
        }
      } else {
        out("  extern __shared__ __align__(16) u64 tab[];\n");
        out("  for (int i = threadIdx.x; i < %d; i += NT) {\n    const int ci = (i / %d) %% %d;\n    u64 idv = 0;\n", copies * n_live * G, G, n_live);
        for (int i = 0; i < n_live; ++i) out("    if (ci == %d) idv = cell_identity<%d>();\n", i, p.cells[live_cells[i]].kind);
        out("    tab[i] = idv;\n  }\n  __syncthreads();\n");
        if (copies > 1 && split > 1) out("  u64* mytab = tab + ((threadIdx.x >> 5) * %d + (threadIdx.x & %d)) * %d;\n", split, split - 1, n_live * G);
        else if (copies > 1) out("  u64* mytab = tab + (threadIdx.x >> 5) * %d;\n", n_live * G);
        else out("  u64* mytab = tab;\n");
      }
    }
    if (p.mode == kSinkLog) {
      gen_log_prologue();
      out("  for (i64 blk = blk_begin; blk < blk_end; ++blk) {\n   if (blk < n_full) {\n");
    } else out("  for (i64 blk = wid; blk < n_blocks; blk += nwarps) {\n   if (blk < n_full) {\n");
    if (p.mode == kSinkArray) gen_array_body(true, st, G); else if (p.mode == kSinkHash) gen_hash_body(true); else if (p.mode == kSinkLog) gen_log_body(true); else gen_fp_body(true);
    out("   } else {\n");
    if (p.mode == kSinkArray) gen_array_body(false, st, G); else if (p.mode == kSinkHash) gen_hash_body(false); else if (p.mode == kSinkLog) gen_log_body(false); else gen_fp_body(false);
    out("   }\n  }\n");
    if (p.mode == kSinkArray) {
      if (st == AS_REGS || st == AS_THREAD) {
        out("  __shared__ u64 scratch[NT / 32];\n");
        for (int i = 0; i < n_live; ++i) {
          const int c = live_cells[i];
          const int kind = p.cells[c].kind;
          if (st == AS_REGS) {
            out("  _Pragma(\"unroll\") for (int gg = 0; gg < %d; ++gg) block_fold<%d, NT / 32>((u64)A%d[gg], scratch, p.partials + ((size_t)blockIdx.x * %d + %d) * %d + gg, accerr);\n", G, kind, c, p.n_cells, c, G);
          } else {
            out("  for (int gg = 0; gg < %d; ++gg) block_fold<%d, NT / 32>((u64)(reinterpret_cast<%s*>(tabc + %d))[gg * NT + threadIdx.x], scratch, p.partials + ((size_t)blockIdx.x * %d + %d) * %d + gg, accerr);\n",
                G, kind, kind == kCellCount ? "u32" : "u64", cell_boff[i], p.n_cells, c, G);
          }
        }
      } else {
        out("  __syncthreads();\n  for (int i = threadIdx.x; i < %d; i += NT) {\n    const int ci = i / %d, gg = i - ci * %d;\n    u64 acc = tab[i];\n", n_live * G, G, G);
        out("    for (int k = 1; k < %d; ++k) {\n      const u64 v = tab[k * %d + i];\n", copies, n_live * G);
        for (int i = 0; i < n_live; ++i) out("      if (ci == %d) acc = cell_comb<%d>(acc, v, accerr);\n", i, p.cells[live_cells[i]].kind);
        out("    }\n    int c = 0;\n");
        for (int i = 0; i < n_live; ++i) out("    if (ci == %d) c = %d;\n", i, live_cells[i]);
        out("    p.partials[((size_t)blockIdx.x * %d + c) * %d + gg] = acc;\n  }\n", p.n_cells, G);
      }
    }
    if (p.mode == kSinkLog) {
      gen_log_epilogue();
      if (send_mode()) {
        // the counts stay with this rank for the next launch and go to each owner's window beside the entries
        out("  send_drain<%s>(scap, sfront, sback, sbase, wl, wbuf, wcnt, sbeg, spend, accerr);\n", send_targs().c_str());
        // the counts stay with this rank for the next launch and go to each owner's window beside the entries
        out("  __syncthreads();\n  if (threadIdx.x < %d) {\n    const unsigned long long c = (unsigned long long)sfront[threadIdx.x] | ((unsigned long long)sback[threadIdx.x] << 32);\n", p.send_world);
        out("    p.send_counts[blockIdx.x * %d + threadIdx.x] = c;\n    p.send_peer_counts[threadIdx.x][blockIdx.x] = c;\n  }\n", p.send_world);
      } else
      out("  __syncthreads();\n  if (threadIdx.x == 0) { p.log_counts[2 * blockIdx.x] = logcur[0]; p.log_counts[2 * blockIdx.x + 1] = logcur[1]; }\n");
      if (p.sorted_probe) out("  if (__any_sync(0xffffffffu, viol) && lane == 0) atomicOr(p.sort_state, 1ull);\n");
    }
    if (p.mode == kSinkHash) {
      out("  _Pragma(\"unroll\") for (int o = 16; o > 0; o >>= 1) inserted += __shfl_xor_sync(0xffffffffu, inserted, o);\n");
      out("  if (lane == 0 && inserted) atomicAdd(p.n_groups_ctr, (u64)inserted);\n");
    }
    out("  if (accerr) atomicOr(p.err_word, accerr);\n}\n");
  }
};

} // namespace

bool fused_gen_source(const PipelineParams& p, FusedPlan& plan, std::string& source, std::string& why) {
  Gen g(p, plan);
  g.generate();
  if (!g.ok) { why = g.why; return false; }
  source.swap(g.s);
  return true;
}

} // namespace vg
