// The fused scan pipeline: stage column tiles -> run the micro-op stream -> sink (selection +
// projections, or group-by accumulation).  One persistent CTA per SM walks tiles of 1024 rows.
// Column tiles are fetched by the bulk-copy engine (cp.async.bulk, SASS UBLKCP) into a ring of
// shared-memory stages guarded by mbarriers, so the bytes in flight do not depend on occupancy.
// The 256 threads then run ONE stream of micro-ops over the staged tile, 4 rows per thread: the
// expression program, the filter, the grouping keys and the accumulator updates are all
// micro-ops whose operands were resolved to shared-memory offsets on the host (lower_mops), so a
// handler is straight-line code.  The running value stays in registers; only values used more
// than once go through shared-memory slots.
//
// Reference behaviour restated here (paths under the reference tree):
//   filter -> selection: velox/exec/FilterProject.cpp:125-191, velox/exec/OperatorUtils.cpp:187-210
//   expression semantics: velox/expression/Expr.cpp:371-441 (default-null), ConjunctExpr.cpp:24-174,
//     velox/functions/prestosql/{Arithmetic.h:45-86,Comparisons.h:43-185},
//     velox/common/base/CheckedArithmetic.h, velox/type/Conversions.h:176-374
//   decoded column access: velox/vector/DecodedVector.h:76-140
//   accumulation: velox/exec/GroupingSet.cpp:240-307,513-559 and the Presto accumulators cited in
//     include/vgpu.h.
#include <cuda_runtime.h>
#include <stdint.h>
#include <algorithm>

#include "internal.h"
#include "device_common.cuh"
#include "fused.h"

namespace vg {

// ------------------------------------------------------------------------------------------
// mbarrier + bulk copy (PTX; see the Blackwell guide: cp.async.bulk -> UBLKCP, expect_tx ->
// SYNCS.ARRIVE.TRANS64)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// Orders prior generic-proxy accesses to shared memory before later async-proxy (bulk copy) writes.
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ int elem_bytes_log2(uint8_t type) {
  switch (type) {
    case VGPU_TINYINT: return 0;
    case VGPU_SMALLINT: return 1;
    case VGPU_INTEGER:
    case VGPU_REAL: return 2;
    case VGPU_VARCHAR: return 4;
    default: return 3;
  }
}

// Bytes of the staged values tile of a column for a full tile.
__device__ __forceinline__ uint32_t col_tile_bytes(const DevCol& c) {
  if (c.elem_log2 == 255) return kTile / 8; // bit-packed BOOLEAN
  return static_cast<uint32_t>(kTile) << c.elem_log2;
}

// Thread 0: start the bulk copies of one FULL tile (row offset `row0`, multiple of kTile).
__device__ void issue_tile(const PipelineParams& p, char* stage, uint64_t* bar, int64_t row0) {
  uint32_t total = 0;
  for (int c = 0; c < p.n_cols; ++c) {
    const DevCol& col = p.cols[c];
    if (!col.bulk_ok) continue;
    if (col.stage_off >= 0) total += col_tile_bytes(col);
    if (col.nulls_off >= 0) total += kTile / 8;
  }
  if (total == 0) return;
  mbar_expect_tx(bar, total);
  for (int c = 0; c < p.n_cols; ++c) {
    const DevCol& col = p.cols[c];
    if (!col.bulk_ok) continue;
    if (col.stage_off >= 0) {
      const char* src;
      if (col.encoding == VGPU_DICTIONARY) {
        src = reinterpret_cast<const char*>(col.indices) + (row0 << 2);
      } else if (col.elem_log2 == 255) {
        src = reinterpret_cast<const char*>(col.values) + (row0 >> 3);
      } else {
        src = reinterpret_cast<const char*>(col.values) + (row0 << col.elem_log2);
      }
      bulk_g2s(stage + col.stage_off, src, col_tile_bytes(col), bar);
    }
    if (col.nulls_off >= 0) {
      bulk_g2s(stage + col.nulls_off, reinterpret_cast<const char*>(col.nulls) + (row0 >> 3),
               kTile / 8, bar);
    }
  }
}

__device__ __forceinline__ bool bit_at(const uint64_t* bits, int64_t i) {
  return (bits[i >> 6] >> (i & 63)) & 1ull;
}

// All threads: element-wise staging of one tile for the columns (or all columns when
// `everything`) that the bulk engine cannot fetch: partial last tile, row indirection through
// in_rows, or unaligned buffers.  Rows past `cnt` are zero-filled.
__device__ void stage_generic(const PipelineParams& p, char* stage, int64_t row0, int cnt,
                              bool everything) {
  const int tid = threadIdx.x;
  for (int c = 0; c < p.n_cols; ++c) {
    const DevCol& col = p.cols[c];
    if (col.bulk_ok && !everything) continue;
#pragma unroll 1
    for (int k = 0; k < kRowsPerThread; ++k) {
      const int j = tid + k * kThreads; // row within the tile; a warp covers 32 adjacent rows
      const bool in = j < cnt;
      int64_t src = 0;
      if (in) src = p.in_rows ? static_cast<int64_t>(p.in_rows[row0 + j]) : row0 + j;
      if (col.stage_off >= 0) {
        char* dst = stage + col.stage_off;
        if (col.encoding == VGPU_DICTIONARY) {
          reinterpret_cast<int32_t*>(dst)[j] = in ? col.indices[src] : 0;
        } else if (col.elem_log2 == 255) {
          const bool b = in && bit_at(reinterpret_cast<const uint64_t*>(col.values), src);
          const uint32_t w = __ballot_sync(0xffffffffu, b);
          if ((tid & 31) == 0) reinterpret_cast<uint32_t*>(dst)[j >> 5] = w;
        } else {
          switch (col.elem_log2) {
            case 0:
              reinterpret_cast<uint8_t*>(dst)[j] =
                  in ? reinterpret_cast<const uint8_t*>(col.values)[src] : 0;
              break;
            case 1:
              reinterpret_cast<uint16_t*>(dst)[j] =
                  in ? reinterpret_cast<const uint16_t*>(col.values)[src] : 0;
              break;
            case 2:
              reinterpret_cast<uint32_t*>(dst)[j] =
                  in ? reinterpret_cast<const uint32_t*>(col.values)[src] : 0;
              break;
            default:
              reinterpret_cast<uint64_t*>(dst)[j] =
                  in ? reinterpret_cast<const uint64_t*>(col.values)[src] : 0;
          }
        }
      }
      if (col.nulls_off >= 0) {
        const bool b = in && bit_at(col.nulls, src);
        const uint32_t w = __ballot_sync(0xffffffffu, b);
        if ((tid & 31) == 0) reinterpret_cast<uint32_t*>(stage + col.nulls_off)[j >> 5] = w;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// Interpreter state.  A thread owns rows pos(p) + {0,1}, p = 0,1 with
//   pos(p) = warp*128 + 64*p + 2*lane
// (the "warp-contiguous pair layout"): a warp covers 128 adjacent rows of the tile, every
// 8-byte operand is one conflict-free LDS.128 per pair, and the selection / null bits of a pair
// group are two ballots.  Row j of the thread is j = 2*p + e.
// Masks (4 bits, bit j = row j): n = value is NULL, e = evaluation raised an error.
// ------------------------------------------------------------------------------------------
struct Val {
  uint64_t v[4]; // int64 (integers, sign-extended) or double bits (REAL is widened to double)
  uint32_t n, e;
};
struct BoolVal {
  uint32_t v, n, e;
};

struct Ctx {
  const PipelineParams* p;
  const char* stage; // current stage (column operands are relative to it)
  char* smem;        // base of dynamic shared memory (slot operands are absolute)
  int pos0, pos1;    // first row of pair 0 / pair 1 inside the tile
  uint32_t valid;    // rows that exist
  uint32_t ecode;    // error bits raised by this thread's ops, one byte per row
  Val acc;
  BoolVal bacc;
  // sink state of the current tile
  uint32_t pass;     // rows that passed the filter (and ignore_null_keys)
  uint32_t raise;    // rows whose latched error must be reported
  uint32_t acc_err;  // checked-sum overflow
  uint32_t goff[4];  // ARRAY: element offset of the row's group inside a cell block
  uint32_t prev;     // ARRAY thread tables: 2 bits per row j: 0 = no earlier equal row, else 1 + i
  uint32_t store;    // ARRAY thread tables: rows that are the last of their group in this thread
  uint64_t key[4];   // HASH: normalized key
  uint32_t keynull;
  int64_t slot[4];   // HASH: table slot of the row's group
};

__device__ __forceinline__ bool is_fp(uint8_t t) { return t == VGPU_REAL || t == VGPU_DOUBLE; }

__device__ __forceinline__ uint64_t load_base(const DevCol& c, int64_t idx) {
  switch (elem_bytes_log2(c.type)) {
    case 0: return (uint64_t)(int64_t)reinterpret_cast<const int8_t*>(c.values)[idx];
    case 1: return (uint64_t)(int64_t)reinterpret_cast<const int16_t*>(c.values)[idx];
    case 2:
      if (c.type == VGPU_REAL) return d2u((double)reinterpret_cast<const float*>(c.values)[idx]);
      return (uint64_t)(int64_t)reinterpret_cast<const int32_t*>(c.values)[idx];
    default: return reinterpret_cast<const uint64_t*>(c.values)[idx];
  }
}

// Two bits (rows pos, pos+1) of a staged 1024-bit tile bitmap.
__device__ __forceinline__ uint32_t pair_bits(const char* tile_bits, int pos) {
  const uint32_t w = reinterpret_cast<const uint32_t*>(tile_bits)[pos >> 5];
  return (w >> (pos & 31)) & 3u;
}
__device__ __forceinline__ uint32_t quad_bits(const Ctx& cx, int off) {
  const char* b = cx.stage + off;
  return pair_bits(b, cx.pos0) | (pair_bits(b, cx.pos1) << 2);
}
__device__ __forceinline__ uint64_t pack64(uint32_t lo, uint32_t hi) { return (uint64_t)lo | ((uint64_t)hi << 32); }
__device__ __forceinline__ uint64_t sext32(uint32_t x) { return (uint64_t)(int64_t)(int32_t)x; }

// Reads a value operand for the thread's 4 rows.
__device__ __forceinline__ void fetch(const Ctx& cx, const XOp& x, Val& out) {
  out.e = 0;
  out.n = 0;
  switch (x.kind) {
    case XK_ACC:
      out = cx.acc;
      return;
    case XK_IMM:
#pragma unroll
      for (int j = 0; j < 4; ++j) out.v[j] = x.imm;
      return;
    case XK_SLOT: {
      const uint4* s = reinterpret_cast<const uint4*>(cx.smem + x.off);
      const uint4 q0 = s[cx.pos0 >> 1];
      const uint4 q1 = s[cx.pos1 >> 1];
      out.v[0] = pack64(q0.x, q0.y); out.v[1] = pack64(q0.z, q0.w);
      out.v[2] = pack64(q1.x, q1.y); out.v[3] = pack64(q1.z, q1.w);
      if (x.has_nulls) { // slot may carry nulls / errors
        const uint32_t m = reinterpret_cast<const uint32_t*>(cx.smem + x.noff)[threadIdx.x];
        out.n = m & 0xFu;
        out.e = (m >> 8) & 0xFu;
      }
      return;
    }
    case XK_COL8: {
      const uint4* s = reinterpret_cast<const uint4*>(cx.stage + x.off);
      const uint4 q0 = s[cx.pos0 >> 1];
      const uint4 q1 = s[cx.pos1 >> 1];
      out.v[0] = pack64(q0.x, q0.y); out.v[1] = pack64(q0.z, q0.w);
      out.v[2] = pack64(q1.x, q1.y); out.v[3] = pack64(q1.z, q1.w);
      break;
    }
    case XK_COL4I: {
      const uint2* s = reinterpret_cast<const uint2*>(cx.stage + x.off);
      const uint2 q0 = s[cx.pos0 >> 1];
      const uint2 q1 = s[cx.pos1 >> 1];
      out.v[0] = sext32(q0.x); out.v[1] = sext32(q0.y);
      out.v[2] = sext32(q1.x); out.v[3] = sext32(q1.y);
      break;
    }
    case XK_COL4F: {
      const float2* s = reinterpret_cast<const float2*>(cx.stage + x.off);
      const float2 q0 = s[cx.pos0 >> 1];
      const float2 q1 = s[cx.pos1 >> 1];
      out.v[0] = d2u((double)q0.x); out.v[1] = d2u((double)q0.y);
      out.v[2] = d2u((double)q1.x); out.v[3] = d2u((double)q1.y);
      break;
    }
    case XK_COL2: {
      const uint32_t* s = reinterpret_cast<const uint32_t*>(cx.stage + x.off);
      const uint32_t q0 = s[cx.pos0 >> 1], q1 = s[cx.pos1 >> 1];
      out.v[0] = (uint64_t)(int64_t)(int16_t)(q0 & 0xffffu); out.v[1] = (uint64_t)(int64_t)(int16_t)(q0 >> 16);
      out.v[2] = (uint64_t)(int64_t)(int16_t)(q1 & 0xffffu); out.v[3] = (uint64_t)(int64_t)(int16_t)(q1 >> 16);
      break;
    }
    case XK_COL1: {
      const uint16_t* s = reinterpret_cast<const uint16_t*>(cx.stage + x.off);
      const uint32_t q0 = s[cx.pos0 >> 1], q1 = s[cx.pos1 >> 1];
      out.v[0] = (uint64_t)(int64_t)(int8_t)(q0 & 0xffu); out.v[1] = (uint64_t)(int64_t)(int8_t)(q0 >> 8);
      out.v[2] = (uint64_t)(int64_t)(int8_t)(q1 & 0xffu); out.v[3] = (uint64_t)(int64_t)(int8_t)(q1 >> 8);
      break;
    }
    case XK_DICT: {
      const DevCol& c = cx.p->cols[x.col];
      const uint2* s = reinterpret_cast<const uint2*>(cx.stage + x.off);
      const uint2 i0 = s[cx.pos0 >> 1];
      const uint2 i1 = s[cx.pos1 >> 1];
      const int32_t idx[4] = {(int32_t)i0.x, (int32_t)i0.y, (int32_t)i1.x, (int32_t)i1.y};
      uint32_t notnull = x.has_nulls ? quad_bits(cx, x.noff) : 0xFu;
      const uint32_t live = cx.valid & notnull;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint64_t raw = 0;
        if ((live >> j) & 1u) {
          raw = load_base(c, idx[j]);
          if (c.base_nulls != nullptr && !bit_at(c.base_nulls, idx[j])) notnull &= ~(1u << j);
        }
        out.v[j] = raw;
      }
      out.n = (~notnull) & 0xFu;
      return;
    }
    default: // XK_NULL
#pragma unroll
      for (int j = 0; j < 4; ++j) out.v[j] = 0;
      out.n = 0xFu;
      return;
  }
  if (x.has_nulls) out.n = (~quad_bits(cx, x.noff)) & 0xFu;
}

__device__ __forceinline__ void store_slot(const Ctx& cx, const MOp& op, const Val& v) {
  uint4* s = reinterpret_cast<uint4*>(cx.smem + op.dst);
  s[cx.pos0 >> 1] = make_uint4((uint32_t)v.v[0], (uint32_t)(v.v[0] >> 32), (uint32_t)v.v[1],
                               (uint32_t)(v.v[1] >> 32));
  s[cx.pos1 >> 1] = make_uint4((uint32_t)v.v[2], (uint32_t)(v.v[2] >> 32), (uint32_t)v.v[3],
                               (uint32_t)(v.v[3] >> 32));
  if (op.dstn >= 0) reinterpret_cast<uint32_t*>(cx.smem + op.dstn)[threadIdx.x] = v.n | (v.e << 8);
}

// Reads a boolean operand.
__device__ __forceinline__ void fetchb(const Ctx& cx, const XOp& x, BoolVal& out) {
  out.e = 0;
  out.n = 0;
  switch (x.kind) {
    case XK_ACC:
      out = cx.bacc;
      return;
    case XK_BSLOT: {
      const uint32_t m = reinterpret_cast<const uint32_t*>(cx.smem + x.off)[threadIdx.x];
      out.v = m & 0xFu;
      out.n = (m >> 8) & 0xFu;
      out.e = (m >> 16) & 0xFu;
      return;
    }
    case XK_IMM:
      out.v = x.imm ? 0xFu : 0u;
      return;
    case XK_BCOL:
      out.v = quad_bits(cx, x.off);
      break;
    case XK_BDICT: {
      const DevCol& c = cx.p->cols[x.col];
      const uint2* s = reinterpret_cast<const uint2*>(cx.stage + x.off);
      const uint2 i0 = s[cx.pos0 >> 1];
      const uint2 i1 = s[cx.pos1 >> 1];
      const int32_t idx[4] = {(int32_t)i0.x, (int32_t)i0.y, (int32_t)i1.x, (int32_t)i1.y};
      uint32_t notnull = x.has_nulls ? quad_bits(cx, x.noff) : 0xFu;
      const uint32_t live = cx.valid & notnull;
      uint32_t v = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if ((live >> j) & 1u) {
          if (bit_at(reinterpret_cast<const uint64_t*>(c.values), idx[j])) v |= 1u << j;
          if (c.base_nulls != nullptr && !bit_at(c.base_nulls, idx[j])) notnull &= ~(1u << j);
        }
      }
      out.v = v;
      out.n = (~notnull) & 0xFu;
      return;
    }
    default: // XK_NULL
      out.v = 0;
      out.n = 0xFu;
      return;
  }
  if (x.has_nulls) out.n = (~quad_bits(cx, x.noff)) & 0xFu;
}

// ------------------------------------------------------------------------------------------
// Scalar kernels
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void int_limits(uint8_t type, int64_t& lo, int64_t& hi) {
  switch (type) {
    case VGPU_TINYINT: lo = -128; hi = 127; break;
    case VGPU_SMALLINT: lo = -32768; hi = 32767; break;
    case VGPU_INTEGER: lo = INT32_MIN; hi = INT32_MAX; break;
    default: lo = INT64_MIN; hi = INT64_MAX;
  }
}

// REAL arithmetic is done in double and rounded back to float: for + - * / the double-rounded
// result equals the float operation (53 >= 2*24+2 bits).
__device__ __forceinline__ double fp_round(uint8_t type, double x) {
  return type == VGPU_REAL ? (double)__double2float_rn(x) : x;
}

// Returns error bits for the row (0 = fine).  Integer ops are the checked ones of
// velox/common/base/CheckedArithmetic.h; FP ops never raise.
__device__ __forceinline__ uint32_t arith(uint8_t opcode, uint8_t type, uint64_t a, uint64_t b,
                                          uint64_t& r) {
  if (is_fp(type)) {
    const double x = u2d(a), y = u2d(b);
    double z;
    switch (opcode) {
      case VGPU_OP_ADD: z = __dadd_rn(x, y); break;
      case VGPU_OP_SUB: z = __dsub_rn(x, y); break;
      case VGPU_OP_MUL: z = __dmul_rn(x, y); break;
      case VGPU_OP_DIV: z = __ddiv_rn(x, y); break;
      case VGPU_OP_NEG: z = -x; break;
      default: // MOD (velox/functions/prestosql/ArithmeticImpl.h:102-110)
        z = (y == 0.0) ? __longlong_as_double(0x7ff8000000000000ll) : fmod(x, y);
    }
    r = d2u(fp_round(type, z));
    return 0;
  }
  const int64_t x = (int64_t)a, y = (int64_t)b;
  int64_t lo, hi;
  int_limits(type, lo, hi);
  int64_t z = 0;
  uint32_t err = 0;
  switch (opcode) {
    case VGPU_OP_ADD:
      z = (int64_t)((uint64_t)x + (uint64_t)y);
      if (((x ^ z) & (y ^ z)) < 0) err = kErrOverflow;
      break;
    case VGPU_OP_SUB:
      z = (int64_t)((uint64_t)x - (uint64_t)y);
      if (((x ^ y) & (x ^ z)) < 0) err = kErrOverflow;
      break;
    case VGPU_OP_MUL: {
      z = (int64_t)((uint64_t)x * (uint64_t)y);
      const int64_t h = __mul64hi(x, y);
      if (h != (z >> 63)) err = kErrOverflow;
      break;
    }
    case VGPU_OP_DIV:
      if (y == 0) {
        err = kErrDivZero;
      } else if (x == lo && y == -1) {
        err = kErrOverflow;
      } else {
        z = x / y;
      }
      break;
    case VGPU_OP_NEG:
      if (x == lo) err = kErrOverflow; else z = -x;
      break;
    default: // MOD
      if (y == 0) err = kErrDivZero; else z = (y == -1) ? 0 : x % y;
  }
  if (!err && (z < lo || z > hi)) err = kErrOverflow;
  r = (uint64_t)z;
  return err;
}

// CAST from -> to (velox/type/Conversions.h, non-truncating flavour).
__device__ __forceinline__ uint32_t cast_value(uint8_t from, uint8_t to, uint64_t a, uint64_t& r) {
  if (is_fp(from)) {
    const double x = u2d(a);
    if (is_fp(to)) {
      r = d2u(fp_round(to, x));
      return 0;
    }
    if (x != x) return kErrCastRange;
    const double y = round(x); // half away from zero, as std::round
    int64_t lo, hi;
    int_limits(to, lo, hi);
    // 2^63 is not representable in int64; compare in double space first.
    if (!(y >= -9223372036854775808.0 && y < 9223372036854775808.0)) return kErrCastRange;
    const int64_t z = (int64_t)y;
    if (z < lo || z > hi) return kErrCastRange;
    r = (uint64_t)z;
    return 0;
  }
  const int64_t x = (int64_t)a;
  if (is_fp(to)) {
    r = d2u(to == VGPU_REAL ? (double)(float)x : (double)x);
    return 0;
  }
  int64_t lo, hi;
  int_limits(to, lo, hi);
  if (x < lo || x > hi) return kErrCastRange;
  r = a;
  return 0;
}

__device__ __forceinline__ bool compare(uint8_t opcode, uint8_t type, uint64_t a, uint64_t b) {
  if (is_fp(type)) {
    const double x = u2d(a), y = u2d(b);
    switch (opcode) {
      case VGPU_OP_EQ: return x == y;
      case VGPU_OP_NE: return x != y;
      case VGPU_OP_LT: return x < y;
      case VGPU_OP_LE: return x <= y;
      case VGPU_OP_GT: return x > y;
      default: return x >= y;
    }
  }
  const int64_t x = (int64_t)a, y = (int64_t)b;
  switch (opcode) {
    case VGPU_OP_EQ: return x == y;
    case VGPU_OP_NE: return x != y;
    case VGPU_OP_LT: return x < y;
    case VGPU_OP_LE: return x <= y;
    case VGPU_OP_GT: return x > y;
    default: return x >= y;
  }
}


// 4-row comparison with the (uniform) opcode/type dispatch hoisted out of the row loop.
__device__ __forceinline__ uint32_t cmp4(uint8_t oc, uint8_t type, const uint64_t* a, const uint64_t* b) {
  uint32_t m = 0;
#define VG_CMP_ROWS(T, CONV, OP)                                           \
  _Pragma("unroll") for (int j = 0; j < 4; ++j) {                          \
    m |= ((CONV(a[j]) OP CONV(b[j])) ? 1u : 0u) << j;                      \
  }
#define VG_CMP_TYPE(T, CONV)                                               \
  switch (oc) {                                                            \
    case VGPU_OP_EQ: VG_CMP_ROWS(T, CONV, ==) break;                       \
    case VGPU_OP_NE: VG_CMP_ROWS(T, CONV, !=) break;                       \
    case VGPU_OP_LT: VG_CMP_ROWS(T, CONV, <) break;                        \
    case VGPU_OP_LE: VG_CMP_ROWS(T, CONV, <=) break;                       \
    case VGPU_OP_GT: VG_CMP_ROWS(T, CONV, >) break;                        \
    default: VG_CMP_ROWS(T, CONV, >=)                                      \
  }
  if (is_fp(type)) {
    VG_CMP_TYPE(double, u2d)
  } else {
    VG_CMP_TYPE(int64_t, (int64_t))
  }
#undef VG_CMP_TYPE
#undef VG_CMP_ROWS
  return m;
}

// ------------------------------------------------------------------------------------------
// Boolean helpers
// ------------------------------------------------------------------------------------------
// SQL three-valued AND / OR of two evaluated operands; an error or NULL in one side is dropped
// when the other side decides the row (velox/expression/ConjunctExpr.cpp:24-174).
__device__ __forceinline__ BoolVal and3(const BoolVal& a, const BoolVal& b) {
  const uint32_t okA = ~a.n & ~a.e, okB = ~b.n & ~b.e;
  const uint32_t decided = (okA & ~a.v) | (okB & ~b.v); // some conjunct is FALSE
  BoolVal r;
  r.v = okA & a.v & okB & b.v & 0xFu;
  r.e = ~decided & (a.e | b.e) & 0xFu;
  r.n = ~decided & ~r.e & (a.n | b.n) & 0xFu;
  return r;
}
__device__ __forceinline__ BoolVal or3(const BoolVal& a, const BoolVal& b) {
  const uint32_t okA = ~a.n & ~a.e, okB = ~b.n & ~b.e;
  const uint32_t decided = (okA & a.v) | (okB & b.v); // some disjunct is TRUE
  BoolVal r;
  r.v = decided & 0xFu;
  r.e = ~decided & (a.e | b.e) & 0xFu;
  r.n = ~decided & ~r.e & (a.n | b.n) & 0xFu;
  return r;
}

__device__ __forceinline__ uint32_t cmp4_i32(uint8_t oc, const int32_t a[4], int32_t k) {
  uint32_t m = 0;
  switch (oc) {
#define VG_ROWS(OP) _Pragma("unroll") for (int j = 0; j < 4; ++j) m |= (a[j] OP k ? 1u : 0u) << j;
    case VGPU_OP_EQ: VG_ROWS(==) break;
    case VGPU_OP_NE: VG_ROWS(!=) break;
    case VGPU_OP_LT: VG_ROWS(<) break;
    case VGPU_OP_LE: VG_ROWS(<=) break;
    case VGPU_OP_GT: VG_ROWS(>) break;
    default: VG_ROWS(>=)
#undef VG_ROWS
  }
  return m;
}

__device__ __forceinline__ uint32_t interleave16(uint32_t even, uint32_t odd) {
  // bit i of even -> bit 2i, bit i of odd -> bit 2i+1 (perfect-shuffle network)
  uint32_t x = (even & 0xffffu) | (odd << 16);
  uint32_t t;
  t = (x ^ (x >> 8)) & 0x0000ff00u; x ^= t ^ (t << 8);
  t = (x ^ (x >> 4)) & 0x00f000f0u; x ^= t ^ (t << 4);
  t = (x ^ (x >> 2)) & 0x0c0c0c0cu; x ^= t ^ (t << 2);
  t = (x ^ (x >> 1)) & 0x22222222u; x ^= t ^ (t << 1);
  return x;
}

// Writes the 4-bit-per-thread mask `m` of the tile as packed words at dst[word_base ...].
__device__ __forceinline__ void store_tile_bits(uint32_t* dst, int64_t word_base, int n_words, uint32_t m) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int pr = 0; pr < 2; ++pr) {
    const uint32_t even = __ballot_sync(0xffffffffu, (m >> (2 * pr)) & 1u);
    const uint32_t odd = __ballot_sync(0xffffffffu, (m >> (2 * pr + 1)) & 1u);
    if ((lane & 15) == 0) {
      const int h = lane >> 4;
      const int w = warp * 4 + pr * 2 + h;
      // words past the data but inside the last 64-bit word are written as zero
      if (w < ((n_words + 1) & ~1)) dst[word_base + w] = interleave16(even >> (16 * h), odd >> (16 * h));
    }
  }
}

// Error bits of the rows in `raise` (cx.ecode keeps one byte of bits per row).
__device__ __forceinline__ uint32_t raised_code(uint32_t ecode, uint32_t raise) {
  uint32_t code = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if ((raise >> j) & 1u) code |= (ecode >> (8 * j)) & 0xffu;
  }
  return code ? code : kErrOverflow;
}

// ------------------------------------------------------------------------------------------
// ARRAY sink: per-thread (or per-warp) partial tables in shared memory.
// Layout of a cell block: [group][copy] elements of 8 bytes (4 for counts); copies = 256
// (thread tables: a warp touches 32 adjacent elements, conflict-free) or 1 (warp tables).
// ------------------------------------------------------------------------------------------
template <int KIND>
__device__ __forceinline__ uint64_t cell_comb(uint64_t a, uint64_t b, uint32_t& err) {
  if (KIND == kCellSumF64) return d2u(__dadd_rn(u2d(a), u2d(b)));
  if (KIND == kCellSumI64) {
    const int64_t x = (int64_t)a, y = (int64_t)b, z = (int64_t)(a + b);
    if (((x ^ z) & (y ^ z)) < 0) err |= kErrOverflow;
    return (uint64_t)z;
  }
  if (KIND == kCellCount) return a + b;
  if (KIND == kCellMinI64) return (int64_t)b < (int64_t)a ? b : a;
  if (KIND == kCellMaxI64) return (int64_t)a < (int64_t)b ? b : a;
  if (KIND == kCellMinF64) return u2d(b) < u2d(a) ? b : a;
  return u2d(a) < u2d(b) ? b : a;
}

// Thread-private tables.  The thread first folds rows of the same group together in registers
// (cx.prev names the previous row of the same group), then does one read-modify-write per
// distinct group with all loads issued before the stores: no two of those touch one address.
template <int KIND>
__device__ __forceinline__ void accum_thread(Ctx& cx, char* cbase, const uint64_t v[4], uint32_t live) {
  const uint64_t idv = cell_identity(KIND);
  uint64_t x[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const bool on = (live >> j) & 1u;
    if (KIND == kCellCount) x[j] = on ? 1u : 0u;
    else if (KIND == kCellMinF64 || KIND == kCellMaxF64) x[j] = (on && u2d(v[j]) == u2d(v[j])) ? v[j] : idv;
    else x[j] = on ? v[j] : idv;
  }
  const uint32_t p1 = (cx.prev >> 2) & 3u, p2 = (cx.prev >> 4) & 3u, p3 = (cx.prev >> 6) & 3u;
  x[1] = cell_comb<KIND>(p1 ? x[0] : idv, x[1], cx.acc_err);
  x[2] = cell_comb<KIND>(p2 == 2 ? x[1] : (p2 == 1 ? x[0] : idv), x[2], cx.acc_err);
  x[3] = cell_comb<KIND>(p3 == 3 ? x[2] : (p3 == 2 ? x[1] : (p3 == 1 ? x[0] : idv)), x[3], cx.acc_err);
  if (KIND == kCellCount) {
    uint32_t* t = reinterpret_cast<uint32_t*>(cbase);
    uint32_t old[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) old[j] = ((cx.store >> j) & 1u) ? t[cx.goff[j]] : 0u;
#pragma unroll
    for (int j = 0; j < 4; ++j) if ((cx.store >> j) & 1u) t[cx.goff[j]] = old[j] + (uint32_t)x[j];
  } else {
    uint64_t* t = reinterpret_cast<uint64_t*>(cbase);
    uint64_t old[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) old[j] = ((cx.store >> j) & 1u) ? t[cx.goff[j]] : idv;
#pragma unroll
    for (int j = 0; j < 4; ++j) if ((cx.store >> j) & 1u) t[cx.goff[j]] = cell_comb<KIND>(old[j], x[j], cx.acc_err);
  }
}

// Warp-private tables (more groups than per-thread tables can hold): shared-memory atomics.
// MIN/MAX read first and only go atomic when the row improves the cell, which is rare once the
// cell has seen a few rows.
template <int KIND>
__device__ __forceinline__ void accum_warp(Ctx& cx, char* cbase, const uint64_t v[4], uint32_t live) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (!((live >> j) & 1u)) continue;
    if (KIND == kCellCount) {
      atomicAdd(reinterpret_cast<uint32_t*>(cbase) + cx.goff[j], 1u);
      continue;
    }
    unsigned long long* a = reinterpret_cast<unsigned long long*>(cbase) + cx.goff[j];
    if (KIND == kCellSumF64) {
      atomicAdd(reinterpret_cast<double*>(a), u2d(v[j]));
    } else if (KIND == kCellSumI64) {
      const int64_t y = (int64_t)v[j];
      const int64_t x = (int64_t)atomicAdd(a, (unsigned long long)v[j]);
      const int64_t z = (int64_t)((uint64_t)x + (uint64_t)y);
      if (((x ^ z) & (y ^ z)) < 0) cx.acc_err |= kErrOverflow;
    } else if (KIND == kCellMinI64) {
      if ((int64_t)v[j] < (int64_t)*reinterpret_cast<volatile unsigned long long*>(a)) atomicMin(reinterpret_cast<long long*>(a), (long long)v[j]);
    } else if (KIND == kCellMaxI64) {
      if ((int64_t)*reinterpret_cast<volatile unsigned long long*>(a) < (int64_t)v[j]) atomicMax(reinterpret_cast<long long*>(a), (long long)v[j]);
    } else {
      unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(a);
      while (KIND == kCellMinF64 ? u2d(v[j]) < u2d(cur) : u2d(cur) < u2d(v[j])) {
        const unsigned long long seen = atomicCAS(a, cur, (unsigned long long)v[j]);
        if (seen == cur) break;
        cur = seen;
      }
    }
  }
}

template <int KIND>
__device__ __forceinline__ void accum_array(Ctx& cx, char* cbase, const uint64_t v[4], uint32_t live) {
  if (cx.p->warp_private) accum_warp<KIND>(cx, cbase, v, live);
  else accum_thread<KIND>(cx, cbase, v, live);
}

// ------------------------------------------------------------------------------------------
// The interpreter: one pass over the micro-op stream for the thread's 4 rows.
// ------------------------------------------------------------------------------------------
template <int MODE>
__device__ __forceinline__ void run_mops(Ctx& cx, char* tab, int64_t tile, int cnt) {
  const PipelineParams& p = *cx.p;
#pragma unroll 1
  for (int i = 0; i < p.n_mops; ++i) {
    const MOp& op = p.mops[i];
    switch (op.h) {
      case H_LD: {
        fetch(cx, op.x, cx.acc);
        if (op.flags & MF_STORE) store_slot(cx, op, cx.acc);
        break;
      }
      case H_ARITH: {
        Val a, r;
        fetch(cx, op.x, a);
        const uint8_t oc = op.sub, type = op.type;
        if (oc == VGPU_OP_NEG) {
          r.n = a.n; r.e = a.e;
          uint32_t bad = 0;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            r.v[j] = 0;
            const uint32_t err = arith(oc, type, a.v[j], 0, r.v[j]);
            if (err && !((a.n >> j) & 1u) && ((cx.valid >> j) & 1u)) { bad |= 1u << j; cx.ecode |= err << (8 * j); }
          }
          r.e |= bad;
        } else {
          Val b;
          fetch(cx, op.y, b);
          r.n = a.n | b.n;
          r.e = a.e | (~a.n & b.e); // a row whose first argument is NULL never evaluates the second (Expr.cpp:371-441)
          if (type == VGPU_DOUBLE && oc <= VGPU_OP_MUL) {
            // hot path of Q1/Q6: plain IEEE double ops, dispatch hoisted out of the row loop
            if (oc == VGPU_OP_MUL) {
#pragma unroll
              for (int j = 0; j < 4; ++j) r.v[j] = d2u(__dmul_rn(u2d(a.v[j]), u2d(b.v[j])));
            } else if (oc == VGPU_OP_ADD) {
#pragma unroll
              for (int j = 0; j < 4; ++j) r.v[j] = d2u(__dadd_rn(u2d(a.v[j]), u2d(b.v[j])));
            } else {
#pragma unroll
              for (int j = 0; j < 4; ++j) r.v[j] = d2u(__dsub_rn(u2d(a.v[j]), u2d(b.v[j])));
            }
          } else {
            uint32_t bad = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              r.v[j] = 0;
              const uint32_t err = arith(oc, type, a.v[j], b.v[j], r.v[j]);
              if (err && !((r.n >> j) & 1u) && ((cx.valid >> j) & 1u)) { bad |= 1u << j; cx.ecode |= err << (8 * j); }
            }
            r.e |= bad;
          }
        }
        cx.acc = r;
        if (op.flags & MF_STORE) store_slot(cx, op, r);
        break;
      }
      case H_CAST: {
        Val a, r;
        fetch(cx, op.x, a);
        r.n = a.n; r.e = a.e;
        uint32_t bad = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          r.v[j] = 0;
          const uint32_t err = cast_value(op.x.type, op.type, a.v[j], r.v[j]);
          if (err && !((a.n >> j) & 1u) && ((cx.valid >> j) & 1u)) { bad |= 1u << j; cx.ecode |= err << (8 * j); }
        }
        r.e |= bad;
        cx.acc = r;
        if (op.flags & MF_STORE) store_slot(cx, op, r);
        break;
      }
      case H_CMP: {
        BoolVal r;
        const uint8_t oc = op.sub;
        if (op.type == VGPU_BOOLEAN) { // boolean = / <> boolean
          BoolVal a, b;
          fetchb(cx, op.x, a);
          fetchb(cx, op.y, b);
          const uint32_t x = a.v ^ b.v;
          r.v = (oc == VGPU_OP_EQ ? ~x : x) & 0xFu;
          r.n = a.n | b.n;
          r.e = a.e | (~a.n & b.e); // a row whose first argument is NULL never evaluates the second (Expr.cpp:371-441)
        } else if (op.x.kind == XK_COL4I && op.y.kind == XK_IMM && (oc != VGPU_OP_BETWEEN || op.z.kind == XK_IMM)) {
          // INTEGER / DATE column against literals: 32-bit compares straight off the staged tile
          const uint2* s = reinterpret_cast<const uint2*>(cx.stage + op.x.off);
          const uint2 q0 = s[cx.pos0 >> 1];
          const uint2 q1 = s[cx.pos1 >> 1];
          const int32_t a[4] = {(int32_t)q0.x, (int32_t)q0.y, (int32_t)q1.x, (int32_t)q1.y};
          if (oc == VGPU_OP_BETWEEN) r.v = cmp4_i32(VGPU_OP_GE, a, (int32_t)op.y.imm) & cmp4_i32(VGPU_OP_LE, a, (int32_t)op.z.imm);
          else r.v = cmp4_i32(oc, a, (int32_t)op.y.imm);
          r.n = op.x.has_nulls ? (~quad_bits(cx, op.x.noff)) & 0xFu : 0u;
          r.e = 0;
        } else {
          Val a, b;
          fetch(cx, op.x, a);
          fetch(cx, op.y, b);
          r.n = a.n | b.n;
          r.e = a.e | (~a.n & b.e); // a row whose first argument is NULL never evaluates the second (Expr.cpp:371-441)
          if (oc == VGPU_OP_BETWEEN) {
            Val c;
            fetch(cx, op.z, c);
            r.e |= ~r.n & c.e;
            r.n |= c.n;
            r.v = cmp4(VGPU_OP_GE, op.type, a.v, b.v) & cmp4(VGPU_OP_LE, op.type, a.v, c.v);
          } else {
            r.v = cmp4(oc, op.type, a.v, b.v);
          }
        }
        r.v &= ~r.n;
        if (op.flags & MF_COMB_AND) r = and3(cx.bacc, r);
        else if (op.flags & MF_COMB_OR) r = or3(cx.bacc, r);
        cx.bacc = r;
        if (op.flags & MF_STORE) reinterpret_cast<uint32_t*>(cx.smem + op.dst)[threadIdx.x] = r.v | (r.n << 8) | (r.e << 16);
        break;
      }
      case H_BOOL: {
        BoolVal r;
        const uint8_t oc = op.sub;
        if (oc == VGPU_OP_IS_NULL) {
          if (op.x.type == VGPU_BOOLEAN) {
            BoolVal a;
            fetchb(cx, op.x, a);
            r.v = a.n; r.e = a.e;
          } else {
            Val a;
            fetch(cx, op.x, a);
            r.v = a.n; r.e = a.e;
          }
          r.n = 0;
        } else {
          BoolVal a;
          fetchb(cx, op.x, a);
          if (oc == VGPU_OP_NOT) {
            r.v = (~a.v) & 0xFu; r.n = a.n; r.e = a.e;
          } else {
            BoolVal b;
            fetchb(cx, op.y, b);
            r = (oc == VGPU_OP_AND) ? and3(a, b) : or3(a, b);
          }
        }
        cx.bacc = r;
        if (op.flags & MF_STORE) reinterpret_cast<uint32_t*>(cx.smem + op.dst)[threadIdx.x] = r.v | (r.n << 8) | (r.e << 16);
        break;
      }
      case H_SELECT: {
        Val b, c, r;
        uint32_t take; // rows that read the first value
        if (op.sub == VGPU_OP_SELECT) {
          BoolVal a;
          fetchb(cx, op.x, a);
          fetch(cx, op.y, b);
          fetch(cx, op.z, c);
          take = a.v & ~a.n & ~a.e & 0xFu;
          r.e = (a.e | (take & b.e) | (~take & ~a.e & c.e)) & 0xFu;
        } else {
          fetch(cx, op.x, b);
          fetch(cx, op.y, c);
          take = (~b.n | b.e) & 0xFu;
          r.e = (b.e | (~take & c.e)) & 0xFu;
        }
        r.n = ((take & b.n) | (~take & c.n)) & ~r.e & 0xFu;
#pragma unroll
        for (int j = 0; j < 4; ++j) r.v[j] = ((take >> j) & 1u) ? b.v[j] : c.v[j];
        cx.acc = r;
        if (op.flags & MF_STORE) store_slot(cx, op, r);
        break;
      }
      case H_FILTER: {
        if (op.x.kind == XK_NONE) {
          cx.pass = cx.valid;
        } else {
          BoolVal f;
          fetchb(cx, op.x, f);
          cx.raise |= f.e & cx.valid;
          cx.pass = f.v & ~f.n & ~f.e & cx.valid;
        }
        break;
      }
      case H_KEY: {
        const DevKey& ks = p.keys[op.dst];
        Val v;
        fetch(cx, op.x, v);
        cx.raise |= v.e & cx.pass;
        if (p.ignore_null_keys) cx.pass &= ~v.n;
        if (MODE == kSinkHash && ks.bits == 64) { // raw 64-bit key
#pragma unroll
          for (int j = 0; j < 4; ++j) cx.key[j] = v.v[j];
          cx.keynull = v.n;
          break;
        }
        uint32_t bad = 0;
        const uint64_t span = (uint64_t)(ks.max - ks.min);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint64_t u = v.v[j] - (uint64_t)ks.min;
          const bool isnull = (v.n >> j) & 1u;
          if (isnull ? !ks.nullable : u > span) bad |= 1u << j;
          const uint64_t id = isnull ? 0ull : u + ks.nullable; // 64-bit: a 33-bit field must not wrap into the NULL id
          if (MODE == kSinkHash) cx.key[j] += id << ks.shift;
          else cx.goff[j] += (uint32_t)id * (uint32_t)ks.mult;
        }
        if (bad & cx.pass) {
          cx.ecode |= kErrKeyRange * 0x01010101u;
          cx.raise |= bad & cx.pass;
          cx.pass &= ~bad;
        }
        break;
      }
      case H_KEYEND: {
        if (MODE == kSinkArray) {
          if (p.warp_private) {
#pragma unroll
            for (int j = 0; j < 4; ++j) cx.goff[j] = ((cx.pass >> j) & 1u) ? cx.goff[j] : 0u;
          } else {
            uint32_t* g = cx.goff;
#pragma unroll
            for (int j = 0; j < 4; ++j) g[j] = (((cx.pass >> j) & 1u) ? g[j] * kThreads : 0u) + threadIdx.x;
            const bool e01 = g[0] == g[1], e02 = g[0] == g[2], e03 = g[0] == g[3];
            const bool e12 = g[1] == g[2], e13 = g[1] == g[3], e23 = g[2] == g[3];
            const uint32_t p1 = e01 ? 1u : 0u;
            const uint32_t p2 = e12 ? 2u : (e02 ? 1u : 0u);
            const uint32_t p3 = e23 ? 3u : (e13 ? 2u : (e03 ? 1u : 0u));
            cx.prev = (p1 << 2) | (p2 << 4) | (p3 << 6);
            cx.store = (!(e01 | e02 | e03) ? 1u : 0u) | (!(e12 | e13) ? 2u : 0u) | (!e23 ? 4u : 0u) | 8u;
          }
        } else if (MODE == kSinkHash) {
          HashRef href;
          href.table = p.table; href.cap_mask = p.cap_mask; href.key_map = p.key_map; href.row_words = p.row_words; href.n_groups_ctr = p.n_groups_ctr;
          // Rows of a thread are adjacent in the input; a row reuses its left neighbour's slot
          // when the keys are equal (sorted / clustered keys such as l_orderkey).
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            cx.slot[j] = -1;
            if (!((cx.pass >> j) & 1u)) continue;
            const bool isnull = (cx.keynull >> j) & 1u;
            if ((j & 1) && ((cx.pass >> (j - 1)) & 1u) && cx.key[j] == cx.key[j - 1] &&
                isnull == (bool)((cx.keynull >> (j - 1)) & 1u)) {
              cx.slot[j] = cx.slot[j - 1];
            } else {
              cx.slot[j] = hash_find_or_insert(href, cx.key[j], isnull, cx.acc_err);
            }
            if (cx.slot[j] < 0) cx.pass &= ~(1u << j);
          }
        }
        break;
      }
      case H_ACCUM: {
        Val v;
        uint32_t live = cx.pass;
        if (op.flags & MF_NO_VALUE) {
#pragma unroll
          for (int j = 0; j < 4; ++j) v.v[j] = 0;
        } else if (op.x.type == VGPU_BOOLEAN) {
          BoolVal b;
          fetchb(cx, op.x, b);
          cx.raise |= b.e & cx.pass;
          live &= ~b.n;
#pragma unroll
          for (int j = 0; j < 4; ++j) v.v[j] = (b.v >> j) & 1u;
        } else {
          fetch(cx, op.x, v);
          cx.raise |= v.e & cx.pass;
          live &= ~v.n;
        }
        const uint8_t kind = op.sub;
        if (kind == kCellSumF64 && !is_fp(op.x.type)) {
#pragma unroll
          for (int j = 0; j < 4; ++j) v.v[j] = d2u((double)(int64_t)v.v[j]);
        }
        if (MODE == kSinkArray) {
          char* cbase = tab + op.dst + (p.warp_private ? (threadIdx.x >> 5) * p.warp_tab_bytes : 0);
          switch (kind) {
            case kCellSumF64: accum_array<kCellSumF64>(cx, cbase, v.v, live); break;
            case kCellCount: accum_array<kCellCount>(cx, cbase, v.v, live); break;
            case kCellSumI64: accum_array<kCellSumI64>(cx, cbase, v.v, live); break;
            case kCellMinI64: accum_array<kCellMinI64>(cx, cbase, v.v, live); break;
            case kCellMaxI64: accum_array<kCellMaxI64>(cx, cbase, v.v, live); break;
            case kCellMinF64: accum_array<kCellMinF64>(cx, cbase, v.v, live); break;
            default: accum_array<kCellMaxF64>(cx, cbase, v.v, live);
          }
        } else if (MODE == kSinkHash) {
          const int c = op.dst; // cell index inside the table row
#pragma unroll
          for (int pr = 0; pr < 2; ++pr) {
            const int j0 = 2 * pr, j1 = 2 * pr + 1;
            const bool a0 = (live >> j0) & 1u, a1 = (live >> j1) & 1u;
            uint64_t x0 = v.v[j0];
            const uint64_t x1 = v.v[j1];
            // pre-combine the pair when both rows fall in the same group (row order kept)
            if (a0 && a1 && cx.slot[j0] == cx.slot[j1] && kind == kCellSumF64) {
              x0 = d2u(__dadd_rn(u2d(x0), u2d(x1)));
              update_hbm_cell(kind, p.table + cx.slot[j0] * p.row_words + 1 + c, x0, cx.acc_err);
              continue;
            }
            if (a0) update_hbm_cell(kind, p.table + cx.slot[j0] * p.row_words + 1 + c, x0, cx.acc_err);
            if (a1) update_hbm_cell(kind, p.table + cx.slot[j1] * p.row_words + 1 + c, x1, cx.acc_err);
          }
        }
        break;
      }
      case H_SEL: {
        if (MODE == kSinkFilterProject) {
          store_tile_bits(p.sel_words, tile * (kTile / 32), (cnt + 31) >> 5, cx.pass);
          int wsum = __popc(cx.pass);
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, o);
          if ((threadIdx.x & 31) == 0 && wsum) atomicAdd(&p.tile_counts[tile], wsum);
        }
        break;
      }
      case H_PROJ: {
        if (MODE == kSinkFilterProject) {
          const DevOut out = p.outs[op.dst];
          const int64_t word_base = tile * (kTile / 32);
          const int n_words = (cnt + 31) >> 5;
          const bool full = cnt == kTile;
          uint32_t nmask;
          const int64_t r0 = tile * kTile + cx.pos0, r1 = tile * kTile + cx.pos1;
          if (op.x.type == VGPU_BOOLEAN) {
            BoolVal b;
            fetchb(cx, op.x, b);
            cx.raise |= b.e & cx.pass;
            nmask = b.n;
            store_tile_bits(reinterpret_cast<uint32_t*>(out.values), word_base, n_words, b.v);
          } else {
            Val v;
            fetch(cx, op.x, v);
            cx.raise |= v.e & cx.pass;
            nmask = v.n;
            const uint8_t t = op.x.type;
            const int lg = elem_bytes_log2(t);
            if (t == VGPU_REAL) {
#pragma unroll
              for (int j = 0; j < 4; ++j) v.v[j] = __float_as_uint((float)u2d(v.v[j]));
            }
            if (full && lg == 3) {
              uint4* d = reinterpret_cast<uint4*>(out.values);
              d[r0 >> 1] = make_uint4((uint32_t)v.v[0], (uint32_t)(v.v[0] >> 32), (uint32_t)v.v[1], (uint32_t)(v.v[1] >> 32));
              d[r1 >> 1] = make_uint4((uint32_t)v.v[2], (uint32_t)(v.v[2] >> 32), (uint32_t)v.v[3], (uint32_t)(v.v[3] >> 32));
            } else if (full && lg == 2) {
              uint2* d = reinterpret_cast<uint2*>(out.values);
              d[r0 >> 1] = make_uint2((uint32_t)v.v[0], (uint32_t)v.v[1]);
              d[r1 >> 1] = make_uint2((uint32_t)v.v[2], (uint32_t)v.v[3]);
            } else {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                if (!((cx.valid >> j) & 1u)) continue;
                const int64_t r = (j < 2 ? r0 : r1) + (j & 1);
                switch (lg) {
                  case 0: reinterpret_cast<uint8_t*>(out.values)[r] = (uint8_t)v.v[j]; break;
                  case 1: reinterpret_cast<uint16_t*>(out.values)[r] = (uint16_t)v.v[j]; break;
                  case 2: reinterpret_cast<uint32_t*>(out.values)[r] = (uint32_t)v.v[j]; break;
                  default: reinterpret_cast<uint64_t*>(out.values)[r] = v.v[j];
                }
              }
            }
          }
          if (out.nulls != nullptr) {
            store_tile_bits(reinterpret_cast<uint32_t*>(out.nulls), word_base, n_words, ~nmask & cx.valid & 0xFu);
          }
        }
        break;
      }
      default:
        break;
    }
  }
  if (cx.raise) atomicOr(p.err_word, raised_code(cx.ecode, cx.raise));
  if (cx.acc_err) atomicOr(p.err_word, cx.acc_err);
}

// ---- per-CTA fold of the partial tables (ARRAY mode), fixed order => run-to-run deterministic ---
__device__ void fold_tables(const PipelineParams& p, char* tab) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int G = p.n_groups;
  uint64_t* out = p.partials + (size_t)blockIdx.x * p.n_cells * G;
  uint32_t err = 0;
  if (p.warp_private) {
    for (int i = threadIdx.x; i < p.n_cells * G; i += kThreads) {
      const int c = i / G, g = i - c * G;
      const DevCell cell = p.cells[c];
      if (cell.skip) continue;
      const int w = cell_width(cell.kind);
      uint64_t acc = cell_identity(cell.kind);
      for (int k = 0; k < kWarps; ++k) {
        const char* a = tab + (size_t)k * p.warp_tab_bytes + cell.smem_off + (size_t)g * w;
        const uint64_t v = (w == 4) ? *reinterpret_cast<const uint32_t*>(a) : *reinterpret_cast<const uint64_t*>(a);
        acc = fold2(cell.kind, acc, v, err);
      }
      out[i] = acc;
    }
  } else {
    for (int i = warp; i < p.n_cells * G; i += kWarps) {
      const int c = i / G, g = i - c * G;
      const DevCell cell = p.cells[c];
      if (cell.skip) continue;
      const int w = cell_width(cell.kind);
      const char* row = tab + cell.smem_off + (size_t)g * kThreads * w;
      uint64_t acc = cell_identity(cell.kind);
#pragma unroll
      for (int k = 0; k < kThreads / 32; ++k) {
        const char* a = row + (size_t)(lane + 32 * k) * w;
        const uint64_t v = (w == 4) ? *reinterpret_cast<const uint32_t*>(a) : *reinterpret_cast<const uint64_t*>(a);
        acc = fold2(cell.kind, acc, v, err);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const uint64_t other = __shfl_down_sync(0xffffffffu, acc, o);
        acc = fold2(cell.kind, acc, other, err);
      }
      if (lane == 0) out[i] = acc;
    }
  }
  if (err) atomicOr(p.err_word, err);
}

// ------------------------------------------------------------------------------------------
// The kernel
// ------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(kThreads, 1)
pipeline_kernel(const __grid_constant__ PipelineParams p) {
  extern __shared__ __align__(128) char smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem); // kMaxStages mbarriers in the first 128 B
  char* stages = smem + 128;
  Ctx cx;
  cx.p = &p;
  cx.smem = smem;
  char* tab = smem + p.smem_tab_off;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  cx.pos0 = warp * 128 + 2 * lane;
  cx.pos1 = cx.pos0 + 64;
  cx.ecode = 0;

  const int64_t n_tiles = (p.n_rows + kTile - 1) / kTile;
  const int64_t n_full = p.in_rows ? 0 : p.n_rows / kTile; // tiles the bulk engine may fetch
  const int S = p.n_stages;

  if (tid == 0) {
    for (int s = 0; s < S; ++s) mbar_init(&bars[s], 1);
    fence_barrier_init();
  }
  if (MODE == kSinkArray) {
    // identity-fill the partial tables
    for (int c = 0; c < p.n_cells; ++c) {
      const DevCell cell = p.cells[c];
      if (cell.skip) continue;
      const int w = cell_width(cell.kind);
      const uint64_t idv = cell_identity(cell.kind);
      const int copies = p.warp_private ? 1 : kThreads;
      const int n = p.n_groups * copies;
      for (int k = 0; k < (p.warp_private ? kWarps : 1); ++k) {
        char* b = tab + (size_t)k * p.warp_tab_bytes + cell.smem_off;
        for (int i = tid; i < n; i += kThreads) {
          if (w == 4) reinterpret_cast<uint32_t*>(b)[i] = 0; else reinterpret_cast<uint64_t*>(b)[i] = idv;
        }
      }
    }
  }
  __syncthreads();

  // prologue: fill the ring
  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      const int64_t t = (int64_t)blockIdx.x + (int64_t)s * gridDim.x;
      if (t < n_full) issue_tile(p, stages + (size_t)s * p.stage_bytes, &bars[s], t * kTile);
    }
  }
  bool any_bulk = false, any_generic = false;
  for (int c = 0; c < p.n_cols; ++c) {
    const bool staged = p.cols[c].stage_off >= 0 || p.cols[c].nulls_off >= 0;
    any_bulk |= staged && p.cols[c].bulk_ok != 0;
    any_generic |= staged && p.cols[c].bulk_ok == 0;
  }

  int s = 0;
  uint32_t parity = 0;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    char* stage = stages + (size_t)s * p.stage_bytes;
    const int64_t row0 = tile * kTile;
    const int cnt = (int)min((int64_t)kTile, p.n_rows - row0);
    const bool bulk = tile < n_full;
    if (bulk) {
      if (any_bulk) mbar_wait(&bars[s], parity);
      if (any_generic) { // columns the bulk engine could not take
        stage_generic(p, stage, row0, cnt, false);
        __syncthreads();
      }
    } else {
      stage_generic(p, stage, row0, cnt, true);
      __syncthreads();
    }
    cx.stage = stage;
    cx.valid = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int pos = (j < 2 ? cx.pos0 : cx.pos1) + (j & 1);
      if (pos < cnt) cx.valid |= 1u << j;
    }
    cx.acc.n = cx.acc.e = 0;
    cx.bacc.v = cx.bacc.n = cx.bacc.e = 0;
    cx.pass = cx.valid;
    cx.raise = 0;
    cx.acc_err = 0;
    cx.prev = 0;
    cx.store = 0xFu;
    cx.keynull = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) { cx.goff[j] = 0; cx.key[j] = 0; cx.slot[j] = -1; }
    run_mops<MODE>(cx, tab, tile, cnt);
    __syncthreads(); // every thread is done reading this stage (and the slots)
    if (tid == 0) {
      const int64_t next = tile + (int64_t)S * gridDim.x;
      if (next < n_full) {
        fence_proxy_async();
        issue_tile(p, stage, &bars[s], next * kTile);
      }
    }
    if (++s == S) { s = 0; parity ^= 1u; }
  }
  if (MODE == kSinkArray) {
    __syncthreads();
    fold_tables(p, tab);
  }
}


// ---- selection bitmap -> ascending row numbers ------------------------------------------------
// tile_counts -> exclusive offsets (single CTA), then one CTA per tile expands its 32 words.
__global__ void scan_tile_counts(const int32_t* counts, int64_t n_tiles, int64_t* offsets,
                                 unsigned long long* total) {
  __shared__ int64_t warp_sums[32];
  __shared__ int64_t carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < n_tiles; base += blockDim.x) {
    const int64_t i = base + threadIdx.x;
    int64_t v = i < n_tiles ? counts[i] : 0;
    int64_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int64_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) warp_sums[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int64_t w = lane < (int)(blockDim.x >> 5) ? warp_sums[lane] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int64_t y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
      }
      warp_sums[lane] = w; // inclusive
    }
    __syncthreads();
    const int64_t before = carry + (warp ? warp_sums[warp - 1] : 0) + x - v;
    if (i < n_tiles) offsets[i] = before;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = (unsigned long long)carry;
}

__global__ void expand_selection(const uint32_t* words, const int64_t* offsets, int64_t n_rows,
                                 int32_t* sel_idx) {
  // one warp per tile: lane w owns word w of the tile
  const int lane = threadIdx.x & 31;
  const int64_t tile = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t n_tiles = (n_rows + kTile - 1) / kTile;
  if (tile >= n_tiles) return;
  const int64_t n_words = (n_rows + 31) >> 5;
  const int64_t wi = tile * 32 + lane;
  uint32_t w = wi < n_words ? words[wi] : 0u;
  const int64_t rem = n_rows - wi * 32;
  if (rem < 32) w &= rem <= 0 ? 0u : ((1u << rem) - 1u);
  int c = __popc(w), x = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  int64_t out = offsets[tile] + (x - c);
  while (w) {
    const int b = __ffs(w) - 1;
    w &= w - 1;
    sel_idx[out++] = (int32_t)(wi * 32 + b);
  }
}

// ------------------------------------------------------------------------------------------
// Host side: planning, lowering to micro-ops and launch
// ------------------------------------------------------------------------------------------
static int host_elem_log2(int type) {
  switch (type) {
    case VGPU_BOOLEAN: return 255;
    case VGPU_TINYINT: return 0;
    case VGPU_SMALLINT: return 1;
    case VGPU_INTEGER:
    case VGPU_REAL: return 2;
    case VGPU_VARCHAR: return 4;
    default: return 3;
  }
}
static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int pipeline_grid(int64_t n_rows) {
  const int64_t tiles = (n_rows + kTile - 1) / kTile;
  const int sms = num_sms();
  return (int)std::max<int64_t>(1, std::min<int64_t>(tiles, sms));
}

static void mark_used(const vgpu_operand& o, bool* used) {
  if (o.kind == VGPU_OPND_COL && o.index < kMaxCols) used[o.index] = true;
}

int plan_pipeline(const Program& prog, const vgpu_column* cols, int32_t ncols, int64_t nrows,
                  const int32_t* in_rows, int64_t n_in_rows, PipelineParams& p) {
  if (ncols < prog.n_cols) {
    set_error("program reads %d columns, %d bound", prog.n_cols, ncols);
    return VGPU_ERR_INVALID;
  }
  if (ncols > kMaxCols) {
    set_error("at most %d input columns per pipeline, got %d", kMaxCols, ncols);
    return VGPU_ERR_UNSUPPORTED;
  }
  p.prog = &prog;
  p.n_outputs = (int)prog.outputs.size();
  for (int i = 0; i < p.n_outputs; ++i) p.outputs[i] = prog.outputs[i];
  p.n_cols = ncols;
  p.n_slots = prog.n_slots;
  p.n_bslots = prog.n_bslots;
  p.in_rows = in_rows;
  p.n_rows = in_rows ? n_in_rows : nrows;
  p.err_word = fused_dry_run() ? nullptr : device_err_word();

  bool used[kMaxCols] = {false};
  for (const auto& op : prog.ops) { mark_used(op.a, used); mark_used(op.b, used); mark_used(op.c, used); }
  for (const auto& o : prog.outputs) mark_used(o, used);
  mark_used(prog.filter, used);

  int off = 0;
  for (int c = 0; c < ncols; ++c) {
    const vgpu_column& in = cols[c];
    DevCol& d = p.cols[c];
    d = DevCol{};
    d.values = in.values; d.indices = in.indices; d.nulls = in.nulls; d.base_nulls = in.base_nulls;
    d.constant = in.constant.i;
    d.type = (uint8_t)in.type; d.encoding = (uint8_t)in.encoding; d.const_null = (uint8_t)(in.constant_is_null != 0);
    d.stage_off = -1; d.nulls_off = -1; d.bulk_ok = 0;
    d.small_dict = (in.encoding == VGPU_DICTIONARY && in.type != VGPU_BOOLEAN && in.base_length >= 1 && in.base_length <= 32) ? (uint8_t)in.base_length : 0;
    if (in.type < VGPU_BOOLEAN || in.type > VGPU_VARCHAR) { set_error("column %d: bad type %d", c, in.type); return VGPU_ERR_INVALID; }
    d.elem_log2 = (uint8_t)(in.encoding == VGPU_DICTIONARY ? 2 : host_elem_log2(in.type));
    if (!used[c] || in.encoding == VGPU_CONSTANT) continue;
    if (in.encoding == VGPU_DICTIONARY) {
      if (!in.indices || !in.values) { set_error("column %d: dictionary without indices/values", c); return VGPU_ERR_INVALID; }
    } else if (in.encoding == VGPU_FLAT) {
      if (!in.values && nrows > 0) { set_error("column %d: flat column without values", c); return VGPU_ERR_INVALID; }
    } else { set_error("column %d: bad encoding %d", c, in.encoding); return VGPU_ERR_INVALID; }
    const int bytes = d.elem_log2 == 255 ? kTile / 8 : (kTile << d.elem_log2);
    d.stage_off = off; off += (bytes + 127) & ~127;
    const void* walk = in.encoding == VGPU_DICTIONARY ? (const void*)in.indices : in.values;
    bool ok = in_rows == nullptr && aligned16(walk);
    if (in.nulls) { d.nulls_off = off; off += 128; ok = ok && aligned16(in.nulls); }
    d.bulk_ok = ok ? 1 : 0;
  }
  p.stage_bytes = std::max(128, (off + 127) & ~127);
  return VGPU_OK;
}

// ---- nullability / error analysis of the ABI program ------------------------------------------
struct Flow {
  bool slot_n[kMaxSlots + 1] = {false}, bslot_n[kMaxSlots + 1] = {false};
  bool acc_n = false, bacc_n = false;
};
static bool col_nullable(const DevCol& c) {
  if (c.encoding == VGPU_CONSTANT) return c.const_null != 0;
  return c.nulls != nullptr || c.base_nulls != nullptr;
}
static bool opnd_dirty(const PipelineParams& p, const Flow& f, const vgpu_operand& o) {
  switch (o.kind) {
    case VGPU_OPND_ACC: return o.type == VGPU_BOOLEAN ? f.bacc_n : f.acc_n;
    case VGPU_OPND_SLOT: return o.type == VGPU_BOOLEAN ? f.bslot_n[std::min<int>(o.index, kMaxSlots)] : f.slot_n[std::min<int>(o.index, kMaxSlots)];
    case VGPU_OPND_COL: return col_nullable(p.cols[o.index]);
    case VGPU_OPND_NULL: return true;
    default: return false;
  }
}
static bool op_is_bool(int oc) { return oc >= VGPU_OP_EQ && oc <= VGPU_OP_IS_NULL; }
// with_errors: also count "may carry a latched error" (integer arithmetic, casts).
static void flow_step(const PipelineParams& p, Flow& f, const vgpu_op& op, bool with_errors) {
  bool n = false;
  if (op.opcode != VGPU_OP_IS_NULL || with_errors) {
    if (op.a.kind != VGPU_OPND_NONE) n |= opnd_dirty(p, f, op.a);
    if (op.b.kind != VGPU_OPND_NONE) n |= opnd_dirty(p, f, op.b);
    if (op.c.kind != VGPU_OPND_NONE) n |= opnd_dirty(p, f, op.c);
  }
  if (with_errors) {
    const bool fp = op.type == VGPU_REAL || op.type == VGPU_DOUBLE;
    if (op.opcode == VGPU_OP_CAST) n = true;
    if (op.opcode >= VGPU_OP_ADD && op.opcode <= VGPU_OP_MOD && !fp) n = true;
  }
  if (op_is_bool(op.opcode)) { f.bacc_n = n; if (op.dst_slot >= 0) f.bslot_n[std::min<int>(op.dst_slot, kMaxSlots)] = n; }
  else { f.acc_n = n; if (op.dst_slot >= 0) f.slot_n[std::min<int>(op.dst_slot, kMaxSlots)] = n; }
}

bool output_nullable(const PipelineParams& p, int idx) {
  Flow f;
  for (const vgpu_op& op : p.prog->ops) flow_step(p, f, op, false);
  return opnd_dirty(p, f, p.outputs[idx]);
}

// ---- lowering: ABI ops + sink -> micro-ops ------------------------------------------------------
namespace {
struct MopInfo {
  int st_v = -1, st_b = -1;          // slot stored by the mop (value / boolean numbering)
  int rd_v[3] = {-1, -1, -1}, rd_b[3] = {-1, -1, -1};
};

struct Lowerer {
  PipelineParams& p;
  Flow f;                 // nullability + latched-error flow ("dirty")
  MopInfo info[kMaxMops];
  int n = 0;
  int acc_mirror = -1;    // value slot whose content equals the value accumulator
  int bacc_mirror = -1;   // boolean slot whose content equals the boolean accumulator
  int bacc_mirror_prev = -1;
  bool overflow = false;
  bool bad_output = false;

  explicit Lowerer(PipelineParams& pp) : p(pp) {}

  MOp* emit(uint8_t h) {
    if (n >= kMaxMops) { overflow = true; n = kMaxMops - 1; }
    MOp& m = p.mops[n];
    m = MOp{};
    m.h = h;
    m.dst = -1; m.dstn = -1;
    info[n] = MopInfo{};
    return &p.mops[n++];
  }

  XOp opnd(const vgpu_operand& o, uint64_t imm, int which) {
    XOp x{};
    x.type = o.type;
    MopInfo& mi = info[n - 1];
    switch (o.kind) {
      case VGPU_OPND_NONE: x.kind = XK_NONE; break;
      case VGPU_OPND_ACC: x.kind = XK_ACC; break;
      case VGPU_OPND_IMM: x.kind = XK_IMM; x.imm = imm; break;
      case VGPU_OPND_NULL: x.kind = XK_NULL; break;
      case VGPU_OPND_SLOT:
        if (o.type == VGPU_BOOLEAN) {
          if (bacc_mirror == (int)o.index) { x.kind = XK_ACC; break; }
          x.kind = XK_BSLOT;
          x.off = p.smem_bmeta_off + o.index * kThreads * 4;
          mi.rd_b[which] = o.index;
        } else {
          if (acc_mirror == (int)o.index) { x.kind = XK_ACC; break; }
          x.kind = XK_SLOT;
          x.off = p.smem_slots_off + o.index * kTile * 8;
          x.noff = p.smem_meta_off + o.index * kThreads * 4;
          x.has_nulls = f.slot_n[o.index] ? 1 : 0;
          mi.rd_v[which] = o.index;
        }
        break;
      default: { // column
        const DevCol& c = p.cols[o.index];
        x.col = (uint8_t)o.index;
        x.type = c.type;
        if (c.encoding == VGPU_CONSTANT) {
          if (c.const_null) x.kind = XK_NULL; else { x.kind = XK_IMM; x.imm = (uint64_t)c.constant; }
          break;
        }
        x.off = c.stage_off;
        x.noff = c.nulls_off;
        x.has_nulls = c.nulls_off >= 0 ? 1 : 0;
        if (c.encoding == VGPU_DICTIONARY) { x.kind = c.type == VGPU_BOOLEAN ? XK_BDICT : XK_DICT; break; }
        switch (c.type) {
          case VGPU_BOOLEAN: x.kind = XK_BCOL; break;
          case VGPU_TINYINT: x.kind = XK_COL1; break;
          case VGPU_SMALLINT: x.kind = XK_COL2; break;
          case VGPU_INTEGER: x.kind = XK_COL4I; break;
          case VGPU_REAL: x.kind = XK_COL4F; break;
          default: x.kind = XK_COL8;
        }
      }
    }
    return x;
  }

  void store(MOp* m, const vgpu_op& op, bool is_bool, bool dirty) {
    if (op.dst_slot < 0) return;
    m->flags |= MF_STORE;
    if (is_bool) {
      m->dst = p.smem_bmeta_off + op.dst_slot * kThreads * 4;
      info[n - 1].st_b = op.dst_slot;
    } else {
      m->dst = p.smem_slots_off + op.dst_slot * kTile * 8;
      m->dstn = dirty ? p.smem_meta_off + op.dst_slot * kThreads * 4 : -1;
      info[n - 1].st_v = op.dst_slot;
    }
  }

  void lower_op(const vgpu_op& op) {
    const int oc = op.opcode;
    const bool is_bool = op_is_bool(oc);
    // AND/OR of (boolean slot the accumulator mirrored before the last compare) with that
    // compare: fold into the compare (Q6's conjunct chain becomes one micro-op per conjunct).
    if ((oc == VGPU_OP_AND || oc == VGPU_OP_OR) && n > 0 && p.mops[n - 1].h == H_CMP &&
        !(p.mops[n - 1].flags & (MF_STORE | MF_COMB_AND | MF_COMB_OR))) {
      const vgpu_operand* slot = nullptr;
      if (op.a.kind == VGPU_OPND_SLOT && op.b.kind == VGPU_OPND_ACC) slot = &op.a;
      if (op.b.kind == VGPU_OPND_SLOT && op.a.kind == VGPU_OPND_ACC) slot = &op.b;
      if (slot && bacc_mirror_prev == (int)slot->index) {
        MOp* m = &p.mops[n - 1];
        m->flags |= (oc == VGPU_OP_AND) ? MF_COMB_AND : MF_COMB_OR;
        flow_step(p, f, op, true);
        store(m, op, true, true);
        bacc_mirror = op.dst_slot;
        return;
      }
    }
    MOp* m;
    if (oc >= VGPU_OP_ADD && oc <= VGPU_OP_MOD) {
      m = emit(H_ARITH);
      m->sub = (uint8_t)oc; m->type = op.type;
      m->x = opnd(op.a, (uint64_t)op.imm.i, 0);
      if (oc != VGPU_OP_NEG) m->y = opnd(op.b, (uint64_t)op.imm.i, 1);
    } else if (oc == VGPU_OP_CAST) {
      m = emit(H_CAST);
      m->type = op.type;
      m->x = opnd(op.a, (uint64_t)op.imm.i, 0);
      m->x.type = op.a.type;
    } else if (oc == VGPU_OP_MOV) {
      m = emit(H_LD);
      m->type = op.type;
      m->x = opnd(op.a, (uint64_t)op.imm.i, 0);
    } else if (oc == VGPU_OP_SELECT || oc == VGPU_OP_COALESCE) {
      m = emit(H_SELECT);
      m->sub = (uint8_t)oc; m->type = op.type;
      m->x = opnd(op.a, (uint64_t)op.imm.i, 0);
      m->y = opnd(op.b, (uint64_t)op.imm.i, 1);
      if (oc == VGPU_OP_SELECT) m->z = opnd(op.c, (uint64_t)op.imm2.i, 2);
    } else if (oc >= VGPU_OP_EQ && oc <= VGPU_OP_BETWEEN) {
      m = emit(H_CMP);
      m->sub = (uint8_t)oc; m->type = op.type;
      m->x = opnd(op.a, (uint64_t)op.imm.i, 0);
      m->y = opnd(op.b, (uint64_t)op.imm.i, 1);
      if (oc == VGPU_OP_BETWEEN) m->z = opnd(op.c, (uint64_t)op.imm2.i, 2);
    } else {
      m = emit(H_BOOL);
      m->sub = (uint8_t)oc; m->type = VGPU_BOOLEAN;
      m->x = opnd(op.a, (uint64_t)op.imm.i, 0);
      if (oc == VGPU_OP_AND || oc == VGPU_OP_OR) m->y = opnd(op.b, (uint64_t)op.imm.i, 1);
    }
    flow_step(p, f, op, true);
    const bool dirty = is_bool ? f.bacc_n : f.acc_n;
    store(m, op, is_bool, dirty);
    if (is_bool) { bacc_mirror_prev = bacc_mirror; bacc_mirror = op.dst_slot; }
    else acc_mirror = op.dst_slot;
  }

  XOp out_opnd(int idx) {
    const vgpu_operand& o = p.outputs[idx];
    if (o.kind == VGPU_OPND_IMM || o.kind == VGPU_OPND_ACC) bad_output = true;
    return opnd(o, 0, 0);
  }

  void lower_sink() {
    const Program& prog = *p.prog;
    MOp* m = emit(H_FILTER);
    m->x = opnd(prog.filter, 1, 0);
    if (prog.filter.kind == VGPU_OPND_ACC) m->x.kind = XK_ACC;
    if (p.mode == kSinkFilterProject) {
      emit(H_SEL);
      for (int k = 0; k < p.n_outputs; ++k) {
        if (p.outs[k].values == nullptr) continue;
        m = emit(H_PROJ);
        m->dst = k;
        m->x = out_opnd(k);
      }
      return;
    }
    for (int k = 0; k < p.n_keys; ++k) {
      m = emit(H_KEY);
      m->dst = k;
      m->x = out_opnd(p.keys[k].output);
    }
    emit(H_KEYEND);
    for (int c = 0; c < p.n_cells; ++c) {
      const DevCell& cell = p.cells[c];
      if (cell.skip) continue;
      m = emit(H_ACCUM);
      m->sub = cell.kind;
      m->dst = p.mode == kSinkArray ? cell.smem_off : c;
      if (cell.input >= 0) m->x = out_opnd(cell.input);
      else m->flags |= MF_NO_VALUE;
    }
  }

  // Drops slot stores nobody reads (after accumulator forwarding most are dead).
  void dead_stores() {
    bool live_v[kMaxSlots + 1] = {false}, live_b[kMaxSlots + 1] = {false};
    for (int i = n - 1; i >= 0; --i) {
      MopInfo& mi = info[i];
      MOp& m = p.mops[i];
      if (mi.st_v >= 0) {
        if (!live_v[mi.st_v]) { m.flags &= ~MF_STORE; }
        live_v[mi.st_v] = false;
      }
      if (mi.st_b >= 0) {
        if (!live_b[mi.st_b]) { m.flags &= ~MF_STORE; }
        live_b[mi.st_b] = false;
      }
      for (int k = 0; k < 3; ++k) {
        if (mi.rd_v[k] >= 0) live_v[mi.rd_v[k]] = true;
        if (mi.rd_b[k] >= 0) live_b[mi.rd_b[k]] = true;
      }
    }
  }
};
} // namespace

static int lower_mops(PipelineParams& p) {
  Lowerer lw(p);
  for (const vgpu_op& op : p.prog->ops) lw.lower_op(op);
  lw.lower_sink();
  lw.dead_stores();
  if (lw.overflow) { set_error("pipeline lowers to more than %d micro-ops", kMaxMops); return VGPU_ERR_UNSUPPORTED; }
  if (lw.bad_output) { set_error("program outputs must be columns, slots or NULL"); return VGPU_ERR_INVALID; }
  p.n_mops = lw.n;
  return VGPU_OK;
}

constexpr int kSmemLimit = 227 * 1024;

int launch_pipeline(PipelineParams& p, int tab_bytes, int grid, cudaStream_t stream) {
  const int slots_bytes = p.n_slots * kTile * 8;
  const int meta_bytes = p.n_slots * kThreads * 4;
  const int bmeta_bytes = p.n_bslots * kThreads * 4;
  const int fixed = 128 + slots_bytes + meta_bytes + bmeta_bytes + ((tab_bytes + 127) & ~127);
  int S = (kSmemLimit - fixed) / p.stage_bytes;
  if (S < 1) {
    set_error("pipeline needs %d B of shared memory per stage plus %d B fixed; exceeds %d", p.stage_bytes, fixed, kSmemLimit);
    return VGPU_ERR_UNSUPPORTED;
  }
  S = std::min(S, kMaxStages);
  p.n_stages = S;
  p.smem_slots_off = 128 + S * p.stage_bytes;
  p.smem_meta_off = p.smem_slots_off + slots_bytes;
  p.smem_bmeta_off = p.smem_meta_off + meta_bytes;
  p.smem_tab_off = (p.smem_bmeta_off + bmeta_bytes + 127) & ~127;
  const int smem = p.smem_tab_off + tab_bytes;
  for (int c = 0; c < p.n_cols; ++c) {
    if (p.cols[c].type == VGPU_VARCHAR && p.cols[c].stage_off >= 0) {
      set_error("column %d: VGPU_VARCHAR columns are read by the plan-specialised kernels only (NVRTC or an AOT plan is required)", c);
      return VGPU_ERR_UNSUPPORTED;
    }
  }
  int rc = lower_mops(p);
  if (rc) return rc;
  auto run = [&](auto kernel) -> int {
    VG_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    kernel_timing_begin(stream);
    kernel<<<grid, kThreads, smem, stream>>>(p);
    kernel_timing_end(stream);
    VG_CUDA(cudaGetLastError());
    count_launch();
    return VGPU_OK;
  };
  switch (p.mode) {
    case kSinkFilterProject: return run(pipeline_kernel<kSinkFilterProject>);
    case kSinkArray: return run(pipeline_kernel<kSinkArray>);
    default: return run(pipeline_kernel<kSinkHash>);
  }
}

// Entry used by api.cu for the FilterProject sink's follow-up kernels.
int launch_selection_indices(const uint32_t* words, int32_t* tile_counts, int64_t* tile_offsets,
                             int64_t n_rows, int32_t* sel_idx, unsigned long long* total,
                             cudaStream_t stream) {
  const int64_t n_tiles = (n_rows + kTile - 1) / kTile;
  scan_tile_counts<<<1, 1024, 0, stream>>>(tile_counts, n_tiles, tile_offsets, total);
  VG_CUDA(cudaGetLastError());
  count_launch();
  if (sel_idx != nullptr && n_tiles > 0) {
    const int warps_per_block = 8;
    const int blocks = (int)((n_tiles + warps_per_block - 1) / warps_per_block);
    expand_selection<<<blocks, warps_per_block * 32, 0, stream>>>(words, tile_offsets, n_rows, sel_idx);
    VG_CUDA(cudaGetLastError());
    count_launch();
  }
  return VGPU_OK;
}

} // namespace vg
