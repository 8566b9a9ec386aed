// The fused scan pipeline: stage column tiles -> evaluate the expression program -> sink
// (selection + projections, or group-by accumulation).  One persistent CTA per SM slot walks
// tiles of 1024 rows.  Column tiles are fetched by the bulk-copy engine (cp.async.bulk, SASS
// UBLKCP) into a ring of shared-memory stages guarded by mbarriers, so the bytes in flight do
// not depend on occupancy; the 256 threads then interpret the program over the staged tile,
// 4 rows per thread, keeping the running value in registers and spilling only named
// intermediates to shared-memory slots.
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
  const char* stage;
  uint64_t* slots;   // [n_slots][kTile]
  uint32_t* meta;    // [n_slots][kThreads]   n | e << 8
  uint32_t* bmeta;   // [n_bslots][kThreads]  v | n << 8 | e << 16
  int pos0, pos1;    // first row of pair 0 / pair 1 inside the tile
  uint32_t valid;    // rows that exist
  uint32_t ecode;    // error bits raised by this thread's ops, one byte per row
  Val acc;
  BoolVal bacc;
};

__device__ __forceinline__ bool is_fp(uint8_t t) { return t == VGPU_REAL || t == VGPU_DOUBLE; }

// Widen a raw staged/base element to the register representation.
__device__ __forceinline__ uint64_t widen(uint8_t type, uint64_t raw) {
  switch (type) {
    case VGPU_TINYINT: return (uint64_t)(int64_t)(int8_t)raw;
    case VGPU_SMALLINT: return (uint64_t)(int64_t)(int16_t)raw;
    case VGPU_INTEGER: return (uint64_t)(int64_t)(int32_t)raw;
    case VGPU_REAL: return d2u((double)__uint_as_float((uint32_t)raw));
    default: return raw;
  }
}

__device__ __forceinline__ uint64_t load_base(const DevCol& c, int64_t idx) {
  switch (elem_bytes_log2(c.type)) {
    case 0: return reinterpret_cast<const uint8_t*>(c.values)[idx];
    case 1: return reinterpret_cast<const uint16_t*>(c.values)[idx];
    case 2: return reinterpret_cast<const uint32_t*>(c.values)[idx];
    default: return reinterpret_cast<const uint64_t*>(c.values)[idx];
  }
}

// Two bits (rows pos, pos+1) of a staged 1024-bit tile bitmap.
__device__ __forceinline__ uint32_t pair_bits(const char* tile_bits, int pos) {
  const uint32_t w = reinterpret_cast<const uint32_t*>(tile_bits)[pos >> 5];
  return (w >> (pos & 31)) & 3u;
}

__device__ __forceinline__ void load_col(const Ctx& cx, const DevCol& c, Val& out) {
  out.e = 0;
  if (c.encoding == VGPU_CONSTANT) {
    const uint64_t k = (uint64_t)c.constant;
#pragma unroll
    for (int j = 0; j < 4; ++j) out.v[j] = k;
    out.n = c.const_null ? 0xFu : 0u;
    return;
  }
  uint32_t notnull = 0xFu;
  if (c.nulls_off >= 0) {
    notnull = pair_bits(cx.stage + c.nulls_off, cx.pos0) | (pair_bits(cx.stage + c.nulls_off, cx.pos1) << 2);
  }
  const char* t = cx.stage + c.stage_off;
  if (c.encoding == VGPU_DICTIONARY) {
    const uint2 i0 = reinterpret_cast<const uint2*>(t)[cx.pos0 >> 1];
    const uint2 i1 = reinterpret_cast<const uint2*>(t)[cx.pos1 >> 1];
    const int32_t idx[4] = {(int32_t)i0.x, (int32_t)i0.y, (int32_t)i1.x, (int32_t)i1.y};
    const uint32_t live = cx.valid & notnull;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint64_t raw = 0;
      if ((live >> j) & 1u) {
        raw = load_base(c, idx[j]);
        if (c.base_nulls != nullptr && !bit_at(c.base_nulls, idx[j])) notnull &= ~(1u << j);
      }
      out.v[j] = widen(c.type, raw);
    }
    out.n = (~notnull) & 0xFu;
    return;
  }
  switch (c.elem_log2) {
    case 3: {
      const uint4 q0 = reinterpret_cast<const uint4*>(t)[cx.pos0 >> 1];
      const uint4 q1 = reinterpret_cast<const uint4*>(t)[cx.pos1 >> 1];
      out.v[0] = (uint64_t)q0.x | ((uint64_t)q0.y << 32);
      out.v[1] = (uint64_t)q0.z | ((uint64_t)q0.w << 32);
      out.v[2] = (uint64_t)q1.x | ((uint64_t)q1.y << 32);
      out.v[3] = (uint64_t)q1.z | ((uint64_t)q1.w << 32);
      break;
    }
    case 2: {
      const uint2 q0 = reinterpret_cast<const uint2*>(t)[cx.pos0 >> 1];
      const uint2 q1 = reinterpret_cast<const uint2*>(t)[cx.pos1 >> 1];
      out.v[0] = widen(c.type, q0.x);
      out.v[1] = widen(c.type, q0.y);
      out.v[2] = widen(c.type, q1.x);
      out.v[3] = widen(c.type, q1.y);
      break;
    }
    case 1: {
      const uint32_t q0 = reinterpret_cast<const uint32_t*>(t)[cx.pos0 >> 1];
      const uint32_t q1 = reinterpret_cast<const uint32_t*>(t)[cx.pos1 >> 1];
      out.v[0] = widen(c.type, q0 & 0xffffu);
      out.v[1] = widen(c.type, q0 >> 16);
      out.v[2] = widen(c.type, q1 & 0xffffu);
      out.v[3] = widen(c.type, q1 >> 16);
      break;
    }
    default: {
      const uint32_t q0 = reinterpret_cast<const uint16_t*>(t)[cx.pos0 >> 1];
      const uint32_t q1 = reinterpret_cast<const uint16_t*>(t)[cx.pos1 >> 1];
      out.v[0] = widen(c.type, q0 & 0xffu);
      out.v[1] = widen(c.type, q0 >> 8);
      out.v[2] = widen(c.type, q1 & 0xffu);
      out.v[3] = widen(c.type, q1 >> 8);
    }
  }
  out.n = (~notnull) & 0xFu;
}

// which: 0 = a/b (reads op.imm), 1 = c (reads op.imm2)
__device__ __forceinline__ void load_value(const Ctx& cx, const vgpu_op& op, const vgpu_operand o,
                                           int which, Val& out) {
  switch (o.kind) {
    case VGPU_OPND_ACC:
      out = cx.acc;
      break;
    case VGPU_OPND_SLOT: {
      const uint4* s = reinterpret_cast<const uint4*>(cx.slots + (size_t)o.index * kTile);
      const uint4 q0 = s[cx.pos0 >> 1];
      const uint4 q1 = s[cx.pos1 >> 1];
      out.v[0] = (uint64_t)q0.x | ((uint64_t)q0.y << 32);
      out.v[1] = (uint64_t)q0.z | ((uint64_t)q0.w << 32);
      out.v[2] = (uint64_t)q1.x | ((uint64_t)q1.y << 32);
      out.v[3] = (uint64_t)q1.z | ((uint64_t)q1.w << 32);
      const uint32_t m = cx.meta[o.index * kThreads + threadIdx.x];
      out.n = m & 0xFu;
      out.e = (m >> 8) & 0xFu;
      break;
    }
    case VGPU_OPND_COL:
      load_col(cx, cx.p->cols[o.index], out);
      break;
    case VGPU_OPND_IMM: {
      const uint64_t k = (uint64_t)(which ? op.imm2.i : op.imm.i);
#pragma unroll
      for (int j = 0; j < 4; ++j) out.v[j] = k;
      out.n = 0;
      out.e = 0;
      break;
    }
    default: // NULL constant
#pragma unroll
      for (int j = 0; j < 4; ++j) out.v[j] = 0;
      out.n = 0xFu;
      out.e = 0;
  }
}

__device__ __forceinline__ void store_slot(const Ctx& cx, int slot, const Val& v) {
  uint4* s = reinterpret_cast<uint4*>(cx.slots + (size_t)slot * kTile);
  s[cx.pos0 >> 1] = make_uint4((uint32_t)v.v[0], (uint32_t)(v.v[0] >> 32), (uint32_t)v.v[1],
                               (uint32_t)(v.v[1] >> 32));
  s[cx.pos1 >> 1] = make_uint4((uint32_t)v.v[2], (uint32_t)(v.v[2] >> 32), (uint32_t)v.v[3],
                               (uint32_t)(v.v[3] >> 32));
  cx.meta[slot * kThreads + threadIdx.x] = v.n | (v.e << 8);
}

__device__ __forceinline__ void load_bool(const Ctx& cx, const vgpu_op& op, const vgpu_operand o,
                                          BoolVal& out) {
  switch (o.kind) {
    case VGPU_OPND_ACC:
      out = cx.bacc;
      break;
    case VGPU_OPND_SLOT: {
      const uint32_t m = cx.bmeta[o.index * kThreads + threadIdx.x];
      out.v = m & 0xFu;
      out.n = (m >> 8) & 0xFu;
      out.e = (m >> 16) & 0xFu;
      break;
    }
    case VGPU_OPND_COL: {
      const DevCol& c = cx.p->cols[o.index];
      out.e = 0;
      if (c.encoding == VGPU_CONSTANT) {
        out.v = c.constant ? 0xFu : 0u;
        out.n = c.const_null ? 0xFu : 0u;
        break;
      }
      uint32_t notnull = 0xFu;
      if (c.nulls_off >= 0) {
        notnull = pair_bits(cx.stage + c.nulls_off, cx.pos0) |
            (pair_bits(cx.stage + c.nulls_off, cx.pos1) << 2);
      }
      if (c.encoding == VGPU_DICTIONARY) {
        const char* t = cx.stage + c.stage_off;
        const uint2 i0 = reinterpret_cast<const uint2*>(t)[cx.pos0 >> 1];
        const uint2 i1 = reinterpret_cast<const uint2*>(t)[cx.pos1 >> 1];
        const int32_t idx[4] = {(int32_t)i0.x, (int32_t)i0.y, (int32_t)i1.x, (int32_t)i1.y};
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
      } else {
        out.v = pair_bits(cx.stage + c.stage_off, cx.pos0) |
            (pair_bits(cx.stage + c.stage_off, cx.pos1) << 2);
      }
      out.n = (~notnull) & 0xFu;
      break;
    }
    case VGPU_OPND_IMM:
      out.v = op.imm.i ? 0xFu : 0u;
      out.n = 0;
      out.e = 0;
      break;
    default:
      out.v = 0;
      out.n = 0xFu;
      out.e = 0;
  }
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
// The interpreter: one pass over the program for the thread's 4 rows.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void run_program(Ctx& cx) {
  const PipelineParams& p = *cx.p;
#pragma unroll 1
  for (int i = 0; i < p.n_ops; ++i) {
    const vgpu_op& op = p.ops[i];
    const uint8_t oc = op.opcode;
    if (oc >= VGPU_OP_AND && oc <= VGPU_OP_IS_NULL) {
      BoolVal r;
      if (oc == VGPU_OP_IS_NULL) {
        if (op.a.type == VGPU_BOOLEAN) {
          BoolVal a;
          load_bool(cx, op, op.a, a);
          r.v = a.n; r.e = a.e;
        } else {
          Val a;
          load_value(cx, op, op.a, 0, a);
          r.v = a.n; r.e = a.e;
        }
        r.n = 0;
      } else {
        BoolVal a;
        load_bool(cx, op, op.a, a);
        if (oc == VGPU_OP_NOT) {
          r.v = (~a.v) & 0xFu; r.n = a.n; r.e = a.e;
        } else {
          BoolVal b;
          load_bool(cx, op, op.b, b);
          const uint32_t okA = ~a.n & ~a.e, okB = ~b.n & ~b.e;
          if (oc == VGPU_OP_AND) {
            const uint32_t decided = (okA & ~a.v) | (okB & ~b.v); // some conjunct is FALSE
            r.v = okA & a.v & okB & b.v & 0xFu;
            r.e = ~decided & (a.e | b.e) & 0xFu;
            r.n = ~decided & ~r.e & (a.n | b.n) & 0xFu;
          } else {
            const uint32_t decided = (okA & a.v) | (okB & b.v); // some disjunct is TRUE
            r.v = decided & 0xFu;
            r.e = ~decided & (a.e | b.e) & 0xFu;
            r.n = ~decided & ~r.e & (a.n | b.n) & 0xFu;
          }
        }
      }
      cx.bacc = r;
      if (op.dst_slot >= 0) {
        cx.bmeta[op.dst_slot * kThreads + threadIdx.x] = r.v | (r.n << 8) | (r.e << 16);
      }
      continue;
    }
    if (oc >= VGPU_OP_EQ && oc <= VGPU_OP_BETWEEN) {
      BoolVal r;
      if (op.type == VGPU_BOOLEAN) { // boolean = / <> boolean
        BoolVal a, b;
        load_bool(cx, op, op.a, a);
        load_bool(cx, op, op.b, b);
        const uint32_t x = a.v ^ b.v;
        r.v = (oc == VGPU_OP_EQ ? ~x : x) & 0xFu;
        r.n = a.n | b.n;
        r.e = a.e | b.e;
      } else {
        Val a, b;
        load_value(cx, op, op.a, 0, a);
        load_value(cx, op, op.b, 0, b);
        r.n = a.n | b.n;
        r.e = a.e | b.e;
        r.v = 0;
        if (oc == VGPU_OP_BETWEEN) {
          Val c;
          load_value(cx, op, op.c, 1, c);
          r.n |= c.n;
          r.e |= c.e;
          r.v = cmp4(VGPU_OP_GE, op.type, a.v, b.v) & cmp4(VGPU_OP_LE, op.type, a.v, c.v);
        } else {
          r.v = cmp4(oc, op.type, a.v, b.v);
        }
      }
      r.v &= ~r.n;
      cx.bacc = r;
      if (op.dst_slot >= 0) {
        cx.bmeta[op.dst_slot * kThreads + threadIdx.x] = r.v | (r.n << 8) | (r.e << 16);
      }
      continue;
    }
    // value ops
    Val a, r;
    load_value(cx, op, op.a, 0, a);
    r.n = a.n;
    r.e = a.e;
    if (oc == VGPU_OP_MOV) {
#pragma unroll
      for (int j = 0; j < 4; ++j) r.v[j] = a.v[j];
    } else if (oc == VGPU_OP_CAST) {
      uint32_t bad = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        r.v[j] = 0;
        const uint32_t err = cast_value(op.a.type, op.type, a.v[j], r.v[j]);
        if (err && !((a.n >> j) & 1u) && ((cx.valid >> j) & 1u)) { bad |= 1u << j; cx.ecode |= err << (8 * j); }
      }
      r.e |= bad;
    } else if (oc == VGPU_OP_NEG) {
      uint32_t bad = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        r.v[j] = 0;
        const uint32_t err = arith(oc, op.type, a.v[j], 0, r.v[j]);
        if (err && !((a.n >> j) & 1u) && ((cx.valid >> j) & 1u)) { bad |= 1u << j; cx.ecode |= err << (8 * j); }
      }
      r.e |= bad;
    } else {
      Val b;
      load_value(cx, op, op.b, 0, b);
      r.n |= b.n;
      r.e |= b.e;
      if (op.type == VGPU_DOUBLE && oc <= VGPU_OP_MUL) {
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
          const uint32_t err = arith(oc, op.type, a.v[j], b.v[j], r.v[j]);
          if (err && !((r.n >> j) & 1u) && ((cx.valid >> j) & 1u)) { bad |= 1u << j; cx.ecode |= err << (8 * j); }
        }
        r.e |= bad;
      }
    }
    cx.acc = r;
    if (op.dst_slot >= 0) store_slot(cx, op.dst_slot, r);
  }
}

// ------------------------------------------------------------------------------------------
// Sinks
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t interleave16(uint32_t even, uint32_t odd) {
  // bit i of even -> bit 2i, bit i of odd -> bit 2i+1
  uint32_t x = (even & 0xffffu) | (odd << 16);
  // spread the two 16-bit halves: classic perfect-shuffle network
  uint32_t t;
  t = (x ^ (x >> 8)) & 0x0000ff00u; x ^= t ^ (t << 8);
  t = (x ^ (x >> 4)) & 0x00f000f0u; x ^= t ^ (t << 4);
  t = (x ^ (x >> 2)) & 0x0c0c0c0cu; x ^= t ^ (t << 2);
  t = (x ^ (x >> 1)) & 0x22222222u; x ^= t ^ (t << 1);
  return x;
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

// Filter operand -> rows that pass; raises latched errors for rows the filter left undecided.
__device__ __forceinline__ uint32_t eval_filter(Ctx& cx, uint32_t& raise) {
  const PipelineParams& p = *cx.p;
  if (p.filter.kind == VGPU_OPND_NONE) return cx.valid;
  BoolVal f;
  vgpu_op dummy;
  dummy.imm.i = 1;
  load_bool(cx, dummy, p.filter, f);
  raise |= f.e & cx.valid;
  return f.v & ~f.n & ~f.e & cx.valid;
}

__device__ void sink_filter_project(Ctx& cx, int64_t tile, int cnt) {
  const PipelineParams& p = *cx.p;
  uint32_t raise = 0;
  const uint32_t pass = eval_filter(cx, raise);
  const int64_t word_base = tile * (kTile / 32);
  const int n_words = (cnt + 31) >> 5; // words of this tile that exist (partial last tile)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool full = cnt == kTile;
  // selection bitmap
  {
#pragma unroll
    for (int pr = 0; pr < 2; ++pr) {
      const uint32_t even = __ballot_sync(0xffffffffu, (pass >> (2 * pr)) & 1u);
      const uint32_t odd = __ballot_sync(0xffffffffu, (pass >> (2 * pr + 1)) & 1u);
      if ((lane & 15) == 0) {
        const int h = lane >> 4;
        const int w = warp * 4 + pr * 2 + h;
        // words past the data but inside the last 64-bit word are written as zero
        if (w < ((n_words + 1) & ~1)) {
          p.sel_words[word_base + w] = interleave16(even >> (16 * h), odd >> (16 * h));
        }
      }
    }
    const int c = __popc(pass);
    int wsum = c;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, o);
    if (lane == 0 && wsum) atomicAdd(&p.tile_counts[tile], wsum);
  }
  // projections
  for (int k = 0; k < p.n_outputs; ++k) {
    const vgpu_operand o = p.outputs[k];
    const DevOut out = p.outs[k];
    if (out.values == nullptr) continue;
    vgpu_op dummy;
    dummy.imm.i = 0;
    uint32_t nmask;
    const int64_t r0 = tile * kTile + cx.pos0, r1 = tile * kTile + cx.pos1;
    if (o.type == VGPU_BOOLEAN) {
      BoolVal b;
      load_bool(cx, dummy, o, b);
      raise |= b.e & pass;
      nmask = b.n;
      // value bits
#pragma unroll
      for (int pr = 0; pr < 2; ++pr) {
        const uint32_t even = __ballot_sync(0xffffffffu, (b.v >> (2 * pr)) & 1u);
        const uint32_t odd = __ballot_sync(0xffffffffu, (b.v >> (2 * pr + 1)) & 1u);
        if ((lane & 15) == 0) {
          const int h = lane >> 4;
          const int w = warp * 4 + pr * 2 + h;
          if (w < ((n_words + 1) & ~1)) {
            reinterpret_cast<uint32_t*>(out.values)[word_base + w] =
                interleave16(even >> (16 * h), odd >> (16 * h));
          }
        }
      }
    } else {
      Val v;
      load_value(cx, dummy, o, 0, v);
      raise |= v.e & pass;
      nmask = v.n;
      const int lg = elem_bytes_log2(o.type);
      if (o.type == VGPU_REAL) {
#pragma unroll
        for (int j = 0; j < 4; ++j) v.v[j] = __float_as_uint((float)u2d(v.v[j]));
      }
      if (full && lg == 3) {
        uint4* d = reinterpret_cast<uint4*>(out.values);
        d[r0 >> 1] = make_uint4((uint32_t)v.v[0], (uint32_t)(v.v[0] >> 32), (uint32_t)v.v[1],
                                (uint32_t)(v.v[1] >> 32));
        d[r1 >> 1] = make_uint4((uint32_t)v.v[2], (uint32_t)(v.v[2] >> 32), (uint32_t)v.v[3],
                                (uint32_t)(v.v[3] >> 32));
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
      const uint32_t notnull = ~nmask & cx.valid & 0xFu;
#pragma unroll
      for (int pr = 0; pr < 2; ++pr) {
        const uint32_t even = __ballot_sync(0xffffffffu, (notnull >> (2 * pr)) & 1u);
        const uint32_t odd = __ballot_sync(0xffffffffu, (notnull >> (2 * pr + 1)) & 1u);
        if ((lane & 15) == 0) {
          const int h = lane >> 4;
          const int w = warp * 4 + pr * 2 + h;
          if (w < ((n_words + 1) & ~1)) {
            reinterpret_cast<uint32_t*>(out.nulls)[word_base + w] =
                interleave16(even >> (16 * h), odd >> (16 * h));
          }
        }
      }
    }
  }
  if (raise) atomicOr(p.err_word, raised_code(cx.ecode, raise));
}

// ---- group-by accumulation -------------------------------------------------------------------
// Converts the raw register value of output operand type `t` into what cell `kind` consumes.
__device__ __forceinline__ uint64_t cell_input(uint8_t kind, uint8_t t, uint64_t v) {
  if (kind == kCellSumF64 && !is_fp(t)) return d2u((double)(int64_t)v);
  return v;
}

// One shared-memory cell update (thread- or leader-exclusive, so no atomics).
__device__ __forceinline__ void update_smem_cell(uint8_t kind, char* addr, uint64_t v, uint32_t& ecode) {
  switch (kind) {
    case kCellSumF64: {
      double* d = reinterpret_cast<double*>(addr);
      *d = __dadd_rn(*d, u2d(v));
      break;
    }
    case kCellSumI64: {
      int64_t* d = reinterpret_cast<int64_t*>(addr);
      const int64_t x = *d, y = (int64_t)v, z = (int64_t)((uint64_t)x + (uint64_t)y);
      if (((x ^ z) & (y ^ z)) < 0) ecode |= kErrOverflow;
      *d = z;
      break;
    }
    case kCellCount: {
      uint32_t* d = reinterpret_cast<uint32_t*>(addr);
      *d = *d + 1;
      break;
    }
    case kCellMinI64: {
      int64_t* d = reinterpret_cast<int64_t*>(addr);
      if ((int64_t)v < *d) *d = (int64_t)v;
      break;
    }
    case kCellMaxI64: {
      int64_t* d = reinterpret_cast<int64_t*>(addr);
      if (*d < (int64_t)v) *d = (int64_t)v;
      break;
    }
    case kCellMinF64: {
      double* d = reinterpret_cast<double*>(addr);
      if (u2d(v) < *d) *d = u2d(v);
      break;
    }
    default: {
      double* d = reinterpret_cast<double*>(addr);
      if (*d < u2d(v)) *d = u2d(v);
    }
  }
}

// Group number / normalized key of the thread's rows; clears rows dropped by ignore_null_keys
// from `pass` and latches kErrKeyRange for values outside the declared key domain.
__device__ __forceinline__ void compute_keys(Ctx& cx, uint32_t& pass, uint32_t& raise, uint64_t key[4],
                                             uint32_t& keynull) {
  const PipelineParams& p = *cx.p;
#pragma unroll
  for (int j = 0; j < 4; ++j) key[j] = 0;
  keynull = 0;
  vgpu_op dummy;
  dummy.imm.i = 0;
  for (int k = 0; k < p.n_keys; ++k) {
    const DevKey& ks = p.keys[k];
    Val v;
    load_value(cx, dummy, p.outputs[ks.output], 0, v);
    raise |= v.e & pass;
    if (p.ignore_null_keys) pass &= ~v.n;
    if (ks.bits == 64) { // raw 64-bit key
#pragma unroll
      for (int j = 0; j < 4; ++j) key[j] = v.v[j];
      keynull = v.n;
      continue;
    }
    uint32_t bad = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t x = (int64_t)v.v[j];
      const bool isnull = (v.n >> j) & 1u;
      uint64_t id;
      if (isnull) {
        id = 0;
        if (!ks.nullable) bad |= 1u << j;
      } else {
        if (x < ks.min || x > ks.max) bad |= 1u << j;
        id = (uint64_t)(x - ks.min) + (ks.nullable ? 1u : 0u);
      }
      key[j] += (p.mode == kSinkArray) ? id * (uint64_t)ks.mult : (id << ks.shift);
    }
    if (bad & pass) {
      cx.ecode |= kErrKeyRange * 0x01010101u;
      raise |= bad & pass;
      pass &= ~bad;
    }
  }
}

__device__ void sink_array(Ctx& cx, char* tab) {
  const PipelineParams& p = *cx.p;
  uint32_t raise = 0;
  uint32_t pass = eval_filter(cx, raise);
  uint64_t key[4];
  uint32_t keynull;
  compute_keys(cx, pass, raise, key, keynull);
  const int lane = threadIdx.x & 31;
  // Conflict schedule: thread-private tables need none; warp-private tables let the lanes that
  // share a group take turns (rank among the peers found with match.any).
  uint32_t rank[4] = {0, 0, 0, 0};
  uint32_t rounds[4] = {1, 1, 1, 1};
  if (p.warp_private) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const bool act = (pass >> j) & 1u;
      const uint32_t todo = __ballot_sync(0xffffffffu, act);
      uint32_t r = 0;
      if (act) {
        const uint32_t peers = __match_any_sync(todo, (uint32_t)key[j]);
        r = __popc(peers & ((1u << lane) - 1u));
      }
      rank[j] = r;
      uint32_t m = act ? r + 1 : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
      rounds[j] = m;
    }
  }
  const int owner = p.warp_private ? 0 : threadIdx.x;
  const int copies = p.warp_private ? 1 : kThreads;
  char* base = tab;
  if (p.warp_private) base += (size_t)(threadIdx.x >> 5) * (size_t)p.warp_tab_bytes;
  vgpu_op dummy;
  dummy.imm.i = 0;
  uint32_t acc_err = 0; // checked-sum overflow of rows that did pass
  for (int c = 0; c < p.n_cells; ++c) {
    const DevCell cell = p.cells[c];
    if (cell.skip) continue;
    Val v;
    uint32_t live = pass;
    uint8_t in_type = VGPU_BIGINT;
    if (cell.input >= 0) {
      const vgpu_operand o = p.outputs[cell.input];
      in_type = o.type;
      if (o.type == VGPU_BOOLEAN) {
        BoolVal b;
        load_bool(cx, dummy, o, b);
        raise |= b.e & pass;
        live &= ~b.n;
#pragma unroll
        for (int j = 0; j < 4; ++j) v.v[j] = (b.v >> j) & 1u;
      } else {
        load_value(cx, dummy, o, 0, v);
        raise |= v.e & pass;
        live &= ~v.n;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) v.v[j] = 0;
    }
    const int w = cell_width(cell.kind);
    char* cbase = base + cell.smem_off;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const bool act = (live >> j) & 1u;
      char* addr = cbase + ((size_t)key[j] * copies + owner) * w;
      const uint64_t x = cell_input(cell.kind, in_type, v.v[j]);
      if (!p.warp_private) {
        if (act) update_smem_cell(cell.kind, addr, x, acc_err);
      } else {
        for (uint32_t r = 0; r < rounds[j]; ++r) {
          if (act && rank[j] == r) update_smem_cell(cell.kind, addr, x, acc_err);
          __syncwarp();
        }
      }
    }
  }
  if (raise) atomicOr(p.err_word, raised_code(cx.ecode, raise));
  if (acc_err) atomicOr(p.err_word, acc_err);
}

__device__ void sink_hash(Ctx& cx) {
  const PipelineParams& p = *cx.p;
  uint32_t raise = 0;
  uint32_t pass = eval_filter(cx, raise);
  uint64_t key[4];
  uint32_t keynull;
  compute_keys(cx, pass, raise, key, keynull);
  uint32_t acc_err = 0;
  HashRef href;
  href.table = p.table; href.cap_mask = p.cap_mask; href.row_words = p.row_words; href.n_groups_ctr = p.n_groups_ctr;
  // Rows of a thread are adjacent in the input; fold a row into its left neighbour's slot lookup
  // when the keys are equal (sorted / clustered keys such as l_orderkey).
  int64_t slot[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    slot[j] = -1;
    if (!((pass >> j) & 1u)) continue;
    const bool isnull = (keynull >> j) & 1u;
    if ((j & 1) && ((pass >> (j - 1)) & 1u) && key[j] == key[j - 1] &&
        isnull == (bool)((keynull >> (j - 1)) & 1u)) {
      slot[j] = slot[j - 1];
    } else {
      slot[j] = hash_find_or_insert(href, key[j], isnull, acc_err);
    }
    if (slot[j] < 0) pass &= ~(1u << j);
  }
  vgpu_op dummy;
  dummy.imm.i = 0;
  for (int c = 0; c < p.n_cells; ++c) {
    const DevCell cell = p.cells[c];
    if (cell.skip) continue;
    Val v;
    uint32_t live = pass;
    uint8_t in_type = VGPU_BIGINT;
    if (cell.input >= 0) {
      const vgpu_operand o = p.outputs[cell.input];
      in_type = o.type;
      if (o.type == VGPU_BOOLEAN) {
        BoolVal b;
        load_bool(cx, dummy, o, b);
        raise |= b.e & pass;
        live &= ~b.n;
#pragma unroll
        for (int j = 0; j < 4; ++j) v.v[j] = (b.v >> j) & 1u;
      } else {
        load_value(cx, dummy, o, 0, v);
        raise |= v.e & pass;
        live &= ~v.n;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) v.v[j] = 0;
    }
#pragma unroll
    for (int pr = 0; pr < 2; ++pr) {
      const int j0 = 2 * pr, j1 = 2 * pr + 1;
      const bool a0 = (live >> j0) & 1u, a1 = (live >> j1) & 1u;
      uint64_t x0 = cell_input(cell.kind, in_type, v.v[j0]);
      const uint64_t x1 = cell_input(cell.kind, in_type, v.v[j1]);
      // pre-combine the pair when both rows fall in the same group (sum / count only)
      if (a0 && a1 && slot[j0] == slot[j1] && cell.kind == kCellSumF64) {
        // Keep row order: (x0 + x1) then one atomic.
        x0 = d2u(__dadd_rn(u2d(x0), u2d(x1)));
        update_hbm_cell(cell.kind, p.table + slot[j0] * p.row_words + 1 + c, x0, acc_err);
        continue;
      }
      if (a0) update_hbm_cell(cell.kind, p.table + slot[j0] * p.row_words + 1 + c, x0, acc_err);
      if (a1) update_hbm_cell(cell.kind, p.table + slot[j1] * p.row_words + 1 + c, x1, acc_err);
    }
  }
  if (raise) atomicOr(p.err_word, raised_code(cx.ecode, raise));
  if (acc_err) atomicOr(p.err_word, acc_err);
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
  cx.slots = reinterpret_cast<uint64_t*>(smem + p.smem_slots_off);
  cx.meta = reinterpret_cast<uint32_t*>(smem + p.smem_meta_off);
  cx.bmeta = reinterpret_cast<uint32_t*>(smem + p.smem_bmeta_off);
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
  bool any_bulk = false;
  for (int c = 0; c < p.n_cols; ++c) any_bulk |= (p.cols[c].bulk_ok != 0) && (p.cols[c].stage_off >= 0 || p.cols[c].nulls_off >= 0);

  int it = 0;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
    const int s = it % S;
    char* stage = stages + (size_t)s * p.stage_bytes;
    const int64_t row0 = tile * kTile;
    const int cnt = (int)min((int64_t)kTile, p.n_rows - row0);
    const bool bulk = tile < n_full;
    if (bulk) {
      if (any_bulk) mbar_wait(&bars[s], (uint32_t)((it / S) & 1));
      stage_generic(p, stage, row0, cnt, false); // columns the bulk engine could not take
    } else {
      stage_generic(p, stage, row0, cnt, true);
    }
    __syncthreads();
    cx.stage = stage;
    cx.valid = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int pos = (j < 2 ? cx.pos0 : cx.pos1) + (j & 1);
      if (pos < cnt) cx.valid |= 1u << j;
    }
    cx.acc.n = cx.acc.e = 0;
    cx.bacc.v = cx.bacc.n = cx.bacc.e = 0;
    run_program(cx);
    if (MODE == kSinkFilterProject) sink_filter_project(cx, tile, cnt);
    else if (MODE == kSinkArray) sink_array(cx, tab);
    else sink_hash(cx);
    __syncthreads(); // every thread is done reading this stage
    if (tid == 0) {
      const int64_t next = tile + (int64_t)S * gridDim.x;
      if (next < n_full) {
        fence_proxy_async();
        issue_tile(p, stage, &bars[s], next * kTile);
      }
    }
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
// Host side: planning and launch
// ------------------------------------------------------------------------------------------
static int host_elem_log2(int type) {
  switch (type) {
    case VGPU_BOOLEAN: return 255;
    case VGPU_TINYINT: return 0;
    case VGPU_SMALLINT: return 1;
    case VGPU_INTEGER:
    case VGPU_REAL: return 2;
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
  p.n_ops = (int)prog.ops.size();
  for (int i = 0; i < p.n_ops; ++i) p.ops[i] = prog.ops[i];
  p.n_outputs = (int)prog.outputs.size();
  for (int i = 0; i < p.n_outputs; ++i) p.outputs[i] = prog.outputs[i];
  p.filter = prog.filter;
  p.n_cols = ncols;
  p.n_slots = prog.n_slots;
  p.n_bslots = prog.n_bslots;
  p.in_rows = in_rows;
  p.n_rows = in_rows ? n_in_rows : nrows;
  p.err_word = device_err_word();

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
    if (in.type < VGPU_BOOLEAN || in.type > VGPU_DOUBLE) { set_error("column %d: bad type %d", c, in.type); return VGPU_ERR_INVALID; }
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

bool output_nullable(const PipelineParams& p, int idx) {
  bool slot_n[kMaxSlots + 1] = {false}, bslot_n[kMaxSlots + 1] = {false};
  bool acc_n = false, bacc_n = false;
  auto opnd = [&](const vgpu_operand& o) -> bool {
    switch (o.kind) {
      case VGPU_OPND_ACC: return o.type == VGPU_BOOLEAN ? bacc_n : acc_n;
      case VGPU_OPND_SLOT: return o.type == VGPU_BOOLEAN ? bslot_n[std::min<int>(o.index, kMaxSlots)] : slot_n[std::min<int>(o.index, kMaxSlots)];
      case VGPU_OPND_COL: {
        const DevCol& c = p.cols[o.index];
        if (c.encoding == VGPU_CONSTANT) return c.const_null != 0;
        return c.nulls != nullptr || c.base_nulls != nullptr;
      }
      case VGPU_OPND_NULL: return true;
      default: return false;
    }
  };
  for (int i = 0; i < p.n_ops; ++i) {
    const vgpu_op& op = p.ops[i];
    bool n = false;
    if (op.opcode != VGPU_OP_IS_NULL) {
      if (op.a.kind != VGPU_OPND_NONE) n |= opnd(op.a);
      if (op.b.kind != VGPU_OPND_NONE) n |= opnd(op.b);
      if (op.c.kind != VGPU_OPND_NONE) n |= opnd(op.c);
    }
    const bool is_bool = op.opcode >= VGPU_OP_EQ && op.opcode <= VGPU_OP_IS_NULL;
    if (is_bool) { bacc_n = n; if (op.dst_slot >= 0) bslot_n[std::min<int>(op.dst_slot, kMaxSlots)] = n; }
    else { acc_n = n; if (op.dst_slot >= 0) slot_n[std::min<int>(op.dst_slot, kMaxSlots)] = n; }
  }
  return opnd(p.outputs[idx]);
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
  auto run = [&](auto kernel) -> int {
    VG_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    kernel<<<grid, kThreads, smem, stream>>>(p);
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
