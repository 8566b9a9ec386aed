// Prelude of the plan-specialised ("fused") kernels.  fused.cpp generates, per pipeline shape, a
// straight-line kernel body over these helpers and compiles it for sm_100a — ahead of time with
// nvcc for the shapes listed in velox-1_b200/aot_plans.py, or at first use with NVRTC — so the
// expression tree, the column encodings, the key layout and the accumulator list are compile-time
// facts of the kernel.  Self-contained: no #include, so the same text compiles under nvcc and
// NVRTC.  Scalar semantics are the interpreter's (pipeline.cu), restating the reference:
//   velox/functions/prestosql/{Arithmetic.h:45-86,Comparisons.h:43-185},
//   velox/common/base/CheckedArithmetic.h, velox/type/Conversions.h:176-374,
//   velox/expression/ConjunctExpr.cpp:24-174, velox/exec/OperatorUtils.cpp:187-210.
#ifndef VGPU_FUSED_PRELUDE_CUH_
#define VGPU_FUSED_PRELUDE_CUH_

namespace vgf {

typedef unsigned long long u64;
typedef long long i64;
typedef unsigned int u32;
typedef int i32;

// vgpu_type (include/vgpu.h)
#define VGF_BOOLEAN 0
#define VGF_TINYINT 1
#define VGF_SMALLINT 2
#define VGF_INTEGER 3
#define VGF_BIGINT 4
#define VGF_REAL 5
#define VGF_DOUBLE 6
// vgpu_opcode
#define VGF_ADD 1
#define VGF_SUB 2
#define VGF_MUL 3
#define VGF_DIV 4
#define VGF_NEG 5
#define VGF_MOD 6
#define VGF_EQ 16
#define VGF_NE 17
#define VGF_LT 18
#define VGF_LE 19
#define VGF_GT 20
#define VGF_GE 21
// error bits (internal.h)
#define VGF_ERR_OVERFLOW 1u
#define VGF_ERR_DIVZERO 2u
#define VGF_ERR_CAST 4u
#define VGF_ERR_KEYRANGE 8u
#define VGF_ERR_TABLEFULL 16u
#define VGF_ERR_STRKEY 32u
// CellKind (internal.h)
#define VGF_SUMF64 0
#define VGF_SUMI64 1
#define VGF_COUNT 2
#define VGF_MINI64 3
#define VGF_MAXI64 4
#define VGF_MINF64 5
#define VGF_MAXF64 6

#define VGF_I64_MIN (-9223372036854775807ll - 1ll)
#define VGF_I64_MAX 9223372036854775807ll
#define VGF_EMPTY_KEY 0xffffffffffffffffull

struct FusedParams {
  const void* values[16];
  const i32* indices[16];
  const u64* nulls[16];
  const u64* base_nulls[16];
  i32 base_len[16]; // DICTIONARY columns with <= 32 base values (else 0): the base lives in one register per lane
  i64 n_rows;
  const i32* in_rows; // optional row selection: the kernel walks in_rows[0 .. n_rows)
  u32* err_word;
  // ARRAY / GLOBAL sink: per-CTA partials [gridDim][n_cells][n_groups]
  u64* partials;
  // HASH sink
  u64* table;
  u64 cap_mask;
  u64* n_groups_ctr;
  u64 key_min, key_span, key_scale; // monotone home slots for a ranged raw key (internal.h KeyMap)
  u32 key_mode, key_pad;
  // FilterProject sink
  u32* sel_words;
  i32* tile_counts;
  void* out_values[24];
  u64* out_nulls[24];
  // HASH sink: parked segments (see hash_probe4 / ovf_reserve)
  u64* ovf;
  u64 ovf_cap;
  u64* ovf_cursor;
  // LOG sink (internal.h PipelineParams::log...)
  u64* log;
  u64 log_region_cap;
  u32* log_counts;
  u64* log_special;
  u64* sort_state;
  const u64* sort_prev_in;
  u64* sort_prev_out;
  // LOG sink, exchange fused in (internal.h PipelineParams::send...)
  u64 send_sub_cap;
  u64* send_base[8];
  unsigned long long* send_peer_counts[8];
  unsigned long long* send_counts;
};

__device__ __forceinline__ u64 d2u(double d) { return (u64)__double_as_longlong(d); }
__device__ __forceinline__ double u2d(u64 u) { return __longlong_as_double((i64)u); }
__device__ __forceinline__ u64 pack64(u32 lo, u32 hi) { return (u64)lo | ((u64)hi << 32); }
__device__ __forceinline__ u64 sext32(u32 x) { return (u64)(i64)(i32)x; }

// ---- loads (streaming: every input byte is read once) -------------------------------------------
__device__ __forceinline__ uint4 ld16(const void* base, i64 byte_off) {
  return __ldcs(reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(base) + byte_off));
}
// A VARCHAR value (16-byte StringView: u32 size, 4 prefix bytes, then 8 inline bytes or a pointer) of at most 7
// bytes as the integer the reference uses for short string keys (velox/exec/VectorHasher.h:324-329 stringAsNumber:
// the bytes as a little-endian word plus 1 << (8 * size), so every length has its own range and "" is 0).
// Bit 63 set = longer than 7 bytes (no such number).
__device__ __forceinline__ u64 sv_number(uint4 sv) {
  const u32 size = sv.x;
  if (size > 7u) return 1ull << 63;
  const u64 mark = 1ull << (8u * size);
  return size == 0u ? 0ull : (pack64(sv.y, sv.z) & (mark - 1ull)) + mark;
}
__device__ __forceinline__ uint2 ld8(const void* base, i64 byte_off) {
  return __ldcs(reinterpret_cast<const uint2*>(reinterpret_cast<const char*>(base) + byte_off));
}
// 256-bit load (sm_100+): four consecutive 8-byte values of one thread in one fully coalesced instruction
struct u64x4 { u64 x, y, z, w; };
__device__ __forceinline__ u64x4 ld32(const void* base, i64 byte_off) {
  u64x4 r;
#if defined(__CUDACC_VER_MAJOR__) && (__CUDACC_VER_MAJOR__ * 100 + __CUDACC_VER_MINOR__ >= 1209)
  asm volatile("ld.global.cs.v4.u64 {%0, %1, %2, %3}, [%4];"
               : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w)
               : "l"(reinterpret_cast<const char*>(base) + byte_off));
#else // an NVRTC older than 12.9 (e.g. the one bundled with the framework): its ptxas has no 256-bit vector loads
  const uint4 a = __ldcs(reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(base) + byte_off));
  const uint4 b = __ldcs(reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(base) + byte_off + 16));
  r.x = (u64)a.x | ((u64)a.y << 32); r.y = (u64)a.z | ((u64)a.w << 32);
  r.z = (u64)b.x | ((u64)b.y << 32); r.w = (u64)b.z | ((u64)b.w << 32);
#endif
  return r;
}
// rows r .. r+3 of a bitmap (r a multiple of 4): 4 bits
__device__ __forceinline__ u32 bits4(const u64* bits, i64 r) {
  const u32 w = __ldg(reinterpret_cast<const u32*>(bits) + (r >> 5));
  return (w >> (r & 31)) & 15u;
}
// pulls one 128-byte line into L2 without occupying a register (next block's input, a key's home slot)
__device__ __forceinline__ void prefetch_l2(const void* ptr) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr));
}
// rows r, r+1 of a bitmap (r even): 2 bits
__device__ __forceinline__ u32 bits2(const u64* bits, i64 r) {
  const u32 w = __ldg(reinterpret_cast<const u32*>(bits) + (r >> 5));
  return (w >> (r & 31)) & 3u;
}
__device__ __forceinline__ bool bit1(const u64* bits, i64 i) {
  return (__ldg(bits + (i >> 6)) >> (i & 63)) & 1ull;
}

// ---- scalar semantics -----------------------------------------------------------------------
__device__ __forceinline__ bool is_fp(int t) { return t == VGF_REAL || t == VGF_DOUBLE; }
__device__ __forceinline__ void int_limits(int type, i64& lo, i64& hi) {
  switch (type) {
    case VGF_TINYINT: lo = -128; hi = 127; break;
    case VGF_SMALLINT: lo = -32768; hi = 32767; break;
    case VGF_INTEGER: lo = -2147483648ll; hi = 2147483647ll; break;
    default: lo = VGF_I64_MIN; hi = VGF_I64_MAX;
  }
}
// REAL arithmetic is done in double and rounded back to float: for + - * / the double-rounded
// result equals the float operation (53 >= 2*24+2 bits).
__device__ __forceinline__ double fp_round(int type, double x) {
  return type == VGF_REAL ? (double)__double2float_rn(x) : x;
}

// Returns error bits (0 = fine).  Integer ops are the checked ones; FP ops never raise.
__device__ __forceinline__ u32 arith(int opcode, int type, u64 a, u64 b, u64& r) {
  if (is_fp(type)) {
    const double x = u2d(a), y = u2d(b);
    double z;
    switch (opcode) {
      case VGF_ADD: z = __dadd_rn(x, y); break;
      case VGF_SUB: z = __dsub_rn(x, y); break;
      case VGF_MUL: z = __dmul_rn(x, y); break;
      case VGF_DIV: z = __ddiv_rn(x, y); break;
      case VGF_NEG: z = -x; break;
      default: z = (y == 0.0) ? __longlong_as_double(0x7ff8000000000000ll) : fmod(x, y);
    }
    r = d2u(fp_round(type, z));
    return 0;
  }
  const i64 x = (i64)a, y = (i64)b;
  i64 lo, hi;
  int_limits(type, lo, hi);
  i64 z = 0;
  u32 err = 0;
  switch (opcode) {
    case VGF_ADD:
      z = (i64)((u64)x + (u64)y);
      if (((x ^ z) & (y ^ z)) < 0) err = VGF_ERR_OVERFLOW;
      break;
    case VGF_SUB:
      z = (i64)((u64)x - (u64)y);
      if (((x ^ y) & (x ^ z)) < 0) err = VGF_ERR_OVERFLOW;
      break;
    case VGF_MUL: {
      z = (i64)((u64)x * (u64)y);
      const i64 h = __mul64hi(x, y);
      if (h != (z >> 63)) err = VGF_ERR_OVERFLOW;
      break;
    }
    case VGF_DIV:
      if (y == 0) err = VGF_ERR_DIVZERO;
      else if (x == lo && y == -1) err = VGF_ERR_OVERFLOW;
      else z = x / y;
      break;
    case VGF_NEG:
      if (x == lo) err = VGF_ERR_OVERFLOW; else z = -x;
      break;
    default:
      if (y == 0) err = VGF_ERR_DIVZERO; else z = (y == -1) ? 0 : x % y;
  }
  if (!err && (z < lo || z > hi)) err = VGF_ERR_OVERFLOW;
  r = (u64)z;
  return err;
}

// CAST from -> to (non-truncating flavour: FP -> integer rounds half away from zero, range-checks).
__device__ __forceinline__ u32 cast_value(int from, int to, u64 a, u64& r) {
  r = 0;
  if (is_fp(from)) {
    const double x = u2d(a);
    if (is_fp(to)) { r = d2u(fp_round(to, x)); return 0; }
    if (x != x) return VGF_ERR_CAST;
    const double y = round(x);
    i64 lo, hi;
    int_limits(to, lo, hi);
    if (!(y >= -9223372036854775808.0 && y < 9223372036854775808.0)) return VGF_ERR_CAST;
    const i64 z = (i64)y;
    if (z < lo || z > hi) return VGF_ERR_CAST;
    r = (u64)z;
    return 0;
  }
  const i64 x = (i64)a;
  if (is_fp(to)) { r = d2u(to == VGF_REAL ? (double)(float)x : (double)x); return 0; }
  i64 lo, hi;
  int_limits(to, lo, hi);
  if (x < lo || x > hi) return VGF_ERR_CAST;
  r = a;
  return 0;
}

__device__ __forceinline__ bool compare(int opcode, int type, u64 a, u64 b) {
  if (is_fp(type)) {
    const double x = u2d(a), y = u2d(b);
    switch (opcode) {
      case VGF_EQ: return x == y;
      case VGF_NE: return x != y;
      case VGF_LT: return x < y;
      case VGF_LE: return x <= y;
      case VGF_GT: return x > y;
      default: return x >= y;
    }
  }
  const i64 x = (i64)a, y = (i64)b;
  switch (opcode) {
    case VGF_EQ: return x == y;
    case VGF_NE: return x != y;
    case VGF_LT: return x < y;
    case VGF_LE: return x <= y;
    case VGF_GT: return x > y;
    default: return x >= y;
  }
}

// A value / boolean of one row with its NULL and latched-error flags.
struct V { u64 v; bool n, e; };
struct B { bool v, n, e; };

// SQL three-valued AND / OR; an error or NULL on one side is dropped when the other side decides.
__device__ __forceinline__ B and3(const B a, const B b) {
  const bool okA = !a.n && !a.e, okB = !b.n && !b.e;
  const bool decided = (okA && !a.v) || (okB && !b.v);
  B r;
  r.v = okA && a.v && okB && b.v;
  r.e = !decided && (a.e || b.e);
  r.n = !decided && !r.e && (a.n || b.n);
  return r;
}
__device__ __forceinline__ B or3(const B a, const B b) {
  const bool okA = !a.n && !a.e, okB = !b.n && !b.e;
  const bool decided = (okA && a.v) || (okB && b.v);
  B r;
  r.v = decided;
  r.e = !decided && (a.e || b.e);
  r.n = !decided && !r.e && (a.n || b.n);
  return r;
}

__device__ __forceinline__ u32 interleave16(u32 even, u32 odd) {
  u32 x = (even & 0xffffu) | (odd << 16);
  u32 t;
  t = (x ^ (x >> 8)) & 0x0000ff00u; x ^= t ^ (t << 8);
  t = (x ^ (x >> 4)) & 0x00f000f0u; x ^= t ^ (t << 4);
  t = (x ^ (x >> 2)) & 0x0c0c0c0cu; x ^= t ^ (t << 2);
  t = (x ^ (x >> 1)) & 0x22222222u; x ^= t ^ (t << 1);
  return x;
}
// Bits of the warp's 128-row block (4 words) from the per-thread flags m[0..3] in pair layout
// (thread rows: 2*lane, 2*lane+1, 64+2*lane, 64+2*lane+1).  n_words = words of the block that
// exist, rounded up to an even count so whole 64-bit words are written.
__device__ __forceinline__ void store_block_bits(u32* dst, i64 word_base, int n_words, const bool m[4]) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int pr = 0; pr < 2; ++pr) {
    const u32 even = __ballot_sync(0xffffffffu, m[2 * pr]);
    const u32 odd = __ballot_sync(0xffffffffu, m[2 * pr + 1]);
    if ((lane & 15) == 0) {
      const int h = lane >> 4;
      const int w = pr * 2 + h;
      if (w < ((n_words + 1) & ~1)) dst[word_base + w] = interleave16(even >> (16 * h), odd >> (16 * h));
    }
  }
}

// ---- accumulator cells --------------------------------------------------------------------------
template <int KIND>
__device__ __forceinline__ u64 cell_identity() {
  if (KIND == VGF_MINI64) return (u64)VGF_I64_MAX;
  if (KIND == VGF_MAXI64) return (u64)VGF_I64_MIN;
  if (KIND == VGF_MINF64) return 0x7fefffffffffffffull;
  if (KIND == VGF_MAXF64) return 0xffefffffffffffffull;
  return 0;
}
template <int KIND>
__device__ __forceinline__ u64 cell_comb(u64 a, u64 b, u32& err) {
  if (KIND == VGF_SUMF64) return d2u(__dadd_rn(u2d(a), u2d(b)));
  if (KIND == VGF_SUMI64) {
    const i64 x = (i64)a, y = (i64)b, z = (i64)(a + b);
    if (((x ^ z) & (y ^ z)) < 0) err |= VGF_ERR_OVERFLOW;
    return (u64)z;
  }
  if (KIND == VGF_COUNT) return a + b;
  if (KIND == VGF_MINI64) return (i64)b < (i64)a ? b : a;
  if (KIND == VGF_MAXI64) return (i64)a < (i64)b ? b : a;
  if (KIND == VGF_MINF64) return u2d(b) < u2d(a) ? b : a;
  return u2d(a) < u2d(b) ? b : a;
}

// Fixed-order fold of one accumulator over the CTA: xor-butterfly inside each warp, then warp 0
// ... in warp order through shared memory.  `scratch` holds NT/32 words per value.
template <int KIND, int NWARPS>
__device__ __forceinline__ void block_fold(u64 v, u64* scratch, u64* out, u32& err) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const u64 other = __shfl_xor_sync(0xffffffffu, v, o);
    v = cell_comb<KIND>(v, other, err);
  }
  if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    u64 acc = scratch[0];
#pragma unroll 1
    for (int w = 1; w < NWARPS; ++w) acc = cell_comb<KIND>(acc, scratch[w], err);
    *out = acc;
  }
  __syncthreads();
}

// Shared-memory table update with atomics (medium cardinality: one table per warp or per CTA).
template <int KIND>
__device__ __forceinline__ void smem_update(u64* cell, u64 v, u32& err) {
  if (KIND == VGF_COUNT) {
    atomicAdd(cell, 1ull);
  } else if (KIND == VGF_SUMF64) {
    atomicAdd(reinterpret_cast<double*>(cell), u2d(v));
  } else if (KIND == VGF_SUMI64) {
    const i64 y = (i64)v;
    const i64 x = (i64)atomicAdd(cell, v);
    const i64 z = (i64)((u64)x + (u64)y);
    if (((x ^ z) & (y ^ z)) < 0) err |= VGF_ERR_OVERFLOW;
  } else if (KIND == VGF_MINI64) {
    if ((i64)v < (i64)*reinterpret_cast<volatile u64*>(cell)) atomicMin(reinterpret_cast<i64*>(cell), (i64)v);
  } else if (KIND == VGF_MAXI64) {
    if ((i64)*reinterpret_cast<volatile u64*>(cell) < (i64)v) atomicMax(reinterpret_cast<i64*>(cell), (i64)v);
  } else {
    u64 cur = *reinterpret_cast<volatile u64*>(cell);
    while (KIND == VGF_MINF64 ? u2d(v) < u2d(cur) : u2d(cur) < u2d(v)) {
      const u64 seen = atomicCAS(cell, cur, v);
      if (seen == cur) break;
      cur = seen;
    }
  }
}

// Claim of an empty slot, executed only when `pred`: CAS(addr, EMPTY -> desired).  `seen` must hold VGF_EMPTY_KEY on
// entry (from empty_marker(), a register of its own) and holds the slot's previous key afterwards.  The register is both
// the compare operand and the result, which is the form the hardware instruction has: no copy after the atomic, so
// several claims issue back to back and their round trips overlap.
__device__ __forceinline__ u64 empty_marker() {
  u64 e;
  asm volatile("mov.b64 %0, 0xFFFFFFFFFFFFFFFF;" : "=l"(e));
  return e;
}
__device__ __forceinline__ void claim_if(bool pred, u64* addr, u64 desired, u64& seen) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %3, 0;\n\t@p atom.global.cas.b64 %0, [%1], %0, %2;\n\t}"
      : "+l"(seen)
      : "l"(addr), "l"(desired), "r"((u32)pred));
}

// ---- HBM hash table (same layout as device_common.cuh) -------------------------------------------
__device__ __forceinline__ u64 mix64(u64 k) {
  k ^= k >> 30; k *= 0xbf58476d1ce4e5b9ull;
  k ^= k >> 27; k *= 0x94d049bb133111ebull;
  k ^= k >> 31;
  return k;
}
__device__ __forceinline__ i64 f64_ordered(u64 bits) {
  const i64 s = (i64)bits;
  return s >= 0 ? s : (i64)(bits ^ 0x7fffffffffffffffull);
}
__device__ __forceinline__ u64 home_slot(const FusedParams& p, u64 key) {
  const u64 d = key - p.key_min;
  if (p.key_mode == 3) { // bucketed placement: see internal.h KeyMap
    const u32 b = p.key_pad;
    const u64 g = mix64(key >> b), lowmask = (1ull << b) - 1ull;
    return (((g << b) & p.cap_mask) & ~lowmask) | ((key + (g >> 56)) & lowmask);
  }
  if (p.key_mode != 0 && d <= p.key_span) return p.key_mode == 1 ? d : __umul64hi(d, p.key_scale);
  return mix64(key) & p.cap_mask;
}
// Next slot of a probe sequence.  Bucketed placement first steps bucket by bucket (keeping the
// position inside the bucket, so a displaced run of neighbouring keys stays together), then
// falls back to slot-by-slot so that every slot is eventually visited.
#define VGF_BUCKET_STEPS 24ull
__device__ __forceinline__ u64 next_slot(const FusedParams& p, u64 h, u64 probes) {
  const u64 step = (p.key_mode == 3 && probes < VGF_BUCKET_STEPS) ? (1ull << p.key_pad) : 1ull;
  return (h + step) & p.cap_mask;
}
template <int ROW_WORDS>
__device__ __forceinline__ i64 hash_find_or_insert(const FusedParams& p, u64 key, bool isnull, u32& err) {
  u64* table = p.table;
  if (isnull || key == VGF_EMPTY_KEY) {
    const i64 slot = (i64)p.cap_mask + (isnull ? 1 : 2);
    u64* kp = table + slot * ROW_WORDS;
    if (*reinterpret_cast<volatile u64*>(kp) == VGF_EMPTY_KEY) {
      if (atomicCAS(kp, VGF_EMPTY_KEY, 0ull) == VGF_EMPTY_KEY) atomicAdd(p.n_groups_ctr, 1ull);
    }
    return slot;
  }
  u64 h = home_slot(p, key);
  for (u64 probes = 0; probes <= p.cap_mask + VGF_BUCKET_STEPS; ++probes) {
    u64* kp = table + h * ROW_WORDS;
    u64 cur = *reinterpret_cast<volatile u64*>(kp);
    if (cur == key) return (i64)h;
    if (cur == VGF_EMPTY_KEY) {
      cur = atomicCAS(kp, VGF_EMPTY_KEY, key);
      if (cur == VGF_EMPTY_KEY) { atomicAdd(p.n_groups_ctr, 1ull); return (i64)h; }
      if (cur == key) return (i64)h;
    }
    h = next_slot(p, h, probes);
  }
  err |= VGF_ERR_TABLEFULL;
  return -1;
}
// Find-or-insert for up to four keys of one thread at once.  Phase 1 reads the home slot of every
// active key (independent loads, all in flight together); phase 2 claims the empty ones with CAS
// (again all in flight together); whatever is still unresolved (collision, lost race, special
// keys) walks the table one key at a time.  `inserted` counts new groups; the caller adds it to
// the group counter once per warp.
template <int ROW_WORDS>
__device__ __forceinline__ i64 hash_walk(const FusedParams& p, u64 key, u64 h, u32& inserted, u32& err) {
  u64* table = p.table;
  for (u64 probes = 0; probes <= p.cap_mask + VGF_BUCKET_STEPS; ++probes) {
    u64* kp = table + h * ROW_WORDS;
    u64 cur = __ldcg(kp);
    if (cur == key) return (i64)h;
    if (cur == VGF_EMPTY_KEY) {
      cur = atomicCAS(kp, VGF_EMPTY_KEY, key);
      if (cur == VGF_EMPTY_KEY) { ++inserted; return (i64)h; }
      if (cur == key) return (i64)h;
    }
    h = next_slot(p, h, probes);
  }
  err |= VGF_ERR_TABLEFULL;
  return -1;
}
template <int ROW_WORDS>
__device__ __forceinline__ void hash_probe4(const FusedParams& p, const u64 key[4], const bool isnull[4],
                                            const bool active[4], const u64 home[4], i64 slot[4], bool parked[4],
                                            u32& inserted, u32& err) {
  u64* table = p.table;
  u64 h[4], seen[4];
  bool pend[4];
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    pend[s] = active[s] && !isnull[s] && key[s] != VGF_EMPTY_KEY;
    h[s] = pend[s] ? home[s] : 0ull; // home_slot(p, key[s]), computed by the caller (it also prefetched that line)
    slot[s] = -1;
  }
  // Round 1 claims the home slots directly (one atomic round trip, all four in flight together;
  // the caller prefetched the lines into L2 while the run aggregation was computing).  The CAS
  // returns the resident key when the slot is taken, which is all a lookup needs: keys never
  // change once set.
  // (predicated PTX, not `if (pend) atomicCAS`: with branches the compiler gave all four CAS the same result
  //  register and each one waited for the previous round trip — 28 % of the kernel's stall samples)
#pragma unroll
  for (int s = 0; s < 4; ++s) seen[s] = empty_marker();
#pragma unroll
  for (int s = 0; s < 4; ++s) claim_if(pend[s], table + h[s] * ROW_WORDS, key[s], seen[s]);
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    if (!pend[s]) continue;
    if (seen[s] == VGF_EMPTY_KEY) ++inserted;
    if (seen[s] == VGF_EMPTY_KEY || seen[s] == key[s]) { slot[s] = (i64)h[s]; pend[s] = false; }
    else h[s] = next_slot(p, h[s], 0ull);
  }
  // NULL key / the key that equals the empty marker: dedicated slots past the table
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    if (!active[s] || !(isnull[s] || key[s] == VGF_EMPTY_KEY)) continue;
    const i64 sl = (i64)p.cap_mask + (isnull[s] ? 1 : 2);
    u64* kp = table + sl * ROW_WORDS;
    if (__ldcg(kp) == VGF_EMPTY_KEY) {
      if (atomicCAS(kp, VGF_EMPTY_KEY, 0ull) == VGF_EMPTY_KEY) ++inserted;
    }
    slot[s] = sl;
  }
  // A warp waits for its slowest lane, and at 50 % load a few of its ~32 claims always need several
  // more round trips.  With a parking buffer those segments leave the scan here: the caller stores
  // {key, partials} and drain_overflow_kernel (agg.cu) inserts them with one segment per thread.
#pragma unroll
  for (int s = 0; s < 4; ++s) parked[s] = pend[s] && p.ovf != nullptr;
  if (p.ovf != nullptr) return;
  // Later rounds (collisions): one CAS per pending segment per round, again all in flight
  // together, so a thread pays one round trip per probe step, not one per segment and step.
  u64 steps = 1;
  while (pend[0] || pend[1] || pend[2] || pend[3]) {
#pragma unroll
    for (int s = 0; s < 4; ++s) seen[s] = empty_marker();
#pragma unroll
    for (int s = 0; s < 4; ++s) claim_if(pend[s], table + h[s] * ROW_WORDS, key[s], seen[s]);
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      if (!pend[s]) continue;
      if (seen[s] == VGF_EMPTY_KEY) ++inserted;
      if (seen[s] == VGF_EMPTY_KEY || seen[s] == key[s]) { slot[s] = (i64)h[s]; pend[s] = false; }
      else h[s] = next_slot(p, h[s], steps);
    }
    if (++steps > p.cap_mask + VGF_BUCKET_STEPS) { err |= VGF_ERR_TABLEFULL; break; }
  }
}

// Entry numbers in the parking buffer for the warp's unresolved segments (one atomic per warp).
// Must be called by all 32 lanes.  Numbers >= p.ovf_cap mean "buffer full": the caller walks the
// table itself for those.
__device__ __forceinline__ void ovf_reserve(const FusedParams& p, const bool parked[4], i64 entry[4]) {
  const u32 lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
  u32 m[4], total = 0;
#pragma unroll
  for (int s = 0; s < 4; ++s) { m[s] = __ballot_sync(0xffffffffu, parked[s]); total += __popc(m[s]); }
#pragma unroll
  for (int s = 0; s < 4; ++s) entry[s] = -1;
  if (total == 0) return;
  u64 base = 0;
  if (lane == 0) base = atomicAdd(p.ovf_cursor, (u64)total);
  base = __shfl_sync(0xffffffffu, base, 0);
  u32 off = 0;
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    if (parked[s]) entry[s] = (i64)(base + off + __popc(m[s] & lt));
    off += __popc(m[s]);
  }
}

// ---- LOG sink ---------------------------------------------------------------------------------------
// Places of the warp's active segments inside the CTA's log region.  kind[s]: 0 = none, 1 = front (a run that cannot
// continue in another block), 2 = back (a run touching a block edge).  Front entries fill the region from its start,
// back entries from its end; off[s] is the entry's offset from the region start.  Per-lane counts of both sides ride
// in one word through one warp scan; one shared-memory atomic per warp and side.
template <int N>
__device__ __forceinline__ void log_reserve(u32* cur /* [2] in shared memory */, u32 cap, const u32 (&kind)[N], u32 (&off)[N]) {
  const u32 lane = threadIdx.x & 31u;
  u32 mine = 0;
#pragma unroll
  for (int s = 0; s < N; ++s) mine += (kind[s] & 1u) + ((kind[s] & 2u) << 15);
  u32 incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const u32 u = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= (u32)d) incl += u;
  }
  const u32 tot = __shfl_sync(0xffffffffu, incl, 31);
  u32 bf = 0, bb = 0;
  if (lane == 0) {
    if (tot & 0xffffu) bf = atomicAdd(&cur[0], tot & 0xffffu);
    if (tot >> 16) bb = atomicAdd(&cur[1], tot >> 16);
  }
  bf = __shfl_sync(0xffffffffu, bf, 0);
  bb = __shfl_sync(0xffffffffu, bb, 0);
  const u32 excl = incl - mine;
  u32 f = bf + (excl & 0xffffu), b = cap - 1u - (bb + (excl >> 16));
#pragma unroll
  for (int s = 0; s < N; ++s) {
    off[s] = (kind[s] & 1u) ? f : b;
    f += kind[s] & 1u;
    b -= kind[s] >> 1;
  }
}

// ---- LOG sink with the exchange fused in ----------------------------------------------------------
// Entries do not enter a local log: FRONT entries (complete runs) are compacted into a per-warp ring in shared memory
// (kSendRing entries), taken from there 32 at a time, hashed to their owner and appended to a per-owner ring
// (kSendBuf = 3 runs of kSendRun entries); a full run leaves as one aligned store of kSendRun entries into the owner's
// peer-mapped window (measured on 8 GPUs: 64-byte runs reach ~1/3 of the NVLink rate, aligned 256-byte runs ~2/3).
// BACK entries (the runs at the two ends of a warp's range: rare) are stored one by one at the end of the sub-region.
// Per CTA and owner: sfront / sback count the entries stored so far, sbase is the CTA's sub-region in that window.
// The last kSendBackRoom entries of a sub-region are kept for back entries, so that both sides check their own
// counter against a constant (no second counter to read on the hot path).
constexpr u32 kSendRing = 192, kSendRun = 16, kSendBuf = 48, kSendBackRoom = 256;

// Places (after `pend` pending entries) of the lane's front segments in the warp's ring; returns how many were added.
template <int N>
__device__ __forceinline__ u32 warp_compact_front(const u32 (&kind)[N], u32 (&off)[N]) {
  const u32 lane = threadIdx.x & 31u;
  u32 mine = 0;
#pragma unroll
  for (int s = 0; s < N; ++s) mine += kind[s] & 1u;
  u32 incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const u32 u = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= (u32)d) incl += u;
  }
  u32 at = incl - mine;
#pragma unroll
  for (int s = 0; s < N; ++s) { off[s] = at; at += kind[s] & 1u; }
  return __shfl_sync(0xffffffffu, incl, 31);
}
// Owner of a group among W ranks: device_common.cuh partition_of with W known at compile time (VELOX: one raw BIGINT
// key hashed like folly::hasher<int64_t>, velox/exec/VectorHasher.cpp:68-108; else splitmix64 of the packed key).
template <int W, bool VELOX>
__device__ __forceinline__ u32 send_dest(u64 key) {
  u64 k = key;
  if (VELOX) {
    k = (~k) + (k << 21);
    k = k ^ (k >> 24);
    k = k + (k << 3) + (k << 8);
    k = k ^ (k >> 14);
    k = k + (k << 2) + (k << 4);
    k = k ^ (k >> 28);
    k = k + (k << 31);
  } else {
    k ^= k >> 30; k *= 0xbf58476d1ce4e5b9ull;
    k ^= k >> 27; k *= 0x94d049bb133111ebull;
    k ^= k >> 31;
  }
  return (u32)(k % (u64)W);
}
// One BACK entry, straight to the end of its owner's sub-region (called by the one lane that holds it).
template <int EW, int W, bool VELOX>
__device__ __noinline__ void send_back(u32 cap, u32* sfront, u32* sback, u64* const* sbase, const u64* ent, u32& err) {
  const u32 d = send_dest<W, VELOX>(ent[0]);
  const u32 back = atomicAdd(&sback[d], 1u);
  if (back >= kSendBackRoom) { err |= VGF_ERR_TABLEFULL; return; } // more range ends than a sub-region keeps room for
  u64* out = sbase[d] + (size_t)(cap - 1u - back) * EW;
#pragma unroll
  for (int w = 0; w < EW; ++w) out[w] = ent[w];
}
// Stores `go` (<= 32) entries of owner d's ring, starting at `head`, behind the front entries sent to d so far.
template <int EW>
__device__ __forceinline__ void send_run(u32 cap, u32* sfront, const u32* sback, u64* const* sbase, const u64* buf, u32 d, u32 head, u32 go, u32& err) {
  const u32 lane = threadIdx.x & 31u;
  u32 front = 0;
  if (lane == 0) front = atomicAdd(&sfront[d], go);
  front = __shfl_sync(0xffffffffu, front, 0);
  if (front + go > cap - kSendBackRoom) { err |= VGF_ERR_TABLEFULL; return; } // the window is too small for this CTA's share
  if (lane < go) {
    u32 src = head + lane;
    if (src >= kSendBuf) src -= kSendBuf;
    u64* out = sbase[d] + (size_t)(front + lane) * EW;
    const u64* e = buf + (size_t)src * EW;
    if (EW == 2) *reinterpret_cast<ulonglong2*>(out) = *reinterpret_cast<const ulonglong2*>(e);
    else { _Pragma("unroll") for (int w = 0; w < EW; ++w) out[w] = e[w]; }
  }
}
// Takes `n` (<= 32) entries from the warp's ring at `beg`, routes them to the owners' rings (wcnt[d] = head | fill << 8)
// and sends every full run.
template <int EW, int W, bool VELOX>
__device__ __forceinline__ void send_step(u32 cap, u32* sfront, const u32* sback, u64* const* sbase, const u64* wl, u64* wbuf, u32* wcnt,
                                          u32 beg, u32 n, u32& err) {
  const u32 lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
  const bool act = lane < n;
  u32 idx = beg + lane;
  if (idx >= kSendRing) idx -= kSendRing;
  u64 ent[EW];
#pragma unroll
  for (int w = 0; w < EW; ++w) ent[w] = 0ull;
  if (act) {
    if (EW == 2) { const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(wl + (size_t)idx * 2); ent[0] = v.x; ent[1] = v.y; }
    else { _Pragma("unroll") for (int w = 0; w < EW; ++w) ent[w] = wl[(size_t)idx * EW + w]; }
  }
  const u32 dest = send_dest<W, VELOX>(ent[0]);
  const u32 m = __match_any_sync(0xffffffffu, act ? dest : 0xffffffffu);
  if (act) {
    const u32 st = wcnt[dest];
    u32 at = (st & 0xffu) + (st >> 8) + (u32)__popc(m & lt);
    if (at >= kSendBuf) at -= kSendBuf;
    u64* b = wbuf + ((size_t)dest * kSendBuf + at) * EW;
    if (EW == 2) *reinterpret_cast<ulonglong2*>(b) = make_ulonglong2(ent[0], ent[1]);
    else { _Pragma("unroll") for (int w = 0; w < EW; ++w) b[w] = ent[w]; }
  }
  __syncwarp();
  if (act && (m & lt) == 0u) wcnt[dest] += (u32)__popc(m) << 8; // the first lane of each owner's group
  __syncwarp();
  const u32 mine = lane < (u32)W ? wcnt[lane] : 0u;
  u32 full = __ballot_sync(0xffffffffu, (mine >> 8) >= kSendRun);
  while (full) {
    const u32 d = (u32)__ffs((int)full) - 1u;
    full &= full - 1u;
    const u32 st = __shfl_sync(0xffffffffu, mine, (int)d), head = st & 0xffu, fill = st >> 8;
    const u32 go = fill & ~(kSendRun - 1u); // whole runs only: the owner's cursor stays aligned to the run size
    send_run<EW>(cap, sfront, sback, sbase, wbuf + (size_t)d * kSendBuf * EW, d, head, go, err);
    if (lane == d) { u32 h = head + go; if (h >= kSendBuf) h -= kSendBuf; wcnt[d] = h | ((fill - go) << 8); }
  }
  __syncwarp();
}
// End of the kernel: the pending entries of the warp's ring, then whatever is left in the owners' rings.
template <int EW, int W, bool VELOX>
__device__ __forceinline__ void send_drain(u32 cap, u32* sfront, const u32* sback, u64* const* sbase, const u64* wl, u64* wbuf, u32* wcnt,
                                           u32 beg, u32 pend, u32& err) {
  const u32 lane = threadIdx.x & 31u;
  if (pend) send_step<EW, W, VELOX>(cap, sfront, sback, sbase, wl, wbuf, wcnt, beg, pend, err);
  __syncwarp();
  const u32 mine = lane < (u32)W ? wcnt[lane] : 0u;
  u32 left = __ballot_sync(0xffffffffu, (mine >> 8) > 0u);
  while (left) {
    const u32 d = (u32)__ffs((int)left) - 1u;
    left &= left - 1u;
    const u32 st = __shfl_sync(0xffffffffu, mine, (int)d), head = st & 0xffu, fill = st >> 8; // < kSendRun after the step
    send_run<EW>(cap, sfront, sback, sbase, wbuf + (size_t)d * kSendBuf * EW, d, head, fill, err);
    if (lane == d) wcnt[d] = 0u;
  }
  __syncwarp();
}

template <int KIND>
__device__ __forceinline__ void hbm_update(u64* addr, u64 v, u32& err) {
  if (KIND == VGF_SUMF64) {
    atomicAdd(reinterpret_cast<double*>(addr), u2d(v));
  } else if (KIND == VGF_SUMI64) {
    const i64 y = (i64)v;
    const i64 x = (i64)atomicAdd(addr, v);
    const i64 z = (i64)((u64)x + (u64)y);
    if (((x ^ z) & (y ^ z)) < 0) err |= VGF_ERR_OVERFLOW;
  } else if (KIND == VGF_COUNT) {
    atomicAdd(addr, v);
  } else if (KIND == VGF_MINI64) {
    atomicMin(reinterpret_cast<i64*>(addr), (i64)v);
  } else if (KIND == VGF_MAXI64) {
    atomicMax(reinterpret_cast<i64*>(addr), (i64)v);
  } else if (KIND == VGF_MINF64) {
    if (u2d(v) == u2d(v)) atomicMin(reinterpret_cast<i64*>(addr), f64_ordered(v));
  } else {
    if (u2d(v) == u2d(v)) atomicMax(reinterpret_cast<i64*>(addr), f64_ordered(v));
  }
}

} // namespace vgf
#endif // VGPU_FUSED_PRELUDE_CUH_
