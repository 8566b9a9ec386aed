// Device helpers shared by the pipeline kernel and the group-state kernels (agg.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "internal.h"

namespace vg {

__device__ __forceinline__ uint64_t d2u(double d) { return (uint64_t)__double_as_longlong(d); }
__device__ __forceinline__ double u2d(uint64_t u) { return __longlong_as_double((long long)u); }

struct HashRef {
  uint64_t* table; // (capacity + 2) rows of row_words words: key, cells...
  uint64_t cap_mask;
  KeyMap key_map;
  int32_t row_words;
  unsigned long long* n_groups_ctr;
};

__device__ __forceinline__ uint64_t mix64(uint64_t k) { // splitmix64 finaliser
  k ^= k >> 30; k *= 0xbf58476d1ce4e5b9ull;
  k ^= k >> 27; k *= 0x94d049bb133111ebull;
  k ^= k >> 31;
  return k;
}
constexpr uint64_t kEmptyKey = 0xffffffffffffffffull;

// folly::hasher<int64_t> = folly::hash::twang_mix64 (Thomas Wang's 64-bit mix; folly is not vendored in the reference
// tree, the same function is restated in velox/experimental/gpu/tests/HashTableTest.cu:51-60).
__host__ __device__ __forceinline__ uint64_t twang_mix64(uint64_t key) {
  key = (~key) + (key << 21);
  key = key ^ (key >> 24);
  key = key + (key << 3) + (key << 8);
  key = key ^ (key >> 14);
  key = key + (key << 2) + (key << 4);
  key = key ^ (key >> 28);
  key = key + (key << 31);
  return key;
}
// Destination of a group between a partial and a final aggregation: hash % partitions
// (velox/exec/HashPartitionFunction.cpp:59-92).  velox_hash: the key is one raw BIGINT column, hashed exactly as
// VectorHasher::hash does (velox/exec/VectorHasher.cpp:68-108: folly::hasher<int64_t> per value, NULL -> kNullHash = 1,
// velox/vector/BaseVector.h:74) so that a GPU partial stage assigns every group to the partition a CPU Velox stage
// would.  Otherwise the key is a packed normalized key / a dense group number, which the reference never hashes as
// such: any deterministic function does, and splitmix64 is used.
__host__ __device__ __forceinline__ uint32_t partition_of(uint64_t key, uint64_t null_flag, uint32_t n_parts, bool velox_hash) {
  if (velox_hash) return (uint32_t)((null_flag ? 1ull : twang_mix64(key)) % (uint64_t)n_parts);
  uint64_t k = key ^ (null_flag << 63);
  k ^= k >> 30; k *= 0xbf58476d1ce4e5b9ull;
  k ^= k >> 27; k *= 0x94d049bb133111ebull;
  k ^= k >> 31;
  return (uint32_t)(k % (uint64_t)n_parts);
}

// Warp-aggregated atomics.  Millions of threads bumping ONE counter serialise in the L2 atomic unit
// (15 M increments ~ 7 ms on B200); the lanes that reach the call together elect a leader that adds
// their count once.  Each lane is in exactly one such group, so the total is exact.
__device__ __forceinline__ void count_one(unsigned long long* ctr) {
  const unsigned m = __activemask();
  if ((int)(threadIdx.x & 31) == __ffs(m) - 1) atomicAdd(ctr, (unsigned long long)__popc(m));
}
// Same, returning this lane's own position in the sequence the counter hands out.
__device__ __forceinline__ unsigned long long take_one(unsigned long long* ctr) {
  const unsigned m = __activemask();
  const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
  unsigned long long base = 0;
  if (lane == leader) base = atomicAdd(ctr, (unsigned long long)__popc(m));
  base = __shfl_sync(m, base, leader);
  return base + __popc(m & ((1u << lane) - 1u));
}

__device__ __forceinline__ uint64_t home_slot(const KeyMap& m, uint64_t cap_mask, uint64_t key) {
  if (m.mode == 3) {
    const uint32_t b = m.pad;
    const uint64_t g = mix64(key >> b), lowmask = (1ull << b) - 1ull;
    return (((g << b) & cap_mask) & ~lowmask) | ((key + (g >> 56)) & lowmask);
  }
  const uint64_t d = key - m.min;
  if (m.mode != 0 && d <= m.span) return m.mode == 1 ? d : __umul64hi(d, m.scale);
  return mix64(key) & cap_mask;
}
constexpr uint64_t kBucketSteps = 24; // same probe sequence as fused_prelude.cuh next_slot()
__device__ __forceinline__ uint64_t next_slot(const KeyMap& m, uint64_t cap_mask, uint64_t h, uint64_t probes) {
  const uint64_t step = (m.mode == 3 && probes < kBucketSteps) ? (1ull << m.pad) : 1ull;
  return (h + step) & cap_mask;
}

// Order-preserving int64 image of a double (for atomicMin/Max on doubles in HBM).
__device__ __forceinline__ int64_t f64_ordered(uint64_t bits) {
  const int64_t s = (int64_t)bits;
  return s >= 0 ? s : (int64_t)(bits ^ 0x7fffffffffffffffull);
}


__device__ __forceinline__ uint64_t cell_identity(uint8_t kind) {
  switch (kind) {
    case kCellMinI64: return (uint64_t)INT64_MAX;
    case kCellMaxI64: return (uint64_t)INT64_MIN;
    case kCellMinF64: return 0x7fefffffffffffffull;  // DBL_MAX  (numeric_limits::max())
    case kCellMaxF64: return 0xffefffffffffffffull;  // -DBL_MAX (numeric_limits::lowest())
    default: return 0;
  }
}
__device__ __forceinline__ int cell_width(uint8_t kind) { return kind == kCellCount ? 4 : 8; }


__device__ __forceinline__ uint64_t fold2(uint8_t kind, uint64_t a, uint64_t b, uint32_t& err) {
  switch (kind) {
    case kCellSumF64: return d2u(__dadd_rn(u2d(a), u2d(b)));
    case kCellSumI64: {
      const int64_t x = (int64_t)a, y = (int64_t)b, z = (int64_t)(a + b);
      if (((x ^ z) & (y ^ z)) < 0) err |= kErrOverflow;
      return (uint64_t)z;
    }
    case kCellCount: return a + b;
    case kCellMinI64: return (int64_t)b < (int64_t)a ? b : a;
    case kCellMaxI64: return (int64_t)a < (int64_t)b ? b : a;
    case kCellMinF64: return u2d(b) < u2d(a) ? b : a;
    default: return u2d(a) < u2d(b) ? b : a;
  }
}


// HBM open-addressing table: rows of row_words 8-byte words {key, cells...}; slots
// [0, capacity) are probed linearly from mix64(key); slot `capacity` holds the NULL-key group
// and slot capacity+1 the group whose raw key equals the empty marker.
__device__ __forceinline__ int64_t hash_find_or_insert(const HashRef& p, uint64_t key, bool isnull,
                                                       uint32_t& err) {
  uint64_t* table = p.table;
  const int rw = p.row_words;
  if (isnull || key == kEmptyKey) {
    const int64_t slot = (int64_t)p.cap_mask + (isnull ? 1 : 2);
    unsigned long long* kp = reinterpret_cast<unsigned long long*>(table + slot * rw);
    if (*reinterpret_cast<volatile unsigned long long*>(kp) == kEmptyKey) {
      if (atomicCAS(kp, kEmptyKey, 0ull) == kEmptyKey) count_one(p.n_groups_ctr);
    }
    return slot;
  }
  uint64_t h = home_slot(p.key_map, p.cap_mask, key);
  for (uint64_t probes = 0; probes <= p.cap_mask + kBucketSteps; ++probes) {
    unsigned long long* kp = reinterpret_cast<unsigned long long*>(table + h * rw);
    unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(kp);
    if (cur == key) return (int64_t)h;
    if (cur == kEmptyKey) {
      cur = atomicCAS(kp, kEmptyKey, (unsigned long long)key);
      if (cur == kEmptyKey) {
        count_one(p.n_groups_ctr);
        return (int64_t)h;
      }
      if (cur == key) return (int64_t)h;
    }
    h = next_slot(p.key_map, p.cap_mask, h, probes);
  }
  err |= kErrTableFull;
  return -1;
}

__device__ __forceinline__ void update_hbm_cell(uint8_t kind, uint64_t* addr, uint64_t v, uint32_t& err) {
  switch (kind) {
    case kCellSumF64:
      atomicAdd(reinterpret_cast<double*>(addr), u2d(v));
      break;
    case kCellSumI64: {
      // The returned old value gives each add its place in one serial order of the adds, so the
      // add that wraps is the one that reports (SumAggregateBase.h:150-158 checks every add).
      const int64_t y = (int64_t)v;
      const int64_t x = (int64_t)atomicAdd(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)v);
      const int64_t z = (int64_t)((uint64_t)x + (uint64_t)y);
      if (((x ^ z) & (y ^ z)) < 0) err |= kErrOverflow;
      break;
    }
    case kCellCount:
      atomicAdd(reinterpret_cast<unsigned long long*>(addr), 1ull);
      break;
    case kCellMinI64:
      atomicMin(reinterpret_cast<long long*>(addr), (long long)v);
      break;
    case kCellMaxI64:
      atomicMax(reinterpret_cast<long long*>(addr), (long long)v);
      break;
    case kCellMinF64:
      if (u2d(v) == u2d(v)) atomicMin(reinterpret_cast<long long*>(addr), (long long)f64_ordered(v));
      break;
    default:
      if (u2d(v) == u2d(v)) atomicMax(reinterpret_cast<long long*>(addr), (long long)f64_ordered(v));
  }
}


} // namespace vg
