// HASH aggregation, LOG engine: high-cardinality group-by without a table in HBM.
//
// The scan kernels (fused.cu, LOG sink) fold runs of equal keys in registers and append one
// {normalized key, cells...} entry per run to per-CTA regions of a log — streaming stores, no global
// atomics.  This file turns the log into groups:
//   1. fin_prepare   reads the "input was sorted by key" flag the scans maintained.  When it holds, every
//                    front entry is a whole group already and only the back entries (runs that touch a block
//                    edge) are deduplicated; otherwise every entry is.
//   2. fin_hist / fin_scan / fin_scatter   radix-partition the entries to deduplicate by hash bits into
//                    P partitions (exact offsets from a histogram: nothing can overflow).
//   3. fin_aggregate one CTA per partition folds its entries in a shared-memory open-addressing table
//                    (shared-memory atomics) and appends the partition's groups to a dense array.
// Every pass is a sequential read and write of 16-byte entries (C3), so the engine is bandwidth bound where the
// HBM open-addressing table (agg.cu, the TABLE engine) paid a random 32-byte sector read-modify-write per group
// plus the initialisation of a table sized for a 25 % load.
//
// Reference behaviour restated: find-or-insert by key equality with NULL as a key value and accumulator
// folding (velox/exec/GroupingSet.cpp:240-307, velox/exec/HashTable.cpp:426-562, Aggregate::addIntermediateResults,
// velox/exec/Aggregate.h:305-337).  The hash function and the group order are unobservable.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "agg.h"
#include "device_common.cuh"
#include "fused.h"

namespace vg {

enum : uint32_t { kRangeFront = 0, kRangeBack = 1, kRangePlain = 2 };
constexpr uint32_t kChunk = 1024; // entries one CTA takes from a range at a time (deduplication and expansion passes):
                                  // small, so that the few thousand edge entries of a sorted input spread over many CTAs

__device__ __forceinline__ uint64_t range_count(const LogRange& r) {
  return r.count64 ? (uint64_t) * reinterpret_cast<const unsigned long long*>(r.count) : (uint64_t) * reinterpret_cast<const uint32_t*>(r.count);
}
__device__ __forceinline__ const uint64_t* range_first(const LogRange& r, uint64_t cnt, int ew) {
  return r.back == kRangeBack ? r.base + (r.cap - cnt) * ew : r.base;
}
__device__ __forceinline__ uint32_t part_of_key(uint64_t key, int log2p) { return (uint32_t)(mix64(key) >> (64 - log2p)); }

struct FinArgs {
  const LogRange* ranges;
  int n_ranges;
  int ew;
  int log2p_max;                 // the arrays hold 2^log2p_max partitions; the device picks log2p <= log2p_max from n_dedupe
  int probed;                    // host: every batch ran the sorted-input probe
  int slots_max;                 // shared-memory table slots of fin_aggregate (power of two)
  const unsigned long long* sort_state;
  const unsigned long long* src_meta; // received streams: {sorted, first key, last key, entries} per source, or nullptr
  int n_src, src_stride, local_entries; // local_entries: this aggregation logged input of its own as well
  FinHeader* hdr;
  unsigned long long* comb;      // Agg::ctr + kCtrCombine (see agg.h)
  int combine_ok;                // host: the pre-combining pass may run (its table fits shared memory)
  uint32_t* hist;                // [P + 1]
  uint32_t* offsets;             // [P + 1]
  uint32_t* cursors;             // [P]
  uint32_t* chunk_prefix;        // [n_ranges + 1]: chunks of kChunk entries before range r, over the ranges to deduplicate
  uint64_t* part;
  uint64_t* groups;
  uint32_t* err;
  CellKinds kinds;               // kinds of the entry's cells, in entry order
};

// Exclusive prefix of the per-range chunk counts (one CTA): chunk_prefix[r] = chunks before range r.
// take(r, count) says whether range r takes part.
template <typename Take>
__device__ __forceinline__ void chunk_prefix_scan(const LogRange* ranges, int n_ranges, uint32_t* chunk_prefix, uint32_t chunk_entries, Take&& take) {
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < n_ranges; base += blockDim.x) {
    const int r = base + threadIdx.x;
    uint32_t v = 0;
    if (r < n_ranges) {
      const uint64_t c = range_count(ranges[r]);
      if (take(ranges[r])) v = (uint32_t)((c + chunk_entries - 1) / chunk_entries);
    }
    uint32_t x = v;
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((int)(threadIdx.x & 31) >= o) x += y; }
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      uint32_t w = threadIdx.x < (blockDim.x >> 5) ? s_warp[threadIdx.x] : 0u;
      for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, w, o); if ((int)threadIdx.x >= o) w += y; }
      s_warp[threadIdx.x] = w;
    }
    __syncthreads();
    const uint32_t before = s_carry + ((threadIdx.x >> 5) ? s_warp[(threadIdx.x >> 5) - 1] : 0u) + x - v;
    if (r < n_ranges) chunk_prefix[r] = before;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) s_carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) chunk_prefix[n_ranges] = s_carry;
}

__device__ __forceinline__ void locate_chunk(const LogRange* ranges, const uint32_t* chunk_prefix, int n_ranges, uint32_t chunk, int ew,
                                             const uint64_t*& first, uint64_t& n);
constexpr int kSampleChunks = 16;

__global__ void fin_prepare_kernel(FinArgs a) {
  __shared__ unsigned long long s_front[32], s_other[32];
  bool sorted = a.probed && *a.sort_state == 0ull;
  if (a.n_src > 0) {
    // received streams: every one sorted, pairwise disjoint key ranges, nothing logged locally
    sorted = a.local_entries == 0;
    for (int s = 0; s < a.n_src && sorted; ++s) {
      const unsigned long long* m = a.src_meta + (size_t)s * a.src_stride;
      if (m[3] == 0ull) continue;
      if (m[0] == 0ull || m[1] > m[2]) { sorted = false; break; }
      for (int t = 0; t < s; ++t) {
        const unsigned long long* o = a.src_meta + (size_t)t * a.src_stride;
        if (o[3] != 0ull && !(m[1] > o[2] || o[1] > m[2])) { sorted = false; break; }
      }
    }
  }
  unsigned long long front = 0, other = 0;
  for (int r = threadIdx.x; r < a.n_ranges; r += blockDim.x) {
    const uint64_t c = range_count(a.ranges[r]);
    if (a.ranges[r].back == kRangeFront) front += c; else other += c;
  }
  for (int o = 16; o > 0; o >>= 1) { front += __shfl_xor_sync(0xffffffffu, front, o); other += __shfl_xor_sync(0xffffffffu, other, o); }
  if ((threadIdx.x & 31) == 0) { s_front[threadIdx.x >> 5] = front; s_other[threadIdx.x >> 5] = other; }
  __syncthreads();
  if (threadIdx.x == 0) {
    front = other = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { front += s_front[w]; other += s_other[w]; }
    a.hdr->sorted = sorted ? 1ull : 0ull;
    a.hdr->n_front = front;
    const unsigned long long n_dedupe = sorted ? other : front + other;
    a.hdr->n_dedupe = n_dedupe;
    a.hdr->n_groups = 0ull;
    a.hdr->overflow = 0ull;
    // partitions: even if every entry were a group of its own the shared-memory tables stay below ~45 % load
    // and at most ~256 entries per partition on average: the histogram / cursor atomics of one 32-byte sector
    // serialise in L2 (measured: 1.9 M entries into 1024 counters took 0.15 ms per pass)
    const unsigned long long per_part = min((unsigned long long)a.slots_max * 45 / 100, 256ull);
    int log2p = 4;
    while (log2p < a.log2p_max && (1ull << log2p) * per_part < n_dedupe) ++log2p;
    if (n_dedupe <= (unsigned long long)a.slots_max * 45 / 100) log2p = 0; // one table holds them all
    a.hdr->log2p = (unsigned long long)log2p;
  }
  chunk_prefix_scan(a.ranges, a.n_ranges, a.chunk_prefix, kChunk, [&](const LogRange& r) { return !(sorted && r.back == kRangeFront); });
  // Do duplicates abound among the entries to deduplicate (unsorted input of low or medium cardinality)?  Sample
  // kSampleChunks chunks spread over the ranges: distinct keys per chunk against entries per chunk.  If fewer than 60 %
  // are distinct, fin_combine_kernel folds every chunk in shared memory first, so that a key with millions of entries
  // does not end up as one partition walked by one CTA (measured: 4 keys over 60 M unsorted rows took 318 ms).
  __syncthreads();
  __shared__ unsigned long long s_keys[2 * kChunk];
  __shared__ unsigned int s_sampled, s_distinct;
  if (threadIdx.x == 0) { s_sampled = 0; s_distinct = 0; }
  const uint32_t total_chunks = a.chunk_prefix[a.n_ranges];
  const bool sample = a.combine_ok && a.hdr->n_dedupe > 65536ull && total_chunks >= (uint32_t)kSampleChunks;
  if (sample) {
    for (int sc = 0; sc < kSampleChunks; ++sc) {
      for (int i = threadIdx.x; i < (int)(2 * kChunk); i += blockDim.x) s_keys[i] = kEmptyKey;
      __syncthreads();
      const uint64_t* first;
      uint64_t n;
      locate_chunk(a.ranges, a.chunk_prefix, a.n_ranges, (uint32_t)((uint64_t)sc * total_chunks / kSampleChunks), a.ew, first, n);
      for (uint64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const uint64_t key = first[i * a.ew];
        uint32_t h = (uint32_t)mix64(key) & (2 * kChunk - 1);
        for (;;) {
          const unsigned long long seen = atomicCAS(&s_keys[h], kEmptyKey, (unsigned long long)key);
          if (seen == kEmptyKey) { atomicAdd(&s_distinct, 1u); break; }
          if (seen == key) break;
          h = (h + 1) & (2 * kChunk - 1);
        }
      }
      if (threadIdx.x == 0) s_sampled += (unsigned int)n;
      __syncthreads();
    }
  }
  if (threadIdx.x == 0) {
    a.comb[0] = sample && (unsigned long long)s_distinct * 10ull < (unsigned long long)s_sampled * 6ull ? 1ull : 0ull;
    a.comb[1] = 0ull; a.comb[2] = 0ull;
  }
}

// The chunk prefix array in shared memory when it fits (a binary search over global memory is ~10 dependent L2
// round trips per chunk, which dominated the deduplication of a sorted input's few edge entries).
constexpr int kPrefixSmem = 2048;
__device__ __forceinline__ const uint32_t* stage_prefix(const uint32_t* chunk_prefix, int n_ranges, uint32_t* s_prefix) {
  if (n_ranges + 1 > kPrefixSmem) return chunk_prefix;
  for (int i = threadIdx.x; i <= n_ranges; i += blockDim.x) s_prefix[i] = chunk_prefix[i];
  __syncthreads();
  return s_prefix;
}

// Range and entry window of chunk `chunk` (every thread of the CTA computes the same answer).
__device__ __forceinline__ void locate_chunk(const LogRange* ranges, const uint32_t* chunk_prefix, int n_ranges, uint32_t chunk, int ew,
                                             const uint64_t*& first, uint64_t& n) {
  int lo = 0, hi = n_ranges;
  while (hi - lo > 1) { // largest r with chunk_prefix[r] <= chunk (ranges without chunks share their successor's prefix)
    const int mid = (lo + hi) >> 1;
    if (chunk_prefix[mid] <= chunk) lo = mid; else hi = mid;
  }
  const LogRange rg = ranges[lo];
  const uint64_t cnt = range_count(rg);
  const uint64_t c0 = (uint64_t)(chunk - chunk_prefix[lo]) * kChunk;
  first = range_first(rg, cnt, ew) + c0 * ew;
  n = min((uint64_t)kChunk, cnt - c0);
}

// Walks every entry that needs deduplication: the CTAs share the chunks of all taking-part ranges.
template <typename F>
__device__ __forceinline__ void for_each_dedupe_entry(const FinArgs& a, F&& fn, uint32_t first_chunk, uint32_t chunk_stride) {
  if (a.comb[0] == 2ull) { // fin_combine_kernel ran: the entries to deduplicate are its output, one plain array
    const uint64_t n_all = a.comb[1];
    const uint32_t total = (uint32_t)((n_all + kChunk - 1) / kChunk);
    for (uint32_t chunk = first_chunk; chunk < total; chunk += chunk_stride) {
      const uint64_t c0 = (uint64_t)chunk * kChunk, n = min((uint64_t)kChunk, n_all - c0);
      const uint64_t* first = a.groups + c0 * a.ew;
      for (uint64_t i = threadIdx.x; i < n; i += blockDim.x) fn(first + i * a.ew);
    }
    return;
  }
  __shared__ uint32_t s_prefix[kPrefixSmem];
  const uint32_t* prefix = stage_prefix(a.chunk_prefix, a.n_ranges, s_prefix);
  const uint32_t total = prefix[a.n_ranges];
  for (uint32_t chunk = first_chunk; chunk < total; chunk += chunk_stride) {
    const uint64_t* first;
    uint64_t n;
    locate_chunk(a.ranges, prefix, a.n_ranges, chunk, a.ew, first, n);
    for (uint64_t i = threadIdx.x; i < n; i += blockDim.x) fn(first + i * a.ew);
  }
}

__device__ __forceinline__ void smem_fold(uint8_t kind, unsigned long long* cell, uint64_t v, uint32_t& err);

// Pre-combining pass (only when fin_prepare_kernel's sample found many duplicates): every chunk of kChunk entries is
// folded in a shared-memory table of 2 * kChunk slots (never full) and its distinct keys are appended to `groups`, which
// serves as the pass's output buffer until fin_aggregate_kernel overwrites it with the final groups (the partitioning
// passes have moved the entries on to `part` by then).  The last CTA to finish publishes the new entry count and
// partition count.  Unsorted input of low cardinality shrinks by orders of magnitude here; a second round is not needed:
// what is left per key is at most one entry per chunk.
__global__ void fin_combine_kernel(FinArgs a) {
  if (a.comb[0] != 1ull) return;
  extern __shared__ __align__(16) unsigned long long smem[];
  __shared__ uint32_t s_prefix[kPrefixSmem];
  __shared__ uint32_t s_warp[32];
  __shared__ unsigned long long s_base;
  const int ew = a.ew, nc = ew - 1;
  const uint32_t T = 2 * kChunk;
  unsigned long long* keys = smem;
  unsigned long long* vals = smem + T;
  const uint32_t* prefix = stage_prefix(a.chunk_prefix, a.n_ranges, s_prefix);
  const uint32_t total = prefix[a.n_ranges];
  uint32_t err = 0;
  for (uint32_t chunk = blockIdx.x; chunk < total; chunk += gridDim.x) {
    for (uint32_t s = threadIdx.x; s < T; s += blockDim.x) {
      keys[s] = kEmptyKey;
      for (int c = 0; c < nc; ++c) vals[(size_t)s * nc + c] = cell_identity(a.kinds.kind[c]);
    }
    __syncthreads();
    const uint64_t* first;
    uint64_t n;
    locate_chunk(a.ranges, prefix, a.n_ranges, chunk, ew, first, n);
    for (uint64_t i = threadIdx.x; i < n; i += blockDim.x) {
      const uint64_t* e = first + i * ew;
      const uint64_t key = e[0];
      uint32_t h = (uint32_t)mix64(key) & (T - 1);
      for (;;) {
        const unsigned long long seen = atomicCAS(&keys[h], kEmptyKey, (unsigned long long)key);
        if (seen == kEmptyKey || seen == key) break;
        h = (h + 1) & (T - 1);
      }
      for (int c = 0; c < nc; ++c) smem_fold(a.kinds.kind[c], &vals[(size_t)h * nc + c], e[1 + c], err);
    }
    __syncthreads();
    uint32_t mine = 0;
    for (uint32_t s = threadIdx.x; s < T; s += blockDim.x) mine += keys[s] != kEmptyKey;
    uint32_t x = mine;
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((int)(threadIdx.x & 31) >= o) x += y; }
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      uint32_t w = threadIdx.x < (blockDim.x >> 5) ? s_warp[threadIdx.x] : 0u;
      for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, w, o); if ((int)threadIdx.x >= o) w += y; }
      s_warp[threadIdx.x] = w;
      if (threadIdx.x == 31) s_base = w ? atomicAdd(&a.comb[1], (unsigned long long)w) : 0ull;
    }
    __syncthreads();
    uint64_t at = s_base + ((threadIdx.x >> 5) ? s_warp[(threadIdx.x >> 5) - 1] : 0u) + x - mine;
    for (uint32_t s = threadIdx.x; s < T; s += blockDim.x) {
      if (keys[s] == kEmptyKey) continue;
      uint64_t* d = a.groups + at * ew;
      d[0] = keys[s];
      for (int c = 0; c < nc; ++c) d[1 + c] = vals[(size_t)s * nc + c];
      ++at;
    }
    __syncthreads();
  }
  if (err) atomicOr(a.err, err);
  // the last CTA to get here publishes the result
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned long long done = atomicAdd(&a.comb[2], 1ull) + 1ull;
    if (done == (unsigned long long)gridDim.x) {
      __threadfence();
      const unsigned long long n_dedupe = *reinterpret_cast<volatile unsigned long long*>(&a.comb[1]);
      a.hdr->n_dedupe = n_dedupe;
      const unsigned long long per_part = min((unsigned long long)a.slots_max * 45 / 100, 256ull);
      int log2p = 4; // (never "tiny": fin_aggregate_kernel must read `part`, not the buffer it writes)
      while (log2p < a.log2p_max && (1ull << log2p) * per_part < n_dedupe) ++log2p;
      a.hdr->log2p = (unsigned long long)log2p;
      __threadfence();
      a.comb[0] = 2ull;
    }
  }
}

// log2p == 0 ("tiny"): so few entries that one shared-memory table holds them all — fin_aggregate_kernel reads them
// straight from their ranges and the partitioning passes have nothing to do.  (A sorted input leaves only the runs at
// the ends of its batches and the runs longer than a block to deduplicate.)
__global__ void fin_hist_kernel(FinArgs a) {
  const int log2p = (int)a.hdr->log2p;
  if (log2p == 0) return;
  // lanes that meet in one partition add once (hot partitions: few distinct keys among many entries)
  for_each_dedupe_entry(a, [&](const uint64_t* e) {
    const uint32_t p = part_of_key(e[0], log2p);
    const unsigned am = __activemask();
    const unsigned m = __match_any_sync(am, p);
    if ((threadIdx.x & 31u) == (unsigned)(__ffs((int)m) - 1)) atomicAdd(&a.hist[p], (uint32_t)__popc(m));
  }, blockIdx.x, gridDim.x);
}

// Exclusive scan of hist[0..P) -> offsets[0..P], cursors = offsets.  One CTA.
__global__ void fin_scan_kernel(FinArgs a) {
  __shared__ uint32_t warp_sum[32];
  __shared__ uint32_t carry;
  if (a.hdr->log2p == 0ull) return;
  const int P = 1 << (int)a.hdr->log2p;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < P; base += blockDim.x) {
    const int i = base + threadIdx.x;
    const uint32_t v = i < P ? a.hist[i] : 0u;
    uint32_t x = v;
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((int)(threadIdx.x & 31) >= o) x += y; }
    if ((threadIdx.x & 31) == 31) warp_sum[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      uint32_t w = threadIdx.x < (blockDim.x >> 5) ? warp_sum[threadIdx.x] : 0u;
      for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, w, o); if ((int)threadIdx.x >= o) w += y; }
      warp_sum[threadIdx.x] = w; // inclusive
    }
    __syncthreads();
    const uint32_t before = carry + ((threadIdx.x >> 5) ? warp_sum[(threadIdx.x >> 5) - 1] : 0u) + x - v;
    if (i < P) { a.offsets[i] = before; a.cursors[i] = before; }
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) a.offsets[P] = carry;
}

__global__ void fin_scatter_kernel(FinArgs a) {
  const int ew = a.ew, log2p = (int)a.hdr->log2p;
  if (log2p == 0) return;
  for_each_dedupe_entry(a, [&](const uint64_t* e) {
    const uint32_t pt = part_of_key(e[0], log2p);
    const unsigned am = __activemask();
    const unsigned m = __match_any_sync(am, pt);
    const int leader = __ffs((int)m) - 1;
    uint32_t pos0 = 0;
    if ((int)(threadIdx.x & 31u) == leader) pos0 = atomicAdd(&a.cursors[pt], (uint32_t)__popc(m));
    const uint32_t pos = __shfl_sync(m, pos0, leader) + (uint32_t)__popc(m & ((1u << (threadIdx.x & 31u)) - 1u));
    if (ew == 2) {
      const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(e);
      *reinterpret_cast<ulonglong2*>(a.part + (uint64_t)pos * 2) = v;
    } else {
      const uint64_t key = e[0];
      uint64_t* d = a.part + (uint64_t)pos * ew;
      d[0] = key;
      for (int w = 1; w < ew; ++w) d[w] = e[w];
    }
  }, blockIdx.x, gridDim.x);
}

// Shared-memory fold of one cell (values of MIN/MAX over doubles are plain double bits here).
__device__ __forceinline__ void smem_fold(uint8_t kind, unsigned long long* cell, uint64_t v, uint32_t& err) {
  switch (kind) {
    case kCellSumF64: atomicAdd(reinterpret_cast<double*>(cell), u2d(v)); break;
    case kCellSumI64: {
      const int64_t y = (int64_t)v;
      const int64_t x = (int64_t)atomicAdd(cell, (unsigned long long)v);
      const int64_t z = (int64_t)((uint64_t)x + (uint64_t)y);
      if (((x ^ z) & (y ^ z)) < 0) err |= kErrOverflow;
      break;
    }
    case kCellCount: atomicAdd(cell, (unsigned long long)v); break;
    case kCellMinI64: atomicMin(reinterpret_cast<long long*>(cell), (long long)v); break;
    case kCellMaxI64: atomicMax(reinterpret_cast<long long*>(cell), (long long)v); break;
    default: {
      unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(cell);
      while (kind == kCellMinF64 ? u2d(v) < u2d(cur) : u2d(cur) < u2d(v)) {
        const unsigned long long seen = atomicCAS(cell, cur, (unsigned long long)v);
        if (seen == cur) break;
        cur = seen;
      }
    }
  }
}

// One CTA per partition (CTA-stride): fold the partition's entries in a shared-memory table sized for it, then
// append the occupied slots to the dense group array.
__global__ void fin_aggregate_kernel(FinArgs a) {
  extern __shared__ __align__(16) unsigned long long smem[];
  __shared__ uint32_t s_warp[32];
  __shared__ unsigned long long s_base;
  const int ew = a.ew, nc = ew - 1, P = 1 << (int)a.hdr->log2p;
  const bool tiny = a.hdr->log2p == 0ull;
  if (tiny && blockIdx.x != 0) return;
  uint32_t err = 0;
  bool overflow = false;
  for (int p = blockIdx.x; p < P; p += gridDim.x) {
    const uint32_t lo = tiny ? 0u : a.offsets[p], cnt = tiny ? (uint32_t)a.hdr->n_dedupe : a.offsets[p + 1] - lo;
    if (cnt == 0) continue;
    uint32_t T = 64;
    while (T < 2 * cnt && T < (uint32_t)a.slots_max) T <<= 1;
    unsigned long long* keys = smem;
    unsigned long long* vals = smem + T;
    for (uint32_t s = threadIdx.x; s < T; s += blockDim.x) {
      keys[s] = kEmptyKey;
      for (int c = 0; c < nc; ++c) vals[(size_t)s * nc + c] = cell_identity(a.kinds.kind[c]);
    }
    __syncthreads();
    auto fold = [&](const uint64_t* e) {
      const uint64_t key = e[0];
      uint32_t h = (uint32_t)mix64(key) & (T - 1);
      bool placed = false;
      for (uint32_t probes = 0; probes < T; ++probes) {
        const unsigned long long seen = atomicCAS(&keys[h], kEmptyKey, (unsigned long long)key);
        if (seen == kEmptyKey || seen == key) { placed = true; break; }
        h = (h + 1) & (T - 1);
      }
      if (!placed) { overflow = true; return; }
      for (int c = 0; c < nc; ++c) smem_fold(a.kinds.kind[c], &vals[(size_t)h * nc + c], e[1 + c], err);
    };
    if (tiny) {
      for_each_dedupe_entry(a, fold, 0u, 1u); // this CTA alone walks every chunk
    } else if (nc <= 4) {
      // A partition whose entries carry few distinct keys (input of medium cardinality: the partition count follows the
      // entries, not the keys) would hammer one table slot with atomics: a thread folds runs of equal keys in registers
      // first (entries of one key sit together often enough: a thread's entries are kNT apart in the partition).
      const uint64_t* src = a.part + (uint64_t)lo * ew;
      uint64_t ck = kEmptyKey, cv[4] = {0, 0, 0, 0};
      for (uint32_t i = threadIdx.x; i < cnt; i += blockDim.x) {
        const uint64_t* e = src + (uint64_t)i * ew;
        const uint64_t key = e[0];
        if (key == ck) {
#pragma unroll
          for (int c = 0; c < 4; ++c) if (c < nc) cv[c] = fold2(a.kinds.kind[c], cv[c], e[1 + c], err);
        } else {
          if (ck != kEmptyKey) { uint64_t held[5] = {ck, cv[0], cv[1], cv[2], cv[3]}; fold(held); }
          ck = key;
#pragma unroll
          for (int c = 0; c < 4; ++c) if (c < nc) cv[c] = e[1 + c];
        }
      }
      if (ck != kEmptyKey) { uint64_t held[5] = {ck, cv[0], cv[1], cv[2], cv[3]}; fold(held); }
    } else {
      const uint64_t* src = a.part + (uint64_t)lo * ew;
      for (uint32_t i = threadIdx.x; i < cnt; i += blockDim.x) fold(src + (uint64_t)i * ew);
    }
    __syncthreads();
    // compaction: thread t owns slots t, t + NT, ...
    uint32_t mine = 0;
    for (uint32_t s = threadIdx.x; s < T; s += blockDim.x) mine += keys[s] != kEmptyKey;
    uint32_t x = mine;
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((int)(threadIdx.x & 31) >= o) x += y; }
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      uint32_t w = threadIdx.x < (blockDim.x >> 5) ? s_warp[threadIdx.x] : 0u;
      for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, w, o); if ((int)threadIdx.x >= o) w += y; }
      s_warp[threadIdx.x] = w;
      if (threadIdx.x == 31) s_base = w ? atomicAdd(&a.hdr->n_groups, (unsigned long long)w) : 0ull;
    }
    __syncthreads();
    uint64_t at = s_base + ((threadIdx.x >> 5) ? s_warp[(threadIdx.x >> 5) - 1] : 0u) + x - mine;
    for (uint32_t s = threadIdx.x; s < T; s += blockDim.x) {
      if (keys[s] == kEmptyKey) continue;
      uint64_t* d = a.groups + at * ew;
      if (ew == 2) {
        *reinterpret_cast<ulonglong2*>(d) = make_ulonglong2(keys[s], vals[s]);
      } else {
        d[0] = keys[s];
        for (int c = 0; c < nc; ++c) d[1 + c] = vals[(size_t)s * nc + c];
      }
      ++at;
    }
    __syncthreads();
  }
  if (err) atomicOr(a.err, err);
  if (overflow) { a.hdr->overflow = 1ull; atomicOr(a.err, kErrTableFull); }
}

// ---- lean entries -> ABI packed rows -------------------------------------------------------------------
// Where each word of the ABI row comes from: an entry word index (>= 1), kFromRows (the tracked row count, else 1)
// or kOne.
enum : int8_t { kFromOne = -1 };
struct LeanMap {
  int32_t n_accs;
  int32_t ew;
  int8_t rows_src;                 // entry word of the rows cell, or kFromOne
  int8_t value_src[kMaxOutputs];
  int8_t nn_src[kMaxOutputs];
  uint8_t value_kind[kMaxOutputs];
};

__device__ __forceinline__ void write_abi_row(uint64_t* out, const LeanMap& m, uint64_t key, uint64_t flags, const uint64_t* e, bool ordered_f64) {
  const uint64_t rows = m.rows_src >= 0 ? e[m.rows_src] : 1ull;
  out[0] = key;
  out[1] = flags;
  out[2] = rows;
  for (int a = 0; a < m.n_accs; ++a) {
    uint64_t v = m.value_src[a] >= 0 ? e[m.value_src[a]] : rows;
    if (ordered_f64 && (m.value_kind[a] == kCellMinF64 || m.value_kind[a] == kCellMaxF64)) v = (uint64_t)f64_ordered(v);
    out[3 + a] = v;
    out[3 + m.n_accs + a] = m.nn_src[a] >= 0 ? e[m.nn_src[a]] : rows;
  }
}

struct ExpandArgs {
  const LogRange* ranges;   // front ranges of the slabs, then one plain range over the deduplicated groups
  int n_ranges;
  const FinHeader* hdr;
  const uint64_t* special;  // 2 rows x ew words: {present, cells...}
  uint32_t* chunk_prefix;   // [n_ranges + 1]
  LeanMap map;
  uint64_t* out;
  unsigned long long* out_cursor;
};

__global__ void expand_prepare_kernel(ExpandArgs a) {
  const bool sorted = a.hdr->sorted != 0ull; // unsorted input: the front entries went through the deduplication
  chunk_prefix_scan(a.ranges, a.n_ranges, a.chunk_prefix, kChunk, [&](const LogRange& r) { return sorted || r.back != kRangeFront; });
}

__global__ void expand_rows_kernel(ExpandArgs a) {
  __shared__ unsigned long long s_base;
  __shared__ uint32_t s_prefix[kPrefixSmem];
  const int ew = a.map.ew, rw = 3 + 2 * a.map.n_accs;
  const uint32_t* prefix = stage_prefix(a.chunk_prefix, a.n_ranges, s_prefix);
  const uint32_t total = prefix[a.n_ranges];
  for (uint32_t chunk = blockIdx.x; chunk < total; chunk += gridDim.x) {
    const uint64_t* first;
    uint64_t n;
    locate_chunk(a.ranges, prefix, a.n_ranges, chunk, ew, first, n);
    __syncthreads();
    if (threadIdx.x == 0) s_base = atomicAdd(a.out_cursor, (unsigned long long)n);
    __syncthreads();
    for (uint64_t i = threadIdx.x; i < n; i += blockDim.x) {
      const uint64_t* e = first + i * ew;
      write_abi_row(a.out + (s_base + i) * rw, a.map, e[0], 0ull, e, false);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < 2 && a.special != nullptr) {
    const uint64_t* sp = a.special + (size_t)threadIdx.x * ew;
    if (sp[0] != 0ull) {
      const unsigned long long at = atomicAdd(a.out_cursor, 1ull);
      write_abi_row(a.out + at * rw, a.map, threadIdx.x == 0 ? 0ull : kEmptyKey, threadIdx.x == 0 ? 1ull : 0ull, sp, true);
    }
  }
}

__global__ void init_special_kernel(uint64_t* special, int ew, CellKinds k) {
  const int i = threadIdx.x;
  if (i >= 2 * ew) return;
  const int w = i % ew;
  uint64_t v = 0;
  if (w > 0) {
    v = cell_identity(k.kind[w - 1]);
    if (k.kind[w - 1] == kCellMinF64 || k.kind[w - 1] == kCellMaxF64) v = (uint64_t)f64_ordered(v); // updated with integer atomics
  }
  special[i] = v;
}

// ------------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------------
CellKinds log_cell_kinds(const Agg& a) {
  CellKinds k{};
  k.n = a.log_n_cells;
  for (int i = 0; i < a.log_n_cells; ++i) k.kind[i] = a.cells[a.log_cells[i]].kind;
  return k;
}

static int entry_word_of(const Agg& a, int cell) {
  for (int i = 0; i < a.log_n_cells; ++i) if (a.log_cells[i] == cell) return 1 + i;
  return -1;
}

static LeanMap lean_map(const Agg& a) {
  LeanMap m{};
  m.n_accs = a.rm.n_accs;
  m.ew = a.ew;
  const int rows_word = entry_word_of(a, a.rm.rows_cell);
  m.rows_src = rows_word >= 0 ? (int8_t)rows_word : kFromOne;
  for (int i = 0; i < a.rm.n_accs; ++i) {
    const int vw = entry_word_of(a, a.rm.value_cell[i]);
    m.value_src[i] = (int8_t)vw;                     // -1: a count cell that equals the row count
    const int nw = a.rm.nn_cell[i] >= 0 ? entry_word_of(a, a.rm.nn_cell[i]) : -1;
    m.nn_src[i] = (int8_t)nw;                        // -1: every row had a non-null input (or the accumulator is a COUNT)
    m.value_kind[i] = a.rm.value_kind[i];
  }
  return m;
}

int log_setup(Agg& a, const PipelineParams& p, cudaStream_t stream) {
  bool rows_needed = false;
  for (int i = 0; i < a.rm.n_accs; ++i) rows_needed = rows_needed || a.rm.value_kind[i] == kCellCount;
  a.log_n_cells = 0;
  if (rows_needed) a.log_cells[a.log_n_cells++] = 0;
  for (int c = 1; c < p.n_cells; ++c) if (!p.cells[c].skip) a.log_cells[a.log_n_cells++] = (uint8_t)c;
  if (a.log_n_cells == 0) a.log_cells[a.log_n_cells++] = 0; // keys only (DISTINCT): carry the row count
  a.ew = 1 + a.log_n_cells;
  if (fused_dry_run()) return VGPU_OK;
  VG_CUDA(cudaMalloc(&a.special, (size_t)2 * a.ew * 8));
  init_special_kernel<<<1, 128, 0, stream>>>(a.special, a.ew, log_cell_kinds(a));
  VG_CUDA(cudaGetLastError());
  count_launch();
  return VGPU_OK;
}

bool log_layout_covers(const Agg& a, const PipelineParams& p) {
  for (int c = 1; c < p.n_cells; ++c) {
    if (p.cells[c].skip) continue;
    if (entry_word_of(a, c) < 0) return false;
  }
  return true;
}

static int slab_alloc(LogSlab& s, int regions, uint64_t region_cap, int ew) {
  s = LogSlab{};
  s.regions = regions;
  s.region_cap = region_cap;
  s.bytes = (size_t)regions * region_cap * ew * 8;
  VG_CUDA(cudaMalloc(&s.base, std::max<size_t>(s.bytes, 16)));
  VG_CUDA(cudaMalloc(&s.counts, (size_t)regions * 2 * sizeof(uint32_t)));
  return VGPU_OK;
}

int log_begin_batch(Agg& a, PipelineParams& p, int regions, uint64_t region_need, cudaStream_t stream, bool no_slab) {
  LogSlab* use = nullptr;
  if (!no_slab && !a.slabs.empty()) {
    LogSlab& s = a.slabs.back();
    if (s.regions >= regions && s.region_used + region_need <= s.region_cap) use = &s;
  }
  if (!use && !no_slab) {
    // a slab kept from before reset() that is large enough, else a new one (at least as large as everything logged
    // so far, so the number of slabs grows with the logarithm of the input)
    int pick = -1;
    for (size_t i = 0; i < a.free_slabs.size(); ++i) {
      const LogSlab& f = a.free_slabs[i];
      if (f.regions >= regions && f.region_cap >= region_need && (pick < 0 || f.bytes < a.free_slabs[pick].bytes)) pick = (int)i;
    }
    LogSlab s;
    if (pick >= 0) {
      s = a.free_slabs[pick];
      a.free_slabs.erase(a.free_slabs.begin() + pick);
    } else {
      uint64_t cap = region_need;
      for (const LogSlab& o : a.slabs) cap = std::max(cap, o.region_cap * (uint64_t)o.regions / (uint64_t)regions);
      int rc = slab_alloc(s, regions, cap, a.ew);
      if (rc) return rc;
    }
    s.region_used = 0;
    VG_CUDA(cudaMemsetAsync(s.counts, 0, (size_t)s.regions * 2 * sizeof(uint32_t), stream));
    a.slabs.push_back(s);
    use = &a.slabs.back();
  }
  if (use) {
    use->region_used += region_need;
    p.log = use->base;
    p.log_region_cap = use->region_cap;
    p.log_counts = use->counts;
  }
  p.log_special = a.special;
  p.sort_state = a.ctr + kCtrSortState;
  p.sort_prev_in = a.ctr + kCtrPrevKey + 2 * ((a.log_batches + 1) & 1);
  p.sort_prev_out = a.ctr + kCtrPrevKey + 2 * (a.log_batches & 1);
  if (!p.sorted_probe) {
    a.log_probed = false;
  } else if (a.log_batches > 0) {
    // the next batch must not see a stale "row before": carry this slot forward when the batch writes nothing (never: a batch has rows)
  }
  ++a.log_batches;
  return VGPU_OK;
}

static int ensure(void** ptr, size_t* cap, size_t want_bytes, cudaStream_t stream) {
  if (*cap >= want_bytes && *ptr) return VGPU_OK;
  if (*ptr) { VG_CUDA(cudaStreamSynchronize(stream)); VG_CUDA(cudaFree(*ptr)); *ptr = nullptr; *cap = 0; }
  VG_CUDA(cudaMalloc(ptr, std::max<size_t>(want_bytes, 256)));
  *cap = want_bytes;
  return VGPU_OK;
}

static void build_ranges(Agg& a, bool for_output) {
  std::vector<LogRange>& out = a.host_ranges[for_output ? 1 : 0];
  out.clear();
  for (const LogSlab& s : a.slabs) {
    for (int c = 0; c < s.regions; ++c) {
      LogRange r{};
      r.base = s.base + (uint64_t)c * s.region_cap * a.ew;
      r.cap = s.region_cap;
      r.count64 = 0;
      r.count = s.counts + 2 * c;
      r.back = kRangeFront;
      out.push_back(r);
      if (for_output) continue;
      r.count = s.counts + 2 * c + 1;
      r.back = kRangeBack;
      out.push_back(r);
    }
  }
  for (const LogRange& r : a.ext_ranges) if (!for_output || r.back == kRangeFront) out.push_back(r);
  if (for_output) {
    LogRange g{};
    g.base = a.groups;
    g.cap = 0;
    g.count = &reinterpret_cast<FinHeader*>(a.ctr + kCtrFin)->n_groups;
    g.count64 = 1;
    g.back = kRangePlain;
    out.push_back(g);
  }
}

// Copies the list to the device unless the device already holds exactly this list (steady state of a repeated
// aggregation: a host-to-device copy from pageable memory may wait for the stream).
static int upload_ranges(Agg& a, int which, cudaStream_t stream) {
  const std::vector<LogRange>& host = a.host_ranges[which];
  std::vector<LogRange>& up = a.uploaded_ranges[which];
  if (a.ranges[which] && up.size() == host.size() && (host.empty() || memcmp(up.data(), host.data(), host.size() * sizeof(LogRange)) == 0)) return VGPU_OK;
  size_t cap_bytes = a.ranges_cap[which] * sizeof(LogRange);
  void* ptr = a.ranges[which];
  int rc = ensure(&ptr, &cap_bytes, host.size() * sizeof(LogRange), stream);
  a.ranges[which] = (LogRange*)ptr;
  a.ranges_cap[which] = cap_bytes / sizeof(LogRange);
  if (rc) return rc;
  VG_CUDA(cudaMemcpyAsync(a.ranges[which], host.data(), host.size() * sizeof(LogRange), cudaMemcpyHostToDevice, stream));
  up = host;
  return VGPU_OK;
}

int log_finalize(Agg& a, cudaStream_t stream) {
  if (a.finalized) return VGPU_OK;
  FinArgs f{};
  f.ew = a.ew;
  f.kinds = log_cell_kinds(a);
  // shared-memory table of a partition: 2048 slots or what fits 32 KB (seven CTAs per SM), at least 256; partitions
  // average <= 256 entries (fin_prepare), so the tables stay far below half full
  int slots = 2048;
  while (slots > 256 && (size_t)slots * a.ew * 8 > 32 * 1024) slots >>= 1;
  f.slots_max = slots;
  // the arrays are sized for the upper bound of the entries; the device picks the partition count from the real count
  const uint64_t bound = a.log_bound + a.ext_bound;
  const uint64_t per_part = (uint64_t)slots * 45 / 100;
  int log2p = 6;
  while (log2p < 20 && ((uint64_t)1 << log2p) * per_part < bound) ++log2p;
  f.log2p_max = log2p;
  a.n_parts = 1 << log2p;
  const size_t P = (size_t)1 << log2p;
  void* ptr = a.hist; size_t cap = a.hist_cap;
  build_ranges(a, false);
  build_ranges(a, true);
  const size_t n_prefix = std::max(a.host_ranges[0].size(), a.host_ranges[1].size()) + 1;
  int rc = ensure(&ptr, &cap, (3 * (P + 1) + n_prefix) * sizeof(uint32_t), stream);
  a.hist = (uint32_t*)ptr; a.hist_cap = cap;
  if (rc) return rc;
  const size_t entry_bytes = (size_t)a.ew * 8;
  ptr = a.part; cap = a.part_cap;
  rc = ensure(&ptr, &cap, std::max<uint64_t>(bound, 1) * entry_bytes, stream);
  a.part = (uint64_t*)ptr; a.part_cap = cap;
  if (rc) return rc;
  ptr = a.groups; cap = a.groups_cap;
  rc = ensure(&ptr, &cap, std::max<uint64_t>(bound, 1) * entry_bytes, stream);
  a.groups = (uint64_t*)ptr; a.groups_cap = cap;
  if (rc) return rc;
  rc = upload_ranges(a, 0, stream);
  if (rc) return rc;
  f.ranges = a.ranges[0];
  f.n_ranges = (int)a.host_ranges[0].size();
  f.probed = a.log_probed ? 1 : 0;
  f.sort_state = a.ctr + kCtrSortState;
  f.src_meta = a.ext_meta; f.n_src = a.ext_sources; f.src_stride = a.ext_meta_stride; f.local_entries = a.slabs.empty() ? 0 : 1;
  f.hdr = reinterpret_cast<FinHeader*>(a.ctr + kCtrFin);
  f.comb = a.ctr + kCtrCombine;
  const size_t combine_smem = (size_t)2 * kChunk * a.ew * 8;
  f.combine_ok = combine_smem <= 96 * 1024 ? 1 : 0;
  f.hist = a.hist;
  f.offsets = a.hist + (P + 1);
  f.cursors = a.hist + 2 * (P + 1);
  f.chunk_prefix = a.hist + 3 * (P + 1);
  f.part = a.part;
  f.groups = a.groups;
  f.err = a.err;
  VG_CUDA(cudaMemsetAsync(a.hist, 0, (P + 1) * sizeof(uint32_t), stream));
  fin_prepare_kernel<<<1, 1024, 0, stream>>>(f);
  if (f.combine_ok) { // returns at once unless the sample found many duplicates
    VG_CUDA(cudaFuncSetAttribute(fin_combine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    fin_combine_kernel<<<num_sms() * 4, 256, combine_smem, stream>>>(f);
    count_launch();
  }
  const int wide = num_sms() * 8;
  fin_hist_kernel<<<wide, 256, 0, stream>>>(f);
  fin_scan_kernel<<<1, 1024, 0, stream>>>(f);
  fin_scatter_kernel<<<wide, 256, 0, stream>>>(f);
  const size_t smem = (size_t)slots * a.ew * 8;
  VG_CUDA(cudaFuncSetAttribute(fin_aggregate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  fin_aggregate_kernel<<<(unsigned)std::min<size_t>(P, (size_t)num_sms() * 7), 256, smem, stream>>>(f);
  VG_CUDA(cudaGetLastError());
  count_launch(5);
  a.finalized = true;
  return VGPU_OK;
}

int log_read_header(Agg& a, cudaStream_t stream, FinHeader* h, uint64_t* n_special) {
  unsigned long long sp[2] = {0, 0};
  VG_CUDA(cudaMemcpyAsync(h, a.ctr + kCtrFin, sizeof(FinHeader), cudaMemcpyDeviceToHost, stream));
  VG_CUDA(cudaMemcpyAsync(&sp[0], a.special, 8, cudaMemcpyDeviceToHost, stream));
  VG_CUDA(cudaMemcpyAsync(&sp[1], a.special + a.ew, 8, cudaMemcpyDeviceToHost, stream));
  VG_CUDA(cudaStreamSynchronize(stream));
  *n_special = (sp[0] != 0) + (sp[1] != 0);
  return VGPU_OK;
}

int log_expand_rows(Agg& a, uint64_t* out, unsigned long long* n_out, cudaStream_t stream) {
  build_ranges(a, true);
  int rc = upload_ranges(a, 1, stream);
  if (rc) return rc;
  ExpandArgs e{};
  e.ranges = a.ranges[1];
  e.n_ranges = (int)a.host_ranges[1].size();
  e.hdr = reinterpret_cast<const FinHeader*>(a.ctr + kCtrFin);
  e.special = a.special;
  e.map = lean_map(a);
  e.out = out;
  e.out_cursor = n_out;
  e.chunk_prefix = a.hist + 3 * ((size_t)a.n_parts + 1); // the deduplication passes are done with it
  expand_prepare_kernel<<<1, 1024, 0, stream>>>(e);
  expand_rows_kernel<<<num_sms() * 8, 256, 0, stream>>>(e);
  VG_CUDA(cudaGetLastError());
  count_launch(2);
  return VGPU_OK;
}

int log_adopt_layout(Agg& a, const Agg& from, cudaStream_t stream) {
  if (a.engine != kEngineUndecided || a.n_cells != from.n_cells) { set_error("cannot adopt the entry layout"); return VGPU_ERR_INVALID; }
  a.log_n_cells = from.log_n_cells;
  for (int i = 0; i < from.log_n_cells; ++i) a.log_cells[i] = from.log_cells[i];
  a.ew = from.ew;
  a.engine = kEngineLog;
  VG_CUDA(cudaMalloc(&a.special, (size_t)2 * a.ew * 8));
  init_special_kernel<<<1, 128, 0, stream>>>(a.special, a.ew, log_cell_kinds(a));
  VG_CUDA(cudaGetLastError());
  count_launch();
  return VGPU_OK;
}

__global__ void chunk_prefix_kernel(const LogRange* ranges, int n, uint32_t chunk, uint32_t* chunk_prefix) {
  chunk_prefix_scan(ranges, n, chunk_prefix, chunk, [](const LogRange&) { return true; });
}
int log_chunk_prefix(const LogRange* ranges, int n, uint32_t chunk, uint32_t* chunk_prefix, cudaStream_t stream) {
  chunk_prefix_kernel<<<1, 1024, 0, stream>>>(ranges, n, chunk, chunk_prefix);
  VG_CUDA(cudaGetLastError());
  count_launch();
  return VGPU_OK;
}

int log_entry_ranges(Agg& a, cudaStream_t stream, const LogRange** ranges, int* n, uint32_t** chunk_prefix_scratch) {
  build_ranges(a, false);
  int rc = upload_ranges(a, 0, stream);
  if (rc) return rc;
  *ranges = a.ranges[0];
  *n = (int)a.host_ranges[0].size();
  if (chunk_prefix_scratch) *chunk_prefix_scratch = nullptr;
  return VGPU_OK;
}

int log_reset(Agg& a, cudaStream_t stream) {
  for (LogSlab& s : a.slabs) a.free_slabs.push_back(s);
  a.slabs.clear();
  a.ext_ranges.clear();
  a.ext_bound = 0;
  a.ext_meta = nullptr; a.ext_sources = 0;
  a.log_bound = 0;
  a.log_probed = true;
  a.log_batches = 0;
  a.finalized = false;
  if (a.ctr) VG_CUDA(cudaMemsetAsync(a.ctr + kCtrSortState, 0, (kCtrWords - kCtrSortState) * sizeof(unsigned long long), stream));
  if (a.special) {
    init_special_kernel<<<1, 128, 0, stream>>>(a.special, a.ew, log_cell_kinds(a));
    VG_CUDA(cudaGetLastError());
    count_launch();
  }
  return VGPU_OK;
}

void log_free(Agg& a) {
  for (LogSlab& s : a.slabs) { cudaFree(s.base); cudaFree(s.counts); }
  for (LogSlab& s : a.free_slabs) { cudaFree(s.base); cudaFree(s.counts); }
  a.slabs.clear();
  a.free_slabs.clear();
  cudaFree(a.special); cudaFree(a.part); cudaFree(a.groups); cudaFree(a.hist); cudaFree(a.ranges[0]); cudaFree(a.ranges[1]);
  a.special = a.part = a.groups = nullptr;
  a.hist = nullptr;
  a.ranges[0] = a.ranges[1] = nullptr;
  a.uploaded_ranges[0].clear(); a.uploaded_ranges[1].clear();
  a.part_cap = a.groups_cap = a.hist_cap = a.ranges_cap[0] = a.ranges_cap[1] = 0;
}

} // namespace vg
