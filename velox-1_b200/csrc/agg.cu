// Group state that outlives one pipeline launch: the dense ARRAY/GLOBAL state, the HBM hash
// table (init / rehash), extraction of packed group rows, merging packed rows
// (Aggregate::addIntermediateResults) and hash partitioning for the multi-GPU exchange.
//
// Reference behaviour restated (paths under the reference tree):
//   velox/exec/GroupingSet.cpp:644-712 (getOutput/extractGroups), velox/exec/RowContainer.cpp
//   (listRows), velox/exec/HashTable.cpp:684-747 (checkSize/rehash),
//   velox/exec/HashPartitionFunction.cpp:59-92 (partition = hash % numPartitions).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <string.h>

#include <algorithm>

#include "agg.h"
#include "fused.h"
#include "device_common.cuh"

namespace vg {

// ---- ARRAY / GLOBAL ---------------------------------------------------------------------------
struct CellMeta {
  uint8_t kind[kMaxCells];
  uint8_t skip[kMaxCells];
  uint8_t alias[kMaxCells];
  int32_t n_cells;
};

__global__ void init_state_kernel(uint64_t* state, CellMeta m, int G) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m.n_cells * G) return;
  state[i] = cell_identity(m.kind[i / G]);
}

// Launch-time ("compact") group numbering -> the aggregation's state numbering.  A launch whose key
// columns cannot be NULL drops the NULL id of those keys so that the per-thread shared-memory
// tables stay small; the state always keeps it (a later batch may carry NULLs).
struct GroupMap {
  int32_t n_keys;
  int32_t card[kMaxKeys];   // compact cardinality of key k
  int32_t shift[kMaxKeys];  // added to the compact digit to get the state digit (0 or 1)
  int64_t mult[kMaxKeys];   // state multiplier of key k
};
__device__ __forceinline__ int state_group(const GroupMap& gm, int g) {
  int64_t out = 0;
  for (int k = 0; k < gm.n_keys; ++k) {
    const int d = g % gm.card[k];
    g /= gm.card[k];
    out += (int64_t)(d + gm.shift[k]) * gm.mult[k];
  }
  return (int)out;
}

// state[c][sg] = fold(state[c][sg], partials[0][c][g], partials[1][c][g], ...): one warp per
// (cell, group).  Lane l folds CTAs l, l+32, ... in order, then a fixed xor-butterfly joins the
// lanes: the order depends only on the launch width, so results are run-to-run deterministic.
// Gc = groups of the launch, G = groups of the state.
__global__ void merge_partials_kernel(uint64_t* state, const uint64_t* partials, int n_blocks,
                                      CellMeta m, int Gc, int G, GroupMap gm, uint32_t* err_word) {
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= m.n_cells * Gc) return;
  const int c = i / Gc, g = i - c * Gc;
  const int src = m.skip[c] ? m.alias[c] : c;
  const uint8_t kind = m.kind[c];
  uint32_t err = 0;
  uint64_t acc = cell_identity(kind);
  const size_t stride = (size_t)m.n_cells * Gc;
  for (int b = lane; b < n_blocks; b += 32) acc = fold2(kind, acc, partials[b * stride + (size_t)src * Gc + g], err);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const uint64_t other = __shfl_xor_sync(0xffffffffu, acc, o);
    acc = fold2(kind, acc, other, err);
  }
  if (lane == 0) {
    const size_t si = (size_t)c * G + state_group(gm, g);
    state[si] = fold2(kind, state[si], acc, err);
  }
  if (err) atomicOr(err_word, err);
}

// Packed group row (the ABI layout): key, flags, rows, value per accumulator, non-null count per
// accumulator.
__device__ __forceinline__ void write_packed(uint64_t* out, uint64_t key, uint64_t flags, uint64_t rows,
                                             const RowMap& rm, const uint64_t* cells, size_t cell_stride,
                                             bool ordered_f64) {
  out[0] = key;
  out[1] = flags;
  out[2] = rows;
  for (int a = 0; a < rm.n_accs; ++a) {
    uint64_t v = cells[(size_t)rm.value_cell[a] * cell_stride];
    const uint8_t k = rm.value_kind[a];
    if (ordered_f64 && (k == kCellMinF64 || k == kCellMaxF64)) v = (uint64_t)f64_ordered(v); // involution
    out[3 + a] = v;
    out[3 + rm.n_accs + a] = rm.nn_cell[a] >= 0 ? cells[(size_t)rm.nn_cell[a] * cell_stride] : rows;
  }
}

// Every group at its own position (absent groups have rows == 0): the fixed-size image of the state
// that the small-domain exchange all-gathers.
__global__ void dense_rows_kernel(const uint64_t* state, int G, RowMap rm, uint64_t* rows_out) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= G) return;
  write_packed(rows_out + (size_t)g * rm.row_words_out, (uint64_t)g, 0, state[(size_t)rm.rows_cell * G + g], rm, state + g, (size_t)G, false);
}

__global__ void extract_array_kernel(const uint64_t* state, int G, RowMap rm, int always,
                                     uint64_t* rows_out, unsigned long long* counter) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= G) return;
  const uint64_t rows = state[(size_t)rm.rows_cell * G + g];
  if (rows == 0 && !always) return;
  const unsigned long long pos = take_one(counter);
  write_packed(rows_out + pos * rm.row_words_out, (uint64_t)g, 0, rows, rm, state + g, (size_t)G, false);
}

// ---- HASH -------------------------------------------------------------------------------------
// One thread writes whole rows (16-byte stores when the row is a multiple of two words), no division
// per word: the wipe of a 1 GB table runs at store bandwidth instead of at the integer-divide rate.
__global__ void init_table_kernel(uint64_t* table, uint64_t n_slots, int row_words, CellMeta m) {
  __shared__ uint64_t image[(1 + kMaxCells + 3) & ~3];
  for (int w = threadIdx.x; w < row_words && w < ((1 + kMaxCells + 3) & ~3); w += blockDim.x) {
    uint64_t v = kEmptyKey;
    if (w > m.n_cells) {
      v = 0; // padding
    } else if (w > 0) {
      v = cell_identity(m.kind[w - 1]);
      if (m.kind[w - 1] == kCellMinF64 || m.kind[w - 1] == kCellMaxF64) v = (uint64_t)f64_ordered(v);
    }
    image[w] = v;
  }
  __syncthreads();
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  if ((row_words & 1) == 0) {
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_slots; r += stride) {
      ulonglong2* dst = reinterpret_cast<ulonglong2*>(table + r * row_words);
      for (int w = 0; w < row_words; w += 2) dst[w >> 1] = make_ulonglong2(image[w], image[w + 1]);
    }
  } else {
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_slots; r += stride) {
      for (int w = 0; w < row_words; ++w) table[r * row_words + w] = image[w];
    }
  }
}

// Moves every occupied row of `from` into `to` (keys are unique, so plain stores after the claim).
__global__ void rehash_kernel(const uint64_t* from, uint64_t from_slots, HashRef to, uint32_t* err_word) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= from_slots) return;
  const uint64_t* src = from + i * to.row_words;
  if (src[0] == kEmptyKey) return;
  const bool special = i >= from_slots - 2;
  uint32_t err = 0;
  int64_t slot;
  if (special) {
    slot = (int64_t)to.cap_mask + 1 + (int64_t)(i - (from_slots - 2));
    to.table[slot * to.row_words] = 0;
    count_one(to.n_groups_ctr);
  } else {
    slot = hash_find_or_insert(to, src[0], false, err);
  }
  if (slot < 0) { atomicOr(err_word, err); return; }
  uint64_t* dst = to.table + slot * to.row_words;
  for (int w = 1; w < to.row_words; ++w) dst[w] = src[w];
}

// One block walks 1024 consecutive slots; the block takes its output range with ONE atomic (a per-warp
// atomic on a single counter was half of this kernel's time at 32 M slots).
constexpr int kExtractPerThread = 4;
__global__ void extract_hash_kernel(const uint64_t* table, uint64_t n_slots, int row_words, RowMap rm,
                                    uint64_t* rows_out, unsigned long long* counter) {
  __shared__ unsigned int warp_count[8];
  __shared__ unsigned long long block_base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint64_t first = (uint64_t)blockIdx.x * (256 * kExtractPerThread);
  bool occ[kExtractPerThread];
  unsigned int mine = 0, before[kExtractPerThread];
#pragma unroll
  for (int k = 0; k < kExtractPerThread; ++k) {
    const uint64_t i = first + (uint64_t)k * 256 + threadIdx.x;
    occ[k] = i < n_slots && table[i * row_words] != kEmptyKey;
    const unsigned m = __ballot_sync(0xffffffffu, occ[k]);
    before[k] = mine + __popc(m & ((1u << lane) - 1u));   // rank inside the warp, pass k after passes < k
    mine += __popc(m);                                      // warp total so far (same in every lane)
  }
  if (lane == 0) warp_count[warp] = mine;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int total = 0;
    for (int w = 0; w < 8; ++w) { const unsigned int c = warp_count[w]; warp_count[w] = total; total += c; }
    block_base = total ? atomicAdd(counter, (unsigned long long)total) : 0ull;
  }
  __syncthreads();
  const unsigned long long base = block_base + warp_count[warp];
#pragma unroll
  for (int k = 0; k < kExtractPerThread; ++k) {
    if (!occ[k]) continue;
    const uint64_t i = first + (uint64_t)k * 256 + threadIdx.x;
    const uint64_t* src = table + i * row_words;
    uint64_t key = src[0], flags = 0;
    if (i == n_slots - 2) { key = 0; flags = 1; }      // NULL key group
    if (i == n_slots - 1) { key = kEmptyKey; }         // the key that equals the empty marker
    write_packed(rows_out + (base + before[k]) * rm.row_words_out, key, flags, src[1 + rm.rows_cell], rm, src + 1, 1, true);
  }
}

// ---- merging packed rows ---------------------------------------------------------------------
// Global-memory fold with atomics (rows from different producers may hit the same group).
__device__ __forceinline__ void atomic_fold(uint8_t kind, uint64_t* addr, uint64_t v, bool ordered_f64,
                                            uint32_t& err) {
  switch (kind) {
    case kCellSumF64: atomicAdd(reinterpret_cast<double*>(addr), u2d(v)); break;
    case kCellSumI64: {
      const int64_t y = (int64_t)v;
      const int64_t x = (int64_t)atomicAdd(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)v);
      const int64_t z = (int64_t)((uint64_t)x + (uint64_t)y);
      if (((x ^ z) & (y ^ z)) < 0) err |= kErrOverflow;
      break;
    }
    case kCellCount: atomicAdd(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)v); break;
    case kCellMinI64: atomicMin(reinterpret_cast<long long*>(addr), (long long)v); break;
    case kCellMaxI64: atomicMax(reinterpret_cast<long long*>(addr), (long long)v); break;
    default: {
      if (u2d(v) != u2d(v)) break;
      if (ordered_f64) {
        if (kind == kCellMinF64) atomicMin(reinterpret_cast<long long*>(addr), (long long)f64_ordered(v));
        else atomicMax(reinterpret_cast<long long*>(addr), (long long)f64_ordered(v));
      } else {
        unsigned long long* p = reinterpret_cast<unsigned long long*>(addr);
        unsigned long long cur = *p;
        while (true) {
          const bool better = kind == kCellMinF64 ? u2d(v) < u2d(cur) : u2d(cur) < u2d(v);
          if (!better) break;
          const unsigned long long seen = atomicCAS(p, cur, (unsigned long long)v);
          if (seen == cur) break;
          cur = seen;
        }
      }
    }
  }
}

// cells: base of the group's cells, `stride` words apart.
__device__ __forceinline__ void merge_packed(const uint64_t* row, const RowMap& rm, uint64_t* cells,
                                             size_t stride, bool ordered_f64, const uint8_t* cell_done,
                                             uint32_t& err) {
  atomic_fold(kCellCount, cells + (size_t)rm.rows_cell * stride, row[2], ordered_f64, err);
  for (int a = 0; a < rm.n_accs; ++a) {
    const uint8_t k = rm.value_kind[a];
    const uint64_t nn = row[3 + rm.n_accs + a];
    if (!cell_done[a * 2]) {
      // a value whose non-null count is 0 is the identity (or garbage for MIN/MAX): skip it
      if (k == kCellCount || rm.nn_cell[a] < 0 || nn != 0) {
        if (rm.value_cell[a] != rm.rows_cell) atomic_fold(k, cells + (size_t)rm.value_cell[a] * stride, row[3 + a], ordered_f64, err);
      }
    }
    if (!cell_done[a * 2 + 1] && rm.nn_cell[a] >= 0 && rm.nn_cell[a] != rm.rows_cell) {
      atomic_fold(kCellCount, cells + (size_t)rm.nn_cell[a] * stride, nn, ordered_f64, err);
    }
  }
}

struct DoneMap { uint8_t d[2 * kMaxOutputs]; }; // 1 = an earlier accumulator already covers the cell

__device__ __forceinline__ int part_of(const uint64_t* row, int n_parts, bool velox_hash);
// n_parts > 0: rows of every rank's dense state arrive here (fixed-size all-gather); a rank folds the
// groups it owns (part_of == my_part) and skips groups that saw no row.
__global__ void merge_rows_array_kernel(const uint64_t* rows, int64_t n, RowMap rm, DoneMap dm,
                                        uint64_t* state, int G, uint32_t* err_word, int n_parts, int my_part) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t* row = rows + i * rm.row_words_out;
  if (n_parts > 0 && (row[2] == 0 || part_of(row, n_parts, false) != my_part)) return;
  uint32_t err = 0;
  if (row[0] >= (uint64_t)G) { atomicOr(err_word, kErrKeyRange); return; }
  merge_packed(row, rm, state + row[0], (size_t)G, false, dm.d, err);
  if (err) atomicOr(err_word, err);
}

__global__ void merge_rows_hash_kernel(const uint64_t* rows, int64_t n, RowMap rm, DoneMap dm, HashRef h,
                                       uint32_t* err_word) {
  const int rw = h.row_words;
  uint32_t err = 0, inserted = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t* row = rows + i * rm.row_words_out;
    const uint64_t key = row[0];
    const bool isnull = (row[1] & 1ull) != 0;
    int64_t slot = -1;
    if (isnull || key == kEmptyKey) {
      slot = hash_find_or_insert(h, key, isnull, err); // the two dedicated slots
    } else { // claim first: one atomic round trip settles a new group and a present one alike
      uint64_t at = home_slot(h.key_map, h.cap_mask, key);
      for (uint64_t probes = 0; probes <= h.cap_mask + kBucketSteps; ++probes) {
        const unsigned long long seen = atomicCAS(reinterpret_cast<unsigned long long*>(h.table + at * rw), kEmptyKey, (unsigned long long)key);
        if (seen == kEmptyKey) { ++inserted; slot = (int64_t)at; break; }
        if (seen == key) { slot = (int64_t)at; break; }
        at = next_slot(h.key_map, h.cap_mask, at, probes);
      }
      if (slot < 0) err |= kErrTableFull;
    }
    if (slot >= 0) merge_packed(row, rm, h.table + slot * rw + 1, 1, true, dm.d, err);
  }
  for (int o = 16; o > 0; o >>= 1) inserted += __shfl_xor_sync(0xffffffffu, inserted, o);
  if ((threadIdx.x & 31) == 0 && inserted) atomicAdd(h.n_groups_ctr, (unsigned long long)inserted);
  if (err) atomicOr(err_word, err);
}

// Inserts the segments the fused HASH kernels parked ({key, cell values...}, row_words words each):
// one segment per thread, so a long probe chain delays 31 neighbours, not a scan warp's 128 rows.
__global__ void drain_overflow_kernel(const uint64_t* ovf, const unsigned long long* cursor, uint64_t cap,
                                      HashRef h, CellMeta m, uint32_t* err_word) {
  const uint64_t n = min((uint64_t)*cursor, cap);
  const int rw = h.row_words;
  uint32_t err = 0, inserted = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t* e = ovf + i * rw;
    const uint64_t key = e[0];
    uint64_t slot = home_slot(h.key_map, h.cap_mask, key);
    bool found = false;
    for (uint64_t probes = 0; probes <= h.cap_mask + kBucketSteps; ++probes) {
      const unsigned long long seen = atomicCAS(reinterpret_cast<unsigned long long*>(h.table + slot * rw), kEmptyKey, (unsigned long long)key);
      if (seen == kEmptyKey) { ++inserted; found = true; break; }
      if (seen == key) { found = true; break; }
      slot = next_slot(h.key_map, h.cap_mask, slot, probes);
    }
    if (!found) { err |= kErrTableFull; continue; }
    for (int c = 0; c < m.n_cells; ++c) {
      const uint64_t v = e[1 + c];
      if (v == cell_identity(m.kind[c])) continue;
      uint64_t* cell = h.table + slot * rw + 1 + c;
      if (m.kind[c] == kCellCount) atomicAdd(reinterpret_cast<unsigned long long*>(cell), (unsigned long long)v); // a run's row count
      else update_hbm_cell(m.kind[c], cell, v, err);
    }
  }
  for (int o = 16; o > 0; o >>= 1) inserted += __shfl_xor_sync(0xffffffffu, inserted, o);
  if ((threadIdx.x & 31) == 0 && inserted) atomicAdd(h.n_groups_ctr, (unsigned long long)inserted);
  if (err) atomicOr(err_word, err);
}

// ---- partitioning packed rows ----------------------------------------------------------------
__device__ __forceinline__ int part_of(const uint64_t* row, int n_parts, bool velox_hash) {
  return (int)partition_of(row[0], row[1] & 1ull, (uint32_t)n_parts, velox_hash);
}
__global__ void partition_count_kernel(const uint64_t* rows, int64_t n, int row_words, int n_parts,
                                       unsigned long long* counts, bool velox_hash) {
  __shared__ unsigned int hist[32];
  if (threadIdx.x < 32) hist[threadIdx.x] = 0;
  __syncthreads();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) atomicAdd(&hist[part_of(rows + i * row_words, n_parts, velox_hash)], 1u);
  __syncthreads();
  if (threadIdx.x < n_parts && hist[threadIdx.x]) atomicAdd(&counts[threadIdx.x], (unsigned long long)hist[threadIdx.x]);
}
__global__ void partition_scatter_kernel(const uint64_t* rows, int64_t n, int row_words, int n_parts,
                                         unsigned long long* cursors, uint64_t* out, bool velox_hash) {
  // positions inside the block from a shared-memory histogram; one global atomic per part and block
  __shared__ unsigned int hist[32];
  __shared__ unsigned long long base[32];
  if (threadIdx.x < 32) hist[threadIdx.x] = 0;
  __syncthreads();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t* row = rows + i * row_words;
  int part = 0;
  unsigned int local = 0;
  if (i < n) { part = part_of(row, n_parts, velox_hash); local = atomicAdd(&hist[part], 1u); }
  __syncthreads();
  if (threadIdx.x < n_parts && hist[threadIdx.x]) base[threadIdx.x] = atomicAdd(&cursors[threadIdx.x], (unsigned long long)hist[threadIdx.x]);
  __syncthreads();
  if (i >= n) return;
  uint64_t* dst = out + (base[part] + local) * row_words;
  for (int w = 0; w < row_words; ++w) dst[w] = row[w];
}

// ------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------
static CellMeta cell_meta(const Agg& a) {
  CellMeta m{};
  m.n_cells = a.n_cells;
  for (int c = 0; c < a.n_cells; ++c) { m.kind[c] = a.cells[c].kind; m.skip[c] = a.cells[c].skip; m.alias[c] = a.cells[c].alias; }
  return m;
}

static int find_or_add_cell(Agg& a, uint8_t kind, int input) {
  for (int c = 0; c < a.n_cells; ++c) if (a.cells[c].kind == kind && a.cells[c].input == input) return c;
  if (a.n_cells >= kMaxCells) return -1;
  DevCell& cell = a.cells[a.n_cells];
  cell = DevCell{};
  cell.kind = kind; cell.input = (int8_t)input;
  return a.n_cells++;
}

// Builds the cell plan from the accumulator list once operand types are known.
int agg_resolve(Agg& a, const Program& prog, cudaStream_t stream) {
  if (a.resolved) return VGPU_OK;
  a.n_cells = 0;
  find_or_add_cell(a, kCellCount, -1); // cell 0: passing rows per group
  a.rm = RowMap{};
  a.rm.n_accs = (int)a.accs.size();
  a.rm.rows_cell = 0;
  a.rm.row_words_out = 3 + 2 * a.rm.n_accs;
  for (size_t i = 0; i < a.accs.size(); ++i) {
    const vgpu_acc_spec& s = a.accs[i];
    if (s.input >= (int)prog.outputs.size()) { set_error("accumulator %zu reads output %d, program has %zu", i, s.input, prog.outputs.size()); return VGPU_ERR_INVALID; }
    const int t = s.input >= 0 ? prog.outputs[s.input].type : VGPU_BIGINT;
    const bool fp = t == VGPU_REAL || t == VGPU_DOUBLE;
    int kind;
    switch (s.kind) {
      case VGPU_ACC_SUM_F64: kind = kCellSumF64; break;
      case VGPU_ACC_SUM_I64: kind = kCellSumI64; if (fp) { set_error("SUM_I64 over a floating point input"); return VGPU_ERR_INVALID; } break;
      case VGPU_ACC_COUNT: kind = kCellCount; break;
      case VGPU_ACC_MIN: kind = fp ? kCellMinF64 : kCellMinI64; break;
      case VGPU_ACC_MAX: kind = fp ? kCellMaxF64 : kCellMaxI64; break;
      default: set_error("bad accumulator kind %d", s.kind); return VGPU_ERR_INVALID;
    }
    if (s.kind != VGPU_ACC_COUNT && s.input < 0) { set_error("accumulator %zu needs an input", i); return VGPU_ERR_INVALID; }
    // BOOLEAN inputs: MIN / MAX over FALSE(0) < TRUE(1) are bool_and / bool_or; sums of booleans are not defined
    if ((s.kind == VGPU_ACC_SUM_F64 || s.kind == VGPU_ACC_SUM_I64) && t == VGPU_BOOLEAN) { set_error("sum over a boolean input is not supported"); return VGPU_ERR_UNSUPPORTED; }
    const int vc = find_or_add_cell(a, (uint8_t)kind, s.input);
    int nc = -1;
    if (s.kind != VGPU_ACC_COUNT) nc = find_or_add_cell(a, kCellCount, s.input);
    if (vc < 0 || (s.kind != VGPU_ACC_COUNT && nc < 0)) { set_error("too many accumulator cells (max %d)", kMaxCells); return VGPU_ERR_UNSUPPORTED; }
    a.rm.value_cell[i] = (int16_t)vc;
    a.rm.nn_cell[i] = (int16_t)nc;
    a.rm.value_kind[i] = (uint8_t)kind;
  }
  // keys
  if (a.mode == VGPU_AGG_ARRAY || a.mode == VGPU_AGG_GLOBAL) {
    int64_t G = 1;
    for (size_t k = 0; k < a.keys.size(); ++k) {
      const vgpu_key_spec& ks = a.keys[k];
      const int64_t card = ks.max - ks.min + 1 + (a.ignore_null_keys ? 0 : 1);
      if (ks.max < ks.min || card > (1 << 20)) { set_error("ARRAY key %zu: bad domain [%lld, %lld]", k, (long long)ks.min, (long long)ks.max); return VGPU_ERR_INVALID; }
      a.mults[k] = G;
      G *= card;
      if (G > (1 << 20)) { set_error("ARRAY mode supports at most 2^20 groups; use HASH"); return VGPU_ERR_UNSUPPORTED; }
    }
    a.n_groups = (int)G;
    const size_t words = (size_t)a.n_cells * a.n_groups;
    if (!fused_dry_run()) {
      VG_CUDA(cudaMalloc(&a.state, words * 8));
      // on the caller's stream: cudaStreamNonBlocking streams have no implicit ordering with stream 0
      init_state_kernel<<<(int)((words + 255) / 256), 256, 0, stream>>>(a.state, cell_meta(a), a.n_groups);
      VG_CUDA(cudaGetLastError());
      count_launch();
    }
  } else {
    int shift = 0;
    for (size_t k = 0; k < a.keys.size(); ++k) {
      const vgpu_key_spec& ks = a.keys[k];
      if (ks.bits == 64) {
        if (a.keys.size() != 1) { set_error("a raw 64-bit key must be the only key"); return VGPU_ERR_UNSUPPORTED; }
        continue;
      }
      if (ks.bits < 1 || ks.bits > 63) { set_error("key %zu: bits %d", k, ks.bits); return VGPU_ERR_INVALID; }
      // ids are value - min + 1 (0 = NULL unless NULL keys are dropped): the largest one must fit the field
      const uint64_t top = (uint64_t)ks.max - (uint64_t)ks.min + (a.ignore_null_keys ? 0u : 1u);
      if (ks.max < ks.min || top > ((1ull << ks.bits) - 1ull)) { set_error("key %zu: domain [%lld, %lld] does not fit %d bits", k, (long long)ks.min, (long long)ks.max, ks.bits); return VGPU_ERR_INVALID; }
      a.shifts[k] = (uint8_t)shift;
      shift += ks.bits;
      if (shift > 63) { set_error("normalized key wider than 63 bits"); return VGPU_ERR_UNSUPPORTED; }
    }
    // key + cells, padded so that a row never straddles more 32-byte sectors than it has to
    int rw = 1 + a.n_cells;
    if (rw == 3) rw = 4; else if (rw > 4 && rw < 8) rw = 8; else if (rw > 8) rw = (rw + 3) & ~3;
    a.row_words = rw;
  }
  if (!fused_dry_run()) {
    a.err = err_word_acquire();
    VG_CUDA(cudaMalloc(&a.ctr, kCtrWords * sizeof(unsigned long long)));
    VG_CUDA(cudaMemsetAsync(a.ctr, 0, kCtrWords * sizeof(unsigned long long), stream));
  }
  a.resolved = true;
  return VGPU_OK;
}

bool agg_raw_bigint_key(const Agg& a) { return a.mode == VGPU_AGG_HASH && a.keys.size() == 1 && a.keys[0].bits == 64; }

static uint64_t next_pow2(uint64_t x) { uint64_t p = 1024; while (p < x) p <<= 1; return p; }

constexpr int kOvfCursor = kCtrOvfCursor;
constexpr int kDefaultBucketBits = 6; // profiles/r01_c3_experiments.md: 3..7 measured within 4 % of each other
static KeyMap key_map_of(const Agg& a) {
  KeyMap m{};
  // Measured on B200 (profiles/r01_notes.md): for sorted l_orderkey the monotone placement is
  // SLOWER than hashing (3.3 ms vs 2.75 ms at SF10) — neighbouring groups land in one 128-byte line
  // and their CAS / RED operations serialise in the L2 atomic unit.  Kept behind
  // VGPU_RANGED_KEYS=1 for experiments; the default is the hashed placement.
  static const bool enabled = getenv("VGPU_RANGED_KEYS") != nullptr && getenv("VGPU_RANGED_KEYS")[0] == '1';
  // Bucketed placement (mode 3): keys that share key >> b land in one aligned bucket of 2^b slots,
  // each at its own (rotated) position, so sorted / clustered keys fill whole lines of the table
  // while random keys still spread uniformly.  VGPU_BUCKET_BITS=0 selects the plain hash.
  static const int bucket_bits = getenv("VGPU_BUCKET_BITS") ? atoi(getenv("VGPU_BUCKET_BITS")) : kDefaultBucketBits;
  if (!enabled && bucket_bits > 0 && bucket_bits <= 8) {
    m.mode = 3;
    m.pad = (uint32_t)bucket_bits;
    return m;
  }
  if (enabled && a.keys.size() == 1 && a.keys[0].bits == 64 && a.keys[0].max > a.keys[0].min && a.capacity > 0) {
    m.min = (uint64_t)a.keys[0].min;
    m.span = (uint64_t)a.keys[0].max - (uint64_t)a.keys[0].min;
    if (m.span < a.capacity) {
      m.mode = 1;
    } else {
      m.mode = 2;
      m.scale = (uint64_t)((((unsigned __int128)a.capacity) << 64) / ((unsigned __int128)m.span + 1));
    }
  }
  return m;
}

static HashRef href_of(const Agg& a) {
  HashRef h; h.table = a.table; h.cap_mask = a.capacity - 1; h.key_map = key_map_of(a); h.row_words = a.row_words; h.n_groups_ctr = a.ctr;
  return h;
}

// (Re)allocates the table with `capacity` slots and moves existing groups over.
static int hash_resize(Agg& a, uint64_t capacity, cudaStream_t stream) {
  uint64_t* old = a.table;
  const uint64_t old_slots = a.capacity ? a.capacity + 2 : 0;
  uint64_t* fresh = nullptr;
  const uint64_t slots = capacity + 2;
  VG_CUDA(cudaMalloc(&fresh, slots * a.row_words * 8));
  const int blocks = std::max(1, std::min(num_sms() * 8, (int)((slots * a.row_words + 255) / 256)));
  init_table_kernel<<<blocks, 256, 0, stream>>>(fresh, slots, a.row_words, cell_meta(a));
  VG_CUDA(cudaGetLastError());
  count_launch();
  a.table = fresh;
  a.capacity = capacity;
  if (old) {
    VG_CUDA(cudaMemsetAsync(a.ctr, 0, sizeof(unsigned long long), stream));
    rehash_kernel<<<(unsigned)((old_slots + 255) / 256), 256, 0, stream>>>(old, old_slots, href_of(a), a.err);
    VG_CUDA(cudaGetLastError());
    count_launch();
    VG_CUDA(cudaStreamSynchronize(stream));
    VG_CUDA(cudaFree(old));
  }
  return VGPU_OK;
}

static int read_group_count(Agg& a, cudaStream_t stream, uint64_t* out) {
  unsigned long long v = 0;
  VG_CUDA(cudaMemcpyAsync(&v, a.ctr, sizeof(v), cudaMemcpyDeviceToHost, stream));
  VG_CUDA(cudaStreamSynchronize(stream));
  *out = v;
  return VGPU_OK;
}

static void slice_columns(const vgpu_column* in, int n, int64_t off, vgpu_column* out) {
  for (int c = 0; c < n; ++c) {
    out[c] = in[c];
    if (off == 0 || in[c].encoding == VGPU_CONSTANT) continue;
    if (in[c].nulls) out[c].nulls = in[c].nulls + (off >> 6);
    if (in[c].encoding == VGPU_DICTIONARY) { out[c].indices = in[c].indices + off; continue; }
    int lg;
    switch (in[c].type) { case VGPU_BOOLEAN: lg = -3; break; case VGPU_TINYINT: lg = 0; break; case VGPU_SMALLINT: lg = 1; break;
      case VGPU_INTEGER: case VGPU_REAL: lg = 2; break; default: lg = 3; }
    out[c].values = (const char*)in[c].values + (lg < 0 ? (off >> 3) : (off << lg));
  }
}

static int fill_sink(Agg& a, PipelineParams& p, GroupMap* gm = nullptr) {
  p.n_keys = (int)a.keys.size();
  p.ignore_null_keys = a.ignore_null_keys ? 1 : 0;
  int64_t compact = 1;
  if (gm) { *gm = GroupMap{}; gm->n_keys = p.n_keys; }
  for (int k = 0; k < p.n_keys; ++k) {
    DevKey& d = p.keys[k];
    d = DevKey{};
    if (a.keys[k].output < 0 || a.keys[k].output >= p.n_outputs) { set_error("key %d reads output %d", k, a.keys[k].output); return VGPU_ERR_INVALID; }
    d.min = a.keys[k].min; d.max = a.keys[k].max; d.mult = a.mults[k];
    d.output = (int8_t)a.keys[k].output; d.bits = (uint8_t)a.keys[k].bits; d.shift = a.shifts[k];
    d.nullable = a.ignore_null_keys ? 0 : 1;
    if (gm) {
      // ARRAY: this launch numbers groups without the NULL id of keys that cannot be NULL here
      const bool can_null = !a.ignore_null_keys && output_nullable(p, a.keys[k].output);
      const int64_t card = a.keys[k].max - a.keys[k].min + 1 + (can_null ? 1 : 0);
      gm->card[k] = (int32_t)card;
      gm->shift[k] = (!a.ignore_null_keys && !can_null) ? 1 : 0;
      gm->mult[k] = a.mults[k];
      d.nullable = can_null ? 1 : 0;
      d.mult = compact;
      compact *= card;
    }
  }
  if (gm) p.n_groups = (int)compact;
  p.n_cells = a.n_cells;
  for (int c = 0; c < a.n_cells; ++c) {
    DevCell& d = p.cells[c];
    d = a.cells[c];
    d.skip = 0; d.alias = 0;
    // a count cell whose input cannot be NULL counts exactly the passing rows: alias cell 0
    if (c > 0 && d.kind == kCellCount && (d.input < 0 || !output_nullable(p, d.input))) { d.skip = 1; d.alias = 0; }
    a.cells[c].skip = d.skip; a.cells[c].alias = d.alias;
  }
  return VGPU_OK;
}

// HASH mode: count cells that have aliased the rows cell so far get their own value before a
// batch (or a merge) that needs them to diverge.
__global__ void materialize_alias_kernel(uint64_t* table, uint64_t n_slots, int row_words, int cell) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_slots) return;
  uint64_t* row = table + i * row_words;
  if (row[0] != kEmptyKey) row[1 + cell] = row[1];
}

static int hash_materialize(Agg& a, int cell, cudaStream_t stream) {
  if (a.table) {
    const uint64_t slots = a.capacity + 2;
    materialize_alias_kernel<<<(unsigned)((slots + 255) / 256), 256, 0, stream>>>(a.table, slots, a.row_words, cell);
    VG_CUDA(cudaGetLastError());
    count_launch();
  }
  a.hash_aliased[cell] = false;
  return VGPU_OK;
}

// LOG -> TABLE: finalize, expand the groups to ABI rows and fold them into the HBM table.
static int log_migrate_to_table(Agg& a, cudaStream_t stream) {
  int rc = agg_extract_all(a, stream); // finalize + expand into a.packed
  if (rc) return rc;
  const int64_t n = a.packed_rows;
  uint64_t* rows = a.packed;
  a.packed = nullptr; a.packed_cap = 0; a.packed_rows = -1;
  a.engine = kEngineTable;
  log_reset(a, stream);
  rc = n > 0 ? agg_merge_rows(a, rows, n, stream) : VGPU_OK;
  cudaStreamSynchronize(stream);
  cudaFree(rows);
  return rc;
}

int agg_wait_pending(Agg& a, cudaStream_t stream) {
  if (!a.pending_set) return VGPU_OK;
  VG_CUDA(cudaStreamWaitEvent(stream, a.pending, 0));
  a.pending_set = false;
  return VGPU_OK;
}

int agg_add_input(Agg& a, const Program& prog, const vgpu_column* cols, int32_t ncols, int64_t nrows,
                  const int32_t* in_rows, int64_t n_in_rows, cudaStream_t stream) {
  { const int wrc = agg_wait_pending(a, stream); if (wrc) return wrc; }
  int rc = agg_resolve(a, prog, stream);
  if (rc) return rc;
  a.packed_rows = -1;
  const int64_t total = in_rows ? n_in_rows : nrows;
  if (total <= 0) return VGPU_OK;
  PipelineParams p{};
  if (a.mode != VGPU_AGG_HASH) {
    rc = plan_pipeline(prog, cols, ncols, nrows, in_rows, n_in_rows, p);
    if (rc) return rc;
    p.mode = kSinkArray;
    p.err_word = a.err;
    GroupMap gm;
    rc = fill_sink(a, p, &gm);
    if (rc) return rc;
    const int Gc = p.n_groups;
    {
      // plan-specialised kernel first; the interpreter takes the shapes it declines
      int fgrid = 0;
      FusedKernel* fk = fused_prepare(p, &fgrid);
      if (fused_dry_run()) return VGPU_OK;
      if (!fk && fused_strict_failure()) return VGPU_ERR_INVALID;
      if (fk) {
        const size_t need = (size_t)fgrid * a.n_cells * Gc;
        if (need > a.partials_cap) {
          if (a.partials) VG_CUDA(cudaFree(a.partials));
          VG_CUDA(cudaMalloc(&a.partials, need * 8));
          a.partials_cap = need;
        }
        p.partials = a.partials;
        rc = fused_launch(fk, p, fgrid, stream);
        if (rc) return rc;
        const int cells = a.n_cells * Gc;
        merge_partials_kernel<<<(cells + 7) / 8, 256, 0, stream>>>(a.state, a.partials, fgrid, cell_meta(a), Gc, a.n_groups, gm, a.err);
        VG_CUDA(cudaGetLastError());
        count_launch();
        return VGPU_OK;
      }
    }
    // shared-memory plan of the partial tables: one table per thread when it fits next to at
    // least two stages, else one per warp.
    auto layout = [&](int copies) {
      int off = 0;
      for (int c = 0; c < p.n_cells; ++c) {
        if (p.cells[c].skip) { p.cells[c].smem_off = 0; continue; }
        const int w = p.cells[c].kind == kCellCount ? 4 : 8;
        off = (off + 7) & ~7;
        p.cells[c].smem_off = off;
        off += Gc * copies * w;
      }
      return (off + 15) & ~15;
    };
    const int fixed = 128 + p.n_slots * (kTile * 8 + kThreads * 4) + p.n_bslots * kThreads * 4;
    const int budget = 227 * 1024 - fixed - 2 * p.stage_bytes - 256;
    int tab = layout(kThreads);
    p.warp_private = 0;
    if (tab > budget) {
      const int one = layout(1);
      p.warp_private = 1;
      p.warp_tab_bytes = one;
      tab = one * kWarps;
      if (tab > budget) { set_error("ARRAY mode: %d groups x %d cells do not fit shared memory; use HASH", Gc, a.n_cells); return VGPU_ERR_UNSUPPORTED; }
    }
    p.tab_bytes = tab;
    const int grid = pipeline_grid(p.n_rows);
    const size_t need = (size_t)grid * a.n_cells * Gc;
    if (need > a.partials_cap) {
      if (a.partials) VG_CUDA(cudaFree(a.partials));
      VG_CUDA(cudaMalloc(&a.partials, need * 8));
      a.partials_cap = need;
    }
    p.partials = a.partials;
    rc = launch_pipeline(p, tab, grid, stream);
    if (rc) return rc;
    const int cells = a.n_cells * Gc;
    merge_partials_kernel<<<(cells + 7) / 8, 256, 0, stream>>>(a.state, a.partials, grid, cell_meta(a), Gc, a.n_groups, gm, a.err);
    VG_CUDA(cudaGetLastError());
    count_launch();
    return VGPU_OK;
  }
  // LOG engine (hashagg.cu): the scan appends run-aggregated entries, finalize deduplicates them.  Used while
  // every batch has a plan-specialised kernel and updates only cells the entry layout holds; the first batch that
  // does not moves the groups into the HBM table (TABLE engine) for good.
  // (the variable is read per call so that tools can compare the two engines in one process)
  const bool log_enabled = !(getenv("VGPU_HASH_ENGINE") && strcmp(getenv("VGPU_HASH_ENGINE"), "table") == 0);
  if ((log_enabled || a.engine == kEngineLog) && a.engine != kEngineTable) {
    rc = plan_pipeline(prog, cols, ncols, nrows, in_rows, n_in_rows, p);
    if (rc) return rc;
    p.err_word = a.err;
    p.mode = kSinkLog;
    rc = fill_sink(a, p);
    if (rc) return rc;
    bool ok = true;
    if (a.engine == kEngineUndecided) { rc = log_setup(a, p, stream); if (rc) return rc; }
    else ok = log_layout_covers(a, p);
    FusedKernel* fk = nullptr;
    int fgrid = 0;
    if (ok) {
      p.log_entry_words = a.ew;
      p.log_n_cells = a.log_n_cells;
      for (int i = 0; i < a.log_n_cells; ++i) p.log_cells[i] = a.log_cells[i];
      // sorted-input probe: one raw 64-bit key read straight from a FLAT BIGINT column without NULLs, no filter
      p.sorted_probe = 0;
      if (prog.filter.kind == VGPU_OPND_NONE && a.keys.size() == 1 && a.keys[0].bits == 64 && in_rows == nullptr) {
        const vgpu_operand& ko = p.outputs[a.keys[0].output];
        if (ko.kind == VGPU_OPND_COL && p.cols[ko.index].encoding == VGPU_FLAT && p.cols[ko.index].type == VGPU_BIGINT && p.cols[ko.index].nulls == nullptr) {
          p.sorted_probe = 1;
          p.sorted_key_col = ko.index;
        }
      }
      // exchange fused in: the entries leave for their owners' windows from inside the scan
      const bool send = send_wanted(a, p);
      fk = fused_prepare(p, &fgrid);
      if (fused_dry_run()) return VGPU_OK;
      if (!fk && fused_strict_failure()) return VGPU_ERR_INVALID;
      if (!fk && send && !a.send_used) { // the streaming variant does not cover this shape (entry too wide for its buffers): log locally
        p.send_world = 0; p.send_hash = 0;
        fk = fused_prepare(p, &fgrid);
        if (!fk && fused_strict_failure()) return VGPU_ERR_INVALID;
      } else if (fk && send) {
        rc = send_bind(a, p, fgrid, stream);
        if (rc == VGPU_ERR_UNSUPPORTED && !a.send_used) { // log locally instead; the exchange partitions afterwards
          p.send_world = 0; p.send_hash = 0;
          fk = fused_prepare(p, &fgrid);
          if (!fk && fused_strict_failure()) return VGPU_ERR_INVALID;
        } else if (rc) return rc;
      } else if (a.send_used) {
        set_error("this batch cannot use the fused exchange the aggregation's earlier batches used");
        return VGPU_ERR_UNSUPPORTED;
      }
    }
    if (fk) {
      a.engine = kEngineLog;
      const int warps = fused_kernel_threads(fk) / 32;
      const uint64_t n_blocks = ((uint64_t)p.n_rows + 127) / 128;
      const uint64_t per_warp = (n_blocks + (uint64_t)fgrid * warps - 1) / ((uint64_t)fgrid * warps);
      rc = log_begin_batch(a, p, fgrid, per_warp * warps * 128, stream, p.send_world > 0);
      if (rc) return rc;
      if (p.send_world > 0) a.send_used = true;
      a.log_bound += (uint64_t)p.n_rows;
      a.finalized = false;
      if (p.sorted_probe && a.log_batches == 1) // the stream's first key, for the exchange's "sources are disjoint" test
        VG_CUDA(cudaMemcpyAsync(a.ctr + kCtrFirstKey, p.cols[p.sorted_key_col].values, 8, cudaMemcpyDeviceToDevice, stream));
      return fused_launch(fk, p, fgrid, stream);
    }
    if (fused_dry_run()) return VGPU_OK;
    if (a.engine == kEngineLog) { rc = log_migrate_to_table(a, stream); if (rc) return rc; }
    a.engine = kEngineTable;
    p = PipelineParams{};
  }
  if (fused_dry_run()) {
    rc = plan_pipeline(prog, cols, ncols, nrows, in_rows, n_in_rows, p);
    if (rc) return rc;
    p.mode = kSinkHash;
    rc = fill_sink(a, p);
    if (rc) return rc;
    p.row_words = a.row_words;
    int fgrid = 0;
    fused_prepare(p, &fgrid);
    return VGPU_OK;
  }
  a.engine = kEngineTable;
  // HASH: walk the batch in pieces small enough that the table cannot fill up.
  int64_t done = 0;
  std::vector<vgpu_column> sliced(ncols);
  while (done < total) {
    const int64_t rem = total - done;
    uint64_t known = 0;
    if (a.table) {
      // the exact count costs a stream synchronise: skip it while even "every row so far was a new group" fits
      if (a.groups_bound + (uint64_t)rem <= a.capacity - a.capacity / 16) known = a.groups_bound;
      else { rc = read_group_count(a, stream, &known); if (rc) return rc; a.groups_bound = known; }
    }
    // every row of a piece may start a group: a piece may take the table to 15/16 full at worst (still
    // terminating; the reference rehashes at 7/8, velox/exec/HashTable.h:562-565).  Tables are sized for <= 25-50 %
    // load from the group-count hint, so real inputs with repeated keys stay far below that; the looser bound only
    // keeps a batch that is slightly larger than 7/8 of the table from being split into a second, tiny piece
    // (one more host read of the group count and two more launches: ~0.04 ms of C3's 0.74).
    const uint64_t limit = a.capacity - a.capacity / 16;
    uint64_t room = limit > known ? limit - known : 0;
    if (room < (uint64_t)std::min<int64_t>(rem, kTile)) {
      uint64_t want = known + (uint64_t)std::min<int64_t>(rem, std::max<int64_t>(4 << 20, (int64_t)known));
      if (a.expected_groups > 0) want = std::max<uint64_t>(want, (uint64_t)a.expected_groups);
      // Slots per expected group.  4 (load <= 25 %) parks half as many segments as 2 did: C3 SF10 1.20 -> 0.93 ms per
      // batch; the price is a table init twice as long (0.34 -> 0.68 ms, once per aggregation) and twice the HBM, so
      // tables that would pass 8 GB keep the factor 2 (profiles/r01_c3_experiments.md).  VGPU_TABLE_FACTOR overrides.
      static const uint64_t env_factor = getenv("VGPU_TABLE_FACTOR") ? std::max(2, atoi(getenv("VGPU_TABLE_FACTOR"))) : 0;
      uint64_t table_factor = env_factor ? env_factor : 4;
      if (!env_factor && next_pow2(table_factor * want) * (uint64_t)a.row_words * 8 > (8ull << 30)) table_factor = 2;
      rc = hash_resize(a, next_pow2(table_factor * want), stream);
      if (rc) return rc;
      room = a.capacity - a.capacity / 16 - known;
    }
    int64_t piece = std::min<int64_t>(rem, (int64_t)room);
    if (piece < rem) piece = std::max<int64_t>(kTile, piece / kTile * kTile); // keep slices tile aligned
    const vgpu_column* use = cols;
    if (in_rows == nullptr && done > 0) { slice_columns(cols, ncols, done, sliced.data()); use = sliced.data(); }
    p = PipelineParams{};
    rc = plan_pipeline(prog, use, ncols, piece, in_rows ? in_rows + done : nullptr, piece, p);
    if (rc) return rc;
    p.mode = kSinkHash;
    p.err_word = a.err;
    rc = fill_sink(a, p);
    if (rc) return rc;
    for (int c = 1; c < a.n_cells; ++c) {
      if (a.cells[c].kind != kCellCount) continue;
      if (!p.cells[c].skip && a.hash_aliased[c]) { rc = hash_materialize(a, c, stream); if (rc) return rc; }
      if (!a.hash_aliased[c]) p.cells[c].skip = 0; // once materialised the cell is always updated
    }
    p.table = a.table; p.cap_mask = a.capacity - 1; p.key_map = key_map_of(a); p.row_words = a.row_words; p.n_groups_ctr = a.ctr;
    int fgrid = 0;
    FusedKernel* fk = fused_prepare(p, &fgrid);
    if (!fk && fused_strict_failure()) return VGPU_ERR_INVALID;
    static const bool parking = !(getenv("VGPU_NO_PARKING") && getenv("VGPU_NO_PARKING")[0] == '1');
    if (fk && parking) {
      const uint64_t want = std::max<uint64_t>(1u << 16, a.capacity / 8); // a full buffer only means the scan warp walks the table itself
      if (a.ovf_cap < want) {
        if (a.ovf) { VG_CUDA(cudaStreamSynchronize(stream)); VG_CUDA(cudaFree(a.ovf)); a.ovf = nullptr; }
        VG_CUDA(cudaMalloc(&a.ovf, want * a.row_words * 8));
        a.ovf_cap = want;
      }
      VG_CUDA(cudaMemsetAsync(a.ctr + kOvfCursor, 0, sizeof(unsigned long long), stream));
      p.ovf = a.ovf; p.ovf_cap = a.ovf_cap; p.ovf_cursor = a.ctr + kOvfCursor;
    }
    rc = fk ? fused_launch(fk, p, fgrid, stream) : launch_pipeline(p, 0, pipeline_grid(p.n_rows), stream);
    if (rc) return rc;
    if (p.ovf) {
      drain_overflow_kernel<<<num_sms() * 8, 256, 0, stream>>>(a.ovf, a.ctr + kOvfCursor, a.ovf_cap, href_of(a), cell_meta(a), a.err);
      VG_CUDA(cudaGetLastError());
      count_launch();
    }
    done += piece;
    a.groups_bound += (uint64_t)piece;
  }
  return VGPU_OK;
}

int agg_extract_all(Agg& a, cudaStream_t stream) {
  { const int wrc = agg_wait_pending(a, stream); if (wrc) return wrc; }
  if (a.packed_rows >= 0) return VGPU_OK;
  if (!a.resolved) { set_error("aggregation has seen no input; call add_input (an empty batch is fine) first"); return VGPU_ERR_INVALID; }
  size_t max_rows;
  if (a.mode == VGPU_AGG_HASH && a.engine == kEngineLog) {
    int rc = log_finalize(a, stream);
    if (rc) return rc;
    FinHeader h{};
    uint64_t n_special = 0;
    rc = log_read_header(a, stream, &h, &n_special);
    if (rc) return rc;
    if (h.overflow) { set_error("internal: a hash partition overflowed its shared-memory table"); return VGPU_ERR_INVALID; }
    max_rows = (size_t)(h.n_groups + (h.sorted ? h.n_front : 0) + n_special);
    if (max_rows > a.packed_cap || a.packed == nullptr) {
      if (a.packed) VG_CUDA(cudaFree(a.packed));
      a.packed_cap = std::max<size_t>(max_rows, 16);
      VG_CUDA(cudaMalloc(&a.packed, a.packed_cap * a.rm.row_words_out * 8));
    }
    VG_CUDA(cudaMemsetAsync(a.ctr + 1, 0, sizeof(unsigned long long), stream));
    rc = log_expand_rows(a, a.packed, a.ctr + 1, stream);
    if (rc) return rc;
    unsigned long long n = 0;
    VG_CUDA(cudaMemcpyAsync(&n, a.ctr + 1, sizeof(n), cudaMemcpyDeviceToHost, stream));
    VG_CUDA(cudaStreamSynchronize(stream));
    a.packed_rows = (int64_t)n;
    return VGPU_OK;
  }
  if (a.mode == VGPU_AGG_HASH) {
    uint64_t known = 0;
    if (a.table) { int rc = read_group_count(a, stream, &known); if (rc) return rc; }
    max_rows = known;
  } else {
    max_rows = a.n_groups;
  }
  if (max_rows > a.packed_cap || a.packed == nullptr) {
    if (a.packed) VG_CUDA(cudaFree(a.packed));
    a.packed_cap = std::max<size_t>(max_rows, 16);
    VG_CUDA(cudaMalloc(&a.packed, a.packed_cap * a.rm.row_words_out * 8));
  }
  VG_CUDA(cudaMemsetAsync(a.ctr + 1, 0, sizeof(unsigned long long), stream));
  if (a.mode == VGPU_AGG_HASH) {
    if (a.table && max_rows > 0) {
      RowMap rm = a.rm;
      for (int i = 0; i < rm.n_accs; ++i) {
        if (rm.nn_cell[i] >= 0 && a.hash_aliased[rm.nn_cell[i]]) rm.nn_cell[i] = (int16_t)rm.rows_cell;
        if (rm.value_kind[i] == kCellCount && a.hash_aliased[rm.value_cell[i]]) rm.value_cell[i] = (int16_t)rm.rows_cell;
      }
      const uint64_t slots = a.capacity + 2;
      extract_hash_kernel<<<(unsigned)((slots + 256 * kExtractPerThread - 1) / (256 * kExtractPerThread)), 256, 0, stream>>>(a.table, slots, a.row_words, rm, a.packed, a.ctr + 1);
      VG_CUDA(cudaGetLastError());
      count_launch();
    }
  } else {
    extract_array_kernel<<<(a.n_groups + 255) / 256, 256, 0, stream>>>(a.state, a.n_groups, a.rm, a.mode == VGPU_AGG_GLOBAL && a.emit_empty_global ? 1 : 0, a.packed, a.ctr + 1);
    VG_CUDA(cudaGetLastError());
    count_launch();
  }
  unsigned long long n = 0;
  VG_CUDA(cudaMemcpyAsync(&n, a.ctr + 1, sizeof(n), cudaMemcpyDeviceToHost, stream));
  VG_CUDA(cudaStreamSynchronize(stream));
  a.packed_rows = (int64_t)n;
  return VGPU_OK;
}

static DoneMap done_map(const RowMap& rm) {
  DoneMap dm{};
  for (int i = 0; i < rm.n_accs; ++i) {
    for (int j = 0; j < i; ++j) {
      if (rm.value_cell[j] == rm.value_cell[i] || rm.nn_cell[j] == rm.value_cell[i]) dm.d[2 * i] = 1;
      if (rm.nn_cell[i] >= 0 && (rm.nn_cell[j] == rm.nn_cell[i] || rm.value_cell[j] == rm.nn_cell[i])) dm.d[2 * i + 1] = 1;
    }
    if (rm.nn_cell[i] >= 0 && rm.nn_cell[i] == rm.value_cell[i]) dm.d[2 * i + 1] = 1;
  }
  return dm;
}

int agg_merge_rows(Agg& a, const void* rows, int64_t n, cudaStream_t stream) {
  { const int wrc = agg_wait_pending(a, stream); if (wrc) return wrc; }
  if (!a.resolved) { set_error("merge_rows before the cell plan is known; call add_input with an empty batch first"); return VGPU_ERR_INVALID; }
  a.packed_rows = -1;
  if (n <= 0) return VGPU_OK;
  const DoneMap dm = done_map(a.rm);
  if (a.mode != VGPU_AGG_HASH) {
    merge_rows_array_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>((const uint64_t*)rows, n, a.rm, dm, a.state, a.n_groups, a.err, 0, 0);
    VG_CUDA(cudaGetLastError());
    count_launch();
    return VGPU_OK;
  }
  if (a.engine == kEngineLog) { int rc = log_migrate_to_table(a, stream); if (rc) return rc; }
  a.engine = kEngineTable;
  uint64_t known = 0;
  if (a.table) { int rc = read_group_count(a, stream, &known); if (rc) return rc; }
  if (a.capacity / 2 < known + (uint64_t)n) {
    int rc = hash_resize(a, next_pow2(2 * (known + (uint64_t)n)), stream);
    if (rc) return rc;
  }
  a.groups_bound = known + (uint64_t)n;
  for (int c = 1; c < a.n_cells; ++c) {
    if (a.cells[c].kind == kCellCount && a.hash_aliased[c]) { int rc = hash_materialize(a, c, stream); if (rc) return rc; }
  }
  merge_rows_hash_kernel<<<(unsigned)std::min<int64_t>((n + 255) / 256, (int64_t)num_sms() * 16), 256, 0, stream>>>((const uint64_t*)rows, n, a.rm, dm, href_of(a), a.err);
  VG_CUDA(cudaGetLastError());
  count_launch();
  return VGPU_OK;
}

// Small-domain (ARRAY / GLOBAL) exchange helpers: no host synchronisation.
int agg_dense_rows(Agg& a, void* rows_out, cudaStream_t stream) {
  { const int wrc = agg_wait_pending(a, stream); if (wrc) return wrc; }
  if (a.mode == VGPU_AGG_HASH || !a.resolved) { set_error("dense rows need a resolved ARRAY / GLOBAL aggregation"); return VGPU_ERR_INVALID; }
  dense_rows_kernel<<<(a.n_groups + 255) / 256, 256, 0, stream>>>(a.state, a.n_groups, a.rm, (uint64_t*)rows_out);
  VG_CUDA(cudaGetLastError());
  count_launch();
  return VGPU_OK;
}
int agg_merge_owned_rows(Agg& a, const void* rows, int64_t n, int n_parts, int my_part, cudaStream_t stream) {
  { const int wrc = agg_wait_pending(a, stream); if (wrc) return wrc; }
  if (a.mode == VGPU_AGG_HASH || !a.resolved) { set_error("owned-row merge needs a resolved ARRAY / GLOBAL aggregation"); return VGPU_ERR_INVALID; }
  a.packed_rows = -1;
  if (n <= 0) return VGPU_OK;
  merge_rows_array_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>((const uint64_t*)rows, n, a.rm, done_map(a.rm), a.state, a.n_groups, a.err, n_parts, my_part);
  VG_CUDA(cudaGetLastError());
  count_launch();
  return VGPU_OK;
}

int agg_partition(Agg& a, int n_parts, void* rows_out, int64_t* counts_out, cudaStream_t stream) {
  { const int wrc = agg_wait_pending(a, stream); if (wrc) return wrc; }
  if (n_parts < 1 || n_parts > 32) { set_error("n_parts %d out of range", n_parts); return VGPU_ERR_INVALID; }
  int rc = agg_extract_all(a, stream);
  if (rc) return rc;
  const int64_t n = a.packed_rows;
  unsigned long long* counts = a.ctr + 2;   // [32]
  VG_CUDA(cudaMemsetAsync(counts, 0, 32 * sizeof(unsigned long long), stream));
  unsigned long long h[32] = {0};
  if (n > 0) {
    const bool velox_hash = agg_raw_bigint_key(a);
    partition_count_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(a.packed, n, a.rm.row_words_out, n_parts, counts, velox_hash);
    VG_CUDA(cudaGetLastError());
    count_launch();
    VG_CUDA(cudaMemcpyAsync(h, counts, n_parts * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
    VG_CUDA(cudaStreamSynchronize(stream));
    unsigned long long cur[32], run = 0;
    for (int i = 0; i < n_parts; ++i) { cur[i] = run; run += h[i]; }
    VG_CUDA(cudaMemcpyAsync(counts, cur, n_parts * sizeof(unsigned long long), cudaMemcpyHostToDevice, stream));
    partition_scatter_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(a.packed, n, a.rm.row_words_out, n_parts, counts, (uint64_t*)rows_out, velox_hash);
    VG_CUDA(cudaGetLastError());
    count_launch();
    VG_CUDA(cudaStreamSynchronize(stream));
  }
  for (int i = 0; i < n_parts; ++i) counts_out[i] = (int64_t)h[i];
  return VGPU_OK;
}

int agg_reset(Agg& a, cudaStream_t stream) {
  { const int wrc = agg_wait_pending(a, stream); if (wrc) return wrc; }
  a.packed_rows = -1;
  a.emit_empty_global = true;
  if (!a.resolved) return VGPU_OK;
  if (a.mode == VGPU_AGG_HASH && a.engine != kEngineTable) {
    // the entry layout and the slabs stay; a later ineligible batch may still move to the TABLE engine
    send_reset(a);
    return log_reset(a, stream);
  }
  if (a.mode == VGPU_AGG_HASH) {
    if (a.table) { // keep the allocation, wipe it in place
      const uint64_t slots = a.capacity + 2;
      const int blocks = std::max(1, std::min(num_sms() * 8, (int)((slots * a.row_words + 255) / 256)));
      init_table_kernel<<<blocks, 256, 0, stream>>>(a.table, slots, a.row_words, cell_meta(a));
      VG_CUDA(cudaGetLastError());
      count_launch();
    }
    for (int c = 0; c < kMaxCells; ++c) a.hash_aliased[c] = true;
    VG_CUDA(cudaMemsetAsync(a.ctr, 0, sizeof(unsigned long long), stream));
    a.groups_bound = 0;
  } else {
    const size_t words = (size_t)a.n_cells * a.n_groups;
    init_state_kernel<<<(int)((words + 255) / 256), 256, 0, stream>>>(a.state, cell_meta(a), a.n_groups);
    VG_CUDA(cudaGetLastError());
    count_launch();
  }
  return VGPU_OK;
}

int agg_finalize(Agg& a, cudaStream_t stream) {
  { const int wrc = agg_wait_pending(a, stream); if (wrc) return wrc; }
  if (a.mode == VGPU_AGG_HASH && a.engine == kEngineLog) return log_finalize(a, stream);
  return VGPU_OK; // the other sinks keep their groups final after every batch
}

void agg_free(Agg& a) {
  if (a.pending) { if (a.pending_set) cudaEventSynchronize(a.pending); cudaEventDestroy(a.pending); a.pending = nullptr; a.pending_set = false; }
  send_free(a);
  log_free(a);
  err_word_release(a.err);
  a.err = nullptr;
  cudaFree(a.state); cudaFree(a.partials); cudaFree(a.table); cudaFree(a.ctr); cudaFree(a.packed); cudaFree(a.ovf);
  a.ovf = nullptr; a.ovf_cap = 0;
  a.state = a.partials = a.table = a.packed = nullptr; a.ctr = nullptr;
}

} // namespace vg
