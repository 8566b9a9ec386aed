// Internal declarations shared by the CUDA translation units of libvgpu.so.
// Nothing here is part of the ABI (include/vgpu.h is).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "vgpu.h"

namespace vg {

// ---- tile geometry of the pipeline kernel ---------------------------------------------------
// One CTA of 256 threads walks tiles of 1024 rows; a thread owns 2 pairs of adjacent rows
// (warp-contiguous pair layout, see pipeline.cu) so every staged 8-byte column is read with
// conflict-free 128-bit shared loads.
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kRowsPerThread = 4;
constexpr int kTile = kThreads * kRowsPerThread; // 1024 rows
constexpr int kMaxStages = 4;

constexpr int kMaxOps = 48;
constexpr int kMaxCols = 16;
constexpr int kMaxOutputs = 24;
constexpr int kMaxKeys = 8;
constexpr int kMaxCells = 40;
constexpr int kMaxSlots = 12;

// Error bits latched by kernels into a device word (vgpu_check_errors reports them).
enum : uint32_t {
  kErrOverflow = 1u,     // "integer overflow" (checked arithmetic / checked sum)
  kErrDivZero = 2u,      // "division by zero"
  kErrCastRange = 4u,    // cast out of range / NaN to integer
  kErrKeyRange = 8u,     // grouping key outside the declared [min,max] or unexpected NULL
  kErrTableFull = 16u,   // HBM hash table could not place a key (internal; host grows + retries)
  kErrStringKey = 32u    // a VGPU_VARCHAR value longer than 7 bytes was used (the host maps those to ids itself)
};

struct DevCol {
  const void* values;
  const int32_t* indices;
  const uint64_t* nulls;
  const uint64_t* base_nulls;
  int64_t constant;
  int32_t stage_off; // byte offset of the staged values (FLAT) / indices (DICTIONARY) tile; -1 none
  int32_t nulls_off; // byte offset of the staged outer-nulls tile (128 B); -1 none
  uint8_t type;
  uint8_t encoding;
  uint8_t const_null;
  uint8_t elem_log2; // log2 bytes of a staged element (BOOLEAN flat: bits, marked 255)
  uint8_t bulk_ok;   // staged tile may be fetched with cp.async.bulk (16 B aligned, flat walk)
  uint8_t small_dict; // DICTIONARY with 1..32 base values: base_len; the generated kernels keep the base in one register per lane
  uint8_t pad[2];
};

// Cell = one 8-byte lane of group state.  An accumulator maps to a value cell and (for
// SUM/MIN/MAX) a "non-null input count" cell; count cells of the same input are shared.
enum CellKind : uint8_t {
  kCellSumF64 = 0,
  kCellSumI64 = 1,
  kCellCount = 2, // counts rows whose `input` is not null (input < 0: all passing rows)
  kCellMinI64 = 3,
  kCellMaxI64 = 4,
  kCellMinF64 = 5,
  kCellMaxF64 = 6
};

struct DevCell {
  uint8_t kind;
  int8_t input;     // program output index feeding the cell, -1 = none (count(*))
  uint8_t skip;     // launch-time: input cannot be NULL => count cell aliases the rows cell
  uint8_t alias;    // cell whose value this one copies at merge time when skip != 0
  int32_t smem_off; // ARRAY mode: byte offset of this cell's block inside the partial tables
};

struct DevKey {
  int64_t min;
  int64_t max;
  int64_t mult;  // ARRAY mode: multiplier of this key's id inside the group number
  int8_t output; // program output index
  uint8_t bits;  // HASH mode: width of the id field (64 = raw value)
  uint8_t shift; // HASH mode: bit position of the id field
  uint8_t nullable;
  uint8_t pad[4];
};

// Home slot of a key in the HBM table.  mode 0: mix64(key) & cap_mask.  When the caller declares
// the value range of a single raw key (vgpu_key_spec min < max — the VectorHasher range statistics
// of the reference, velox/exec/VectorHasher.h:118-578), keys inside it map MONOTONICALLY onto the
// table (mode 1: key - min when the range fits; mode 2: (key - min) * capacity / range), so sorted
// or clustered inputs walk the table sequentially instead of scattering one sector per group.
// Keys outside the range fall back to mode 0.  Every kernel that probes the table uses this.
struct KeyMap {
  uint64_t min;
  uint64_t span;  // max - min
  uint64_t scale; // mode 2: floor(2^64 * capacity / (span + 1))
  uint32_t mode;
  uint32_t pad;
};

// kSinkLog: HASH aggregation, log engine — the scan appends run-aggregated {key, cells...} entries to
// per-CTA regions of a log (no table, no global atomics); hashagg.cu deduplicates them afterwards.
enum SinkMode : int32_t { kSinkFilterProject = 0, kSinkArray = 1, kSinkHash = 2, kSinkLog = 3 };

struct DevOut {
  void* values;
  uint64_t* nulls;
};

// ---- micro-ops --------------------------------------------------------------------------------
// The ABI program (vgpu_op list + filter + outputs) and the sink (keys, cells, projections) are
// lowered per launch into ONE stream of micro-ops whose operands are already resolved to
// shared-memory offsets, so that a device handler is straight-line code over the thread's 4 rows.
enum XKind : uint8_t {
  XK_NONE = 0,
  XK_ACC,      // value / boolean accumulator registers
  XK_IMM,      // XOp::imm
  XK_NULL,     // typed NULL constant
  XK_SLOT,     // value slot (off = smem byte offset of the 8 KB tile, noff = its meta words)
  XK_COL8,     // staged flat column of 8-byte elements (BIGINT / DOUBLE)
  XK_COL4I,    // staged INTEGER
  XK_COL4F,    // staged REAL
  XK_COL2,     // staged SMALLINT
  XK_COL1,     // staged TINYINT
  XK_DICT,     // staged int32 indices + base values in global memory (col = column number)
  XK_BSLOT,    // boolean slot (off = smem byte offset of its meta words)
  XK_BCOL,     // staged bit-packed BOOLEAN column
  XK_BDICT     // staged indices over a bit-packed BOOLEAN base
};

struct XOp {
  uint8_t kind;
  uint8_t has_nulls; // staged outer null bitmap present (noff valid)
  uint8_t type;      // vgpu_type of the value
  uint8_t col;       // XK_DICT / XK_BDICT: column number (base pointer, base nulls)
  int32_t off;       // stage-relative (columns) or absolute (slots) shared-memory byte offset
  int32_t noff;      // stage-relative null bitmap offset / absolute slot meta offset
  int32_t pad;
  uint64_t imm;      // XK_IMM: the literal (int64 or double bits)
};

enum Handler : uint8_t {
  H_END = 0,
  H_LD,      // acc = x
  H_ARITH,   // acc = x (sub) y          sub = vgpu_opcode ADD..MOD, NEG (y unused)
  H_CAST,    // acc = cast(x) to `type`
  H_CMP,     // bacc = combine(bacc, x (sub) y [, z for BETWEEN])
  H_BOOL,    // bacc = x AND/OR y, NOT x, IS_NULL x
  H_SELECT,  // acc = if(x, y, z) (sub = VGPU_OP_SELECT) or coalesce(x, y) (sub = VGPU_OP_COALESCE)
  H_FILTER,  // pass = x is TRUE
  H_KEY,     // fold value x into the group number / normalized key (dst = key index)
  H_KEYEND,  // ARRAY: table offsets + in-thread duplicate plan; HASH: find-or-insert
  H_ACCUM,   // fold value x into cell `dst` (sub = CellKind)
  H_SEL,     // FilterProject: selection bitmap + per-tile count
  H_PROJ     // FilterProject: write output `dst` from x
};

enum MFlags : uint8_t {
  MF_STORE = 1,     // also store the result to slot (dst / dstn)
  MF_COMB_AND = 2,  // H_CMP: bacc = bacc AND cmp (three-valued)
  MF_COMB_OR = 4,   // H_CMP: bacc = bacc OR cmp
  MF_NO_VALUE = 8   // H_ACCUM: count(*) style cell without an input operand
};

struct MOp {
  uint8_t h, sub, type, flags;
  int32_t dst;   // slot tile offset (MF_STORE), cell offset (H_ACCUM), key / output index
  int32_t dstn;  // slot meta offset (MF_STORE)
  int32_t pad;
  XOp x, y, z;
};
constexpr int kMaxMops = 112;

struct PipelineParams {
  // program (ABI form; lowered to mops by launch_pipeline)
  MOp mops[kMaxMops];
  int32_t n_mops;
  vgpu_operand outputs[kMaxOutputs];
  int32_t n_outputs, n_cols, n_slots, n_bslots;
  DevCol cols[kMaxCols];
  // rows
  int64_t n_rows;         // rows to walk (n_in_rows when in_rows != nullptr)
  const int32_t* in_rows; // optional indirection
  int32_t n_stages, stage_bytes;
  int32_t smem_slots_off, smem_meta_off, smem_bmeta_off, smem_tab_off;
  uint32_t* err_word;
  // sink
  int32_t mode;
  int32_t n_keys, n_cells, n_groups; // ARRAY: n_groups = product of key cardinalities of this launch
  int32_t ignore_null_keys;
  int32_t warp_private; // ARRAY: 1 = one table per warp (shared-memory atomics), 0 = one per thread
  int32_t warp_tab_bytes; // ARRAY, warp_private: bytes of one warp's table
  int32_t tab_bytes;      // ARRAY: bytes of all partial tables of the CTA
  DevKey keys[kMaxKeys];
  DevCell cells[kMaxCells];
  // ARRAY sink: per-block partials [gridDim][n_cells][n_groups] of 8-byte values
  uint64_t* partials;
  // HASH sink
  uint64_t* table; // (capacity + 2) rows of row_words 8-byte words: key, cells...
  uint64_t cap_mask;
  KeyMap key_map;
  int32_t row_words;
  int32_t raw_key; // single 64-bit raw key
  unsigned long long* n_groups_ctr;
  // HASH sink, fused kernels: segments whose first claim met another key are parked here as
  // {key, cell values...} (row_words words each) and inserted by drain_overflow_kernel afterwards
  uint64_t* ovf;
  uint64_t ovf_cap; // entries
  unsigned long long* ovf_cursor;
  // LOG sink (fused kernels only): CTA c owns entries [c * log_region_cap, (c + 1) * log_region_cap) of `log`
  // (log_entry_words 8-byte words each): runs that cannot continue into a neighbouring block grow from the front,
  // runs touching a block edge from the back; log_counts[2c], [2c+1] receive the two counts.
  uint64_t* log;
  uint64_t log_region_cap;
  uint32_t* log_counts;
  int32_t log_entry_words;
  int32_t log_n_cells;          // cells stored in an entry, in order: log_cells[0..n)
  uint8_t log_cells[kMaxCells];
  uint64_t* log_special;        // rows of the NULL-key group and of the key that equals the empty marker: {present, cells...}
  // sorted-input probe (single raw 64-bit FLAT key column, no filter): sort_state[0] != 0 once any adjacent pair of
  // rows was out of order; prev_in = {key, valid} of the row before this batch, prev_out receives this batch's last row
  int32_t sorted_probe;
  int32_t sorted_key_col;
  unsigned long long* sort_state;
  const unsigned long long* sort_prev_in;
  unsigned long long* sort_prev_out;
  // LOG sink with the exchange fused in (exchange.cu, vgpu_agg_attach_exchange): entries do not enter the local log but
  // leave for their owners' windows as they are produced.  CTA c owns sub-region c of the region this rank has in every
  // destination's window (send_sub_cap entries; front entries from its start, back entries from its end) and keeps
  // its counts in send_counts[c * send_world + d] (low 32 bits front, high 32 back; carried from launch to launch);
  // at its end it also stores them into the destination's window (send_peer_counts[d][c]).
  int32_t send_world;           // 0 = off
  int32_t send_hash;            // 1: one raw BIGINT key, owner = HashPartitionFunction's hash (device_common.cuh partition_of)
  uint64_t send_sub_cap;
  uint64_t* send_base[8];
  unsigned long long* send_peer_counts[8];
  unsigned long long* send_counts;
  // FILTER_PROJECT sink
  uint32_t* sel_words;
  int32_t* tile_counts;
  unsigned long long* num_selected;
  DevOut outs[kMaxOutputs];
  // host only: the program being lowered
  const struct Program* prog;
};

// ---- host-side objects behind the opaque ABI handles ---------------------------------------
struct Program {
  std::vector<vgpu_op> ops;
  std::vector<vgpu_operand> outputs;
  vgpu_operand filter;
  int n_cols = 0, n_slots = 0, n_bslots = 0;
  void* scratch = nullptr; // FilterProject: per-tile counts/offsets (grow-only)
  size_t scratch_bytes = 0;
};

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);
#define VG_CUDA(x)                                   \
  do {                                               \
    cudaError_t e_ = (x);                            \
    if (e_ != cudaSuccess) return cuda_fail(e_, #x); \
  } while (0)

uint32_t* device_err_word();  // latched error bits of kernels that belong to no aggregation (FilterProject)
// Error words are slots of one per-device slab, so vgpu_check_errors can read them all with one copy while every
// aggregation reports only its own data errors (N drivers = N operator instances per device).
uint32_t* err_word_acquire();
void err_word_release(uint32_t* w);
// Optional timing of the scan kernels (vgpu_debug_kernel_timing): events on the launching stream
// right before / after the kernel, summed by vgpu_debug_kernel_time.
void kernel_timing_begin(cudaStream_t stream);
void kernel_timing_end(cudaStream_t stream);
void count_launch(int n = 1);
int num_sms();

// Fills the program/column part of PipelineParams and the shared-memory plan.  Returns status.
int plan_pipeline(const Program& prog, const vgpu_column* cols, int32_t ncols, int64_t nrows,
                  const int32_t* in_rows, int64_t n_in_rows, PipelineParams& p);
// Launches the pipeline kernel for p.mode; `tab_bytes` = shared memory wanted by the sink.
int launch_pipeline(PipelineParams& p, int tab_bytes, int grid, cudaStream_t stream);
int pipeline_grid(int64_t n_rows);
// Whether program output `idx` can be NULL given the columns bound in `p`.
bool output_nullable(const PipelineParams& p, int idx);
int launch_selection_indices(const uint32_t* words, int32_t* tile_counts, int64_t* tile_offsets,
                             int64_t n_rows, int32_t* sel_idx, unsigned long long* total,
                             cudaStream_t stream);

} // namespace vg
