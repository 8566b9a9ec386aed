// Host-side aggregation object behind vgpu_agg_t and the launchers of agg.cu.
#pragma once
#include <vector>

#include "internal.h"

namespace vg {

// How accumulators map to cells of the group state / to words of a packed row.
struct RowMap {
  int32_t n_accs;
  int32_t rows_cell;      // cell counting passing rows (always cell 0)
  int32_t row_words_out;  // 3 + 2 * n_accs
  int16_t value_cell[kMaxOutputs];
  int16_t nn_cell[kMaxOutputs]; // -1: accumulator never NULL (COUNT)
  uint8_t value_kind[kMaxOutputs];
};

// Indices into Agg::ctr (device words).
enum : int {
  kCtrGroups = 0,      // TABLE engine: groups in the table
  kCtrScratch = 1,     // extract / partition output cursor
  kCtrPartCounts = 2,  // [32] partition counts
  kCtrOvfCursor = 40,  // parked-segment cursor of the fused HASH kernels
  // LOG engine (hashagg.cu)
  kCtrSortState = 48,  // != 0: some adjacent rows were out of key order (the front entries are not unique)
  kCtrPrevKey = 49,    // [2][2]: {key, valid} of the last row of the previous batch, double-buffered by batch parity
  kCtrFin = 53,        // finalize header: see FinHeader (6 words)
  kCtrFirstKey = 59,   // key of the very first logged row (meaningful while the input is sorted)
  kCtrCombine = 60,    // finalize's pre-combining pass: [0] 0 = off, 1 = wanted (sample says duplicates abound), 2 = done;
                       // [1] combined entries, [2] CTAs finished (ticket), [3] spare
  kCtrWords = 64
};

// LOG engine.  One slab = per-CTA regions that the scan kernels of consecutive batches append to: region c holds
// region_cap entries; entries that can only be whole groups when the input is sorted grow from the region's front,
// the others (runs touching a block edge) from its back; counts[2c], counts[2c+1] are the two fill counts.
struct LogSlab {
  uint64_t* base = nullptr;
  uint32_t* counts = nullptr;
  uint64_t region_cap = 0;   // entries per region
  uint64_t region_used = 0;  // host-side upper bound of the entries appended to any one region so far
  int regions = 0;
  size_t bytes = 0;
};

// A source of entries for the deduplication passes (device-resident description).
struct LogRange {
  const uint64_t* base;   // first word of the region / array
  uint64_t cap;           // entries the region holds (back ranges end at base + cap * ew)
  const void* count;      // device address of the entry count
  uint32_t count64;       // the count is a 64-bit integer (else 32-bit)
  uint32_t back;          // 1: the entries are the LAST `count` entries of the region
};

// What finalize leaves in device memory (Agg::ctr + kCtrFin).
struct FinHeader {
  unsigned long long sorted;    // 1: front ranges are whole groups and were not deduplicated
  unsigned long long n_front;   // entries in front ranges (groups, when sorted)
  unsigned long long n_dedupe;  // entries that went through the deduplication
  unsigned long long n_groups;  // groups the deduplication produced (in Agg::groups)
  unsigned long long overflow;  // a partition held more distinct keys than its shared-memory table
  unsigned long long log2p;     // hash partitions used (chosen on the device from n_dedupe)
};

constexpr int kSrcMetaWords = 4;
enum HashEngine : int { kEngineUndecided = 0, kEngineLog = 1, kEngineTable = 2 };

struct Agg {
  int mode = 0;
  bool ignore_null_keys = false;
  std::vector<vgpu_key_spec> keys;
  std::vector<vgpu_acc_spec> accs;
  int64_t expected_groups = 0;
  // cell plan (resolved on the first add_input, when operand types are known)
  bool resolved = false;
  int n_cells = 0;
  DevCell cells[kMaxCells];
  RowMap rm{};
  // ARRAY / GLOBAL
  int n_groups = 1;
  int64_t mults[kMaxKeys] = {0};
  uint64_t* state = nullptr;   // [n_cells][n_groups]
  uint64_t* partials = nullptr;
  size_t partials_cap = 0;     // in words
  // HASH
  uint64_t* table = nullptr;
  uint64_t capacity = 0;       // power of two; table has capacity + 2 rows
  uint64_t groups_bound = 0;   // host-side upper bound of the groups in the table (rows added so far): lets add_input
                               // skip the device read of the exact count while the bound still fits the table
  int row_words = 0;
  uint8_t shifts[kMaxKeys] = {0};
  bool hash_aliased[kMaxCells]; // count cell still equals the rows cell (never updated on its own)
  unsigned long long* ctr = nullptr; // device words, see the kCtr* indices
  uint64_t* ovf = nullptr;     // parked segments of the fused HASH kernels (see PipelineParams::ovf)
  uint64_t ovf_cap = 0;        // entries
  // HASH, LOG engine (hashagg.cu)
  int engine = kEngineUndecided;
  int ew = 0;                        // words per entry: key + log_n_cells
  int log_n_cells = 0;
  uint8_t log_cells[kMaxCells];      // cells stored in an entry, in entry order
  std::vector<LogSlab> slabs;        // slabs.back() is the one being appended to
  std::vector<LogSlab> free_slabs;   // kept across reset()
  std::vector<LogRange> ext_ranges;  // entries in foreign buffers (received from other GPUs), deduplicated with the slabs'
  uint64_t ext_bound = 0;            // upper bound of the entries in ext_ranges
  // what the senders said about their streams: per source kSrcMetaWords words {sorted, first key, last key, entries}
  // at ext_meta + s * ext_meta_stride (device).  When every source was sorted and their key ranges are disjoint, the
  // received front entries are whole groups and skip the deduplication like a local sorted stream's.
  const unsigned long long* ext_meta = nullptr;
  int ext_sources = 0, ext_meta_stride = 0;
  uint64_t log_bound = 0;            // upper bound of the entries in the slabs
  bool log_probed = true;            // every batch so far ran the sorted-input probe
  int log_batches = 0;
  uint64_t* special = nullptr;       // 2 rows x (1 + log_n_cells) words: NULL-key group, empty-marker-key group
  bool finalized = false;
  uint64_t* part = nullptr;  size_t part_cap = 0;      // scatter target (entries)
  uint64_t* groups = nullptr; size_t groups_cap = 0;   // deduplicated groups (entries)
  uint32_t* hist = nullptr;  size_t hist_cap = 0;      // [3][P + 1]: histogram, offsets, cursors
  // device copies of the range lists ([0] entries to deduplicate, [1] where the groups are) and what they hold
  LogRange* ranges[2] = {nullptr, nullptr}; size_t ranges_cap[2] = {0, 0};
  std::vector<LogRange> host_ranges[2], uploaded_ranges[2];
  int n_parts = 0;
  uint32_t* err = nullptr;           // this aggregation's error word
  // extraction cache
  uint64_t* packed = nullptr;
  int64_t packed_rows = -1;    // -1 = stale
  size_t packed_cap = 0;       // rows
  int device = 0;
  // Exchange fused into the scan (exchange.cu, vgpu_agg_attach_exchange): the LOG-sink kernels of this (partial)
  // aggregation store their entries straight into the owners' windows.
  struct vgpu_comm_s* send_comm = nullptr;
  unsigned long long* send_counts = nullptr; // [send_grid][world], device: entries per CTA and destination so far
  size_t send_counts_cap = 0;                // words
  int send_grid = 0;                         // CTAs of the sending kernels (fixed while send_used)
  uint64_t send_sub_cap = 0;                 // entries per CTA and destination
  bool send_used = false;                    // entries of the current state have left through the window
  uint64_t send_epoch = 0;                   // the exchange's epoch (window half) those entries were stored under
  // A dense exchange folds into this (final) aggregation on the communicator's side stream (exchange.cu): `pending` is
  // recorded there after the fold, and the next operation on the aggregation makes its own stream wait for it.
  cudaEvent_t pending = nullptr;
  bool pending_set = false;
  bool emit_empty_global = true;  // GLOBAL: this rank reports the (possibly empty) single group
  Agg() { for (int c = 0; c < kMaxCells; ++c) hash_aliased[c] = true; }
};

int agg_resolve(Agg& a, const Program& prog, cudaStream_t stream);
int agg_add_input(Agg& a, const Program& prog, const vgpu_column* cols, int32_t ncols, int64_t nrows,
                  const int32_t* in_rows, int64_t n_in_rows, cudaStream_t stream);
int agg_extract_all(Agg& a, cudaStream_t stream); // fills the packed cache
int agg_merge_rows(Agg& a, const void* rows, int64_t n, cudaStream_t stream);
int agg_partition(Agg& a, int n_parts, void* rows_out, int64_t* counts_out, cudaStream_t stream);
int agg_dense_rows(Agg& a, void* rows_out, cudaStream_t stream);
int agg_merge_owned_rows(Agg& a, const void* rows, int64_t n, int n_parts, int my_part, cudaStream_t stream);
int agg_reset(Agg& a, cudaStream_t stream);
int agg_wait_pending(Agg& a, cudaStream_t stream); // orders `stream` after a fold still running on an exchange's side stream
void agg_free(Agg& a);
bool agg_raw_bigint_key(const Agg& a); // one raw BIGINT key: partitions are assigned with the reference's own hash
int agg_finalize(Agg& a, cudaStream_t stream);   // LOG engine: deduplicate what the scans logged (asynchronous)

// ---- hashagg.cu --------------------------------------------------------------------------------
struct CellKinds { uint8_t kind[kMaxCells]; int32_t n; };
// Picks the entry layout from the first eligible batch and allocates the side structures.
int log_setup(Agg& a, const PipelineParams& p, cudaStream_t stream);
// Whether the cells this launch updates fit the entry layout chosen earlier.
bool log_layout_covers(const Agg& a, const PipelineParams& p);
// Makes room for one more batch of `region_need` entries per region in `regions` regions; fills the sink fields of p.
// no_slab: the entries go elsewhere (fused exchange): only the bookkeeping fields.
int log_begin_batch(Agg& a, PipelineParams& p, int regions, uint64_t region_need, cudaStream_t stream, bool no_slab = false);

// ---- exchange.cu -------------------------------------------------------------------------------
// Whether the aggregation's next LOG batch can leave through its attached exchange (window mapped, world <= 8, same
// stream of batches); sets p.send_world / p.send_hash for the kernel's signature.
bool send_wanted(Agg& a, PipelineParams& p);
// Binds the window addresses for a launch of `grid` CTAs.  VGPU_ERR_UNSUPPORTED: the window is too small / the grid changed.
int send_bind(Agg& a, PipelineParams& p, int grid, cudaStream_t stream);
void send_reset(Agg& a);
void send_free(Agg& a);
int log_finalize(Agg& a, cudaStream_t stream);
// Group rows in the ABI layout (key, flags, rows, values, non-null counts) from the finalized state into `out`;
// *n_out (device) receives the count.  `max_rows`: capacity of out.
int log_expand_rows(Agg& a, uint64_t* out, unsigned long long* n_out, cudaStream_t stream);
int log_read_header(Agg& a, cudaStream_t stream, FinHeader* h, uint64_t* n_special); // synchronises
int log_reset(Agg& a, cudaStream_t stream);
// The entry layout of `from` for an aggregation of the same description that has seen no input yet.
int log_adopt_layout(Agg& a, const Agg& from, cudaStream_t stream);
// Device list of every logged entry (front and back ranges): for kernels outside hashagg.cu.  *n receives the count.
int log_entry_ranges(Agg& a, cudaStream_t stream, const LogRange** ranges, int* n, uint32_t** chunk_prefix_scratch);
// chunk_prefix[r] = chunks of `chunk` entries before range r, over all ranges (asynchronous, one small kernel)
int log_chunk_prefix(const LogRange* ranges, int n, uint32_t chunk, uint32_t* chunk_prefix, cudaStream_t stream);
void log_free(Agg& a);
CellKinds log_cell_kinds(const Agg& a);

} // namespace vg

// the opaque ABI handles
struct vgpu_program_s { vg::Program prog; };
struct vgpu_agg_s { vg::Agg agg; };
