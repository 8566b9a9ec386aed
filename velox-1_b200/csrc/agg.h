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
  unsigned long long* ctr = nullptr; // device: [0] groups, [1] scratch counter, [2] overflow cursor
  uint64_t* ovf = nullptr;     // parked segments of the fused HASH kernels (see PipelineParams::ovf)
  uint64_t ovf_cap = 0;        // entries
  // extraction cache
  uint64_t* packed = nullptr;
  int64_t packed_rows = -1;    // -1 = stale
  size_t packed_cap = 0;       // rows
  int device = 0;
  Agg() { for (int c = 0; c < kMaxCells; ++c) hash_aliased[c] = true; }
};

int agg_resolve(Agg& a, const Program& prog);
int agg_add_input(Agg& a, const Program& prog, const vgpu_column* cols, int32_t ncols, int64_t nrows,
                  const int32_t* in_rows, int64_t n_in_rows, cudaStream_t stream);
int agg_extract_all(Agg& a, cudaStream_t stream); // fills the packed cache
int agg_merge_rows(Agg& a, const void* rows, int64_t n, cudaStream_t stream);
int agg_partition(Agg& a, int n_parts, void* rows_out, int64_t* counts_out, cudaStream_t stream);
int agg_dense_rows(Agg& a, void* rows_out, cudaStream_t stream);
int agg_merge_owned_rows(Agg& a, const void* rows, int64_t n, int n_parts, int my_part, cudaStream_t stream);
int agg_reset(Agg& a, cudaStream_t stream);
void agg_free(Agg& a);

} // namespace vg
