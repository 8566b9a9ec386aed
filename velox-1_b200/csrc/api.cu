// extern "C" entry points of libvgpu.so (include/vgpu.h) + small runtime: error strings, device
// error word, statistics, memory/stream helpers and the NCCL exchange (loaded with dlopen so the
// library has no link-time dependency on a particular NCCL build).
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <new>
#include <utility>
#include <vector>

#include "agg.h"
#include "fused.h"
#include "internal.h"

namespace vg {

static thread_local char g_error[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}
int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
  return VGPU_ERR_CUDA;
}

static std::mutex g_mu;
static uint32_t* g_err_words[64] = {nullptr};
static int g_sms[64] = {0};
static vgpu_stats g_stats = {0, 0, 0, 0};

uint32_t* device_err_word() {
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> l(g_mu);
  if (!g_err_words[dev]) {
    cudaMalloc(&g_err_words[dev], 64);
    cudaMemset(g_err_words[dev], 0, 64);
  }
  return g_err_words[dev];
}
// ---- kernel timing (diagnostics) -------------------------------------------------------------
static bool g_timing = false;
static std::vector<std::pair<cudaEvent_t, cudaEvent_t>> g_timing_events;
void kernel_timing_begin(cudaStream_t stream) {
  if (!g_timing) return;
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  cudaEventRecord(a, stream);
  std::lock_guard<std::mutex> l(g_mu);
  g_timing_events.emplace_back(a, b);
}
void kernel_timing_end(cudaStream_t stream) {
  if (!g_timing) return;
  std::lock_guard<std::mutex> l(g_mu);
  if (!g_timing_events.empty()) cudaEventRecord(g_timing_events.back().second, stream);
}

void count_launch(int n) { __atomic_fetch_add(&g_stats.kernel_launches, (int64_t)n, __ATOMIC_RELAXED); }
int num_sms() {
  int dev = 0;
  cudaGetDevice(&dev);
  if (!g_sms[dev]) {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) == cudaSuccess) g_sms[dev] = prop.multiProcessorCount;
    else g_sms[dev] = 148;
  }
  return g_sms[dev];
}

static int validate_operand(const vgpu_operand& o, bool allow_none, int n_cols, int& max_slot, int& max_bslot, int& max_col, const char* what, int i) {
  switch (o.kind) {
    case VGPU_OPND_NONE:
      if (!allow_none) { set_error("op %d: missing operand %s", i, what); return VGPU_ERR_INVALID; }
      return VGPU_OK;
    case VGPU_OPND_ACC: case VGPU_OPND_IMM: case VGPU_OPND_NULL: break;
    case VGPU_OPND_SLOT:
      if (o.index >= kMaxSlots) { set_error("op %d: slot %d out of range (max %d)", i, o.index, kMaxSlots); return VGPU_ERR_UNSUPPORTED; }
      if (o.type == VGPU_BOOLEAN) max_bslot = std::max(max_bslot, (int)o.index); else max_slot = std::max(max_slot, (int)o.index);
      break;
    case VGPU_OPND_COL:
      if (o.index >= n_cols) { set_error("op %d: column %d out of range (%d columns)", i, o.index, n_cols); return VGPU_ERR_INVALID; }
      max_col = std::max(max_col, (int)o.index);
      break;
    default: set_error("op %d: bad operand kind %d", i, o.kind); return VGPU_ERR_INVALID;
  }
  if (o.type > VGPU_DOUBLE) { set_error("op %d: bad operand type %d", i, o.type); return VGPU_ERR_INVALID; }
  return VGPU_OK;
}

} // namespace vg

using namespace vg;

struct vgpu_program_s { Program prog; };
struct vgpu_agg_s { Agg agg; };

extern "C" {

int vgpu_abi_version(void) { return VGPU_ABI_VERSION; }

int vgpu_last_error(char* buf, size_t cap) {
  if (!buf || cap == 0) return VGPU_ERR_INVALID;
  strncpy(buf, g_error, cap - 1);
  buf[cap - 1] = 0;
  return VGPU_OK;
}

int vgpu_device_count(int* count) {
  if (!count) return VGPU_ERR_INVALID;
  VG_CUDA(cudaGetDeviceCount(count));
  return VGPU_OK;
}

int vgpu_init(int device) {
  VG_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  VG_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error("device %d is sm_%d%d; this library only ships sm_100a code", device, prop.major, prop.minor);
    return VGPU_ERR_UNSUPPORTED;
  }
  VG_CUDA(cudaFree(0));
  device_err_word();
  return VGPU_OK;
}

int vgpu_shutdown(void) { return VGPU_OK; }

int vgpu_alloc(size_t bytes, void** dptr) {
  if (!dptr) return VGPU_ERR_INVALID;
  VG_CUDA(cudaMalloc(dptr, bytes ? bytes : 16));
  return VGPU_OK;
}
int vgpu_free(void* dptr) { VG_CUDA(cudaFree(dptr)); return VGPU_OK; }
int vgpu_alloc_host(size_t bytes, void** hptr) {
  if (!hptr) return VGPU_ERR_INVALID;
  VG_CUDA(cudaMallocHost(hptr, bytes ? bytes : 16));
  return VGPU_OK;
}
int vgpu_free_host(void* hptr) { VG_CUDA(cudaFreeHost(hptr)); return VGPU_OK; }
int vgpu_h2d_async(void* dst, const void* src, size_t bytes, void* stream) {
  VG_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  __atomic_fetch_add(&g_stats.bytes_h2d, (int64_t)bytes, __ATOMIC_RELAXED);
  return VGPU_OK;
}
int vgpu_d2h_async(void* dst, const void* src, size_t bytes, void* stream) {
  VG_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  __atomic_fetch_add(&g_stats.bytes_d2h, (int64_t)bytes, __ATOMIC_RELAXED);
  return VGPU_OK;
}
int vgpu_memset_async(void* dst, int value, size_t bytes, void* stream) {
  VG_CUDA(cudaMemsetAsync(dst, value, bytes, (cudaStream_t)stream));
  return VGPU_OK;
}
int vgpu_stream_create(void** stream) {
  if (!stream) return VGPU_ERR_INVALID;
  cudaStream_t s;
  VG_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  *stream = (void*)s;
  return VGPU_OK;
}
int vgpu_stream_destroy(void* stream) { VG_CUDA(cudaStreamDestroy((cudaStream_t)stream)); return VGPU_OK; }
int vgpu_stream_sync(void* stream) { VG_CUDA(cudaStreamSynchronize((cudaStream_t)stream)); return VGPU_OK; }

int vgpu_get_stats(vgpu_stats* out) {
  if (!out) return VGPU_ERR_INVALID;
  *out = g_stats;
  return VGPU_OK;
}
int vgpu_reset_stats(void) { g_stats = vgpu_stats{0, 0, 0, 0}; return VGPU_OK; }

int vgpu_check_errors(void* stream) {
  uint32_t* w = device_err_word();
  uint32_t bits = 0;
  VG_CUDA(cudaMemcpyAsync(&bits, w, 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  VG_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  if (!bits) return VGPU_OK;
  VG_CUDA(cudaMemsetAsync(w, 0, 4, (cudaStream_t)stream));
  if (bits & kErrDivZero) set_error("division by zero");
  else if (bits & kErrOverflow) set_error("integer overflow");
  else if (bits & kErrCastRange) set_error("Cannot cast value: out of range or NaN");
  else if (bits & kErrKeyRange) { set_error("grouping key outside the declared domain"); return VGPU_ERR_INVALID; }
  else { set_error("internal: hash table full"); return VGPU_ERR_INVALID; }
  return VGPU_ERR_USER;
}

// ---- programs -------------------------------------------------------------------------------
int vgpu_program_create(const vgpu_program_desc* desc, vgpu_program_t* out) {
  if (!desc || !out) { set_error("null argument"); return VGPU_ERR_INVALID; }
  if (desc->n_ops < 0 || desc->n_ops > kMaxOps) { set_error("program has %d ops, at most %d supported", desc->n_ops, kMaxOps); return VGPU_ERR_UNSUPPORTED; }
  if (desc->n_outputs < 0 || desc->n_outputs > kMaxOutputs) { set_error("program has %d outputs, at most %d supported", desc->n_outputs, kMaxOutputs); return VGPU_ERR_UNSUPPORTED; }
  if (desc->n_columns < 0 || desc->n_columns > kMaxCols) { set_error("program reads %d columns, at most %d supported", desc->n_columns, kMaxCols); return VGPU_ERR_UNSUPPORTED; }
  int max_slot = -1, max_bslot = -1, max_col = -1;
  bool have_acc = false, have_bacc = false;
  for (int i = 0; i < desc->n_ops; ++i) {
    const vgpu_op& op = desc->ops[i];
    const int oc = op.opcode;
    const bool is_bool = oc >= VGPU_OP_EQ && oc <= VGPU_OP_IS_NULL;
    const bool known = (oc >= VGPU_OP_ADD && oc <= VGPU_OP_MOD) || oc == VGPU_OP_CAST || oc == VGPU_OP_MOV || oc == VGPU_OP_SELECT || oc == VGPU_OP_COALESCE ||
        (oc >= VGPU_OP_EQ && oc <= VGPU_OP_BETWEEN) || (oc >= VGPU_OP_AND && oc <= VGPU_OP_IS_NULL);
    if (!known) { set_error("op %d: unknown opcode %d", i, oc); return VGPU_ERR_INVALID; }
    const bool unary = oc == VGPU_OP_NEG || oc == VGPU_OP_CAST || oc == VGPU_OP_MOV || oc == VGPU_OP_NOT || oc == VGPU_OP_IS_NULL;
    int rc = validate_operand(op.a, false, desc->n_columns, max_slot, max_bslot, max_col, "a", i);
    if (rc) return rc;
    rc = validate_operand(op.b, unary, desc->n_columns, max_slot, max_bslot, max_col, "b", i);
    if (rc) return rc;
    rc = validate_operand(op.c, oc != VGPU_OP_BETWEEN && oc != VGPU_OP_SELECT, desc->n_columns, max_slot, max_bslot, max_col, "c", i);
    if (rc) return rc;
    if (oc == VGPU_OP_SELECT || oc == VGPU_OP_COALESCE) {
      const vgpu_operand& first = oc == VGPU_OP_SELECT ? op.b : op.a;
      const vgpu_operand& second = oc == VGPU_OP_SELECT ? op.c : op.b;
      if (oc == VGPU_OP_SELECT && op.a.type != VGPU_BOOLEAN) { set_error("op %d: SELECT condition must be BOOLEAN", i); return VGPU_ERR_INVALID; }
      if (oc == VGPU_OP_SELECT && (op.a.kind == VGPU_OPND_IMM || op.a.kind == VGPU_OPND_NULL)) { set_error("op %d: constant SELECT condition; fold it on the host", i); return VGPU_ERR_INVALID; }
      if (op.type == VGPU_BOOLEAN) { set_error("op %d: BOOLEAN-valued SELECT / COALESCE is not supported", i); return VGPU_ERR_UNSUPPORTED; }
      if (first.type != op.type || second.type != op.type) { set_error("op %d: branch types differ from op type %d (insert a CAST)", i, op.type); return VGPU_ERR_INVALID; }
    }
    for (const vgpu_operand* o : {&op.a, &op.b, &op.c}) {
      if (o->kind == VGPU_OPND_ACC && !(o->type == VGPU_BOOLEAN ? have_bacc : have_acc)) { set_error("op %d reads ACC before any op wrote it", i); return VGPU_ERR_INVALID; }
    }
    if (oc >= VGPU_OP_AND && oc <= VGPU_OP_NOT) {
      if (op.a.type != VGPU_BOOLEAN || (oc != VGPU_OP_NOT && op.b.type != VGPU_BOOLEAN)) { set_error("op %d: boolean op over non-boolean operands", i); return VGPU_ERR_INVALID; }
    } else if (oc == VGPU_OP_SELECT || oc == VGPU_OP_COALESCE) {
      // checked above
    } else if (!is_bool || (oc >= VGPU_OP_EQ && oc <= VGPU_OP_BETWEEN)) {
      if (oc != VGPU_OP_CAST && op.type != VGPU_BOOLEAN && oc != VGPU_OP_MOV) {
        for (const vgpu_operand* o : {&op.a, &op.b, &op.c}) {
          if (o->kind != VGPU_OPND_NONE && o->type != op.type) { set_error("op %d: operand type %d differs from op type %d (insert a CAST)", i, o->type, op.type); return VGPU_ERR_INVALID; }
        }
      }
      if (!is_bool && op.type == VGPU_BOOLEAN) { set_error("op %d: value op of BOOLEAN type", i); return VGPU_ERR_INVALID; }
    }
    if (op.dst_slot >= kMaxSlots) { set_error("op %d: dst slot %d out of range", i, op.dst_slot); return VGPU_ERR_UNSUPPORTED; }
    if (op.dst_slot >= 0) { if (is_bool) max_bslot = std::max(max_bslot, (int)op.dst_slot); else max_slot = std::max(max_slot, (int)op.dst_slot); }
    if (is_bool) have_bacc = true; else have_acc = true;
  }
  for (int i = 0; i < desc->n_outputs; ++i) {
    int rc = validate_operand(desc->outputs[i], false, desc->n_columns, max_slot, max_bslot, max_col, "output", i);
    if (rc) return rc;
    if (desc->outputs[i].kind == VGPU_OPND_ACC) { set_error("output %d: ACC operands are not allowed in outputs; store to a slot", i); return VGPU_ERR_INVALID; }
  }
  int rc = validate_operand(desc->filter, true, desc->n_columns, max_slot, max_bslot, max_col, "filter", -1);
  if (rc) return rc;
  if (desc->filter.kind != VGPU_OPND_NONE && desc->filter.type != VGPU_BOOLEAN) { set_error("filter must be BOOLEAN"); return VGPU_ERR_INVALID; }
  if (desc->filter.kind == VGPU_OPND_ACC && !have_bacc) { set_error("filter reads boolean ACC but no boolean op exists"); return VGPU_ERR_INVALID; }
  vgpu_program_s* p = new (std::nothrow) vgpu_program_s();
  if (!p) { set_error("out of host memory"); return VGPU_ERR_INVALID; }
  p->prog.ops.assign(desc->ops, desc->ops + desc->n_ops);
  p->prog.outputs.assign(desc->outputs, desc->outputs + desc->n_outputs);
  p->prog.filter = desc->filter;
  p->prog.n_cols = desc->n_columns;
  p->prog.n_slots = max_slot + 1;
  p->prog.n_bslots = max_bslot + 1;
  *out = p;
  return VGPU_OK;
}

int vgpu_program_destroy(vgpu_program_t program) {
  if (program) {
    cudaFree(program->prog.scratch);
    delete program;
  }
  return VGPU_OK;
}

int vgpu_filter_project(vgpu_program_t program, const vgpu_column* cols, int32_t ncols, int64_t nrows,
                        const int32_t* in_rows, int64_t n_in_rows, uint64_t* sel_bits, int32_t* sel_idx,
                        int64_t* num_selected, const vgpu_out_column* outs, void* stream_) {
  if (!program || (!cols && ncols > 0) || !sel_bits || !num_selected) { set_error("null argument"); return VGPU_ERR_INVALID; }
  cudaStream_t stream = (cudaStream_t)stream_;
  Program& prog = program->prog;
  const int64_t total = in_rows ? n_in_rows : nrows;
  if (total <= 0) {
    VG_CUDA(cudaMemsetAsync(num_selected, 0, 8, stream));
    return VGPU_OK;
  }
  if (total > (int64_t)INT32_MAX) { set_error("a batch holds at most 2^31-1 rows (int32 selection indices)"); return VGPU_ERR_UNSUPPORTED; }
  PipelineParams p{};
  int rc = plan_pipeline(prog, cols, ncols, nrows, in_rows, n_in_rows, p);
  if (rc) return rc;
  p.mode = kSinkFilterProject;
  const int64_t n_tiles = (total + kTile - 1) / kTile;
  const size_t need = (size_t)n_tiles * 12 + 64;
  if (need > prog.scratch_bytes && !fused_dry_run()) {
    if (prog.scratch) { VG_CUDA(cudaStreamSynchronize(stream)); VG_CUDA(cudaFree(prog.scratch)); }
    VG_CUDA(cudaMalloc(&prog.scratch, need));
    prog.scratch_bytes = need;
  }
  int64_t* tile_offsets = (int64_t*)prog.scratch;
  int32_t* tile_counts = (int32_t*)((char*)prog.scratch + (size_t)n_tiles * 8);
  if (!fused_dry_run()) VG_CUDA(cudaMemsetAsync(tile_counts, 0, (size_t)n_tiles * 4, stream));
  p.sel_words = (uint32_t*)sel_bits;
  p.tile_counts = tile_counts;
  p.num_selected = (unsigned long long*)num_selected;
  for (int i = 0; i < p.n_outputs; ++i) {
    p.outs[i].values = outs ? outs[i].values : nullptr;
    p.outs[i].nulls = outs ? outs[i].nulls : nullptr;
    if (p.outs[i].values && (((uintptr_t)p.outs[i].values) & 15u)) { set_error("output %d: values must be 16-byte aligned", i); return VGPU_ERR_INVALID; }
  }
  {
    int fgrid = 0;
    FusedKernel* fk = fused_prepare(p, &fgrid);
    if (fused_dry_run()) return VGPU_OK;
    rc = fk ? fused_launch(fk, p, fgrid, stream) : launch_pipeline(p, 0, pipeline_grid(p.n_rows), stream);
  }
  if (rc) return rc;
  __atomic_fetch_add(&g_stats.rows_in, total, __ATOMIC_RELAXED);
  return launch_selection_indices(p.sel_words, tile_counts, tile_offsets, total, sel_idx, (unsigned long long*)num_selected, stream);
}

// ---- aggregation ----------------------------------------------------------------------------
int vgpu_agg_create(const vgpu_agg_desc* d, vgpu_agg_t* out) {
  if (!d || !out) { set_error("null argument"); return VGPU_ERR_INVALID; }
  if (d->mode < VGPU_AGG_GLOBAL || d->mode > VGPU_AGG_HASH) { set_error("bad aggregation mode %d", d->mode); return VGPU_ERR_INVALID; }
  if (d->n_keys < 0 || d->n_keys > kMaxKeys) { set_error("at most %d grouping keys", kMaxKeys); return VGPU_ERR_UNSUPPORTED; }
  if (d->n_accs < 0 || d->n_accs > kMaxOutputs) { set_error("at most %d accumulators", kMaxOutputs); return VGPU_ERR_UNSUPPORTED; }
  if (d->mode == VGPU_AGG_GLOBAL && d->n_keys != 0) { set_error("GLOBAL aggregation takes no keys"); return VGPU_ERR_INVALID; }
  if (d->mode != VGPU_AGG_GLOBAL && d->n_keys == 0) { set_error("grouped aggregation needs keys"); return VGPU_ERR_INVALID; }
  vgpu_agg_s* a = new (std::nothrow) vgpu_agg_s();
  if (!a) { set_error("out of host memory"); return VGPU_ERR_INVALID; }
  a->agg.mode = d->mode;
  a->agg.ignore_null_keys = d->ignore_null_keys != 0;
  a->agg.keys.assign(d->keys, d->keys + d->n_keys);
  a->agg.accs.assign(d->accs, d->accs + d->n_accs);
  a->agg.expected_groups = d->expected_groups;
  *out = a;
  return VGPU_OK;
}
int vgpu_agg_destroy(vgpu_agg_t agg) {
  if (agg) { agg_free(agg->agg); delete agg; }
  return VGPU_OK;
}
int vgpu_agg_reset(vgpu_agg_t agg, void* stream) {
  if (!agg) return VGPU_ERR_INVALID;
  return agg_reset(agg->agg, (cudaStream_t)stream);
}
int vgpu_agg_add_input(vgpu_agg_t agg, vgpu_program_t program, const vgpu_column* cols, int32_t ncols,
                       int64_t nrows, const int32_t* in_rows, int64_t n_in_rows, void* stream) {
  if (!agg || !program) { set_error("null argument"); return VGPU_ERR_INVALID; }
  __atomic_fetch_add(&g_stats.rows_in, in_rows ? n_in_rows : nrows, __ATOMIC_RELAXED);
  return agg_add_input(agg->agg, program->prog, cols, ncols, nrows, in_rows, n_in_rows, (cudaStream_t)stream);
}
int vgpu_agg_num_groups(vgpu_agg_t agg, void* stream, int64_t* n) {
  if (!agg || !n) return VGPU_ERR_INVALID;
  int rc = agg_extract_all(agg->agg, (cudaStream_t)stream);
  if (rc) return rc;
  *n = agg->agg.packed_rows;
  return VGPU_OK;
}
int vgpu_agg_row_bytes(vgpu_agg_t agg, int32_t* bytes) {
  if (!agg || !bytes) return VGPU_ERR_INVALID;
  *bytes = (3 + 2 * (int)agg->agg.accs.size()) * 8;
  return VGPU_OK;
}
int vgpu_agg_extract(vgpu_agg_t agg, int64_t offset, int64_t max_groups, void* rows, int64_t* n_written, void* stream) {
  if (!agg || !n_written || offset < 0) return VGPU_ERR_INVALID;
  int rc = agg_extract_all(agg->agg, (cudaStream_t)stream);
  if (rc) return rc;
  const Agg& a = agg->agg;
  const int64_t n = std::max<int64_t>(0, std::min(max_groups, a.packed_rows - offset));
  if (n > 0) {
    if (!rows) return VGPU_ERR_INVALID;
    const size_t rb = (size_t)a.rm.row_words_out * 8;
    VG_CUDA(cudaMemcpyAsync(rows, (const char*)a.packed + offset * rb, n * rb, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    VG_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  }
  *n_written = n;
  return VGPU_OK;
}
int vgpu_agg_merge_rows(vgpu_agg_t agg, const void* rows, int64_t n_rows, void* stream) {
  if (!agg || (!rows && n_rows > 0)) return VGPU_ERR_INVALID;
  return agg_merge_rows(agg->agg, rows, n_rows, (cudaStream_t)stream);
}
int vgpu_agg_partition(vgpu_agg_t agg, int32_t n_parts, void* rows_out, int64_t* counts_out, void* stream) {
  if (!agg || !counts_out) return VGPU_ERR_INVALID;
  return agg_partition(agg->agg, n_parts, rows_out, counts_out, (cudaStream_t)stream);
}

// ---- diagnostics of the plan-specialised kernels ------------------------------------------------
int vgpu_debug_set_dry_run(int on) { fused_set_dry_run(on != 0); return VGPU_OK; }
int vgpu_debug_last_kernel_source(char* buf, size_t cap, uint64_t* hash) {
  const std::string& s = fused_last_source();
  if (hash) *hash = fused_source_hash(s);
  if (buf && cap) { strncpy(buf, s.c_str(), cap - 1); buf[cap - 1] = 0; }
  return s.size() + 1 > cap ? VGPU_ERR_INVALID : VGPU_OK;
}
int vgpu_debug_last_kernel_note(char* buf, size_t cap) {
  if (!buf || cap == 0) return VGPU_ERR_INVALID;
  strncpy(buf, fused_last_note().c_str(), cap - 1);
  buf[cap - 1] = 0;
  return VGPU_OK;
}

int vgpu_debug_kernel_timing(int on) {
  std::lock_guard<std::mutex> l(g_mu);
  g_timing = on != 0;
  for (auto& e : g_timing_events) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
  g_timing_events.clear();
  return VGPU_OK;
}
int vgpu_debug_kernel_time(double* ms_total, int64_t* launches) {
  if (!ms_total || !launches) return VGPU_ERR_INVALID;
  std::lock_guard<std::mutex> l(g_mu);
  double sum = 0;
  int64_t n = 0;
  for (auto& e : g_timing_events) {
    VG_CUDA(cudaEventSynchronize(e.second));
    float ms = 0;
    VG_CUDA(cudaEventElapsedTime(&ms, e.first, e.second));
    sum += ms;
    ++n;
    cudaEventDestroy(e.first);
    cudaEventDestroy(e.second);
  }
  g_timing_events.clear();
  *ms_total = sum;
  *launches = n;
  return VGPU_OK;
}

// ---- multi-GPU exchange over NCCL (dlopen'ed) -------------------------------------------------
namespace {
typedef struct { char internal[128]; } nccl_unique_id;
typedef void* nccl_comm;
typedef int (*fn_get_unique_id)(nccl_unique_id*);
typedef int (*fn_comm_init_rank)(nccl_comm*, int, nccl_unique_id, int);
typedef int (*fn_comm_destroy)(nccl_comm);
typedef int (*fn_send)(const void*, size_t, int, int, nccl_comm, cudaStream_t);
typedef int (*fn_recv)(void*, size_t, int, int, nccl_comm, cudaStream_t);
typedef int (*fn_group)(void);
typedef int (*fn_allgather)(const void*, void*, size_t, int, nccl_comm, cudaStream_t);
typedef const char* (*fn_errstr)(int);
struct NcclApi {
  void* handle = nullptr;
  fn_get_unique_id get_unique_id; fn_comm_init_rank comm_init_rank; fn_comm_destroy comm_destroy;
  fn_send send; fn_recv recv; fn_group group_start, group_end; fn_allgather allgather; fn_errstr errstr;
};
NcclApi g_nccl;
constexpr int kNcclUint8 = 1; // ncclUint8 in nccl.h's ncclDataType_t

int load_nccl() {
  if (g_nccl.handle) return VGPU_OK;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* n : names) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
  if (!h) { set_error("cannot load libnccl.so.2: %s", dlerror()); return VGPU_ERR_NCCL; }
#define VG_SYM(field, name) g_nccl.field = (decltype(g_nccl.field))dlsym(h, name); if (!g_nccl.field) { set_error("libnccl lacks %s", name); return VGPU_ERR_NCCL; }
  VG_SYM(get_unique_id, "ncclGetUniqueId") VG_SYM(comm_init_rank, "ncclCommInitRank") VG_SYM(comm_destroy, "ncclCommDestroy")
  VG_SYM(send, "ncclSend") VG_SYM(recv, "ncclRecv") VG_SYM(group_start, "ncclGroupStart") VG_SYM(group_end, "ncclGroupEnd")
  VG_SYM(allgather, "ncclAllGather") VG_SYM(errstr, "ncclGetErrorString")
#undef VG_SYM
  g_nccl.handle = h;
  return VGPU_OK;
}
int nccl_fail(int rc, const char* what) { set_error("NCCL error %d (%s) in %s", rc, g_nccl.errstr(rc), what); return VGPU_ERR_NCCL; }
#define VG_NCCL(x) do { int r_ = (x); if (r_ != 0) return nccl_fail(r_, #x); } while (0)
} // namespace

struct vgpu_comm_s {
  nccl_comm comm; int rank, world;
  // grow-only device scratch of vgpu_comm_exchange_groups
  void* send = nullptr; size_t send_cap = 0;
  void* recv = nullptr; size_t recv_cap = 0;
  int64_t* dcounts = nullptr; // [world] + [world * world]
};
static int comm_reserve(void** p, size_t* cap, size_t want) {
  if (*cap >= want) return VGPU_OK;
  if (*p) { VG_CUDA(cudaDeviceSynchronize()); VG_CUDA(cudaFree(*p)); *p = nullptr; *cap = 0; }
  want = std::max<size_t>(want + want / 4, 1 << 16);
  VG_CUDA(cudaMalloc(p, want));
  *cap = want;
  return VGPU_OK;
}

int vgpu_comm_unique_id(uint8_t id[VGPU_COMM_ID_BYTES]) {
  int rc = load_nccl();
  if (rc) return rc;
  nccl_unique_id u;
  VG_NCCL(g_nccl.get_unique_id(&u));
  memcpy(id, &u, VGPU_COMM_ID_BYTES);
  return VGPU_OK;
}
int vgpu_comm_create(const uint8_t id[VGPU_COMM_ID_BYTES], int32_t rank, int32_t world, vgpu_comm_t* out) {
  if (!id || !out || rank < 0 || rank >= world) return VGPU_ERR_INVALID;
  int rc = load_nccl();
  if (rc) return rc;
  nccl_unique_id u;
  memcpy(&u, id, VGPU_COMM_ID_BYTES);
  vgpu_comm_s* c = new vgpu_comm_s();
  c->rank = rank; c->world = world;
  int r = g_nccl.comm_init_rank(&c->comm, world, u, rank);
  if (r != 0) { delete c; return nccl_fail(r, "ncclCommInitRank"); }
  *out = c;
  return VGPU_OK;
}
int vgpu_comm_destroy(vgpu_comm_t comm) {
  if (comm) { g_nccl.comm_destroy(comm->comm); cudaFree(comm->send); cudaFree(comm->recv); cudaFree(comm->dcounts); delete comm; }
  return VGPU_OK;
}
int vgpu_comm_alltoallv(vgpu_comm_t comm, const void* send, const int64_t* send_counts, const int64_t* send_offsets,
                        void* recv, const int64_t* recv_counts, const int64_t* recv_offsets, void* stream) {
  if (!comm || !send_counts || !recv_counts || !send_offsets || !recv_offsets) return VGPU_ERR_INVALID;
  VG_NCCL(g_nccl.group_start());
  for (int r = 0; r < comm->world; ++r) {
    if (send_counts[r] > 0) VG_NCCL(g_nccl.send((const char*)send + send_offsets[r], (size_t)send_counts[r], kNcclUint8, r, comm->comm, (cudaStream_t)stream));
    if (recv_counts[r] > 0) VG_NCCL(g_nccl.recv((char*)recv + recv_offsets[r], (size_t)recv_counts[r], kNcclUint8, r, comm->comm, (cudaStream_t)stream));
  }
  VG_NCCL(g_nccl.group_end());
  return VGPU_OK;
}
int vgpu_comm_allgather(vgpu_comm_t comm, const void* send, void* recv, size_t bytes, void* stream) {
  if (!comm) return VGPU_ERR_INVALID;
  VG_NCCL(g_nccl.allgather(send, recv, bytes, kNcclUint8, comm->comm, (cudaStream_t)stream));
  return VGPU_OK;
}

int vgpu_comm_exchange_groups(vgpu_comm_t comm, vgpu_agg_t partial_agg, vgpu_agg_t final_agg, void* stream_, int64_t* bytes_sent) {
  if (!comm || !partial_agg || !final_agg) return VGPU_ERR_INVALID;
  cudaStream_t stream = (cudaStream_t)stream_;
  Agg& pa = partial_agg->agg;
  Agg& fa = final_agg->agg;
  if (pa.mode != fa.mode || pa.accs.size() != fa.accs.size()) { set_error("partial and final aggregations differ"); return VGPU_ERR_INVALID; }
  const size_t row_bytes = (3 + 2 * pa.accs.size()) * 8;
  const int W = comm->world;
  int rc;
  if (pa.mode != VGPU_AGG_HASH) {
    if (pa.n_groups != fa.n_groups) { set_error("partial and final aggregations differ in their key domains"); return VGPU_ERR_INVALID; }
    const size_t mine = (size_t)pa.n_groups * row_bytes;
    if ((rc = comm_reserve(&comm->send, &comm->send_cap, mine))) return rc;
    if ((rc = comm_reserve(&comm->recv, &comm->recv_cap, mine * W))) return rc;
    if ((rc = agg_dense_rows(pa, comm->send, stream))) return rc;
    VG_NCCL(g_nccl.allgather(comm->send, comm->recv, mine, kNcclUint8, comm->comm, stream));
    if (bytes_sent) *bytes_sent = (int64_t)(mine * (W - 1));
    return agg_merge_owned_rows(fa, comm->recv, (int64_t)pa.n_groups * W, W, comm->rank, stream);
  }
  if ((rc = agg_extract_all(pa, stream))) return rc;
  const int64_t n = pa.packed_rows;
  if ((rc = comm_reserve(&comm->send, &comm->send_cap, (size_t)std::max<int64_t>(n, 1) * row_bytes))) return rc;
  if (!comm->dcounts) VG_CUDA(cudaMalloc(&comm->dcounts, sizeof(int64_t) * (32 + 32 * 32)));
  int64_t counts[32] = {0}, matrix[32 * 32];
  if ((rc = agg_partition(pa, W, comm->send, counts, stream))) return rc;
  VG_CUDA(cudaMemcpyAsync(comm->dcounts, counts, sizeof(int64_t) * W, cudaMemcpyHostToDevice, stream));
  VG_NCCL(g_nccl.allgather(comm->dcounts, comm->dcounts + 32, sizeof(int64_t) * W, kNcclUint8, comm->comm, stream));
  VG_CUDA(cudaMemcpyAsync(matrix, comm->dcounts + 32, sizeof(int64_t) * W * W, cudaMemcpyDeviceToHost, stream));
  VG_CUDA(cudaStreamSynchronize(stream));
  int64_t sc[32], so[32], rcv[32], ro[32], srun = 0, rrun = 0;
  for (int r = 0; r < W; ++r) { // matrix[src][dst]
    sc[r] = counts[r] * (int64_t)row_bytes; so[r] = srun; srun += sc[r];
    rcv[r] = matrix[r * W + comm->rank] * (int64_t)row_bytes; ro[r] = rrun; rrun += rcv[r];
  }
  if ((rc = comm_reserve(&comm->recv, &comm->recv_cap, (size_t)std::max<int64_t>(rrun, 8)))) return rc;
  if ((rc = vgpu_comm_alltoallv(comm, comm->send, sc, so, comm->recv, rcv, ro, stream_))) return rc;
  if (bytes_sent) *bytes_sent = srun - sc[comm->rank];
  return agg_merge_rows(fa, comm->recv, rrun / (int64_t)row_bytes, stream);
}

} // extern "C"
