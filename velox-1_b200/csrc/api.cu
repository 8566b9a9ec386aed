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
// Error words: one slab of kErrSlots 32-bit words per device.  Slot 0 is the device-wide word (kernels that belong to
// no aggregation, i.e. FilterProject); every aggregation takes a slot of its own, so with N drivers on a device an
// overflow in one operator is raised by that operator (vgpu_agg_check_errors), not by whoever checks first.
constexpr int kErrSlots = 4096;
static uint32_t* g_err_slab[64] = {nullptr};
static std::vector<bool> g_err_used[64];
static int g_sms[64] = {0};
static vgpu_stats g_stats = {0, 0, 0, 0};

static uint32_t* err_slab_locked(int dev) {
  if (!g_err_slab[dev]) {
    cudaMalloc(&g_err_slab[dev], kErrSlots * sizeof(uint32_t));
    cudaMemset(g_err_slab[dev], 0, kErrSlots * sizeof(uint32_t));
    cudaDeviceSynchronize(); // streams created with cudaStreamNonBlocking are not ordered after the legacy stream
    g_err_used[dev].assign(kErrSlots, false);
    g_err_used[dev][0] = true;
  }
  return g_err_slab[dev];
}
uint32_t* device_err_word() {
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> l(g_mu);
  return err_slab_locked(dev);
}
uint32_t* err_word_acquire() {
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> l(g_mu);
  uint32_t* base = err_slab_locked(dev);
  if (!base) return nullptr;
  for (int i = 1; i < kErrSlots; ++i) {
    if (!g_err_used[dev][i]) { g_err_used[dev][i] = true; return base + i; }
  }
  return base; // out of slots: share the device-wide word
}
void err_word_release(uint32_t* w) {
  if (!w) return;
  std::lock_guard<std::mutex> l(g_mu);
  for (int dev = 0; dev < 64; ++dev) {
    if (g_err_slab[dev] && w > g_err_slab[dev] && w < g_err_slab[dev] + kErrSlots) {
      // a latched error nobody collected must not leak into the slot's next owner
      cudaMemset(w, 0, sizeof(uint32_t));
      g_err_used[dev][w - g_err_slab[dev]] = false;
    }
  }
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

namespace vg {
// dst[i] = table[src[i]]: a dictionary-encoded VARCHAR key column (int32 indices into a batch's base strings) becomes
// the operator-wide value ids of those strings while the batch is gathered on the device (VectorHasher::makeValueIds
// per base value, velox/exec/VectorHasher.cpp:241-289, then one gather per row — on the GPU instead of the host).
__global__ void translate_i32_kernel(int32_t* dst, const int32_t* src, const int32_t* table, int32_t table_len, int64_t n, uint32_t* err) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t ix = src[i];
  if (ix < 0 || ix >= table_len) { atomicOr(err, kErrKeyRange); dst[i] = 0; return; }
  dst[i] = __ldg(table + ix);
}
} // namespace vg

using namespace vg;


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

int vgpu_event_create(void** event) {
  if (!event) return VGPU_ERR_INVALID;
  cudaEvent_t e;
  VG_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  *event = (void*)e;
  return VGPU_OK;
}
int vgpu_event_destroy(void* event) { VG_CUDA(cudaEventDestroy((cudaEvent_t)event)); return VGPU_OK; }
int vgpu_event_record(void* event, void* stream) { VG_CUDA(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream)); return VGPU_OK; }
int vgpu_stream_wait_event(void* stream, void* event) { VG_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)event, 0)); return VGPU_OK; }
int vgpu_event_sync(void* event) { VG_CUDA(cudaEventSynchronize((cudaEvent_t)event)); return VGPU_OK; }
int vgpu_host_is_pinned(const void* hptr, int32_t* pinned) {
  if (!pinned) return VGPU_ERR_INVALID;
  cudaPointerAttributes a;
  const cudaError_t e = cudaPointerGetAttributes(&a, hptr);
  if (e != cudaSuccess) { cudaGetLastError(); *pinned = 0; return VGPU_OK; }
  *pinned = a.type == cudaMemoryTypeHost ? 1 : 0;
  return VGPU_OK;
}

int vgpu_translate_i32(int32_t* dst, const int32_t* src, const int32_t* table, int32_t table_len, int64_t n, void* stream) {
  if (n <= 0) return VGPU_OK;
  if (!dst || !src || !table) return VGPU_ERR_INVALID;
  translate_i32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(dst, src, table, table_len, n, device_err_word());
  VG_CUDA(cudaGetLastError());
  count_launch();
  return VGPU_OK;
}

int vgpu_get_stats(vgpu_stats* out) {
  if (!out) return VGPU_ERR_INVALID;
  *out = g_stats;
  return VGPU_OK;
}
int vgpu_reset_stats(void) { g_stats = vgpu_stats{0, 0, 0, 0}; return VGPU_OK; }

static int report_error_bits(uint32_t bits) {
  if (!bits) return VGPU_OK;
  if (bits & kErrDivZero) set_error("division by zero");
  else if (bits & kErrOverflow) set_error("integer overflow");
  else if (bits & kErrCastRange) set_error("Cannot cast value: out of range or NaN");
  else if (bits & kErrKeyRange) { set_error("grouping key outside the declared domain"); return VGPU_ERR_INVALID; }
  else if (bits & kErrStringKey) { set_error("VARCHAR value longer than 7 bytes read on the device (map long strings to ids on the host)"); return VGPU_ERR_UNSUPPORTED; }
  else { set_error("internal: a device table or an exchange window region is full (hash table, log region, or the per-CTA share of vgpu_comm_reserve_window's window)"); return VGPU_ERR_INVALID; }
  return VGPU_ERR_USER;
}

// Every error word of the device (the device-wide one and those of all aggregations).
int vgpu_check_errors(void* stream) {
  uint32_t* slab = device_err_word();
  static thread_local std::vector<uint32_t> host;
  host.resize(kErrSlots);
  VG_CUDA(cudaMemcpyAsync(host.data(), slab, kErrSlots * sizeof(uint32_t), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  VG_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  uint32_t bits = 0;
  for (uint32_t w : host) bits |= w;
  if (!bits) return VGPU_OK;
  VG_CUDA(cudaMemsetAsync(slab, 0, kErrSlots * sizeof(uint32_t), (cudaStream_t)stream));
  return report_error_bits(bits);
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
    if (!fk && fused_strict_failure()) return VGPU_ERR_INVALID;
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
int vgpu_agg_finalize(vgpu_agg_t agg, void* stream) {
  if (!agg) return VGPU_ERR_INVALID;
  return agg_finalize(agg->agg, (cudaStream_t)stream);
}
int vgpu_agg_check_errors(vgpu_agg_t agg, void* stream) {
  if (!agg) return VGPU_ERR_INVALID;
  if (!agg->agg.err) return VGPU_OK; // no kernel ran yet
  { const int wrc = agg_wait_pending(agg->agg, (cudaStream_t)stream); if (wrc) return wrc; }
  uint32_t bits = 0;
  VG_CUDA(cudaMemcpyAsync(&bits, agg->agg.err, 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  VG_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  if (!bits) return VGPU_OK;
  VG_CUDA(cudaMemsetAsync(agg->agg.err, 0, 4, (cudaStream_t)stream));
  return report_error_bits(bits);
}
int vgpu_agg_memory_bytes(vgpu_agg_t agg, int64_t* bytes) {
  if (!agg || !bytes) return VGPU_ERR_INVALID;
  const Agg& a = agg->agg;
  int64_t b = 0;
  if (a.mode == VGPU_AGG_HASH) {
    b += (int64_t)(a.capacity ? (a.capacity + 2) * a.row_words * 8 : 0);          // TABLE engine
    b += (int64_t)(a.log_bound * (uint64_t)a.ew * 8);                               // LOG engine: entries logged so far (upper bound)
  } else {
    b += (int64_t)a.n_cells * a.n_groups * 8;
  }
  *bytes = b;
  return VGPU_OK;
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

// info[0] = engine of a HASH aggregation (0 undecided, 1 log, 2 table), then — log engine, after finalize —
// [1] sorted input detected, [2] front entries, [3] entries deduplicated, [4] groups the deduplication produced,
// [5] hash partitions, [6] words per entry, [7] upper bound of the logged entries.  Synchronises the stream.
int vgpu_debug_agg_info(vgpu_agg_t agg, void* stream, int64_t info[8]) {
  if (!agg || !info) return VGPU_ERR_INVALID;
  Agg& a = agg->agg;
  for (int i = 0; i < 8; ++i) info[i] = 0;
  info[0] = a.mode == VGPU_AGG_HASH ? a.engine : -1;
  if (a.mode == VGPU_AGG_HASH && a.engine == kEngineLog && a.finalized) {
    FinHeader h{};
    uint64_t nsp = 0;
    int rc = log_read_header(a, (cudaStream_t)stream, &h, &nsp);
    if (rc) return rc;
    info[1] = (int64_t)h.sorted; info[2] = (int64_t)h.n_front; info[3] = (int64_t)h.n_dedupe; info[4] = (int64_t)h.n_groups;
    info[5] = (int64_t)1 << (int)h.log2p; info[6] = a.ew; info[7] = (int64_t)a.log_bound;
  }
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

} // extern "C"
