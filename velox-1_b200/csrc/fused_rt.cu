// Plan-specialised kernels: compile, cache and launch.
//
// Lookup order for the kernel of a pipeline shape:
//   1. in-memory cache keyed by the shape signature,
//   2. <library dir>/aot/<hash>.cubin   — prebuilt by build() with nvcc for the shapes of
//      velox-1_b200/aot_plans.py (the configurations of BASELINE.json),
//   3. $VGPU_CACHE_DIR/<hash>.cubin     — earlier NVRTC results (default /tmp/vgpu_kernel_cache),
//   4. NVRTC (libnvrtc, loaded with dlopen) for sm_100a, -fmad=false, then stored in 3.
// <hash> = FNV-1a 64 of prelude + generated body.  Modules are loaded and launched through the
// driver API (libcuda, dlopen'ed so the library still loads on a box without a GPU).
// VGPU_FUSED=0 disables the whole path (the interpreter runs everything).
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "fused.h"
#include "fused_prelude.cuh"
#include "build/prelude_str.h" // const char kFusedPrelude[] = the text of fused_prelude.cuh

namespace vg {

static_assert(VGF_BOOLEAN == VGPU_BOOLEAN && VGF_DOUBLE == VGPU_DOUBLE && VGF_INTEGER == VGPU_INTEGER, "type codes");
static_assert(VGF_ADD == VGPU_OP_ADD && VGF_MOD == VGPU_OP_MOD && VGF_EQ == VGPU_OP_EQ && VGF_GE == VGPU_OP_GE, "opcodes");
static_assert(VGF_SUMF64 == kCellSumF64 && VGF_COUNT == kCellCount && VGF_MAXF64 == kCellMaxF64, "cell kinds");
static_assert(VGF_ERR_OVERFLOW == kErrOverflow && VGF_ERR_KEYRANGE == kErrKeyRange && VGF_ERR_TABLEFULL == kErrTableFull && VGF_ERR_STRKEY == kErrStringKey, "error bits");

struct FusedKernel {
  void* module = nullptr;   // CUmodule
  void* function = nullptr; // CUfunction
  int threads = 512;
  int smem_bytes = 0;
  int blocks_per_sm = 1;
  std::string origin;       // "aot", "disk", "nvrtc"
};

namespace {

// ---- driver API / NVRTC through dlopen ----------------------------------------------------------
typedef int (*cu_module_load_data_t)(void**, const void*);
typedef int (*cu_module_get_function_t)(void**, void*, const char*);
typedef int (*cu_launch_kernel_t)(void*, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, void*, void**, void**);
typedef int (*cu_func_set_attribute_t)(void*, int, int);
typedef int (*cu_occupancy_t)(int*, void*, int, size_t);
typedef int (*cu_get_error_string_t)(int, const char**);
struct CuApi {
  bool tried = false, ok = false;
  cu_module_load_data_t module_load_data;
  cu_module_get_function_t module_get_function;
  cu_launch_kernel_t launch_kernel;
  cu_func_set_attribute_t func_set_attribute;
  cu_occupancy_t occupancy;
  cu_get_error_string_t get_error_string;
} g_cu;

typedef int (*nvrtc_create_t)(void**, const char*, const char*, int, const char* const*, const char* const*);
typedef int (*nvrtc_compile_t)(void*, int, const char* const*);
typedef int (*nvrtc_size_t)(void*, size_t*);
typedef int (*nvrtc_get_t)(void*, char*);
typedef int (*nvrtc_destroy_t)(void**);
struct NvrtcApi {
  bool tried = false, ok = false;
  nvrtc_create_t create;
  nvrtc_compile_t compile;
  nvrtc_size_t cubin_size, log_size;
  nvrtc_get_t get_cubin, get_log;
  nvrtc_destroy_t destroy;
} g_rtc;

std::mutex g_mu;
std::map<std::string, FusedKernel*> g_cache; // signature bytes -> kernel (nullptr = declined)
thread_local std::string g_last_source, g_last_note;
thread_local bool g_strict_failed = false;
bool g_dry = false;

bool load_cu() {
  if (g_cu.tried) return g_cu.ok;
  g_cu.tried = true;
  void* h = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libcuda.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return false;
#define VG_SYM(field, type, name) g_cu.field = (type)dlsym(h, name); if (!g_cu.field) return false;
  VG_SYM(module_load_data, cu_module_load_data_t, "cuModuleLoadData")
  VG_SYM(module_get_function, cu_module_get_function_t, "cuModuleGetFunction")
  VG_SYM(launch_kernel, cu_launch_kernel_t, "cuLaunchKernel")
  VG_SYM(func_set_attribute, cu_func_set_attribute_t, "cuFuncSetAttribute")
  VG_SYM(occupancy, cu_occupancy_t, "cuOccupancyMaxActiveBlocksPerMultiprocessor")
  VG_SYM(get_error_string, cu_get_error_string_t, "cuGetErrorString")
#undef VG_SYM
  g_cu.ok = true;
  return true;
}

bool load_rtc() {
  if (g_rtc.tried) return g_rtc.ok;
  g_rtc.tried = true;
  // the toolkit's own NVRTC first: the one a framework bundles may be older than the installed ptxas features
  const char* names[] = {"/usr/local/cuda/lib64/libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so", "libnvrtc.so.12", "libnvrtc.so"};
  void* h = nullptr;
  for (const char* n : names) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
  if (!h) return false;
#define VG_SYM(field, type, name) g_rtc.field = (type)dlsym(h, name); if (!g_rtc.field) return false;
  VG_SYM(create, nvrtc_create_t, "nvrtcCreateProgram")
  VG_SYM(compile, nvrtc_compile_t, "nvrtcCompileProgram")
  VG_SYM(cubin_size, nvrtc_size_t, "nvrtcGetCUBINSize")
  VG_SYM(get_cubin, nvrtc_get_t, "nvrtcGetCUBIN")
  VG_SYM(log_size, nvrtc_size_t, "nvrtcGetProgramLogSize")
  VG_SYM(get_log, nvrtc_get_t, "nvrtcGetProgramLog")
  VG_SYM(destroy, nvrtc_destroy_t, "nvrtcDestroyProgram")
#undef VG_SYM
  g_rtc.ok = true;
  return true;
}

uint64_t fnv(const void* data, size_t n, uint64_t h = 0xcbf29ce484222325ull) {
  const unsigned char* p = (const unsigned char*)data;
  for (size_t i = 0; i < n; ++i) { h ^= p[i]; h *= 0x100000001b3ull; }
  return h;
}

std::string lib_dir() {
  Dl_info info;
  if (dladdr((void*)&load_cu, &info) && info.dli_fname) {
    std::string s(info.dli_fname);
    const size_t k = s.rfind('/');
    return k == std::string::npos ? "." : s.substr(0, k);
  }
  return ".";
}
// Per-user cache directory (mode 0700): $VGPU_CACHE_DIR, else $XDG_CACHE_HOME/vgpu, else ~/.cache/vgpu, else
// /tmp/vgpu_kernel_cache.<uid>.  A directory that is not ours or is writable by others is not used.
std::string cache_dir() {
  const char* e = getenv("VGPU_CACHE_DIR");
  if (e && *e) return e;
  const char* x = getenv("XDG_CACHE_HOME");
  if (x && *x) return std::string(x) + "/vgpu";
  const char* h = getenv("HOME");
  if (h && *h) return std::string(h) + "/.cache/vgpu";
  return "/tmp/vgpu_kernel_cache." + std::to_string((long)getuid());
}
bool cache_dir_trusted(const std::string& dir) {
  struct stat st;
  if (stat(dir.c_str(), &st) != 0) return false;
  return S_ISDIR(st.st_mode) && st.st_uid == getuid() && (st.st_mode & (S_IWGRP | S_IWOTH)) == 0;
}

bool read_file(const std::string& path, std::vector<char>& out) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  const long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  out.resize(n > 0 ? n : 0);
  const bool ok = n > 0 && fread(out.data(), 1, n, f) == (size_t)n;
  fclose(f);
  return ok;
}
void write_file(const std::string& dir, const std::string& name, const std::vector<char>& data) {
  const size_t slash = dir.rfind('/');
  if (slash != std::string::npos && slash > 0) mkdir(dir.substr(0, slash).c_str(), 0700); // e.g. ~/.cache
  mkdir(dir.c_str(), 0700);
  if (!cache_dir_trusted(dir)) return;
  const std::string tmp = dir + "/." + name + "." + std::to_string((long)getpid());
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) return;
  const bool ok = fwrite(data.data(), 1, data.size(), f) == data.size();
  fclose(f);
  if (ok) rename(tmp.c_str(), (dir + "/" + name).c_str()); else remove(tmp.c_str());
}

bool nvrtc_compile(const std::string& full_source, std::vector<char>& cubin, std::string& log) {
  if (!load_rtc()) { log = "libnvrtc not found"; return false; }
  void* prog = nullptr;
  if (g_rtc.create(&prog, full_source.c_str(), "vgpu_fused.cu", 0, nullptr, nullptr) != 0) { log = "nvrtcCreateProgram failed"; return false; }
  const char* opts[] = {"--gpu-architecture=sm_100a", "-fmad=false", "-lineinfo", "--std=c++17"};
  const int rc = g_rtc.compile(prog, 4, opts);
  size_t ls = 0;
  if (g_rtc.log_size(prog, &ls) == 0 && ls > 1) { log.resize(ls); g_rtc.get_log(prog, &log[0]); }
  bool ok = false;
  if (rc == 0) {
    size_t n = 0;
    if (g_rtc.cubin_size(prog, &n) == 0 && n > 0) { cubin.resize(n); ok = g_rtc.get_cubin(prog, cubin.data()) == 0; }
  }
  g_rtc.destroy(&prog);
  return ok;
}

// The facts of `p` the generated text depends on.
std::string signature(const PipelineParams& p) {
  std::string s;
  auto put = [&](const void* d, size_t n) { s.append((const char*)d, n); };
  const Program& prog = *p.prog;
  put(&p.mode, 4);
  const int n_ops = (int)prog.ops.size();
  put(&n_ops, 4);
  if (n_ops) put(prog.ops.data(), sizeof(vgpu_op) * n_ops);
  put(&prog.filter, sizeof(prog.filter));
  put(&p.n_outputs, 4);
  put(p.outputs, sizeof(vgpu_operand) * p.n_outputs);
  put(&p.n_cols, 4);
  for (int c = 0; c < p.n_cols; ++c) {
    const DevCol& d = p.cols[c];
    const void* walk = d.encoding == VGPU_DICTIONARY ? (const void*)d.indices : d.values;
    const uint8_t f[6] = {d.type, d.encoding, (uint8_t)(d.nulls != nullptr), (uint8_t)(d.base_nulls != nullptr), d.const_null,
                          (uint8_t)(((reinterpret_cast<uintptr_t>(walk) & 15u) != 0) | (((reinterpret_cast<uintptr_t>(d.nulls) & 7u) != 0) << 1) |
                                    (((reinterpret_cast<uintptr_t>(walk) & 31u) != 0) << 2) | (((reinterpret_cast<uintptr_t>(walk) & 7u) != 0) << 3) |
                                    (((reinterpret_cast<uintptr_t>(walk) & 3u) != 0) << 4) | (((reinterpret_cast<uintptr_t>(walk) & 1u) != 0) << 5))};
    put(f, 6);
    const uint8_t sd = d.small_dict != 0 && fused_small_dict_enabled();
    put(&sd, 1);
    if (d.encoding == VGPU_CONSTANT) put(&d.constant, 8);
  }
  const uint8_t ind = p.in_rows != nullptr;
  put(&ind, 1);
  if (p.mode != kSinkFilterProject) {
    put(&p.n_keys, 4); put(&p.n_cells, 4); put(&p.n_groups, 4); put(&p.ignore_null_keys, 4); put(&p.row_words, 4);
    for (int k = 0; k < p.n_keys; ++k) {
      const DevKey& ks = p.keys[k];
      put(&ks.min, 8); put(&ks.max, 8); put(&ks.mult, 8); put(&ks.output, 1); put(&ks.bits, 1); put(&ks.shift, 1); put(&ks.nullable, 1);
    }
    for (int c = 0; c < p.n_cells; ++c) { put(&p.cells[c].kind, 1); put(&p.cells[c].input, 1); put(&p.cells[c].skip, 1); }
    if (p.mode == kSinkLog) {
      put(&p.log_entry_words, 4); put(&p.log_n_cells, 4); put(p.log_cells, (size_t)p.log_n_cells);
      put(&p.sorted_probe, 4); put(&p.sorted_key_col, 4); put(&p.send_world, 4); put(&p.send_hash, 4);
    }
  } else {
    for (int k = 0; k < p.n_outputs; ++k) {
      const uint8_t f[2] = {(uint8_t)(p.outs[k].values != nullptr), (uint8_t)(p.outs[k].nulls != nullptr)};
      put(f, 2);
    }
  }
  return s;
}

} // namespace

void fused_set_dry_run(bool on) { g_dry = on; }
bool fused_dry_run() { return g_dry; }
bool fused_small_dict_enabled() {
  static const bool on = !(getenv("VGPU_SMALL_DICT") && getenv("VGPU_SMALL_DICT")[0] == '0');
  return on;
}
bool fused_strict_failure() { const bool f = g_strict_failed; g_strict_failed = false; return f; }
const std::string& fused_last_source() { return g_last_source; }
const std::string& fused_last_note() { return g_last_note; }
uint64_t fused_source_hash(const std::string& source) {
  return fnv(source.data(), source.size(), fnv(kFusedPrelude, strlen(kFusedPrelude)));
}

FusedKernel* fused_prepare(const PipelineParams& p, int* grid) {
  const char* env = getenv("VGPU_FUSED");
  if (env && env[0] == '0') { g_last_note = "disabled by VGPU_FUSED=0"; return nullptr; }
  int dev = 0;
  cudaGetDevice(&dev);
  const std::string sig = std::string(1, (char)dev) + signature(p);
  FusedKernel* k = nullptr;
  {
    std::lock_guard<std::mutex> l(g_mu);
    auto it = g_cache.find(sig);
    if (it != g_cache.end()) {
      k = it->second;
      if (!k) return nullptr;
    }
  }
  if (!k) {
    FusedPlan plan;
    std::string body, why;
    if (!fused_gen_source(p, plan, body, why)) {
      g_last_note = "declined: " + why;
      std::lock_guard<std::mutex> l(g_mu);
      g_cache[sig] = nullptr;
      return nullptr;
    }
    g_last_source = body;
    if (g_dry) { g_last_note = "dry run"; return nullptr; }
    if (!load_cu()) { g_last_note = "libcuda not available"; return nullptr; }
    char name[32];
    snprintf(name, sizeof(name), "%016llx.cubin", (unsigned long long)fused_source_hash(body));
    std::vector<char> cubin;
    std::string origin;
    if (read_file(lib_dir() + "/aot/" + name, cubin)) origin = "aot";
    else if (cache_dir_trusted(cache_dir()) && read_file(cache_dir() + "/" + name, cubin)) origin = "disk";
    else {
      std::string log;
      const std::string full = std::string(kFusedPrelude) + "\n" + body;
      if (!nvrtc_compile(full, cubin, log)) {
        g_last_note = "nvrtc failed: " + log.substr(0, 2000);
        // VGPU_STRICT=1 (the test suites): a generated kernel that does not compile is an error, not a reason to run the
        // shape on the interpreter
        if (getenv("VGPU_STRICT") && getenv("VGPU_STRICT")[0] == '1') { g_strict_failed = true; set_error("%s", g_last_note.c_str()); }
        if (getenv("VGPU_FUSED_VERBOSE")) fprintf(stderr, "[vgpu] %s\n", g_last_note.c_str());
        std::lock_guard<std::mutex> l(g_mu);
        g_cache[sig] = nullptr;
        return nullptr;
      }
      origin = "nvrtc";
      write_file(cache_dir(), name, cubin);
    }
    k = new FusedKernel();
    k->threads = plan.threads;
    k->smem_bytes = plan.smem_bytes;
    k->origin = origin;
    int rc = g_cu.module_load_data(&k->module, cubin.data());
    if (rc == 0) rc = g_cu.module_get_function(&k->function, k->module, "vgpu_fused");
    if (rc == 0 && k->smem_bytes > 48 * 1024) rc = g_cu.func_set_attribute(k->function, 8 /*MAX_DYNAMIC_SHARED_SIZE_BYTES*/, k->smem_bytes);
    int occ = 0;
    if (rc == 0) rc = g_cu.occupancy(&occ, k->function, k->threads, (size_t)k->smem_bytes);
    if (rc != 0 || occ < 1) {
      const char* msg = "?";
      g_cu.get_error_string(rc, &msg);
      g_last_note = std::string("module load failed (") + origin + "): " + (msg ? msg : "?");
      if (getenv("VGPU_FUSED_VERBOSE")) fprintf(stderr, "[vgpu] %s\n", g_last_note.c_str());
      delete k;
      std::lock_guard<std::mutex> l(g_mu);
      g_cache[sig] = nullptr;
      return nullptr;
    }
    k->blocks_per_sm = occ;
    g_last_note = "built from " + origin;
    if (getenv("VGPU_FUSED_VERBOSE")) fprintf(stderr, "[vgpu] fused kernel %s from %s, %d CTA/SM\n", name, origin.c_str(), occ);
    std::lock_guard<std::mutex> l(g_mu);
    g_cache[sig] = k;
  }
  const int64_t blocks = (p.n_rows + 127) / 128;
  const int64_t want = (blocks + k->threads / 32 - 1) / (k->threads / 32);
  *grid = (int)std::max<int64_t>(1, std::min<int64_t>(want, (int64_t)num_sms() * k->blocks_per_sm));
  // exchange fused in: the launch shape names the sub-regions of the peers' windows, so it must not follow the batch size
  if (p.mode == kSinkLog && p.send_world > 0) *grid = num_sms() * k->blocks_per_sm;
  return k;
}

int fused_kernel_threads(const FusedKernel* k) { return k->threads; }

int fused_launch(FusedKernel* k, const PipelineParams& p, int grid, cudaStream_t stream) {
  vgf::FusedParams fp;
  memset(&fp, 0, sizeof(fp));
  for (int c = 0; c < p.n_cols && c < 16; ++c) {
    fp.values[c] = p.cols[c].values;
    fp.indices[c] = p.cols[c].indices;
    fp.nulls[c] = (const vgf::u64*)p.cols[c].nulls;
    fp.base_nulls[c] = (const vgf::u64*)p.cols[c].base_nulls;
    fp.base_len[c] = p.cols[c].small_dict;
  }
  fp.n_rows = p.n_rows;
  fp.in_rows = p.in_rows;
  fp.err_word = p.err_word;
  fp.partials = (vgf::u64*)p.partials;
  fp.table = (vgf::u64*)p.table;
  fp.cap_mask = p.cap_mask;
  fp.key_min = p.key_map.min; fp.key_span = p.key_map.span; fp.key_scale = p.key_map.scale; fp.key_mode = p.key_map.mode; fp.key_pad = p.key_map.pad;
  fp.n_groups_ctr = (vgf::u64*)p.n_groups_ctr;
  fp.ovf = (vgf::u64*)p.ovf; fp.ovf_cap = p.ovf_cap; fp.ovf_cursor = (vgf::u64*)p.ovf_cursor;
  fp.log = (vgf::u64*)p.log; fp.log_region_cap = p.log_region_cap; fp.log_counts = p.log_counts; fp.log_special = (vgf::u64*)p.log_special;
  fp.sort_state = (vgf::u64*)p.sort_state; fp.sort_prev_in = (const vgf::u64*)p.sort_prev_in; fp.sort_prev_out = (vgf::u64*)p.sort_prev_out;
  fp.send_sub_cap = p.send_sub_cap; fp.send_counts = p.send_counts;
  for (int d = 0; d < 8; ++d) { fp.send_base[d] = (vgf::u64*)p.send_base[d]; fp.send_peer_counts[d] = p.send_peer_counts[d]; }
  fp.sel_words = p.sel_words;
  fp.tile_counts = p.tile_counts;
  for (int i = 0; i < p.n_outputs && i < 24; ++i) { fp.out_values[i] = p.outs[i].values; fp.out_nulls[i] = (vgf::u64*)p.outs[i].nulls; }
  void* args[] = {&fp};
  kernel_timing_begin(stream);
  const int rc = g_cu.launch_kernel(k->function, (unsigned)grid, 1, 1, (unsigned)k->threads, 1, 1, (unsigned)k->smem_bytes, (void*)stream, args, nullptr);
  if (rc != 0) {
    const char* msg = "?";
    g_cu.get_error_string(rc, &msg);
    set_error("cuLaunchKernel failed: %s", msg ? msg : "?");
    return VGPU_ERR_CUDA;
  }
  kernel_timing_end(stream);
  count_launch();
  return VGPU_OK;
}

} // namespace vg
