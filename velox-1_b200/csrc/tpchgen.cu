// TPC-H lineitem generated directly into HBM: the device analogue of the reference's TableScan over the TPC-H connector
//   velox/connectors/tpch/TpchConnector.cpp:139-163 (TpchDataSource::next), velox/tpch/gen/TpchGen.cpp:401-489
//   (genTpchLineItem, decimalToDouble :79-81, toDate :83-85), velox/tpch/gen/DBGenIterator.cpp:70-73 (initOrder),
//   velox/tpch/gen/dbgen/build.cpp:110-223 (mk_order), :43-51 (rpb_routine), :29-35 (PART_SUPP_BRIDGE), :85-108 (mk_sparse),
//   velox/tpch/gen/dbgen/rnd.cpp (NextRand, UnifInt, row_stop_h), speed_seed.cpp:67-88 (NthElement), :120-159 (sd_line / sd_order),
//   velox/tpch/gen/dbgen/include/dbgen/dss.h:491-540 (the seed table), dists_dss.h (instruct / smode / rflag).
// Bit-exact with the reference's generator (tests compare with oracle/_ref/libdbgen_ref.so, the reference's own dbgen):
// dbgen's streams are Lehmer generators x -> 16807 x mod (2^31 - 1), one per column, and every stream advances a FIXED number
// of draws per order (its "boundary": row_stop_h skips the draws an order did not use), so the state of stream s at order o
// is seed_s * 16807^(o * boundary_s) — computed here with three table look-ups per stream instead of a sequential walk.
// One thread generates one order (its 1..7 lineitems); a block scan of the line counts gives the row positions.
// String columns are produced as the int32 dictionary ids the rest of this library uses for VARCHAR (the dictionaries
// are returned by vgpu_tpch_dictionary, sorted like tests/lineitem.py's).
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <mutex>

#include "internal.h"

namespace vg {

namespace {
constexpr uint64_t kMod = 2147483647ull;
constexpr uint64_t kMul = 16807ull;
// dss.h:491-540
enum Stream : int { S_ODATE = 0, S_LCNT, S_QTY, S_DCNT, S_TAX, S_SHIP, S_SMODE, S_PKEY, S_SKEY, S_SDTE, S_CDTE, S_RDTE, S_RFLG, S_COUNT };
__constant__ uint32_t c_seed0[S_COUNT] = {1066728069u, 1434868289u, 209208115u, 554590007u, 721958466u, 1371272478u, 675466456u,
                                          1808217256u, 2095021727u, 1769349045u, 904914315u, 373135028u, 717419739u};
constexpr int kStartDay = 8035;      // 1992-01-01 (STARTDATE 92001) in days since the epoch
constexpr int kCurrentDay = 9298;    // 1995-06-17 (CURRENTDATE 95168)
constexpr int kTotDate = 2557;
constexpr int kODateSpan = kTotDate - (121 + 30) - 1 + 1; // O_ODATE_MAX - O_ODATE_MIN + 1 = TOTDATE - (L_SDTE_MAX + L_RDTE_MAX)

// 16807^e mod M as T0[e & 1023] * T1[(e >> 10) & 1023] * T2[(e >> 20) & 1023] * T3[e >> 30]
struct PowTables { uint32_t t[4][1024]; };
PowTables* g_tables[64] = {nullptr};
std::mutex g_tables_mu;

__host__ __device__ __forceinline__ uint64_t mulmod(uint64_t x, uint64_t y) {
  uint64_t p = x * y; // < 2^62
  p = (p & kMod) + (p >> 31);
  p = (p & kMod) + (p >> 31);
  return p >= kMod ? p - kMod : p;
}
__device__ __forceinline__ uint64_t pow_mul(const PowTables* T, uint64_t e) {
  uint64_t r = mulmod(T->t[0][e & 1023], T->t[1][(e >> 10) & 1023]);
  r = mulmod(r, T->t[2][(e >> 20) & 1023]);
  if (e >> 30) r = mulmod(r, T->t[3][(e >> 30) & 1023]);
  return r;
}
// UnifInt (rnd.cpp): low + (int64)(((double)seed / dM) * (double)(high - low + 1)), seed advanced first
__device__ __forceinline__ int64_t draw(uint64_t& seed, int64_t low, int64_t high) {
  seed = mulmod(seed, kMul);
  const double range = (double)(high - low + 1);
  return low + (int64_t)__dmul_rn(__ddiv_rn((double)seed, 2147483647.0), range);
}

struct GenArgs {
  const PowTables* tables;
  int64_t order_first;   // absolute 0-based index of the slice's first order
  int64_t n_orders;      // orders in this slice
  int64_t scale;         // dbgen's integer scale factor (1 for sf < 1)
  uint8_t* lines;        // [n_orders] line count of each order
  uint32_t* intra;       // [n_orders] rows before the order inside its block
  unsigned long long* block_rows;  // [blocks] rows of each block; after the scan: rows before the block (in the slice)
  unsigned long long* row_base;    // device: rows generated before this slice (running total)
  int64_t capacity;      // rows the output arrays hold
  vgpu_lineitem_columns out;
  uint32_t* err;
};

constexpr int kGenThreads = 256;

__global__ void __launch_bounds__(kGenThreads) tpch_count_kernel(GenArgs a) {
  __shared__ uint32_t s_warp[kGenThreads / 32];
  const int64_t o = (int64_t)blockIdx.x * kGenThreads + threadIdx.x;
  uint32_t n = 0;
  if (o < a.n_orders) {
    uint64_t s = mulmod(c_seed0[S_LCNT], pow_mul(a.tables, (uint64_t)(a.order_first + o)));
    n = (uint32_t)draw(s, 1, 7);
    a.lines[o] = (uint8_t)n;
  }
  uint32_t x = n;
  for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if ((int)(threadIdx.x & 31) >= d) x += y; }
  if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = x;
  __syncthreads();
  uint32_t before = 0;
  for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) before += s_warp[w];
  if (o < a.n_orders) a.intra[o] = before + x - n;
  if (threadIdx.x == kGenThreads - 1) a.block_rows[blockIdx.x] = before + x;
}

// Exclusive scan of the block totals (one CTA); the slice's rows are added to the running total.
__global__ void tpch_scan_kernel(unsigned long long* block_rows, int n_blocks, unsigned long long* row_base, unsigned long long* slice_base) {
  __shared__ unsigned long long s_warp[32];
  __shared__ unsigned long long carry;
  if (threadIdx.x == 0) { carry = 0; *slice_base = *row_base; }
  __syncthreads();
  for (int base = 0; base < n_blocks; base += blockDim.x) {
    const int i = base + threadIdx.x;
    const unsigned long long v = i < n_blocks ? block_rows[i] : 0ull;
    unsigned long long x = v;
    for (int d = 1; d < 32; d <<= 1) { const unsigned long long y = __shfl_up_sync(0xffffffffu, x, d); if ((int)(threadIdx.x & 31) >= d) x += y; }
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      unsigned long long w = threadIdx.x < (blockDim.x >> 5) ? s_warp[threadIdx.x] : 0ull;
      for (int d = 1; d < 32; d <<= 1) { const unsigned long long y = __shfl_up_sync(0xffffffffu, w, d); if ((int)threadIdx.x >= d) w += y; }
      s_warp[threadIdx.x] = w;
    }
    __syncthreads();
    const unsigned long long before = carry + ((threadIdx.x >> 5) ? s_warp[(threadIdx.x >> 5) - 1] : 0ull) + x - v;
    if (i < n_blocks) block_rows[i] = before;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) *row_base += carry;
}

__global__ void __launch_bounds__(kGenThreads) tpch_fill_kernel(GenArgs a, const unsigned long long* slice_base) {
  const int64_t o = (int64_t)blockIdx.x * kGenThreads + threadIdx.x;
  if (o >= a.n_orders) return;
  const int lines = a.lines[o];
  const int64_t row0 = (int64_t)(*slice_base + a.block_rows[blockIdx.x] + a.intra[o]);
  if (row0 + lines > a.capacity) { atomicOr(a.err, kErrTableFull); return; }
  const uint64_t abs_o = (uint64_t)(a.order_first + o);
  const vgpu_lineitem_columns& out = a.out;
  // mk_sparse (build.cpp:85-108) of the 1-based order index
  const uint64_t idx = abs_o + 1;
  const int64_t okey = (int64_t)(((idx >> 3) << 5) | (idx & 7));
  uint64_t s_odate = mulmod(c_seed0[S_ODATE], pow_mul(a.tables, abs_o));
  const int64_t odate = draw(s_odate, 0, kODateSpan - 1); // days after STARTDATE
  const uint64_t p7 = pow_mul(a.tables, abs_o * 7);        // every line stream has used 7 draws per earlier order
  const bool want_price = out.l_extendedprice || out.l_partkey || out.l_suppkey;
  const bool want_rflag = out.l_returnflag != nullptr;
  uint64_t s_qty = mulmod(c_seed0[S_QTY], p7), s_dcnt = mulmod(c_seed0[S_DCNT], p7), s_tax = mulmod(c_seed0[S_TAX], p7);
  uint64_t s_ship = mulmod(c_seed0[S_SHIP], p7), s_smode = mulmod(c_seed0[S_SMODE], p7), s_pkey = mulmod(c_seed0[S_PKEY], p7);
  uint64_t s_skey = mulmod(c_seed0[S_SKEY], p7), s_sdte = mulmod(c_seed0[S_SDTE], p7), s_cdte = mulmod(c_seed0[S_CDTE], p7);
  uint64_t s_rdte = mulmod(c_seed0[S_RDTE], p7), s_rflg = mulmod(c_seed0[S_RFLG], p7);
  const int64_t tot_scnt = 10000 * a.scale;
  for (int l = 0; l < lines; ++l) {
    const int64_t r = row0 + l;
    if (out.l_orderkey) out.l_orderkey[r] = okey;
    if (out.l_linenumber) out.l_linenumber[r] = l + 1;
    int64_t qty = 0;
    if (out.l_quantity || out.l_extendedprice) { qty = draw(s_qty, 1, 50); if (out.l_quantity) out.l_quantity[r] = __dmul_rn((double)qty, 0.01); }
    if (out.l_discount) out.l_discount[r] = __dmul_rn((double)draw(s_dcnt, 0, 10), 0.01);
    if (out.l_tax) out.l_tax[r] = __dmul_rn((double)draw(s_tax, 0, 8), 0.01);
    if (out.l_shipinstruct) { // pick_str over instruct: DELIVER IN PERSON, COLLECT COD, TAKE BACK RETURN, NONE -> ids in sorted order
      const int j = (int)draw(s_ship, 1, 4);
      out.l_shipinstruct[r] = j == 1 ? 1 : j == 2 ? 0 : j == 3 ? 3 : 2;
    }
    if (out.l_shipmode) {     // REG AIR, AIR, RAIL, TRUCK, MAIL, FOB, SHIP -> ids in sorted order
      const int j = (int)draw(s_smode, 1, 7);
      const int map[8] = {0, 4, 0, 3, 6, 2, 1, 5};
      out.l_shipmode[r] = map[j];
    }
    if (want_price) {
      const int64_t pkey = draw(s_pkey, 1, 200000 * a.scale);
      if (out.l_partkey) out.l_partkey[r] = pkey;
      if (out.l_suppkey) {
        const int64_t snum = draw(s_skey, 0, 3);
        out.l_suppkey[r] = (pkey + snum * (tot_scnt / 4 + (pkey - 1) / tot_scnt)) % tot_scnt + 1;
      }
      if (out.l_extendedprice) {
        const int64_t rprice = 90000 + (pkey / 10) % 20001 + (pkey % 1000) * 100; // rpb_routine
        out.l_extendedprice[r] = __dmul_rn((double)(rprice * qty), 0.01);
      }
    }
    const int64_t sdate = odate + draw(s_sdte, 1, 121);
    if (out.l_shipdate) out.l_shipdate[r] = (int32_t)(kStartDay + sdate);
    if (out.l_commitdate) out.l_commitdate[r] = (int32_t)(kStartDay + odate + draw(s_cdte, 30, 90));
    if (out.l_receiptdate || want_rflag) {
      const int64_t rdate = sdate + draw(s_rdte, 1, 30);
      if (out.l_receiptdate) out.l_receiptdate[r] = (int32_t)(kStartDay + rdate);
      if (want_rflag) { // the R/A draw happens only for lines received by CURRENTDATE; ids of {"A","N","R"}
        int32_t id = 1;
        if (kStartDay + rdate <= kCurrentDay) id = draw(s_rflg, 1, 2) == 1 ? 2 : 0;
        out.l_returnflag[r] = id;
      }
    }
    if (out.l_linestatus) out.l_linestatus[r] = kStartDay + sdate <= kCurrentDay ? 0 : 1; // ids of {"F","O"}
  }
}

const PowTables* pow_tables() {
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> l(g_tables_mu);
  if (g_tables[dev]) return g_tables[dev];
  PowTables* h = new PowTables();
  uint64_t step = kMul;
  for (int level = 0; level < 4; ++level) {
    uint64_t v = 1;
    for (int j = 0; j < 1024; ++j) { h->t[level][j] = (uint32_t)v; v = mulmod(v, step); }
    step = v; // step^1024
  }
  PowTables* d = nullptr;
  if (cudaMalloc(&d, sizeof(PowTables)) == cudaSuccess && cudaMemcpy(d, h, sizeof(PowTables), cudaMemcpyHostToDevice) == cudaSuccess) g_tables[dev] = d;
  delete h;
  return g_tables[dev];
}

struct Scratch {
  uint8_t* lines = nullptr; uint32_t* intra = nullptr; unsigned long long* block_rows = nullptr; unsigned long long* ctr = nullptr;
  int64_t cap_orders = 0;
};
thread_local Scratch g_scratch;
constexpr int64_t kSliceOrders = 8 << 20;

int scratch_for(int64_t orders) {
  Scratch& s = g_scratch;
  if (s.cap_orders >= orders) return VGPU_OK;
  cudaFree(s.lines); cudaFree(s.intra); cudaFree(s.block_rows);
  s.cap_orders = 0;
  VG_CUDA(cudaMalloc(&s.lines, (size_t)orders));
  VG_CUDA(cudaMalloc(&s.intra, (size_t)orders * 4));
  VG_CUDA(cudaMalloc(&s.block_rows, ((size_t)orders / kGenThreads + 2) * 8));
  if (!s.ctr) VG_CUDA(cudaMalloc(&s.ctr, 64));
  s.cap_orders = orders;
  return VGPU_OK;
}

int generate(double sf, int64_t order_offset, int64_t max_orders, const vgpu_lineitem_columns* out, int64_t capacity, int64_t* rows_host, cudaStream_t stream) {
  if (sf <= 0 || order_offset < 0 || max_orders < 0) { set_error("bad TPC-H split"); return VGPU_ERR_INVALID; }
  const int64_t total = (int64_t)(1500000 * sf);         // TpchGen.cpp:142-143
  const int64_t n = order_offset >= total ? 0 : std::min(total - order_offset, max_orders);
  const PowTables* T = pow_tables();
  if (!T) return cuda_fail(cudaGetLastError(), "TPC-H power tables");
  int rc = scratch_for(std::min<int64_t>(std::max<int64_t>(n, 1), kSliceOrders));
  if (rc) return rc;
  Scratch& s = g_scratch;
  VG_CUDA(cudaMemsetAsync(s.ctr, 0, 64, stream));
  GenArgs a{};
  a.tables = T;
  a.scale = (sf < 1.0) ? 1 : (int64_t)sf;                // DBGenIterator.cpp:51-61
  a.lines = s.lines; a.intra = s.intra; a.block_rows = s.block_rows; a.row_base = s.ctr;
  a.capacity = capacity;
  if (out) a.out = *out;
  a.err = device_err_word();
  for (int64_t done = 0; done < n; done += kSliceOrders) {
    a.order_first = order_offset + done;
    a.n_orders = std::min(kSliceOrders, n - done);
    const int blocks = (int)((a.n_orders + kGenThreads - 1) / kGenThreads);
    tpch_count_kernel<<<blocks, kGenThreads, 0, stream>>>(a);
    tpch_scan_kernel<<<1, 1024, 0, stream>>>(s.block_rows, blocks, s.ctr, s.ctr + 1);
    if (out) tpch_fill_kernel<<<blocks, kGenThreads, 0, stream>>>(a, s.ctr + 1);
    VG_CUDA(cudaGetLastError());
    count_launch(out ? 3 : 2);
  }
  if (rows_host) {
    unsigned long long r = 0;
    VG_CUDA(cudaMemcpyAsync(&r, s.ctr, 8, cudaMemcpyDeviceToHost, stream));
    VG_CUDA(cudaStreamSynchronize(stream));
    *rows_host = (int64_t)r;
  }
  return VGPU_OK;
}
} // namespace

} // namespace vg

extern "C" {

int vgpu_tpch_lineitem_rows(double sf, int64_t order_offset, int64_t max_orders, int64_t* rows, void* stream) {
  if (!rows) return VGPU_ERR_INVALID;
  return vg::generate(sf, order_offset, max_orders, nullptr, 0, rows, (cudaStream_t)stream);
}

int vgpu_tpch_gen_lineitem(double sf, int64_t order_offset, int64_t max_orders, const vgpu_lineitem_columns* out,
                           int64_t capacity_rows, int64_t* rows, void* stream) {
  if (!out) return VGPU_ERR_INVALID;
  return vg::generate(sf, order_offset, max_orders, out, capacity_rows, rows, (cudaStream_t)stream);
}

int vgpu_tpch_dictionary(int32_t which, const char** out, int32_t cap) {
  static const char* rf[] = {"A", "N", "R"};
  static const char* ls[] = {"F", "O"};
  static const char* si[] = {"COLLECT COD", "DELIVER IN PERSON", "NONE", "TAKE BACK RETURN"};
  static const char* sm[] = {"AIR", "FOB", "MAIL", "RAIL", "REG AIR", "SHIP", "TRUCK"};
  const char** d = nullptr;
  int32_t n = 0;
  switch (which) {
    case 0: d = rf; n = 3; break;
    case 1: d = ls; n = 2; break;
    case 2: d = si; n = 4; break;
    case 3: d = sm; n = 7; break;
    default: return -1;
  }
  for (int32_t i = 0; i < n && i < cap && out; ++i) out[i] = d[i];
  return n;
}

} // extern "C"
