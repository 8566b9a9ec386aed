// Multi-GPU exchange between a partial and a final aggregation (one process per GPU):
// partialAggregation -> PartitionedOutput (HashPartitionFunction) -> Exchange -> finalAggregation of the reference
// (velox/exec/tests/MultiFragmentTest.cpp:180-212, velox/exec/HashPartitionFunction.cpp:59-92).
//
// B200-native shape.  Every rank owns a WINDOW of device memory that all peers map through CUDA IPC (NVLink 5 /
// NVSwitch peer stores).  The window has two halves used by alternate exchanges, and per half one region per source
// rank, so a sender owns its regions on every peer and needs no remote atomics:
//   * HASH aggregations (log engine): ONE kernel reads the partial stage's logged entries, hash-partitions them by
//     destination in shared memory and stores every destination's share straight into that peer's window over NVLink —
//     partitioning and transfer are the same pass, tile by tile.  The per-destination counts then travel in one tiny
//     ncclAllGather, which is also the only synchronisation between the ranks; the final stage deduplicates what
//     it received with the radix passes of hashagg.cu.  No host synchronisation anywhere in the step.
//   * GLOBAL / ARRAY aggregations (a few groups): every rank stores its dense group rows into all peers' windows and
//     raises a per-source epoch flag; the same kernel waits for the peers' flags and folds the groups this rank owns —
//     one launch instead of pack + ncclAllGather + merge.
// NCCL (dlopen'ed: the library has no link-time dependency on one NCCL build) carries the count matrix, the IPC handles
// at set-up and the whole payload when no window is available (vgpu_comm_alltoallv path, kept as the baseline).
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "agg.h"
#include "device_common.cuh"
#include "fused.h"
#include "internal.h"

namespace vg {

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

constexpr int kMaxWorld = 32;
constexpr int kMsgWords = kMaxWorld + 8;  // one rank's count message
constexpr int kWinParts = 3;          // window parts used by exchanges e, e+1, e+2 in turn (see exchange_log_fused for why three)
constexpr size_t kDcWords = (size_t)kMsgWords + 32 * kMsgWords + 32 * 32; // one part's message block
constexpr size_t kWinHeader = 4096;   // per half: epoch flags of the dense exchange, one 64-byte line per source
} // namespace

} // namespace vg

using namespace vg;

struct vgpu_comm_s {
  nccl_comm comm; int rank, world;
  // grow-only device scratch of the NCCL payload path
  void* send = nullptr; size_t send_cap = 0;
  void* recv = nullptr; size_t recv_cap = 0;
  int64_t* dcounts = nullptr;          // kWinParts blocks (one per window part) of: [kMsgWords] this rank's message + [32 * kMsgWords] the all-gathered messages [src]
                                       // message: [world] entries per destination (low 32 bits front, high 32 back), then kSrcMetaWords of stream facts
  // peer-mapped window
  char* win = nullptr;                 // this rank's window
  size_t win_bytes = 0;                // both halves
  char* peer[kMaxWorld] = {nullptr};   // peer[r] = rank r's window in this process' address space (peer[rank] == win)
  uint64_t epoch = 0;                  // exchanges done through the window
  uint32_t* chunk_prefix = nullptr; size_t chunk_prefix_cap = 0;
  // dense exchange: the wait-for-peers + fold half runs on a side stream, so the caller's stream goes on (next scan)
  cudaStream_t side = nullptr;
  cudaEvent_t stored = nullptr;                    // the caller's stream reached the point where the side stream takes over
  cudaEvent_t folded = nullptr, folded_prev = nullptr; // end of the latest exchange's tail, and of the one before it
  bool folded_valid = false, folded_prev_valid = false;
  bool local = false;                  // vgpu_debug_local_comms: all ranks on one device, peer[] are plain local allocations
};

namespace vg {

static int comm_reserve(void** p, size_t* cap, size_t want) {
  if (*cap >= want) return VGPU_OK;
  if (*p) { VG_CUDA(cudaDeviceSynchronize()); VG_CUDA(cudaFree(*p)); *p = nullptr; *cap = 0; }
  want = std::max<size_t>(want + want / 4, 1 << 16);
  VG_CUDA(cudaMalloc(p, want));
  *cap = want;
  return VGPU_OK;
}

// ---- destination of a group ------------------------------------------------------------------------------------
// partition_of (device_common.cuh): the same function as agg.cu's part_of, so the stepwise NCCL path and the fused
// path give every group the same owner — and, for one raw BIGINT key, the owner the reference's HashPartitionFunction
// would pick.
__device__ __forceinline__ uint32_t dest_of(uint64_t key, uint64_t null_flag, uint32_t world, bool velox_hash) {
  return partition_of(key, null_flag, world, velox_hash);
}

// ---- HASH: partition + peer store ---------------------------------------------------------------------------------
struct SendArgs {
  const LogRange* ranges;
  const uint32_t* chunk_prefix;
  int n_ranges, ew;
  uint32_t chunk;                 // entries per chunk (= what the shared-memory stage holds)
  uint32_t world, rank;
  int velox_hash;                 // one raw BIGINT key: HashPartitionFunction's own hash
  uint64_t* peer_region[kMaxWorld]; // destination d: first entry of the region this rank owns in d's window
  uint64_t region_cap;            // entries
  unsigned long long* cursor;     // [world]: entries stored per destination so far: low 32 bits front, high 32 bits back
  uint32_t* err;
};

constexpr int kSendThreads = 256;
constexpr int kSendPerThread = 8;

__global__ void __launch_bounds__(kSendThreads) xchg_send_kernel(SendArgs a) {
  extern __shared__ __align__(16) uint64_t stage[];
  __shared__ uint32_t cnt[kMaxWorld], off[kMaxWorld + 1];
  __shared__ unsigned long long base[kMaxWorld];
  const int ew = a.ew;
  const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
  const uint32_t total = a.chunk_prefix[a.n_ranges];
  bool overflow = false;
  for (uint32_t chunk = blockIdx.x; chunk < total; chunk += gridDim.x) {
    // locate the chunk (every thread computes the same answer)
    int lo = 0, hi = a.n_ranges;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (a.chunk_prefix[mid] <= chunk) lo = mid; else hi = mid; }
    const LogRange rg = a.ranges[lo];
    const uint64_t rcnt = rg.count64 ? (uint64_t) * reinterpret_cast<const unsigned long long*>(rg.count) : (uint64_t) * reinterpret_cast<const uint32_t*>(rg.count);
    const uint64_t c0 = (uint64_t)(chunk - a.chunk_prefix[lo]) * a.chunk;
    const uint64_t* first = (rg.back == 1 ? rg.base + (rg.cap - rcnt) * ew : rg.base) + c0 * ew;
    const uint32_t n = (uint32_t)min((uint64_t)a.chunk, rcnt - c0);
    if (threadIdx.x < kMaxWorld) cnt[threadIdx.x] = 0;
    __syncthreads();
    // pass 1: destination and rank inside the chunk's share for that destination (one shared atomic per warp and destination)
    uint32_t dst[kSendPerThread], rk[kSendPerThread];
#pragma unroll
    for (int k = 0; k < kSendPerThread; ++k) {
      const uint32_t i = k * kSendThreads + threadIdx.x;
      dst[k] = i < n ? dest_of(first[(uint64_t)i * ew], 0ull, a.world, a.velox_hash != 0) : 0xffffffffu;
      const uint32_t m = __match_any_sync(0xffffffffu, dst[k]);
      const int leader = __ffs(m) - 1;
      uint32_t b = 0;
      if ((int)lane == leader && dst[k] != 0xffffffffu) b = atomicAdd(&cnt[dst[k]], (uint32_t)__popc(m));
      b = __shfl_sync(0xffffffffu, b, leader);
      rk[k] = b + __popc(m & lt);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t run = 0;
      for (uint32_t d = 0; d < a.world; ++d) { off[d] = run; run += cnt[d]; }
      off[a.world] = run;
    }
    // front entries (whole groups if the stream is sorted) fill a region from its start, back entries from its end;
    // one packed cursor per destination keeps "front + back <= capacity" checkable with a single atomic
    const bool is_back = rg.back == 1;
    if (threadIdx.x < a.world) base[threadIdx.x] = cnt[threadIdx.x] ? atomicAdd(&a.cursor[threadIdx.x], (unsigned long long)cnt[threadIdx.x] << (is_back ? 32 : 0)) : 0ull;
    __syncthreads();
    // pass 2: entries grouped by destination in shared memory
#pragma unroll
    for (int k = 0; k < kSendPerThread; ++k) {
      if (dst[k] == 0xffffffffu) continue;
      const uint32_t i = k * kSendThreads + threadIdx.x;
      const uint64_t* e = first + (uint64_t)i * ew;
      uint64_t* s = stage + (uint64_t)(off[dst[k]] + rk[k]) * ew;
      if (ew == 2) *reinterpret_cast<ulonglong2*>(s) = *reinterpret_cast<const ulonglong2*>(e);
      else for (int w = 0; w < ew; ++w) s[w] = e[w];
    }
    __syncthreads();
    // pass 3: every destination's share leaves as one contiguous, coalesced run of peer stores
    for (uint32_t i = threadIdx.x; i < n; i += kSendThreads) {
      uint32_t d = 0;
      while (d + 1 < a.world && off[d + 1] <= i) ++d;
      const unsigned long long front = base[d] & 0xffffffffull, back = base[d] >> 32;
      const unsigned long long j = i - off[d];
      if (front + back + cnt[d] > a.region_cap) { overflow = true; continue; }
      uint64_t* out = is_back ? a.peer_region[d] + (a.region_cap - 1ull - (back + j)) * ew : a.peer_region[d] + (front + j) * ew;
      const uint64_t* s = stage + (uint64_t)i * ew;
      if (ew == 2) *reinterpret_cast<ulonglong2*>(out) = *reinterpret_cast<const ulonglong2*>(s);
      else for (int w = 0; w < ew; ++w) out[w] = s[w];
    }
    __syncthreads();
  }
  if (overflow) atomicOr(a.err, kErrTableFull);
}

// Facts about the sender's stream for the receivers' "all sources sorted and disjoint" shortcut.
__global__ void xchg_meta_kernel(unsigned long long* msg, uint32_t world, const unsigned long long* ctr, int probed, int last_parity) {
  if (threadIdx.x != 0) return;
  unsigned long long n = 0;
  for (uint32_t d = 0; d < world; ++d) n += (msg[d] & 0xffffffffull) + (msg[d] >> 32);
  msg[kMaxWorld + 0] = probed && ctr[kCtrSortState] == 0ull ? 1ull : 0ull;
  msg[kMaxWorld + 1] = ctr[kCtrFirstKey];
  msg[kMaxWorld + 2] = ctr[kCtrPrevKey + 2 * last_parity];
  msg[kMaxWorld + 3] = n;
}

// Fused exchange: the per-CTA counts summed per destination (the message's first `world` words), and the launch shape.
__global__ void xchg_send_totals_kernel(unsigned long long* msg, const unsigned long long* counts, int grid, uint32_t world,
                                        unsigned long long sub_cap) {
  __shared__ unsigned long long s_sum[kMaxWorld];
  if (threadIdx.x < kMaxWorld) s_sum[threadIdx.x] = 0ull;
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < (uint32_t)grid * world; i += blockDim.x) atomicAdd(&s_sum[i % world], counts[i]);
  __syncthreads();
  if (threadIdx.x < world) msg[threadIdx.x] = s_sum[threadIdx.x];
  if (threadIdx.x == 0) { msg[kMaxWorld + 4] = (unsigned long long)grid; msg[kMaxWorld + 5] = sub_cap; }
}

// The two dedicated groups (NULL key, key equal to the empty marker) go to their owners' slots in front of the region;
// every other destination gets "absent" so that a stale slot of an earlier exchange cannot be read.
struct SpecialArgs {
  const uint64_t* special;   // sender's rows: 2 x ew words {present, cells...}
  uint64_t* peer_slot[kMaxWorld];
  uint32_t world;
  int ew;
  int velox_hash;
};
__global__ void xchg_send_special_kernel(SpecialArgs a) {
  const uint32_t d = blockIdx.x;
  if (d >= a.world) return;
  const uint32_t owner[2] = {dest_of(0ull, 1ull, a.world, a.velox_hash != 0), dest_of(kEmptyKey, 0ull, a.world, a.velox_hash != 0)};
  for (int i = threadIdx.x; i < 2 * a.ew; i += blockDim.x) {
    const int r = i / a.ew, w = i % a.ew;
    uint64_t v = a.special[i];
    if (w == 0 && owner[r] != d) v = 0ull;
    a.peer_slot[d][i] = v;
  }
}

// Final stage: fold the received dedicated groups (one thread: at most 2 * world rows).
struct FoldSpecialArgs {
  const uint64_t* slot[kMaxWorld]; // source s: its 2 x ew words in this rank's window
  uint64_t* special;               // final aggregation's rows
  CellKinds kinds;
  uint32_t world;
  int ew;
  uint32_t* err;
  // fused exchange: every source that sent entries must have used the launch shape this rank laid its ranges out for
  const unsigned long long* meta;  // source s: meta + s * meta_stride -> {.., [3] entries, [4] CTAs, [5] entries per CTA and owner}
  int meta_stride;
  unsigned long long expect_grid, expect_sub_cap; // 0 = not a fused exchange
};
__global__ void xchg_fold_special_kernel(FoldSpecialArgs a) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  uint32_t err = 0;
  if (a.expect_grid) {
    for (uint32_t s = 0; s < a.world; ++s) {
      const unsigned long long* m = a.meta + (size_t)s * a.meta_stride;
      if (m[3] != 0ull && (m[4] != a.expect_grid || m[5] != a.expect_sub_cap)) err |= kErrTableFull;
    }
  }
  for (uint32_t s = 0; s < a.world; ++s) {
    for (int r = 0; r < 2; ++r) {
      const uint64_t* in = a.slot[s] + (size_t)r * a.ew;
      if (in[0] == 0ull) continue;
      uint64_t* out = a.special + (size_t)r * a.ew;
      out[0] = 1ull;
      for (int c = 0; c + 1 < a.ew; ++c) {
        uint8_t kind = a.kinds.kind[c];
        // MIN / MAX over doubles are kept as order-preserving integers in these rows (hashagg.cu init_special_kernel)
        if (kind == kCellMinF64) kind = kCellMinI64; else if (kind == kCellMaxF64) kind = kCellMaxI64;
        out[1 + c] = fold2(kind, out[1 + c], in[1 + c], err);
      }
    }
  }
  if (err) atomicOr(a.err, err);
}

// ---- GLOBAL / ARRAY: store to all peers, flag, wait, fold — one launch ---------------------------------------------
struct DenseArgs {
  const uint64_t* state;     // partial aggregation: [n_cells][G]
  uint64_t* final_state;     // final aggregation:   [n_cells][G]
  int G, n_cells;
  uint8_t kind[kMaxCells];
  uint32_t world, rank;
  uint64_t* peer_data[kMaxWorld];            // peer d: where this rank's [n_cells][G] image goes
  unsigned long long* peer_flag[kMaxWorld];  // peer d: this rank's epoch flag
  const uint64_t* my_data[kMaxWorld];        // this rank's window: source s' image
  const unsigned long long* my_flag[kMaxWorld];
  unsigned long long epoch;
  uint32_t* err;
  int phase;                 // 3 = the whole exchange (one rank per GPU); 1 = store + flag only, 2 = wait + fold only
                             // (all ranks emulated on ONE GPU must not wait for each other inside a kernel)
};

// Launched with ONE CTA (the images are small: G <= 4096 groups).  Each rank runs its own copy on its own GPU.
__global__ void xchg_dense_kernel(DenseArgs a) {
  const int words = a.G * a.n_cells;
  // 1. this rank's image -> every peer (including itself), then the flag, ordered by a system-scope fence
  if (a.phase & 1) {
    for (uint32_t d = 0; d < a.world; ++d) {
      for (int i = threadIdx.x; i < words; i += blockDim.x) a.peer_data[d][i] = a.state[i];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < a.world) {
      *reinterpret_cast<volatile unsigned long long*>(a.peer_flag[threadIdx.x]) = a.epoch;
    }
  }
  if (!(a.phase & 2)) return;
  // 2. wait until every source's flag carries this epoch (bounded: a missing peer must not hang the GPU)
  __shared__ int s_ok;
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if (threadIdx.x < a.world) {
    const volatile unsigned long long* f = a.my_flag[threadIdx.x];
    const long long t0 = clock64();
    while (*f < a.epoch) {
      if (clock64() - t0 > 20000000000ll) { s_ok = 0; break; } // ~10 s
      __nanosleep(200);
    }
  }
  __syncthreads();
  if (!s_ok) { if (threadIdx.x == 0) atomicOr(a.err, kErrTableFull); return; }
  __threadfence_system();
  // 3. fold the groups this rank owns, sources in rank order (deterministic)
  for (int g = threadIdx.x; g < a.G; g += blockDim.x) {
    if (dest_of((uint64_t)g, 0ull, a.world, false) != a.rank) continue;
    uint32_t err = 0;
    for (int c = 0; c < a.n_cells; ++c) {
      uint64_t acc = a.final_state[(size_t)c * a.G + g];
      for (uint32_t s = 0; s < a.world; ++s) {
        const volatile uint64_t* src = a.my_data[s];
        const uint64_t rows = src[g]; // cell 0 = rows of the group at source s
        if (rows == 0) continue;
        acc = fold2(a.kind[c], acc, src[(size_t)c * a.G + g], err);
      }
      a.final_state[(size_t)c * a.G + g] = acc;
    }
    if (err) atomicOr(a.err, err);
  }
}

} // namespace vg

extern "C" {

int vgpu_partition_of(int64_t key, int32_t is_null, int32_t n_parts, int32_t raw_bigint_key, int32_t* part) {
  if (!part || n_parts < 1) return VGPU_ERR_INVALID;
  *part = (int32_t)partition_of((uint64_t)key, is_null ? 1ull : 0ull, (uint32_t)n_parts, raw_bigint_key != 0);
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
  if (!id || !out || rank < 0 || rank >= world || world > kMaxWorld) return VGPU_ERR_INVALID;
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
static void window_close(vgpu_comm_s* c) {
  for (int r = 0; r < c->world; ++r) {
    if (r != c->rank && c->peer[r] && !c->local) cudaIpcCloseMemHandle(c->peer[r]);
    c->peer[r] = nullptr;
  }
  if (c->win) cudaFree(c->win);
  c->win = nullptr;
  c->win_bytes = 0;
}
int vgpu_comm_destroy(vgpu_comm_t comm) {
  if (comm) {
    if (comm->side) { cudaStreamSynchronize(comm->side); cudaStreamDestroy(comm->side); }
    if (comm->stored) cudaEventDestroy(comm->stored);
    if (comm->folded) cudaEventDestroy(comm->folded);
    if (comm->folded_prev) cudaEventDestroy(comm->folded_prev);
    window_close(comm);
    if (comm->comm) g_nccl.comm_destroy(comm->comm);
    cudaFree(comm->send); cudaFree(comm->recv); cudaFree(comm->dcounts); cudaFree(comm->chunk_prefix);
    delete comm;
  }
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

// Collective: every rank allocates a window of `bytes` (two halves), publishes its CUDA IPC handle through an
// ncclAllGather and maps the peers' windows.  All ranks must pass the same size.  When any rank cannot map a peer the
// window is dropped on every rank (*mapped = 0) and the exchanges use the NCCL payload path.
int vgpu_comm_reserve_window(vgpu_comm_t comm, size_t bytes, int32_t* mapped) {
  if (!comm) return VGPU_ERR_INVALID;
  if (mapped) *mapped = 0;
  if (comm->win && comm->win_bytes >= bytes) { if (mapped) *mapped = 1; return VGPU_OK; }
  VG_CUDA(cudaDeviceSynchronize());
  window_close(comm);
  const int W = comm->world;
  bytes = (bytes + kWinParts * kWinHeader + 4095) & ~(size_t)4095;
  struct Slot { cudaIpcMemHandle_t handle; int32_t ok; int32_t pad[15]; };
  static_assert(sizeof(Slot) == 128, "slot size");
  Slot mine{};
  mine.ok = 0;
  if (cudaMalloc(&comm->win, bytes) == cudaSuccess) {
    cudaMemset(comm->win, 0, bytes);
    mine.ok = W == 1 || cudaIpcGetMemHandle(&mine.handle, comm->win) == cudaSuccess ? 1 : 0;
  }
  cudaGetLastError();
  Slot* dslots = nullptr;
  VG_CUDA(cudaMalloc(&dslots, sizeof(Slot) * (W + 1)));
  std::vector<Slot> all(W);
  auto gather = [&]() -> int {
    VG_CUDA(cudaMemcpy(dslots + W, &mine, sizeof(Slot), cudaMemcpyHostToDevice));
    VG_NCCL(g_nccl.allgather(dslots + W, dslots, sizeof(Slot), kNcclUint8, comm->comm, (cudaStream_t)0));
    VG_CUDA(cudaDeviceSynchronize());
    VG_CUDA(cudaMemcpy(all.data(), dslots, sizeof(Slot) * W, cudaMemcpyDeviceToHost));
    return VGPU_OK;
  };
  int rc = gather();
  bool ok = rc == VGPU_OK;
  for (int r = 0; ok && r < W; ++r) ok = all[r].ok != 0;
  if (ok) {
    for (int r = 0; r < W; ++r) {
      if (r == comm->rank) { comm->peer[r] = comm->win; continue; }
      void* p = nullptr;
      if (cudaIpcOpenMemHandle(&p, all[r].handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
      comm->peer[r] = (char*)p;
    }
  }
  // second round: did everybody map everybody?
  mine.ok = ok ? 1 : 0;
  if (rc == VGPU_OK) rc = gather();
  for (int r = 0; rc == VGPU_OK && r < W; ++r) ok = ok && all[r].ok != 0;
  cudaFree(dslots);
  if (rc) { window_close(comm); return rc; }
  if (!ok) { window_close(comm); return VGPU_OK; }
  comm->win_bytes = bytes;
  comm->epoch = 0;
  if (mapped) *mapped = 1;
  return VGPU_OK;
}

// Layout of one half of a window: [kWinHeader bytes of flags][world regions].
static size_t half_bytes(const vgpu_comm_s* c) { return (c->win_bytes / kWinParts) & ~(size_t)255; } // one part (the name is from the two-part days)
static int64_t* dc_of(const vgpu_comm_s* c) { return c->dcounts + (size_t)(c->epoch % kWinParts) * kDcWords; } // the current exchange's message block
static size_t region_bytes(const vgpu_comm_s* c) { return ((half_bytes(c) - kWinHeader) / c->world) & ~(size_t)255; }
static char* region_of(const vgpu_comm_s* c, int owner_rank, int half, int source) {
  return c->peer[owner_rank] + (size_t)half * half_bytes(c) + kWinHeader + (size_t)source * region_bytes(c);
}

static int exchange_nccl(vgpu_comm_s* comm, Agg& pa, Agg& fa, cudaStream_t stream, int64_t* bytes_sent) {
  const size_t row_bytes = (3 + 2 * pa.accs.size()) * 8;
  const int W = comm->world;
  int rc;
  if (pa.mode != VGPU_AGG_HASH) {
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
  int64_t counts[32] = {0}, matrix[32 * 32];
  if ((rc = agg_partition(pa, W, comm->send, counts, stream))) return rc;
  VG_CUDA(cudaMemcpyAsync(dc_of(comm), counts, sizeof(int64_t) * W, cudaMemcpyHostToDevice, stream));
  int64_t* gathered = dc_of(comm) + kMsgWords + 32 * kMsgWords;
  VG_NCCL(g_nccl.allgather(dc_of(comm), gathered, sizeof(int64_t) * W, kNcclUint8, comm->comm, stream));
  VG_CUDA(cudaMemcpyAsync(matrix, gathered, sizeof(int64_t) * W * W, cudaMemcpyDeviceToHost, stream));
  VG_CUDA(cudaStreamSynchronize(stream));
  int64_t sc[32], so[32], rcv[32], ro[32], srun = 0, rrun = 0;
  for (int r = 0; r < W; ++r) { // matrix[src][dst]
    sc[r] = counts[r] * (int64_t)row_bytes; so[r] = srun; srun += sc[r];
    rcv[r] = matrix[r * W + comm->rank] * (int64_t)row_bytes; ro[r] = rrun; rrun += rcv[r];
  }
  if ((rc = comm_reserve(&comm->recv, &comm->recv_cap, (size_t)std::max<int64_t>(rrun, 8)))) return rc;
  if ((rc = vgpu_comm_alltoallv(comm, comm->send, sc, so, comm->recv, rcv, ro, stream))) return rc;
  if (bytes_sent) *bytes_sent = srun - sc[comm->rank];
  return agg_merge_rows(fa, comm->recv, rrun / (int64_t)row_bytes, stream);
}

// HASH, log engine on both sides, window mapped: partition + peer stores in one kernel, counts in one all-gather,
// deduplication at the receiver.
static size_t special_slot_bytes(int ew) { return ((size_t)2 * ew * 8 + 255) & ~(size_t)255; }
static uint64_t region_cap_entries(const vgpu_comm_s* comm, int ew) { return (region_bytes(comm) - special_slot_bytes(ew)) / ((size_t)ew * 8); }

static int log_p2p_send(vgpu_comm_s* comm, Agg& pa, cudaStream_t stream) {
  const int W = comm->world, ew = pa.ew;
  const int half = (int)(comm->epoch % kWinParts);
  const size_t special_bytes = special_slot_bytes(ew);
  int rc;
  const LogRange* ranges = nullptr;
  int n_ranges = 0;
  if ((rc = log_entry_ranges(pa, stream, &ranges, &n_ranges, nullptr))) return rc;
  SendArgs s{};
  s.ranges = ranges; s.n_ranges = n_ranges; s.ew = ew;
  uint32_t chunk = kSendThreads * kSendPerThread;
  while (chunk > kSendThreads && (size_t)chunk * ew * 8 > 64 * 1024) chunk -= kSendThreads;
  s.chunk = chunk;
  s.world = (uint32_t)W; s.rank = (uint32_t)comm->rank;
  s.velox_hash = agg_raw_bigint_key(pa) ? 1 : 0;
  for (int d = 0; d < W; ++d) s.peer_region[d] = (uint64_t*)(region_of(comm, d, half, comm->rank) + special_bytes);
  s.region_cap = region_cap_entries(comm, ew);
  s.cursor = (unsigned long long*)dc_of(comm);
  s.err = pa.err;
  {
    void* p = comm->chunk_prefix; size_t cap = comm->chunk_prefix_cap;
    if ((rc = comm_reserve(&p, &cap, ((size_t)n_ranges + 1) * sizeof(uint32_t)))) return rc;
    comm->chunk_prefix = (uint32_t*)p; comm->chunk_prefix_cap = cap;
  }
  s.chunk_prefix = comm->chunk_prefix;
  if ((rc = log_chunk_prefix(ranges, n_ranges, chunk, comm->chunk_prefix, stream))) return rc;
  VG_CUDA(cudaMemsetAsync(dc_of(comm), 0, sizeof(int64_t) * kMsgWords, stream));
  const size_t smem = (size_t)chunk * ew * 8;
  VG_CUDA(cudaFuncSetAttribute(xchg_send_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  xchg_send_kernel<<<num_sms() * 4, kSendThreads, smem, stream>>>(s);
  SpecialArgs sp{};
  sp.special = pa.special; sp.world = (uint32_t)W; sp.ew = ew; sp.velox_hash = agg_raw_bigint_key(pa) ? 1 : 0;
  for (int d = 0; d < W; ++d) sp.peer_slot[d] = (uint64_t*)region_of(comm, d, half, comm->rank);
  xchg_send_special_kernel<<<W, 64, 0, stream>>>(sp);
  xchg_meta_kernel<<<1, 32, 0, stream>>>((unsigned long long*)dc_of(comm), (uint32_t)W, pa.ctr, pa.log_probed && pa.ext_ranges.empty() ? 1 : 0,
                                         (pa.log_batches + 1) & 1);
  VG_CUDA(cudaGetLastError());
  count_launch(4);
  return VGPU_OK;
}

static int log_p2p_receive(vgpu_comm_s* comm, Agg& pa, Agg& fa, cudaStream_t stream) {
  const int W = comm->world, ew = pa.ew;
  const int half = (int)(comm->epoch % kWinParts);
  const size_t special_bytes = special_slot_bytes(ew);
  const uint64_t region_cap = region_cap_entries(comm, ew);
  int rc;
  if (fa.engine == kEngineUndecided) { if ((rc = log_adopt_layout(fa, pa, stream))) return rc; }
  if (fa.engine != kEngineLog || fa.ew != ew) { set_error("the final aggregation cannot take the partial stage's entries"); return VGPU_ERR_INVALID; }
  FoldSpecialArgs fs{};
  fs.special = fa.special; fs.kinds = log_cell_kinds(fa); fs.world = (uint32_t)W; fs.ew = ew; fs.err = fa.err;
  for (int src = 0; src < W; ++src) {
    fs.slot[src] = (const uint64_t*)region_of(comm, comm->rank, half, src);
    LogRange r{};
    r.base = (const uint64_t*)(region_of(comm, comm->rank, half, src) + special_bytes);
    r.cap = region_cap;
    const uint32_t* packed = reinterpret_cast<const uint32_t*>(dc_of(comm) + kMsgWords + (size_t)src * kMsgWords + comm->rank); // message[src][dst = me]
    r.count64 = 0;
    r.count = packed;       // low half: front entries
    r.back = 0;
    fa.ext_ranges.push_back(r);
    r.count = packed + 1;   // high half: back entries
    r.back = 1;
    fa.ext_ranges.push_back(r);
  }
  fa.ext_bound += region_cap * (uint64_t)W;
  fa.ext_meta = (const unsigned long long*)(dc_of(comm) + kMsgWords + kMaxWorld);
  fa.ext_sources = W;
  fa.ext_meta_stride = kMsgWords;
  fa.finalized = false;
  fa.packed_rows = -1;
  xchg_fold_special_kernel<<<1, 32, 0, stream>>>(fs);
  VG_CUDA(cudaGetLastError());
  count_launch();
  ++comm->epoch;
  return log_finalize(fa, stream);
}

// ---- side stream -------------------------------------------------------------------------------------------------------
static int comm_side_init(vgpu_comm_s* comm) {
  if (comm->side) return VGPU_OK;
  VG_CUDA(cudaStreamCreateWithFlags(&comm->side, cudaStreamNonBlocking));
  VG_CUDA(cudaEventCreateWithFlags(&comm->stored, cudaEventDisableTiming));
  VG_CUDA(cudaEventCreateWithFlags(&comm->folded, cudaEventDisableTiming));
  VG_CUDA(cudaEventCreateWithFlags(&comm->folded_prev, cudaEventDisableTiming));
  return VGPU_OK;
}
// The current exchange's last work was just issued on `stream` (the side stream, or the caller's for the paths that
// do not use one): remember it as "latest", the former latest as "the one before".
static int comm_tail_issued(vgpu_comm_s* comm, cudaStream_t stream) {
  int rc = comm_side_init(comm);
  if (rc) return rc;
  std::swap(comm->folded, comm->folded_prev);
  comm->folded_prev_valid = comm->folded_valid;
  VG_CUDA(cudaEventRecord(comm->folded, stream));
  comm->folded_valid = true;
  return VGPU_OK;
}

// ---- fused exchange (entries already left from inside the scan: fused_prelude.cuh send_entries) --------------------
static constexpr size_t kFusedCountsBytes = 1024 * sizeof(unsigned long long); // = vg::kSendCountsBytes (below)

// Sender: what is left to do after the scans — the count message, the dedicated groups, the stream facts.
static int log_fused_send(vgpu_comm_s* comm, Agg& pa, cudaStream_t stream) {
  const int W = comm->world, ew = pa.ew;
  const int half = (int)(comm->epoch % kWinParts);
  VG_CUDA(cudaMemsetAsync(dc_of(comm), 0, sizeof(int64_t) * kMsgWords, stream));
  xchg_send_totals_kernel<<<1, 256, 0, stream>>>((unsigned long long*)dc_of(comm), pa.send_counts, pa.send_grid, (uint32_t)W, pa.send_sub_cap);
  SpecialArgs sp{};
  sp.special = pa.special; sp.world = (uint32_t)W; sp.ew = ew; sp.velox_hash = agg_raw_bigint_key(pa) ? 1 : 0;
  for (int d = 0; d < W; ++d) sp.peer_slot[d] = (uint64_t*)region_of(comm, d, half, comm->rank);
  xchg_send_special_kernel<<<W, 64, 0, stream>>>(sp);
  xchg_meta_kernel<<<1, 32, 0, stream>>>((unsigned long long*)dc_of(comm), (uint32_t)W, pa.ctr, pa.log_probed ? 1 : 0, (pa.log_batches + 1) & 1);
  VG_CUDA(cudaGetLastError());
  count_launch(3);
  return VGPU_OK;
}

// Receiver: one front and one back range per source and sending CTA, counted by the words the senders left beside them.
static int log_fused_receive(vgpu_comm_s* comm, Agg& pa, Agg& fa, cudaStream_t stream) {
  const int W = comm->world, ew = pa.ew, grid = pa.send_grid;
  const int half = (int)(comm->epoch % kWinParts);
  const size_t special_bytes = special_slot_bytes(ew);
  const uint64_t sub_cap = pa.send_sub_cap;
  int rc;
  if (fa.engine == kEngineUndecided) { if ((rc = log_adopt_layout(fa, pa, stream))) return rc; }
  if (fa.engine != kEngineLog || fa.ew != ew) { set_error("the final aggregation cannot take the partial stage's entries"); return VGPU_ERR_INVALID; }
  FoldSpecialArgs fs{};
  fs.special = fa.special; fs.kinds = log_cell_kinds(fa); fs.world = (uint32_t)W; fs.ew = ew; fs.err = fa.err;
  fs.meta = (const unsigned long long*)(dc_of(comm) + kMsgWords + kMaxWorld);
  fs.meta_stride = kMsgWords;
  fs.expect_grid = (unsigned long long)grid; fs.expect_sub_cap = sub_cap;
  fa.ext_ranges.reserve(fa.ext_ranges.size() + (size_t)2 * W * grid);
  for (int src = 0; src < W; ++src) {
    const char* region = region_of(comm, comm->rank, half, src);
    fs.slot[src] = (const uint64_t*)region;
    const uint32_t* counts = reinterpret_cast<const uint32_t*>(region + special_bytes);
    const uint64_t* entries = reinterpret_cast<const uint64_t*>(region + special_bytes + kFusedCountsBytes);
    for (int c = 0; c < grid; ++c) {
      LogRange r{};
      r.base = entries + (uint64_t)c * sub_cap * ew;
      r.cap = sub_cap;
      r.count64 = 0;
      r.count = counts + 2 * c;      // low half: front entries
      r.back = 0;
      fa.ext_ranges.push_back(r);
      r.count = counts + 2 * c + 1;  // high half: back entries
      r.back = 1;
      fa.ext_ranges.push_back(r);
    }
  }
  fa.ext_bound += sub_cap * (uint64_t)grid * (uint64_t)W;
  fa.ext_meta = (const unsigned long long*)(dc_of(comm) + kMsgWords + kMaxWorld);
  fa.ext_sources = W;
  fa.ext_meta_stride = kMsgWords;
  fa.finalized = false;
  fa.packed_rows = -1;
  xchg_fold_special_kernel<<<1, 32, 0, stream>>>(fs);
  VG_CUDA(cudaGetLastError());
  count_launch();
  ++comm->epoch;
  return log_finalize(fa, stream);
}

// The caller's stream only issues the three small sender kernels; the count all-gather (the rendezvous with the peers),
// the receiver's fold and its deduplication run on the communicator's side stream, so the caller's next scan overlaps
// them.  That is why the window has THREE parts: scan e+1 stores into part (e+1) % 3 while the peers may still read
// parts (e-1) % 3 and e % 3 in the tails of exchanges e-1 and e.  Scan e+1 has to know that every peer finished reading
// part (e+1) % 3 = (e-2) % 3, i.e. its tail e-2: this rank's own tail e-1 is past its all-gather only when every peer
// has entered that all-gather on its side stream, behind its tail e-2 — so scan e+1 waits for this rank's tail e-1
// ("the one before the latest", send_bind).  The final aggregation remembers the tail's event (agg_wait_pending).
static int exchange_log_fused(vgpu_comm_s* comm, Agg& pa, Agg& fa, cudaStream_t stream, int64_t* bytes_sent) {
  static const bool serial = getenv("VGPU_EXCHANGE_OVERLAP") && getenv("VGPU_EXCHANGE_OVERLAP")[0] == '0'; // tuning aid: everything on the caller's stream
  int rc = comm_side_init(comm);
  if (rc) return rc;
  rc = log_fused_send(comm, pa, stream);
  if (rc) return rc;
  cudaStream_t tail = serial ? stream : comm->side;
  if (!serial) {
    VG_CUDA(cudaEventRecord(comm->stored, stream));
    VG_CUDA(cudaStreamWaitEvent(comm->side, comm->stored, 0));
  }
  // the barrier between "everybody's scan has stored" and "anybody reads", carrying the stream facts
  VG_NCCL(g_nccl.allgather(dc_of(comm), dc_of(comm) + kMsgWords, sizeof(int64_t) * kMsgWords, kNcclUint8, comm->comm, tail));
  if (bytes_sent) *bytes_sent = -1; // known on the device only; see vgpu_comm_last_counts
  rc = log_fused_receive(comm, pa, fa, tail);
  if (rc) return rc;
  rc = comm_tail_issued(comm, tail);
  if (rc) return rc;
  if (!serial) {
    if (!fa.pending) VG_CUDA(cudaEventCreateWithFlags(&fa.pending, cudaEventDisableTiming));
    VG_CUDA(cudaEventRecord(fa.pending, comm->side));
    fa.pending_set = true;
  }
  return VGPU_OK;
}

static int exchange_log_p2p(vgpu_comm_s* comm, Agg& pa, Agg& fa, cudaStream_t stream, int64_t* bytes_sent) {
  int rc = log_p2p_send(comm, pa, stream);
  if (rc) return rc;
  // the counts, and the barrier between "everybody has stored" and "anybody reads"
  VG_NCCL(g_nccl.allgather(dc_of(comm), dc_of(comm) + kMsgWords, sizeof(int64_t) * kMsgWords, kNcclUint8, comm->comm, stream));
  if (bytes_sent) *bytes_sent = -1; // known on the device only; see vgpu_comm_last_counts
  return log_p2p_receive(comm, pa, fa, stream);
}

// GLOBAL / ARRAY through the window.
static int dense_p2p(vgpu_comm_s* comm, Agg& pa, Agg& fa, cudaStream_t stream, int phase, bool advance) {
  const int W = comm->world;
  const int half = (int)(comm->epoch % kWinParts);
  DenseArgs d{};
  d.state = pa.state; d.final_state = fa.state; d.G = pa.n_groups; d.n_cells = pa.n_cells;
  for (int c = 0; c < pa.n_cells; ++c) d.kind[c] = pa.cells[c].kind;
  d.world = (uint32_t)W; d.rank = (uint32_t)comm->rank;
  for (int r = 0; r < W; ++r) {
    d.peer_data[r] = (uint64_t*)region_of(comm, r, half, comm->rank);
    d.peer_flag[r] = (unsigned long long*)(comm->peer[r] + (size_t)half * half_bytes(comm) + (size_t)comm->rank * 64);
    d.my_data[r] = (const uint64_t*)region_of(comm, comm->rank, half, r);
    d.my_flag[r] = (const unsigned long long*)(comm->win + (size_t)half * half_bytes(comm) + (size_t)r * 64);
  }
  d.epoch = comm->epoch / kWinParts + 1; // per part: strictly increasing
  d.err = fa.err;
  d.phase = phase;
  // the waiting half is one small CTA that must fit beside the resident CTAs of a scan running on another stream: it
  // asks for the same (maximum) shared-memory carve-out as the scans, otherwise the SM it sits on cannot take a scan CTA
  // until it exits (measured at N = 8: the scan kernel 0.38 -> 0.44 ms)
  static bool carveout_set = false;
  if (!carveout_set) { cudaFuncSetAttribute(xchg_dense_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared); cudaGetLastError(); carveout_set = true; }
  xchg_dense_kernel<<<1, phase == 2 ? 64 : 512, 0, stream>>>(d);
  VG_CUDA(cudaGetLastError());
  count_launch();
  if (advance) ++comm->epoch;
  fa.packed_rows = -1;
  return VGPU_OK;
}
// Two launches: store + flag on the caller's stream, wait + fold on the communicator's side stream.  The caller's stream
// is free again after the store, so the next batch's scan overlaps the rendezvous with the peers (and absorbs the skew
// between ranks); the final aggregation remembers the fold's event and its next operation waits for it (agg_wait_pending).
// Window halves stay safe: store(j) waits for this rank's fold(j-1), so a peer's flag(j) implies that peer finished
// reading half (j-1) & 1, and this rank's store(j+1) — after its own fold(j), which saw every peer's flag(j) — may
// overwrite it.
static int exchange_dense_p2p(vgpu_comm_s* comm, Agg& pa, Agg& fa, cudaStream_t stream, int64_t* bytes_sent) {
  if (bytes_sent) *bytes_sent = (int64_t)((size_t)pa.n_groups * pa.n_cells * 8 * (comm->world - 1));
  static const bool serial = getenv("VGPU_DENSE_OVERLAP") && getenv("VGPU_DENSE_OVERLAP")[0] == '0'; // tuning aid: one launch, no side stream
  if (serial) { const int src = dense_p2p(comm, pa, fa, stream, 3, true); return src ? src : comm_tail_issued(comm, stream); }
  { const int irc = comm_side_init(comm); if (irc) return irc; }
  if (!fa.pending) VG_CUDA(cudaEventCreateWithFlags(&fa.pending, cudaEventDisableTiming));
  int rc = dense_p2p(comm, pa, fa, stream, 1, false);
  if (rc) return rc;
  VG_CUDA(cudaEventRecord(comm->stored, stream));
  VG_CUDA(cudaStreamWaitEvent(comm->side, comm->stored, 0));
  rc = dense_p2p(comm, pa, fa, comm->side, 2, true);
  if (rc) return rc;
  rc = comm_tail_issued(comm, comm->side);
  if (rc) return rc;
  VG_CUDA(cudaEventRecord(fa.pending, comm->side));
  fa.pending_set = true;
  return VGPU_OK;
}

static bool dense_fits(const vgpu_comm_s* comm, const Agg& pa) {
  return (size_t)pa.n_groups * pa.n_cells * 8 <= region_bytes(comm) && pa.n_groups <= 4096;
}
static bool log_fits(const vgpu_comm_s* comm, const Agg& pa, const Agg& fa) {
  return pa.engine == kEngineLog && fa.engine != kEngineTable && region_bytes(comm) > special_slot_bytes(pa.ew) + (size_t)pa.ew * 8 * 1024;
}

int vgpu_comm_exchange_groups(vgpu_comm_t comm, vgpu_agg_t partial_agg, vgpu_agg_t final_agg, void* stream_, int64_t* bytes_sent) {
  if (!comm || !partial_agg || !final_agg) return VGPU_ERR_INVALID;
  cudaStream_t stream = (cudaStream_t)stream_;
  Agg& pa = partial_agg->agg;
  Agg& fa = final_agg->agg;
  if (pa.mode != fa.mode || pa.accs.size() != fa.accs.size()) { set_error("partial and final aggregations differ"); return VGPU_ERR_INVALID; }
  if (!pa.resolved || !fa.resolved) { set_error("both aggregations must have seen an (empty) vgpu_agg_add_input"); return VGPU_ERR_INVALID; }
  if (!comm->dcounts) VG_CUDA(cudaMalloc(&comm->dcounts, sizeof(int64_t) * kDcWords * kWinParts));
  { int wrc = agg_wait_pending(pa, stream); if (wrc) return wrc; wrc = agg_wait_pending(fa, stream); if (wrc) return wrc; }
  if (pa.mode == VGPU_AGG_HASH && pa.send_used) {
    if (pa.send_comm != comm) { set_error("the partial aggregation's entries left through another exchange"); return VGPU_ERR_INVALID; }
    if (pa.send_epoch != comm->epoch) { set_error("another exchange ran through this communicator since the partial aggregation's scan"); return VGPU_ERR_INVALID; }
    return exchange_log_fused(comm, pa, fa, stream, bytes_sent);
  }
  // every other path starts behind the latest exchange's tail, which may still run on the side stream (exchange_dense_p2p,
  // exchange_log_fused), and leaves its own end as the new "latest"
  if (comm->folded_valid) VG_CUDA(cudaStreamWaitEvent(stream, comm->folded, 0));
  auto done = [&](int rc) -> int { if (rc) return rc; return comm_tail_issued(comm, stream); };
  const bool no_p2p = getenv("VGPU_EXCHANGE") && strcmp(getenv("VGPU_EXCHANGE"), "nccl") == 0;
  if (pa.mode != VGPU_AGG_HASH) {
    if (pa.n_groups != fa.n_groups || pa.n_cells != fa.n_cells) { set_error("partial and final aggregations differ in their key domains"); return VGPU_ERR_INVALID; }
    // a global aggregation's single row (present even without input) is reported by the rank that owns group 0 only
    if (pa.mode == VGPU_AGG_GLOBAL) fa.emit_empty_global = (int)partition_of(0, 0, (uint32_t)comm->world, false) == comm->rank;
    if (comm->win && !no_p2p && dense_fits(comm, pa)) return exchange_dense_p2p(comm, pa, fa, stream, bytes_sent);
    return done(exchange_nccl(comm, pa, fa, stream, bytes_sent));
  }
  if (comm->win && !no_p2p && log_fits(comm, pa, fa)) return done(exchange_log_p2p(comm, pa, fa, stream, bytes_sent));
  return done(exchange_nccl(comm, pa, fa, stream, bytes_sent));
}

int vgpu_debug_local_comms(int32_t world, size_t window_bytes, vgpu_comm_t* out) {
  if (!out || world < 1 || world > kMaxWorld) return VGPU_ERR_INVALID;
  window_bytes = (window_bytes + kWinParts * kWinHeader + 4095) & ~(size_t)4095;
  int rc = VGPU_OK;
  for (int r = 0; r < world; ++r) out[r] = nullptr;
  for (int r = 0; r < world && rc == VGPU_OK; ++r) {
    vgpu_comm_s* c = new vgpu_comm_s();
    out[r] = c;
    c->comm = nullptr; c->rank = r; c->world = world; c->local = true;
    if (cudaMalloc(&c->win, window_bytes) != cudaSuccess || cudaMemset(c->win, 0, window_bytes) != cudaSuccess ||
        cudaMalloc(&c->dcounts, sizeof(int64_t) * kDcWords * kWinParts) != cudaSuccess) rc = cuda_fail(cudaGetLastError(), "local window");
    c->win_bytes = window_bytes;
  }
  if (rc) { for (int r = 0; r < world; ++r) { if (out[r]) vgpu_comm_destroy(out[r]); out[r] = nullptr; } return rc; }
  for (int r = 0; r < world; ++r) for (int q = 0; q < world; ++q) out[r]->peer[q] = out[q]->win;
  return VGPU_OK;
}

int vgpu_debug_exchange_local_comms(vgpu_comm_t* comms, vgpu_agg_t* partials, vgpu_agg_t* finals, int32_t world, void* stream_) {
  if (!comms || !partials || !finals || world < 1 || world > kMaxWorld) return VGPU_ERR_INVALID;
  cudaStream_t stream = (cudaStream_t)stream_;
  for (int r = 0; r < world; ++r) {
    if (!comms[r] || !comms[r]->local || comms[r]->world != world || !partials[r] || !finals[r]) { set_error("rank %d: not a local exchange of %d ranks", r, world); return VGPU_ERR_INVALID; }
    Agg& pa = partials[r]->agg;
    Agg& fa = finals[r]->agg;
    if (!pa.resolved || !fa.resolved || pa.mode != fa.mode) { set_error("rank %d: aggregations not ready", r); return VGPU_ERR_INVALID; }
    if (pa.send_used) { if (pa.send_comm != comms[r]) { set_error("rank %d: entries left through another exchange", r); return VGPU_ERR_INVALID; } continue; }
    if (pa.mode == VGPU_AGG_HASH ? !log_fits(comms[r], pa, fa) : !dense_fits(comms[r], pa)) { set_error("rank %d: the window path does not apply", r); return VGPU_ERR_UNSUPPORTED; }
  }
  const bool hash = partials[0]->agg.mode == VGPU_AGG_HASH;
  for (int r = 0; r < world; ++r) {
    Agg& pa = partials[r]->agg;
    int rc = !hash ? dense_p2p(comms[r], pa, finals[r]->agg, stream, 1, false) : pa.send_used ? log_fused_send(comms[r], pa, stream) : log_p2p_send(comms[r], pa, stream);
    if (rc) return rc;
  }
  if (hash) { // the all-gather of the counts
    for (int r = 0; r < world; ++r) for (int src = 0; src < world; ++src)
      VG_CUDA(cudaMemcpyAsync(dc_of(comms[r]) + kMsgWords + (size_t)src * kMsgWords, dc_of(comms[src]), sizeof(int64_t) * kMsgWords, cudaMemcpyDeviceToDevice, stream));
  }
  for (int r = 0; r < world; ++r) {
    Agg& pa = partials[r]->agg;
    Agg& fa = finals[r]->agg;
    if (pa.mode == VGPU_AGG_GLOBAL) fa.emit_empty_global = (int)partition_of(0, 0, (uint32_t)world, false) == r;
    int rc = !hash ? dense_p2p(comms[r], pa, fa, stream, 2, true) : pa.send_used ? log_fused_receive(comms[r], pa, fa, stream) : log_p2p_receive(comms[r], pa, fa, stream);
    if (rc) return rc;
  }
  return VGPU_OK;
}

// All `world` ranks of an exchange emulated on the CURRENT device, for boxes with fewer GPUs than ranks (tests,
// smoke()): rank r owns partials[r] / finals[r]; the windows are plain local allocations, the count all-gather is a
// device copy, and the phases run rank after rank (no kernel waits for another one).  The kernels, the region layout,
// the ownership function and the receiver's deduplication are the ones vgpu_comm_exchange_groups uses.
int vgpu_debug_exchange_local(vgpu_agg_t* partials, vgpu_agg_t* finals, int32_t world, size_t window_bytes, void* stream_) {
  if (!partials || !finals || world < 1 || world > kMaxWorld) return VGPU_ERR_INVALID;
  cudaStream_t stream = (cudaStream_t)stream_;
  std::vector<vgpu_comm_t> comms(world, nullptr);
  int rc = vgpu_debug_local_comms(world, window_bytes, comms.data());
  if (rc) return rc;
  rc = vgpu_debug_exchange_local_comms(comms.data(), partials, finals, world, stream_);
  // the final aggregations must not keep pointers into the temporary windows: materialise their groups now
  for (int r = 0; r < world && rc == VGPU_OK; ++r) if (finals[r]->agg.mode == VGPU_AGG_HASH) rc = agg_extract_all(finals[r]->agg, stream);
  cudaStreamSynchronize(stream);
  for (int r = 0; r < world; ++r) {
    if (partials[r] && finals[r]) { Agg& fa = finals[r]->agg; fa.ext_ranges.clear(); fa.ext_bound = 0; fa.ext_meta = nullptr; fa.ext_sources = 0; }
    vgpu_comm_destroy(comms[r]);
  }
  return rc;
}

int vgpu_agg_attach_exchange(vgpu_agg_t agg, vgpu_comm_t comm) {
  if (!agg) return VGPU_ERR_INVALID;
  Agg& a = agg->agg;
  if (a.send_used) { set_error("entries of the current state already left through an exchange: reset the aggregation first"); return VGPU_ERR_INVALID; }
  if (comm && a.mode != VGPU_AGG_HASH) { set_error("only HASH aggregations stream their entries (ARRAY / GLOBAL states are exchanged whole)"); return VGPU_ERR_UNSUPPORTED; }
  a.send_comm = comm;
  return VGPU_OK;
}

// Entries this rank sent to each destination in the last window exchange of a HASH aggregation (and the whole
// [src][dst] matrix): synchronises the stream.
int vgpu_comm_last_counts(vgpu_comm_t comm, int64_t* matrix, void* stream) {
  if (!comm || !matrix || !comm->dcounts) return VGPU_ERR_INVALID;
  const int W = comm->world;
  std::vector<int64_t> msgs((size_t)W * kMsgWords);
  if (comm->side) VG_CUDA(cudaStreamSynchronize(comm->side));
  const int64_t* last = comm->dcounts + (size_t)((comm->epoch + kWinParts - 1) % kWinParts) * kDcWords; // the exchange that just ran
  VG_CUDA(cudaMemcpyAsync(msgs.data(), last + kMsgWords, sizeof(int64_t) * W * kMsgWords, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  VG_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  for (int s = 0; s < W; ++s) for (int d = 0; d < W; ++d) {
    const uint64_t w = (uint64_t)msgs[(size_t)s * kMsgWords + d];
    matrix[s * W + d] = (int64_t)((w & 0xffffffffull) + (w >> 32));
  }
  return VGPU_OK;
}

} // extern "C"

// ---- the exchange fused into the scan ------------------------------------------------------------------------------
// A partial aggregation with an attached exchange does not log its entries locally: the LOG-sink kernel stores each
// entry into the window of the rank that owns its key while it scans (fused_prelude.cuh send_entries), so the NVLink
// transfer overlaps the scan instead of following it.  Layout of the region source s has in owner o's window:
//   [special slots][kSendCountsBytes: one packed count per sending CTA][grid sub-regions of sub_cap entries]
// Sub-region c belongs to CTA c of the sender: front entries from its start, back entries from its end.
namespace vg {
namespace {
constexpr int kSendMaxGrid = 1024;
constexpr size_t kSendCountsBytes = (size_t)kSendMaxGrid * sizeof(unsigned long long);
static_assert(kSendCountsBytes == kFusedCountsBytes, "one layout");
constexpr uint64_t kSendMinSubCap = 1024; // of which fused_prelude.cuh kSendBackRoom (256) are kept for back entries

uint64_t send_sub_cap_for(const vgpu_comm_s* comm, int ew, int grid) {
  const size_t head = special_slot_bytes(ew) + kSendCountsBytes;
  if (region_bytes(comm) <= head) return 0;
  return ((region_bytes(comm) - head) / ((size_t)ew * 8) / (size_t)grid) & ~(size_t)15; // whole runs of 16 entries
}
} // namespace

bool send_wanted(Agg& a, PipelineParams& p) {
  p.send_world = 0; p.send_hash = 0;
  if (fused_dry_run()) { // ahead-of-time plans: "world,hash" names the variant to generate
    const char* e = getenv("VGPU_DRY_SEND");
    int w = 0, h = 0;
    if (e && sscanf(e, "%d,%d", &w, &h) == 2 && w >= 1 && w <= 8) { p.send_world = w; p.send_hash = h ? 1 : 0; return true; }
    return false;
  }
  vgpu_comm_s* c = a.send_comm;
  if (!c || !c->win || c->world > 8) return false;
  static const bool off = getenv("VGPU_FUSED_EXCHANGE") && getenv("VGPU_FUSED_EXCHANGE")[0] == '0'; // tuning aid
  if (off) return false;
  if (!a.slabs.empty() || !a.ext_ranges.empty()) return false; // entries were already logged locally: stay local
  p.send_world = c->world;
  p.send_hash = agg_raw_bigint_key(a) ? 1 : 0;
  return true;
}

int send_bind(Agg& a, PipelineParams& p, int grid, cudaStream_t stream) {
  vgpu_comm_s* c = a.send_comm;
  const int W = c->world, ew = a.ew;
  if (grid > kSendMaxGrid) { set_error("fused exchange: %d CTAs exceed the %d count slots of a window region", grid, kSendMaxGrid); return VGPU_ERR_UNSUPPORTED; }
  const uint64_t sub_cap = send_sub_cap_for(c, ew, grid);
  if (sub_cap < kSendMinSubCap) { set_error("fused exchange: the window is too small (vgpu_comm_reserve_window)"); return VGPU_ERR_UNSUPPORTED; }
  if (a.send_used && (grid != a.send_grid || sub_cap != a.send_sub_cap)) {
    set_error("fused exchange: the launch shape changed between batches (%d -> %d CTAs)", a.send_grid, grid);
    return VGPU_ERR_UNSUPPORTED;
  }
  const size_t words = (size_t)grid * W;
  if (words > a.send_counts_cap) {
    if (a.send_counts) { VG_CUDA(cudaStreamSynchronize(stream)); VG_CUDA(cudaFree(a.send_counts)); a.send_counts = nullptr; a.send_counts_cap = 0; }
    VG_CUDA(cudaMalloc(&a.send_counts, words * sizeof(unsigned long long)));
    a.send_counts_cap = words;
  }
  if (a.send_used && a.send_epoch != c->epoch) {
    set_error("fused exchange: another exchange ran through this communicator between two batches of the aggregation");
    return VGPU_ERR_INVALID;
  }
  if (!a.send_used) { VG_CUDA(cudaMemsetAsync(a.send_counts, 0, words * sizeof(unsigned long long), stream)); a.send_epoch = c->epoch; }
  a.send_grid = grid;
  a.send_sub_cap = sub_cap;
  // the part this scan stores into was read by the peers in the tail of the exchange three back: every peer is past
  // that once this rank's tail "before the latest" is complete (exchange_log_fused)
  if (c->folded_prev_valid) VG_CUDA(cudaStreamWaitEvent(stream, c->folded_prev, 0));
  const int half = (int)(c->epoch % kWinParts);
  const size_t special_bytes = special_slot_bytes(ew);
  p.send_sub_cap = sub_cap;
  p.send_counts = a.send_counts;
  for (int d = 0; d < 8; ++d) { p.send_base[d] = nullptr; p.send_peer_counts[d] = nullptr; }
  for (int d = 0; d < W; ++d) {
    char* region = region_of(c, d, half, c->rank);
    p.send_peer_counts[d] = (unsigned long long*)(region + special_bytes);
    p.send_base[d] = (uint64_t*)(region + special_bytes + kSendCountsBytes);
  }
  return VGPU_OK;
}

void send_reset(Agg& a) { a.send_used = false; }
void send_free(Agg& a) {
  cudaFree(a.send_counts);
  a.send_counts = nullptr; a.send_counts_cap = 0; a.send_comm = nullptr; a.send_used = false;
}
} // namespace vg
