// vgpu_selection_indices_noalloc: selection bitmap -> selected row numbers without any allocation.
//
// vgpu_filter_project turns its selection bitmap into the index list (the SelectivityVector -> indices step of
// velox/exec/FilterProject.cpp:125-174) with two follow-up kernels that need scratch memory, the per-tile selected-row
// counts and their exclusive prefix, which it keeps in a grow-only cudaMalloc'ed buffer.  Some callers cannot
// allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): for them this file
// holds the same step with its scratch in a file-scope __device__ array of a fixed, modest size.
//
// - Every GPU that loads the library holds the array (kScratchBytes, ~192 KiB) until the process exits.
// - All calls on a device share the array and it keeps the previous call's contents.  Nothing in it has to start at
//   zero: every call overwrites the counts and offsets of all tiles it reads before reading them.  Calls that overlap
//   on the GPU (different streams, no ordering between them) race on it: order them.
// - In host code the array's NAME is a host-side stand-in, not a device address; the address kernels get comes from
//   cudaGetSymbolAddress, looked up once per device with that device current.
// - No cudaMalloc / cudaFree / allocating library call on this path.
#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>

#include "internal.h"

namespace vg {
namespace {

constexpr int64_t kNoAllocMaxRows = VGPU_NOALLOC_MAX_ROWS;           // 16 Mi rows per call
constexpr int64_t kNoAllocMaxTiles = kNoAllocMaxRows / kTile;        // 16384 tiles of 1024 rows
constexpr size_t kOffsetsBytes = (size_t)kNoAllocMaxTiles * sizeof(int64_t);
constexpr size_t kCountsBytes = (size_t)kNoAllocMaxTiles * sizeof(int32_t);
constexpr size_t kScratchBytes = kOffsetsBytes + kCountsBytes;       // 192 KiB
static_assert(kNoAllocMaxRows % kTile == 0, "the row limit is a whole number of tiles");

// [kNoAllocMaxTiles] int64 tile offsets, then [kNoAllocMaxTiles] int32 tile counts
static __device__ __align__(256) unsigned char g_sel_scratch[kScratchBytes];

// One warp per tile of 1024 rows (32 words of 32 bits): the tile's selected-row count.  Bits past n_rows do not count.
__global__ void noalloc_count_tiles(const uint32_t* words, int64_t n_rows, int32_t* tile_counts) {
  const int lane = threadIdx.x & 31;
  const int64_t tile = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t n_tiles = (n_rows + kTile - 1) / kTile;
  if (tile >= n_tiles) return;
  const int64_t n_words = (n_rows + 31) >> 5;
  const int64_t wi = tile * 32 + lane;
  uint32_t w = wi < n_words ? words[wi] : 0u;
  const int64_t rem = n_rows - wi * 32;
  if (rem < 32) w &= rem <= 0 ? 0u : ((1u << rem) - 1u);
  int c = __popc(w);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane == 0) tile_counts[tile] = c;
}

// Exclusive prefix of the tile counts (one CTA) and the total.
__global__ void noalloc_scan_tiles(const int32_t* counts, int64_t n_tiles, int64_t* offsets, unsigned long long* total) {
  __shared__ int64_t warp_sums[32];
  __shared__ int64_t carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < n_tiles; base += blockDim.x) {
    const int64_t i = base + threadIdx.x;
    const int64_t v = i < n_tiles ? counts[i] : 0;
    int64_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int64_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) warp_sums[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int64_t w = lane < (int)(blockDim.x >> 5) ? warp_sums[lane] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int64_t y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
      }
      warp_sums[lane] = w;
    }
    __syncthreads();
    const int64_t before = carry + (warp ? warp_sums[warp - 1] : 0) + x - v;
    if (i < n_tiles) offsets[i] = before;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = (unsigned long long)carry;
}

// One warp per tile: lane w expands word w of the tile behind the lanes before it.
__global__ void noalloc_expand_tiles(const uint32_t* words, const int64_t* offsets, int64_t n_rows, int32_t* sel_idx) {
  const int lane = threadIdx.x & 31;
  const int64_t tile = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t n_tiles = (n_rows + kTile - 1) / kTile;
  if (tile >= n_tiles) return;
  const int64_t n_words = (n_rows + 31) >> 5;
  const int64_t wi = tile * 32 + lane;
  uint32_t w = wi < n_words ? words[wi] : 0u;
  const int64_t rem = n_rows - wi * 32;
  if (rem < 32) w &= rem <= 0 ? 0u : ((1u << rem) - 1u);
  const int c = __popc(w);
  int x = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  int64_t out = offsets[tile] + (x - c);
  while (w) {
    const int b = __ffs(w) - 1;
    w &= w - 1;
    sel_idx[out++] = (int32_t)(wi * 32 + b);
  }
}

// Device address of g_sel_scratch on the current device: looked up once per device (with that device current) and kept.
constexpr int kMaxDevices = 64;
std::atomic<unsigned char*> g_scratch_addr[kMaxDevices];

int scratch_address(unsigned char** out) {
  int dev = -1;
  VG_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) { set_error("device ordinal %d outside the supported range", dev); return VGPU_ERR_UNSUPPORTED; }
  unsigned char* p = g_scratch_addr[dev].load(std::memory_order_acquire);
  if (!p) {
    void* sym = nullptr;
    VG_CUDA(cudaGetSymbolAddress(&sym, g_sel_scratch));
    if (!sym) { set_error("cudaGetSymbolAddress returned no address for the selection scratch"); return VGPU_ERR_CUDA; }
    p = static_cast<unsigned char*>(sym);
    g_scratch_addr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return VGPU_OK;
}

} // namespace
} // namespace vg

using namespace vg;

extern "C" int vgpu_selection_indices_noalloc(const uint64_t* sel_bits, int64_t nrows, int32_t* sel_idx, int64_t* num_selected,
                                              void* stream_) {
  if (!sel_bits || !sel_idx || !num_selected) { set_error("null argument"); return VGPU_ERR_INVALID; }
  if (nrows < 0) { set_error("negative row count"); return VGPU_ERR_INVALID; }
  if (nrows > kNoAllocMaxRows) {
    set_error("vgpu_selection_indices_noalloc handles at most %lld rows per call (fixed scratch), got %lld: split the batch or use vgpu_filter_project",
              (long long)kNoAllocMaxRows, (long long)nrows);
    return VGPU_ERR_UNSUPPORTED;
  }
  cudaStream_t stream = (cudaStream_t)stream_;
  unsigned char* scratch = nullptr;
  int rc = scratch_address(&scratch);
  if (rc) return rc;
  int64_t* tile_offsets = reinterpret_cast<int64_t*>(scratch);
  int32_t* tile_counts = reinterpret_cast<int32_t*>(scratch + kOffsetsBytes);
  const uint32_t* words = reinterpret_cast<const uint32_t*>(sel_bits);
  const int64_t n_tiles = (nrows + kTile - 1) / kTile;
  const int warps_per_block = 8;
  const int blocks = (int)((n_tiles + warps_per_block - 1) / warps_per_block);
  if (n_tiles > 0) {
    noalloc_count_tiles<<<blocks, warps_per_block * 32, 0, stream>>>(words, nrows, tile_counts);
    VG_CUDA(cudaGetLastError());
  }
  noalloc_scan_tiles<<<1, 1024, 0, stream>>>(tile_counts, n_tiles, tile_offsets, reinterpret_cast<unsigned long long*>(num_selected));
  VG_CUDA(cudaGetLastError());
  if (n_tiles > 0) {
    noalloc_expand_tiles<<<blocks, warps_per_block * 32, 0, stream>>>(words, tile_offsets, nrows, sel_idx);
    VG_CUDA(cudaGetLastError());
  }
  count_launch(n_tiles > 0 ? 3 : 1);
  return VGPU_OK;
}
