// Small RAII layer over the C ABI for the operators: status -> exception, device buffers, and the
// device image of the input columns a program reads (the DecodedVector triple per column,
// velox/vector/DecodedVector.cpp:60-187).
#pragma once
#include "GpuOperators.h"

namespace vb200 {
namespace gpu {

struct DeviceBatch {
  std::vector<vgpu_column> cols;
  std::vector<void*> owned;
  void* stream = nullptr;
  ~DeviceBatch() { release(); }
  void release() {
    for (void* p : owned) vgpu_free(p);
    owned.clear();
    cols.clear();
  }
  void* upload(const void* host, size_t bytes) {
    void* d = nullptr;
    check(vgpu_alloc(bytes ? bytes : 8, &d), "vgpu_alloc");
    owned.push_back(d);
    if (bytes) check(vgpu_h2d_async(d, host, bytes, stream), "vgpu_h2d_async");
    return d;
  }
  void* scratch(size_t bytes, bool zero) {
    void* d = nullptr;
    check(vgpu_alloc(bytes ? bytes : 8, &d), "vgpu_alloc");
    owned.push_back(d);
    if (zero) check(vgpu_memset_async(d, 0, bytes ? bytes : 8, stream), "vgpu_memset_async");
    return d;
  }
};

size_t deviceWidth(int vgpuType); // bytes per value (0 for bit-packed BOOLEAN)

// Appends the device image of `v` (numeric / boolean, any stack of dictionary / constant wrappers) to `b`.
// `idsOfBase` (optional): replaces the base values by these int32 value ids (VARCHAR keys).
void bindColumn(DeviceBatch& b, const VectorPtr& v, const std::vector<int32_t>* idsOfBase = nullptr);

// Flattened view of a vector: base flat vector + per-row index (or -1 for "row itself") + null test.
struct Decoded {
  const BaseVector* base = nullptr;
  std::vector<int32_t> indices;  // empty = identity
  bool constant = false;         // every row reads base[0]
  bool constantNull = false;
  std::vector<uint64_t> nulls;   // combined OUTER nulls from the wrappers (empty = none); base nulls live in base
  size_t size = 0;
};
Decoded decode(const VectorPtr& v);

} // namespace gpu
} // namespace vb200
