// Small RAII layer over the C ABI for the operators: status -> exception, the host -> device staging ring, and the
// device image of the input columns a program reads (the DecodedVector triple per column,
// velox/vector/DecodedVector.cpp:60-187).
#pragma once
#include "GpuOperators.h"

namespace vb200 {
namespace gpu {

// Vectors built after usePinnedHostMemory(true) keep their value arrays in page-locked memory (Vectors.h
// memory::hostAllocator): the staging ring then copies them to HBM with one DMA per column, straight out of the vector.
void usePinnedHostMemory(bool on);

// Host -> device staging of one operator.  kSlots slots are used round-robin; a slot is a device arena plus a pinned host
// area of the same size.  Uploads run on a copy stream of their own, so the copies of the batch being gathered overlap
// the kernels of the previous one; events order the two streams without blocking the host:
//   copy stream   waits for the slot's "free" event (recorded on the compute stream after the kernels that read it),
//   compute stream waits for the slot's "ready" event (recorded on the copy stream after the batch's last copy).
// The host only blocks when it is a full ring ahead of the GPU (back-pressure), never per batch — this replaces the
// cudaMalloc / pageable cudaMemcpy / stream synchronise per addInput of round 1.
class Staging {
 public:
  static constexpr int kSlots = 3;
  explicit Staging(void* computeStream);
  ~Staging();
  Staging(const Staging&) = delete;
  Staging& operator=(const Staging&) = delete;

  // Starts the next slot with at least `bytes` of device (and pinned) room; waits until its previous user is done.
  void begin(size_t bytes);
  bool active() const { return active_; }
  size_t room() const { return active_ ? slots_[cur_].cap - slots_[cur_].used : 0; }
  // Device memory inside the current slot (256-byte aligned).  Throws when the slot is full.
  void* reserve(size_t bytes);
  // Copies host memory into slot memory.  Page-locked sources are copied directly (asynchronously: `owner` keeps the
  // source alive until the slot is reused; without an owner the source is staged like pageable memory); pageable
  // sources go through the slot's pinned area.
  void copyIn(void* dst, const void* host, size_t bytes, const std::shared_ptr<const void>& owner = nullptr);
  void* upload(const void* host, size_t bytes, const std::shared_ptr<const void>& owner = nullptr) {
    void* d = reserve(bytes);
    if (bytes) copyIn(d, host, bytes, owner);
    return d;
  }
  void* scratch(size_t bytes, bool zero);
  // The pinned host address that mirrors device address `dev` of the current slot (same offset in the slot's pinned
  // area), and the copy of `bytes` from there to `dev`: callers gather many small batches on the host, then move
  // them with one DMA.  Mirror space and the staged copies of copyIn share the pinned area: copyIn stages behind
  // everything reserved so far.
  void* mirror(void* dev);
  void copyMirror(void* dev, size_t bytes);
  // All copies of the slot are issued: the compute stream will wait for them.
  void ready();
  // The kernels reading the slot are launched: the slot may be reused once they finish.
  void release();
  void* copyStream() const { return copy_; }
  int64_t h2dBytes = 0, stagedBytes = 0;

 private:
  struct Slot {
    char* dev = nullptr;
    char* pin = nullptr;
    size_t cap = 0, used = 0, pinUsed = 0;
    void* freeEv = nullptr;
    void* readyEv = nullptr;
    bool inFlight = false;
    std::vector<std::shared_ptr<const void>> keep;
  };
  Slot slots_[kSlots];
  int cur_ = -1;
  bool active_ = false;
  void* compute_ = nullptr;
  void* copy_ = nullptr;
};

struct DeviceBatch {
  std::vector<vgpu_column> cols;
  std::vector<void*> owned;
  void* stream = nullptr;
  Staging* staging = nullptr; // when set, memory comes from the staging slot and copies run on its copy stream
  ~DeviceBatch() { release(); }
  void release() {
    for (void* p : owned) vgpu_free(p);
    owned.clear();
    cols.clear();
  }
  void* upload(const void* host, size_t bytes) {
    if (staging) return staging->upload(host, bytes);
    void* d = nullptr;
    check(vgpu_alloc(bytes ? bytes : 8, &d), "vgpu_alloc");
    owned.push_back(d);
    if (bytes) check(vgpu_h2d_async(d, host, bytes, stream), "vgpu_h2d_async");
    return d;
  }
  void* scratch(size_t bytes, bool zero) {
    if (staging) return staging->scratch(bytes, zero);
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
// `valuesOverride` (optional): a flat BOOLEAN or BIGINT vector of the base's length whose values replace the base's
// (string predicates evaluated per base value; VARCHAR keys mapped to numbers on the host); nulls, indices and wrappers
// stay those of `v`.
// `stringViews`: the base is a StringViewVector and travels as VGPU_VARCHAR (its 16-byte views, read on the device as
// the stringAsNumber value of strings of at most 7 bytes).
void bindColumn(DeviceBatch& b, const VectorPtr& v, const std::vector<int32_t>* idsOfBase = nullptr, const BaseVector* valuesOverride = nullptr,
                bool stringViews = false);
// Upper bound of the device bytes bindColumn needs for `v`.
size_t bindBytes(const VectorPtr& v);

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
// Raw value array of a flat vector (bit words for BOOLEAN).
const void* rawValuesOf(const BaseVector* flat);

} // namespace gpu
} // namespace vb200
