// Host -> device staging ring and the pinned host allocator of the GPU operators (GpuDevice.h).
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "GpuDevice.h"

namespace vb200 {

namespace memory {
namespace {
void* mallocAlloc(size_t bytes) { return std::malloc(bytes ? bytes : 1); }
void mallocFree(void* p) { std::free(p); }
void* pinnedAlloc(size_t bytes) {
  void* p = nullptr;
  return vgpu_alloc_host(bytes ? bytes : 16, &p) == VGPU_OK ? p : nullptr;
}
void pinnedFree(void* p) { vgpu_free_host(p); }
HostAllocator g_allocator{mallocAlloc, mallocFree};
} // namespace
HostAllocator& hostAllocator() { return g_allocator; }
} // namespace memory

namespace gpu {

void usePinnedHostMemory(bool on) {
  memory::hostAllocator() = on ? memory::HostAllocator{memory::pinnedAlloc, memory::pinnedFree} : memory::HostAllocator{memory::mallocAlloc, memory::mallocFree};
}

// Device arenas and pinned areas outlive the operators: cudaMalloc + cudaMallocHost of a 60 MB slot cost ~25 ms, more
// than a whole Q1 over 20 M rows.  Closed operators hand their slots back; the next operator (the next query, or the
// next driver of this one) takes them from here.
namespace {
struct Arena { char* dev; char* pin; size_t cap; };
std::mutex g_arenas_mu;
std::vector<Arena> g_arenas;
bool takeArena(size_t bytes, Arena& out) {
  std::lock_guard<std::mutex> l(g_arenas_mu);
  int best = -1;
  for (size_t i = 0; i < g_arenas.size(); ++i) {
    if (g_arenas[i].cap >= bytes && (best < 0 || g_arenas[i].cap < g_arenas[best].cap)) best = (int)i;
  }
  if (best < 0) return false;
  out = g_arenas[best];
  g_arenas.erase(g_arenas.begin() + best);
  return true;
}
void giveArena(const Arena& a) {
  std::lock_guard<std::mutex> l(g_arenas_mu);
  if (g_arenas.size() < 64) { g_arenas.push_back(a); return; }
  vgpu_free(a.dev);
  vgpu_free_host(a.pin);
}
} // namespace

Staging::Staging(void* computeStream) : compute_(computeStream) {
  check(vgpu_stream_create(&copy_), "vgpu_stream_create");
  for (auto& s : slots_) {
    check(vgpu_event_create(&s.freeEv), "vgpu_event_create");
    check(vgpu_event_create(&s.readyEv), "vgpu_event_create");
  }
}

Staging::~Staging() {
  for (auto& s : slots_) {
    if (s.inFlight) vgpu_event_sync(s.freeEv);
    if (s.dev) giveArena(Arena{s.dev, s.pin, s.cap});
    if (s.freeEv) vgpu_event_destroy(s.freeEv);
    if (s.readyEv) vgpu_event_destroy(s.readyEv);
  }
  if (copy_) vgpu_stream_destroy(copy_);
}

void Staging::begin(size_t bytes) {
  VB_CHECK(!active_, "Staging::begin while a slot is open");
  cur_ = (cur_ + 1) % kSlots;
  Slot& s = slots_[cur_];
  if (s.inFlight) { // the only place the host waits: a full ring ahead of the GPU
    check(vgpu_event_sync(s.freeEv), "vgpu_event_sync");
    s.inFlight = false;
  }
  s.keep.clear();
  bytes = (bytes + 4095) & ~(size_t)4095;
  if (bytes > s.cap) {
    if (s.dev) giveArena(Arena{s.dev, s.pin, s.cap});
    s.dev = s.pin = nullptr;
    s.cap = 0;
    Arena a;
    if (takeArena(bytes, a)) {
      s.dev = a.dev; s.pin = a.pin; s.cap = a.cap;
    } else {
      void* d = nullptr;
      void* h = nullptr;
      check(vgpu_alloc(bytes, &d), "vgpu_alloc");
      check(vgpu_alloc_host(bytes, &h), "vgpu_alloc_host");
      s.dev = (char*)d;
      s.pin = (char*)h;
      s.cap = bytes;
    }
  }
  s.used = s.pinUsed = 0;
  active_ = true;
}

void* Staging::reserve(size_t bytes) {
  VB_CHECK(active_, "Staging::reserve without a slot");
  Slot& s = slots_[cur_];
  const size_t at = (s.used + 255) & ~(size_t)255;
  VB_CHECK(at + bytes <= s.cap, "staging slot overflow (internal: the caller sized the slot too small)");
  s.used = at + bytes;
  return s.dev + at;
}

void Staging::copyIn(void* dst, const void* host, size_t bytes, const std::shared_ptr<const void>& owner) {
  if (!bytes) return;
  Slot& s = slots_[cur_];
  int32_t pinned = 0;
  if (owner) vgpu_host_is_pinned(host, &pinned);
  if (pinned) {
    s.keep.push_back(owner);
    check(vgpu_h2d_async(dst, host, bytes, copy_), "vgpu_h2d_async");
  } else {
    // the pinned area mirrors the device arena offset for offset: stage at the destination's own mirror
    const size_t at = (size_t)((char*)dst - s.dev);
    VB_CHECK(at + bytes <= s.cap, "staging slot overflow (pinned area)");
    std::memcpy(s.pin + at, host, bytes);
    check(vgpu_h2d_async(dst, s.pin + at, bytes, copy_), "vgpu_h2d_async");
    stagedBytes += (int64_t)bytes;
  }
  h2dBytes += (int64_t)bytes;
}

void* Staging::mirror(void* dev) {
  Slot& s = slots_[cur_];
  return s.pin + ((char*)dev - s.dev);
}

void Staging::copyMirror(void* dev, size_t bytes) {
  if (!bytes) return;
  Slot& s = slots_[cur_];
  check(vgpu_h2d_async(dev, s.pin + ((char*)dev - s.dev), bytes, copy_), "vgpu_h2d_async");
  h2dBytes += (int64_t)bytes;
  stagedBytes += (int64_t)bytes;
}

void* Staging::scratch(size_t bytes, bool zero) {
  void* d = reserve(bytes ? bytes : 8);
  if (zero) check(vgpu_memset_async(d, 0, bytes ? bytes : 8, copy_), "vgpu_memset_async");
  return d;
}

void Staging::ready() {
  VB_CHECK(active_, "Staging::ready without a slot");
  Slot& s = slots_[cur_];
  check(vgpu_event_record(s.readyEv, copy_), "vgpu_event_record");
  check(vgpu_stream_wait_event(compute_, s.readyEv), "vgpu_stream_wait_event");
}

void Staging::release() {
  VB_CHECK(active_, "Staging::release without a slot");
  Slot& s = slots_[cur_];
  check(vgpu_event_record(s.freeEv, compute_), "vgpu_event_record");
  // the slot's next copies must not overtake the kernels that still read it
  check(vgpu_stream_wait_event(copy_, s.freeEv), "vgpu_stream_wait_event");
  s.inFlight = true;
  active_ = false;
}

} // namespace gpu
} // namespace vb200
