// Aligned host memory with leak accounting and size-class reuse: the native side of DnnStorage (reference:
// S/tensor/DnnStorage.scala:99-125 — Memory.AlignedMalloc(bytes, 64) per storage, a global set of live pointers that
// release() checks so a double free or a leak is detected). Blocks are rounded up to a power of two >= 256 bytes and
// returned to a per-class free list, so the per-iteration DnnTensors of an inference loop do not hit malloc.
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <mutex>
#include <unordered_map>
#include <vector>

namespace {

struct Pool {
  std::mutex mu;
  std::unordered_map<void*, size_t> live;            // pointer -> class size
  std::unordered_map<size_t, std::vector<void*>> free_lists;
  size_t live_bytes = 0, cached_bytes = 0, total_allocs = 0, reused = 0;
};

Pool& pool() {
  static Pool* p = new Pool();                       // never destroyed: blocks may outlive static destruction order
  return *p;
}

size_t class_size(size_t bytes) {
  size_t c = 256;
  while (c < bytes) c <<= 1;
  return c;
}

}  // namespace

extern "C" void* bdlh_aligned_alloc(size_t bytes, size_t align) {
  if (align < 64) align = 64;
  if (align & (align - 1)) return nullptr;           // power of two only
  const size_t cls = class_size(bytes ? bytes : 1);
  Pool& p = pool();
  std::lock_guard<std::mutex> lock(p.mu);
  void* ptr = nullptr;
  auto it = p.free_lists.find(cls);
  if (it != p.free_lists.end() && !it->second.empty() && align <= 4096) {
    ptr = it->second.back();
    it->second.pop_back();
    p.cached_bytes -= cls;
    ++p.reused;
  } else {
    const size_t a = align < 4096 ? 4096 : align;    // page alignment for every block: any smaller request is satisfied
    if (posix_memalign(&ptr, a, cls < a ? a : cls) != 0) return nullptr;
  }
  p.live[ptr] = cls;
  p.live_bytes += cls;
  ++p.total_allocs;
  return ptr;
}

// 0 = released, 1 = not a live pointer of this pool (double release or foreign pointer)
extern "C" int bdlh_aligned_free(void* ptr) {
  if (ptr == nullptr) return 1;
  Pool& p = pool();
  std::lock_guard<std::mutex> lock(p.mu);
  auto it = p.live.find(ptr);
  if (it == p.live.end()) return 1;
  const size_t cls = it->second;
  p.live.erase(it);
  p.live_bytes -= cls;
  if (p.cached_bytes + cls <= (size_t(1) << 30)) {   // keep at most 1 GiB cached
    p.free_lists[cls].push_back(ptr);
    p.cached_bytes += cls;
  } else {
    free(ptr);
  }
  return 0;
}

// what: 0 live blocks, 1 live bytes, 2 cached bytes, 3 allocations so far, 4 allocations served from the cache
extern "C" size_t bdlh_aligned_stat(int what) {
  Pool& p = pool();
  std::lock_guard<std::mutex> lock(p.mu);
  switch (what) {
    case 0: return p.live.size();
    case 1: return p.live_bytes;
    case 2: return p.cached_bytes;
    case 3: return p.total_allocs;
    default: return p.reused;
  }
}

extern "C" void bdlh_aligned_trim() {
  Pool& p = pool();
  std::lock_guard<std::mutex> lock(p.mu);
  for (auto& kv : p.free_lists)
    for (void* q : kv.second) free(q);
  p.free_lists.clear();
  p.cached_bytes = 0;
}
