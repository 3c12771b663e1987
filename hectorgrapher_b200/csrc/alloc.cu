// Small caching allocator and per-thread stream for the short-lived objects of the
// per-pair path (one cost function per cloud per (node, submap) pair, scratch of one
// scan match). cudaMalloc / cudaFree / cudaMallocHost take a process-wide driver lock and
// cudaFree waits for the whole device, which serialises host threads that each drive
// their own search; blocks released by a thread are therefore kept in that thread's
// free lists (power-of-two size classes) and handed out again. Blocks above 32 MiB (the
// outputs of a whole optimisation window) bypass the cache.
#include <map>
#include <mutex>
#include <unordered_map>

#include "hg_internal.cuh"

namespace hg {
namespace {

constexpr size_t kMaxCachedBytes = size_t(32) << 20;

struct BlockInfo {
  int device;  // -1 = pinned host memory
  int cls;     // size class (log2 of the capacity), -1 = not cached
};

std::mutex g_live_mutex;
std::unordered_map<void*, BlockInfo> g_live;

int ClassOf(const size_t bytes) {
  int cls = 8;  // 256 B
  while ((size_t(1) << cls) < bytes) ++cls;
  return cls;
}

struct ThreadCache {
  std::map<std::pair<int, int>, std::vector<void*>> free_lists;  // (device or -1, class) -> blocks
  std::map<int, cudaStream_t> streams;                           // device -> stream
  ~ThreadCache() {
    for (auto& kv : free_lists) {
      for (void* p : kv.second) {
        if (kv.first.first < 0) {
          cudaFreeHost(p);
        } else {
          cudaSetDevice(kv.first.first);
          cudaFree(p);
        }
      }
    }
    for (auto& kv : streams) cudaStreamDestroy(kv.second);
  }
};

ThreadCache& Cache() {
  static thread_local ThreadCache cache;
  return cache;
}

void* Allocate(const size_t bytes, const bool pinned) {
  int device = -1;
  if (!pinned && cudaGetDevice(&device) != cudaSuccess) return nullptr;
  const size_t want = bytes == 0 ? 1 : bytes;
  BlockInfo info{device, want > kMaxCachedBytes ? -1 : ClassOf(want)};
  void* p = nullptr;
  if (info.cls >= 0) {
    std::vector<void*>& list = Cache().free_lists[{device, info.cls}];
    if (!list.empty()) {
      p = list.back();
      list.pop_back();
    }
  }
  if (p == nullptr) {
    const size_t cap = info.cls >= 0 ? (size_t(1) << info.cls) : want;
    const cudaError_t e = pinned ? cudaMallocHost(&p, cap) : cudaMalloc(&p, cap);
    if (e != cudaSuccess) {
      SetError(std::string(pinned ? "cudaMallocHost: " : "cudaMalloc: ") + cudaGetErrorString(e));
      return nullptr;
    }
  }
  std::lock_guard<std::mutex> lock(g_live_mutex);
  g_live[p] = info;
  return p;
}

void Release(void* p) {
  if (p == nullptr) return;
  BlockInfo info;
  {
    std::lock_guard<std::mutex> lock(g_live_mutex);
    auto it = g_live.find(p);
    if (it == g_live.end()) return;
    info = it->second;
    g_live.erase(it);
  }
  if (info.cls < 0) {
    if (info.device < 0) cudaFreeHost(p); else cudaFree(p);
    return;
  }
  Cache().free_lists[{info.device, info.cls}].push_back(p);
}

}  // namespace

void* CachedDeviceAlloc(const size_t bytes) { return Allocate(bytes, false); }
void* CachedPinnedAlloc(const size_t bytes) { return Allocate(bytes, true); }
void CachedFree(void* p) { Release(p); }

cudaStream_t ThreadStream() {
  int device = 0;
  if (cudaGetDevice(&device) != cudaSuccess) return nullptr;
  std::map<int, cudaStream_t>& streams = Cache().streams;
  auto it = streams.find(device);
  if (it != streams.end()) return it->second;
  cudaStream_t s = nullptr;
  if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
  streams[device] = s;
  return s;
}

}  // namespace hg
