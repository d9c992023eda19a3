// Size-bucketed caching allocator for device scratch (see common.cuh).
#include "common.cuh"
#include <map>
#include <mutex>

namespace efgp {

namespace {
std::mutex g_mu;
std::multimap<size_t, void*> g_free;      // cached blocks by capacity
std::map<void*, size_t> g_live;           // capacity of blocks handed out
size_t g_cached_bytes = 0;
const size_t kMaxCached = (size_t)24 << 30;
}  // namespace

void* pool_alloc(size_t bytes) {
  if (bytes == 0) bytes = 256;
  bytes = (bytes + 511) & ~(size_t)511;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_free.lower_bound(bytes);
    if (it != g_free.end() && it->first <= bytes + bytes / 4 + (1 << 20)) {
      void* p = it->second;
      g_cached_bytes -= it->first;
      g_live[p] = it->first;
      g_free.erase(it);
      return p;
    }
  }
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) {
    cudaGetLastError();
    pool_trim();
    e = cudaMalloc(&p, bytes);
  }
  if (e != cudaSuccess)
    throw Error(2, std::string("cudaMalloc(") + std::to_string(bytes) + "): " + cudaGetErrorString(e));
  std::lock_guard<std::mutex> lk(g_mu);
  g_live[p] = bytes;
  return p;
}

void pool_free(void* p) {
  if (!p) return;
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_live.find(p);
  if (it == g_live.end()) {
    cudaFree(p);
    return;
  }
  size_t cap = it->second;
  g_live.erase(it);
  if (g_cached_bytes + cap > kMaxCached) {
    cudaFree(p);
    return;
  }
  g_free.emplace(cap, p);
  g_cached_bytes += cap;
}

void pool_trim() {
  std::lock_guard<std::mutex> lk(g_mu);
  for (auto& kv : g_free) cudaFree(kv.second);
  g_free.clear();
  g_cached_bytes = 0;
}

}  // namespace efgp
