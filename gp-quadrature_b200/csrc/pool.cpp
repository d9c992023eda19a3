// Size-bucketed caching allocator for device scratch (see common.cuh).
//
// Stream awareness: a cached block remembers the stream of the library call that released it.  A call on the SAME
// stream may take it at once (stream order protects the data: the usual case, one torch stream per process); a call
// on ANOTHER stream first waits for the releasing stream to drain, so a block is never handed to work that can run
// concurrently with its previous user.  The C-ABI entry points publish their stream with set_current_stream().
//
// One GPU per process: the pool, the cuFFT plan cache and the deconvolution tables are process-wide and belong to the
// device of the first call; a call with another current device is refused with a clear error instead of handing a
// cuda:0 block to work on cuda:1 (one process per GPU).
#include "common.cuh"
#include <map>
#include <mutex>

namespace efgp {

namespace {
struct Block {
  size_t cap;
  cudaStream_t last;      // stream of the call that released (or last allocated) it
};
std::mutex g_mu;
std::multimap<size_t, std::pair<void*, cudaStream_t>> g_free;   // cached blocks by capacity
std::map<void*, Block> g_live;                                    // blocks handed out
size_t g_cached_bytes = 0;
const size_t kMaxCached = (size_t)16 << 30;
int g_device = -1;
thread_local cudaStream_t t_stream = nullptr;
}  // namespace

void set_current_stream(cudaStream_t st) { t_stream = st; }
cudaStream_t current_stream() { return t_stream; }

void check_device() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) throw Error(2, "cudaGetDevice failed: no usable CUDA device");
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_device < 0) {
    g_device = dev;
    return;
  }
  if (dev != g_device)
    throw Error(4, "libefgp_b200 drives ONE GPU per process (its scratch pool, cuFFT plans and tables live on cuda:" +
                       std::to_string(g_device) + "); this call arrived with cuda:" + std::to_string(dev) +
                       " current -- run one process per GPU (torchrun) as INTEGRATION.md describes");
}

void* pool_alloc(size_t bytes) {
  if (bytes == 0) bytes = 256;
  bytes = (bytes + 511) & ~(size_t)511;
  void* p = nullptr;
  cudaStream_t prev = nullptr;
  bool hit = false;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_free.lower_bound(bytes);
    if (it != g_free.end() && it->first <= bytes + bytes / 4 + (1 << 20)) {
      p = it->second.first;
      prev = it->second.second;
      g_cached_bytes -= it->first;
      g_live[p] = Block{it->first, t_stream};
      g_free.erase(it);
      hit = true;
    }
  }
  if (hit) {
    // released on another stream: its last user may still be running there
    if (prev != t_stream) cudaStreamSynchronize(prev);
    return p;
  }
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) {
    cudaGetLastError();
    pool_trim();
    e = cudaMalloc(&p, bytes);
  }
  if (e != cudaSuccess)
    throw Error(2, std::string("cudaMalloc(") + std::to_string(bytes) + "): " + cudaGetErrorString(e));
  std::lock_guard<std::mutex> lk(g_mu);
  g_live[p] = Block{bytes, t_stream};
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
  size_t cap = it->second.cap;
  g_live.erase(it);
  if (g_cached_bytes + cap > kMaxCached) {
    cudaFree(p);      // (synchronises the device: safe for any stream)
    return;
  }
  g_free.emplace(cap, std::make_pair(p, t_stream));
  g_cached_bytes += cap;
}

void pool_trim() {
  std::lock_guard<std::mutex> lk(g_mu);
  for (auto& kv : g_free) cudaFree(kv.second.first);
  g_free.clear();
  g_cached_bytes = 0;
}

}  // namespace efgp
