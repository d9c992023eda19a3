// efgp_bbox_static: the bounding-box reduction of efgp_bbox_async (fit.cu) for callers that may not allocate --
// code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  It has no counterpart in the
// reference; no other code path of the library uses this file.
//
// The per-CTA partial extrema live in a file-scope `static __device__` array instead of the scratch pool:
//   * every GPU that loads the library holds kStaticBytes = 786,688 bytes (768 KiB of slots + one 256-byte header) until the process exits;
//   * one CTA reduces one tile of kTilePoints points and owns one slot of the array, so the array's size is the
//     entry point's maximum: npts <= kStaticMaxCtas * kTilePoints = 134,217,728 (EFGP_BBOX_STATIC_MAX_POINTS);
//   * EVERY call shares the array.  Two calls that overlap on the GPU (different streams, or replays of two captured
//     graphs at once) race on the partials and on the ticket -- serialise them.  The array keeps the last call's
//     contents: the ticket that elects the finishing CTA must start at zero and is cleared by a cudaMemsetAsync at
//     the head of every call; the partial slots are all overwritten before they are read.
// In host code the NAME of a __device__ array is a host-side stand-in, not a device address, so the kernel is never
// given `g_bbox_scratch` itself: scratch_for_current_device() asks cudaGetSymbolAddress once per device, with that
// device current, and the kernel receives that pointer.
//
// The call enqueues one memset and ONE kernel (the last CTA to finish, elected by the ticket, folds the partials), and
// calls no cudaMalloc / cudaFree / pool / cuFFT / NCCL routine.
#include "common.cuh"
#include <mutex>

namespace efgp {

constexpr int kStaticThreads = 256;
constexpr int kStaticPerThread = 32;                                  // points per thread
constexpr int kTilePoints = kStaticThreads * kStaticPerThread;        // 8192 points per CTA
constexpr int kStaticMaxCtas = 16384;
constexpr int64_t kStaticMaxPoints = (int64_t)kStaticMaxCtas * kTilePoints;   // 134,217,728
constexpr int kStaticHeader = 32;                                     // doubles (256 bytes): the ticket lives here
constexpr size_t kStaticDoubles = (size_t)kStaticHeader + (size_t)kStaticMaxCtas * 2 * kMaxDim;
constexpr size_t kStaticBytes = kStaticDoubles * sizeof(double);      // 786,688 bytes
constexpr int kMaxDevices = 64;

static __device__ __align__(256) double g_bbox_scratch[kStaticDoubles];

// part = scratch + kStaticHeader: [ctas][2 D] partial extrema; ticket = (unsigned*)scratch
template <typename T, int D>
__global__ void __launch_bounds__(kStaticThreads)
bbox_static_kernel(const T* __restrict__ x, int64_t N, double* scratch, double* __restrict__ ext) {
  unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
  double* part = scratch + kStaticHeader;
  double hi[D], nlo[D];
#pragma unroll
  for (int a = 0; a < D; ++a) hi[a] = nlo[a] = -1.0e308;
  const int64_t base = (int64_t)blockIdx.x * kTilePoints + threadIdx.x;
#pragma unroll 8
  for (int j = 0; j < kStaticPerThread; ++j) {
    const int64_t i = base + (int64_t)j * kStaticThreads;
    if (i < N) {
#pragma unroll
      for (int a = 0; a < D; ++a) {
        const double v = (double)x[i * D + a];
        hi[a] = v > hi[a] ? v : hi[a];
        nlo[a] = -v > nlo[a] ? -v : nlo[a];
      }
    }
  }
  __shared__ double sh[2 * D][kStaticThreads / 32];
  __shared__ bool last;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int a = 0; a < D; ++a) {
    double u = hi[a], l = nlo[a];
    for (int o = 16; o > 0; o >>= 1) {
      u = fmax(u, __shfl_xor_sync(0xffffffffu, u, o));
      l = fmax(l, __shfl_xor_sync(0xffffffffu, l, o));
    }
    if (lane == 0) {
      sh[a][w] = u;
      sh[D + a][w] = l;
    }
  }
  __syncthreads();
  if (threadIdx.x < 2 * D) {
    double m = -1.0e308;
    for (int q = 0; q < kStaticThreads / 32; ++q) m = fmax(m, sh[threadIdx.x][q]);
    part[(int64_t)blockIdx.x * 2 * D + threadIdx.x] = m;
    __threadfence();                       // the partial is visible device-wide before this CTA takes its ticket
  }
  __syncthreads();
  if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!last) return;
  // the finishing CTA: fold all partials (read past L1: they were written by other SMs)
  __threadfence();
  double m[2 * D];
#pragma unroll
  for (int c = 0; c < 2 * D; ++c) m[c] = -1.0e308;
  for (int b = threadIdx.x; b < (int)gridDim.x; b += kStaticThreads) {
#pragma unroll
    for (int c = 0; c < 2 * D; ++c) m[c] = fmax(m[c], __ldcg(part + (int64_t)b * 2 * D + c));
  }
#pragma unroll
  for (int c = 0; c < 2 * D; ++c) {
    double u = m[c];
    for (int o = 16; o > 0; o >>= 1) u = fmax(u, __shfl_xor_sync(0xffffffffu, u, o));
    if (lane == 0) sh[c][w] = u;
  }
  __syncthreads();
  if (threadIdx.x < 2 * D) {
    double r = -1.0e308;
    for (int q = 0; q < kStaticThreads / 32; ++q) r = fmax(r, sh[threadIdx.x][q]);
    ext[threadIdx.x] = r;
  }
}

// Device address of g_bbox_scratch on the CURRENT device, resolved once per device.
static double* scratch_for_current_device() {
  static std::mutex mu;
  static double* ptrs[kMaxDevices] = {};
  int dev = -1;
  EFGP_CUDA(cudaGetDevice(&dev));
  EFGP_CHECK(dev >= 0 && dev < kMaxDevices, "efgp_bbox_static: device ordinal out of range");
  std::lock_guard<std::mutex> lk(mu);
  if (ptrs[dev] == nullptr) {
    double* p = nullptr;
    EFGP_CUDA(cudaGetSymbolAddress((void**)&p, g_bbox_scratch));   // `dev` is current here
    EFGP_CHECK(p != nullptr, "efgp_bbox_static: cudaGetSymbolAddress returned a null address");
    ptrs[dev] = p;
  }
  return ptrs[dev];
}

static void bbox_static(const void* x, int64_t N, int dim, int dtype, double* ext_dev, cudaStream_t st) {
  EFGP_CHECK(x != nullptr && ext_dev != nullptr, "efgp_bbox_static: null argument");
  EFGP_CHECK(dtype == 0 || dtype == 1, "efgp_bbox_static: dtype must be 0 (f32) or 1 (f64)");
  EFGP_CHECK(dim >= 1 && dim <= kMaxDim, "efgp_bbox_static: dim must be 1..3");
  EFGP_CHECK(N >= 1, "efgp_bbox_static: bounding box of an empty point set");
  if (N > kStaticMaxPoints)
    throw Error(1, "efgp_bbox_static: npts = " + std::to_string(N) + " exceeds the static scratch array's maximum of " +
                       std::to_string(kStaticMaxPoints) + " points (" + std::to_string(kStaticMaxCtas) + " CTA slots x " +
                       std::to_string(kTilePoints) + " points); use efgp_bbox_async, which takes scratch from the pool");
  double* scratch = scratch_for_current_device();
  const int ctas = (int)((N + kTilePoints - 1) / kTilePoints);
  EFGP_CUDA(cudaMemsetAsync(scratch, 0, kStaticHeader * sizeof(double), st));   // the ticket starts at zero
#define EFGP_BOXS(T, D) bbox_static_kernel<T, D><<<ctas, kStaticThreads, 0, st>>>((const T*)x, N, scratch, ext_dev)
  if (dtype == 0) {
    if (dim == 1) EFGP_BOXS(float, 1); else if (dim == 2) EFGP_BOXS(float, 2); else EFGP_BOXS(float, 3);
  } else {
    if (dim == 1) EFGP_BOXS(double, 1); else if (dim == 2) EFGP_BOXS(double, 2); else EFGP_BOXS(double, 3);
  }
#undef EFGP_BOXS
  EFGP_CUDA(cudaGetLastError());
}

}  // namespace efgp

extern "C" {

int efgp_bbox_static(const void* x, int64_t npts, int dim, int dtype, double* ext_dev, void* stream) {
  try {
    efgp::bbox_static(x, npts, dim, dtype, ext_dev, (cudaStream_t)stream);
    return 0;
  } catch (const efgp::Error& e) {
    efgp::set_last_error(e.what());
    return e.code;
  } catch (const std::exception& e) {
    efgp::set_last_error(e.what());
    return 99;
  }
}

}  // extern "C"
