// Bin-sorting of nonuniform points by fine-grid cell.
//
// Replaces the point handling FINUFFT does inside every pff.finufft_type1/type2 call
// (reference call sites efgpnd.py:1496-1499, 1533-1549; phi = 2*pi*h*(x - xcen), efgpnd.py:1451).
// Points are folded to one period, scaled to fine-grid units, keyed by cell and radix-sorted
// once per (point set, fine grid); every later spread/interp streams them in cell order.
#include "nufft.cuh"
#include <cub/device/device_radix_sort.cuh>

namespace efgp {

SortedPoints::~SortedPoints() {
  for (int a = 0; a < kMaxDim; ++a)
    if (xs[a]) pool_free(xs[a]);
  if (perm) pool_free(perm);
}

void* DevBuf::get(size_t bytes) {
  if (bytes > cap) {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) {
      p = nullptr;
      throw Error(2, std::string("cudaMalloc(") + std::to_string(want) + "): " + cudaGetErrorString(e));
    }
    cap = want;
  }
  return p;
}
DevBuf::~DevBuf() {
  if (p) cudaFree(p);
}

template <typename T>
struct FoldParams {
  T h;
  T xcen[kMaxDim];
  T nf[kMaxDim];
  int inf[kMaxDim];
  T shift;
};

// fine-grid coordinate of one coordinate value, shifted by 1/2 for odd widths, in [0, nf)
template <typename T>
__device__ __forceinline__ T fold_coord(T xv, T h, T xc, T nf, T shift) {
  T t = h * (xv - xc);
  t -= floor(t);
  T X = t * nf + shift;
  if (X >= nf) X -= nf;
  if (!(X >= T(0)) || X >= nf) X = T(0);  // rounding at the period edge (and NaN guard)
  return X;
}

template <typename T, int D>
__global__ void __launch_bounds__(256) key_kernel(const T* __restrict__ x, int64_t N, FoldParams<T> p,
                                                  uint32_t* __restrict__ keys, uint32_t* __restrict__ idx) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  uint32_t key = 0;
#pragma unroll
  for (int a = 0; a < D; ++a) {
    T X = fold_coord(x[i * D + a], p.h, p.xcen[a], p.nf[a], p.shift);
    int c = (int)X;
    if (c >= p.inf[a]) c = p.inf[a] - 1;
    key = key * (uint32_t)p.inf[a] + (uint32_t)c;
  }
  keys[i] = key;
  idx[i] = (uint32_t)i;
}

// One thread per SLOT (coalesced writes); the slot's sorted rank follows from the tile-transposed
// layout described in nufft.cuh.
template <typename T, int D>
__global__ void __launch_bounds__(256) gather_kernel(const T* __restrict__ x, int64_t N, int P, FoldParams<T> p,
                                                     const uint32_t* __restrict__ sorted_idx,
                                                     uint32_t* __restrict__ perm, T* __restrict__ o0,
                                                     T* __restrict__ o1, T* __restrict__ o2, int64_t npos) {
  int64_t pos = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= npos) return;
  const int64_t tile_len = (int64_t)kTileThreads * P;
  int64_t tile = pos / tile_len;
  int r = (int)(pos - tile * tile_len);
  int j = r / kTileThreads, t = r - j * kTileThreads;
  int64_t rank = tile * tile_len + (int64_t)t * P + j;
  T* outs[3] = {o0, o1, o2};
  if (rank >= N) {
    outs[0][pos] = T(-1);
#pragma unroll
    for (int a = 1; a < D; ++a) outs[a][pos] = T(0);
    perm[pos] = 0;
    return;
  }
  int64_t src = sorted_idx[rank];
  perm[pos] = (uint32_t)src;
#pragma unroll
  for (int a = 0; a < D; ++a) {
    T X = fold_coord(x[src * D + a], p.h, p.xcen[a], p.nf[a], p.shift);
    if ((int)X >= p.inf[a]) X = T(0);
    outs[a][pos] = X;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) permute_kernel(const T* __restrict__ c, int64_t cstride,
                                                      const uint32_t* __restrict__ perm, int64_t npos,
                                                      T* __restrict__ cs) {
  int64_t pos = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= npos) return;
  cs[pos] = c[(int64_t)perm[pos] * cstride];
}

void permute_strengths(const SortedPoints& sp, const void* c, int64_t cstride, void* cs, cudaStream_t st) {
  if (sp.npos == 0) return;
  int nb = div_up(sp.npos, 256);
  ScopedStage stg("permute_strengths", st);
  if (sp.dtype == 0)
    permute_kernel<float><<<nb, 256, 0, st>>>((const float*)c, cstride, sp.perm, sp.npos, (float*)cs);
  else
    permute_kernel<double><<<nb, 256, 0, st>>>((const double*)c, cstride, sp.perm, sp.npos, (double*)cs);
  EFGP_CUDA(cudaGetLastError());
}

static int pick_points_per_thread(int64_t N) {
  // >= 4 waves of 148 SMs x 256 threads when N allows, and long register runs when N is large
  int64_t P = N / (148 * 256 * 4);
  if (P < 1) P = 1;
  if (P > 128) P = 128;
  int p2 = 1;
  while (p2 * 2 <= P) p2 *= 2;   // power of two: the gather kernel reads the sort output in 8-wide chunks
  return p2;
}

template <typename T>
static void build_t(SortedPoints& sp, const T* x, int64_t N, int dim, double h, const double* xcen,
                    const int64_t* nf, int odd, cudaStream_t st) {
  FoldParams<T> p;
  p.h = (T)h;
  p.shift = odd ? T(0.5) : T(0);
  double cells = 1;
  for (int a = 0; a < kMaxDim; ++a) {
    p.xcen[a] = a < dim ? (T)xcen[a] : T(0);
    p.nf[a] = a < dim ? (T)nf[a] : T(1);
    p.inf[a] = a < dim ? (int)nf[a] : 1;
    if (a < dim) cells *= (double)nf[a];
  }
  EFGP_CHECK(cells < 4.0e9, "fine grid has too many cells for 32-bit sort keys");
  EFGP_CHECK(N < (int64_t)2147483000, "N must be < 2^31");
  int bits = 1;
  while ((1ull << bits) < (unsigned long long)cells) ++bits;

  sp.P = dim == 3 ? 1 : pick_points_per_thread(N);  // 3-D uses the per-point kernels: keep rank order
  const int64_t tile_len = (int64_t)kTileThreads * sp.P;
  sp.ntiles = (N + tile_len - 1) / tile_len;
  sp.npos = sp.ntiles * tile_len;

  uint32_t* keys_in = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)N);
  uint32_t* keys_out = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)N);
  uint32_t* idx_in = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)N);
  uint32_t* idx_out = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)N);
  sp.perm = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)sp.npos);
  int nb = div_up(N, 256);
  {
  ScopedStage stg("sort.key_kernel", st);
  if (dim == 1) key_kernel<T, 1><<<nb, 256, 0, st>>>(x, N, p, keys_in, idx_in);
  if (dim == 2) key_kernel<T, 2><<<nb, 256, 0, st>>>(x, N, p, keys_in, idx_in);
  if (dim == 3) key_kernel<T, 3><<<nb, 256, 0, st>>>(x, N, p, keys_in, idx_in);
  }
  EFGP_CUDA(cudaGetLastError());
  size_t tmp_bytes = 0;
  EFGP_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys_in, keys_out, idx_in, idx_out, (int)N, 0,
                                            bits, st));
  void* tmp = pool_alloc(tmp_bytes);
  {
    ScopedStage stg("sort.cub_radix_sort_pairs", st, 0);
    EFGP_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys_in, keys_out, idx_in, idx_out, (int)N, 0, bits,
                                              st));
  }
  for (int a = 0; a < dim; ++a) sp.xs[a] = pool_alloc(sizeof(T) * (size_t)sp.npos);
  T* o0 = (T*)sp.xs[0];
  T* o1 = (T*)sp.xs[1];
  T* o2 = (T*)sp.xs[2];
  int nbp = div_up(sp.npos, 256);
  {
  ScopedStage stg("sort.gather_kernel", st);
  if (dim == 1) gather_kernel<T, 1><<<nbp, 256, 0, st>>>(x, N, sp.P, p, idx_out, sp.perm, o0, o1, o2, sp.npos);
  if (dim == 2) gather_kernel<T, 2><<<nbp, 256, 0, st>>>(x, N, sp.P, p, idx_out, sp.perm, o0, o1, o2, sp.npos);
  if (dim == 3) gather_kernel<T, 3><<<nbp, 256, 0, st>>>(x, N, sp.P, p, idx_out, sp.perm, o0, o1, o2, sp.npos);
  }
  EFGP_CUDA(cudaGetLastError());
  // scratch goes back to the pool; later users run on the same stream, so ordering is preserved
  pool_free(tmp);
  pool_free(keys_in);
  pool_free(keys_out);
  pool_free(idx_in);
  pool_free(idx_out);
}

void build_sorted_points(SortedPoints& sp, const void* x, int64_t N, int dim, int dtype, double h,
                         const double* xcen, const int64_t* nf, int odd, cudaStream_t st) {
  sp.dim = dim;
  sp.dtype = dtype;
  sp.N = N;
  sp.odd = odd;
  for (int a = 0; a < kMaxDim; ++a) sp.nf[a] = a < dim ? nf[a] : 1;
  if (N == 0) return;
  if (dtype == 0)
    build_t<float>(sp, (const float*)x, N, dim, h, xcen, nf, odd, st);
  else
    build_t<double>(sp, (const double*)x, N, dim, h, xcen, nf, odd, st);
}

}  // namespace efgp
