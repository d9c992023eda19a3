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

// Scratch comes from the caching pool (pool.cpp): handles are rebuilt every hyperparameter step and a
// cudaMalloc/cudaFree pair per buffer would synchronise the device each time.
void* DevBuf::get(size_t bytes) {
  if (bytes > cap) {
    if (p) pool_free(p);
    p = nullptr;
    cap = 0;
    p = pool_alloc(bytes);
    cap = bytes;
  }
  return p;
}
DevBuf::~DevBuf() {
  if (p) pool_free(p);
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
// Slots [pos0, pos1) belong to the chunk whose points are the original indices [lo, lo + n): ranks are
// chunk-local sort positions, `sorted_idx` holds chunk-local indices.
template <typename T, int D>
__global__ void __launch_bounds__(256) gather_kernel(const T* __restrict__ x, int64_t N, int P, FoldParams<T> p,
                                                     const uint32_t* __restrict__ sorted_idx,
                                                     uint32_t* __restrict__ perm, T* __restrict__ o0,
                                                     T* __restrict__ o1, T* __restrict__ o2, int64_t pos0,
                                                     int64_t pos1, int64_t lo) {
  int64_t pos = pos0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= pos1) return;
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
  int64_t src = lo + sorted_idx[rank - pos0];
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
  ScopedStage stg("permute_strengths", st, 1, sp.N);
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

// Points per sort chunk: a whole number of tiles, sized so that one chunk's coordinates plus the sort's
// key/index ping-pong buffers (d*word + 16 bytes per point) fit comfortably in L2.  EFGP_SORT_CHUNK
// (points) overrides it; 0 or a value >= N gives one global sort.
static int64_t pick_sort_chunk(int64_t N, int64_t tile_len, int dim, size_t word) {
  int64_t target = (int64_t)((48ull << 20) / (dim * word + 16));   // ~48 MB working set
  if (const char* e = getenv("EFGP_SORT_CHUNK")) {
    long long v = atoll(e);
    target = v <= 0 ? N : (int64_t)v;
  }
  int64_t ntile_total = (N + tile_len - 1) / tile_len;
  int64_t tiles = std::max<int64_t>(1, target / tile_len);
  if (tiles * 3 >= ntile_total * 2) tiles = ntile_total;            // not worth splitting
  return tiles * tile_len;
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

  // Chunked sort.  The points are cut into chunks of consecutive ORIGINAL indices (a whole number of
  // tiles each) and every chunk is cell-sorted on its own: key -> radix sort -> gather run back to back
  // per chunk, so the chunk's coordinates, keys and indices stay resident in the 126 MB L2 between the
  // three kernels, and -- because a chunk's slots only ever reference original indices of the same chunk --
  // the strength gathers (permute_strengths) and the type-2 result scatters (interp) of later transforms
  // hit an L2-sized window instead of issuing one 32-byte DRAM sector per 8-byte element.
  const int64_t chunk = pick_sort_chunk(N, tile_len, dim, sizeof(T));
  sp.chunk = chunk;
  uint32_t* keys_in = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)chunk);
  uint32_t* keys_out = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)chunk);
  uint32_t* idx_in = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)chunk);
  uint32_t* idx_out = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)chunk);
  sp.perm = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)sp.npos);
  for (int a = 0; a < dim; ++a) sp.xs[a] = pool_alloc(sizeof(T) * (size_t)sp.npos);
  T* o0 = (T*)sp.xs[0];
  T* o1 = (T*)sp.xs[1];
  T* o2 = (T*)sp.xs[2];
  size_t tmp_bytes = 0;
  EFGP_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys_in, keys_out, idx_in, idx_out,
                                            (int)std::min<int64_t>(chunk, N), 0, bits, st));
  void* tmp = pool_alloc(tmp_bytes);
  for (int64_t lo = 0; lo < sp.npos; lo += chunk) {
    const int64_t n = std::min(chunk, N - lo);          // points of this chunk (> 0: npos - N < tile_len)
    const int64_t pos1 = std::min(lo + chunk, sp.npos);
    const T* xc = x + lo * dim;
    int nb = div_up(n, 256);
    {
      ScopedStage stg("sort.key_kernel", st, 1, n);
      if (dim == 1) key_kernel<T, 1><<<nb, 256, 0, st>>>(xc, n, p, keys_in, idx_in);
      if (dim == 2) key_kernel<T, 2><<<nb, 256, 0, st>>>(xc, n, p, keys_in, idx_in);
      if (dim == 3) key_kernel<T, 3><<<nb, 256, 0, st>>>(xc, n, p, keys_in, idx_in);
    }
    {
      ScopedStage stg("sort.cub_radix_sort_pairs", st, 0, n);
      size_t tb = tmp_bytes;
      EFGP_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tb, keys_in, keys_out, idx_in, idx_out, (int)n, 0, bits, st));
    }
    int nbp = div_up(pos1 - lo, 256);
    {
      ScopedStage stg("sort.gather_kernel", st, 1, n);
      if (dim == 1) gather_kernel<T, 1><<<nbp, 256, 0, st>>>(x, N, sp.P, p, idx_out, sp.perm, o0, o1, o2, lo, pos1, lo);
      if (dim == 2) gather_kernel<T, 2><<<nbp, 256, 0, st>>>(x, N, sp.P, p, idx_out, sp.perm, o0, o1, o2, lo, pos1, lo);
      if (dim == 3) gather_kernel<T, 3><<<nbp, 256, 0, st>>>(x, N, sp.P, p, idx_out, sp.perm, o0, o1, o2, lo, pos1, lo);
    }
    EFGP_CUDA(cudaGetLastError());
  }
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
