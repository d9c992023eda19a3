// Type-1 spreading: real strengths at bin-sorted nonuniform points -> real fine grid.
//
// This is the arithmetic FINUFFT performs inside pff.finufft_type1 (reference call site
// efgpnd.py:1496-1499); the algorithm here is this library's own.
//
// Design (B200): points are sorted by fine-grid cell, so long runs of consecutive points share
// one W^d footprint.  Each thread walks a contiguous slice of the sorted points and keeps the
// footprint's partial sums in REGISTERS (W*W doubles for d=2), so a point costs only the
// piecewise-polynomial kernel evaluation plus W^d FMAs -- no shared-memory or global atomics in
// the inner loop (sm_100a has no native shared fp atomics: ATOMS.CAST.SPIN loops).  Only when the
// cell changes (or the slice ends) is the footprint flushed with native REDG.ADD.F64/F32 into
// the L2-resident fine grid.  EFGP grids are tiny next to N (e.g. 282^2 cells for 1e8 points),
// so flushes are ~1/100 of the FMA work at the headline size and degrade gracefully to one
// flush per point for sparse cases.
#include "nufft.cuh"
#include "stream_ring.cuh"

namespace efgp {

template <typename T, int W>
__device__ __forceinline__ void es_eval(const ESDev<T>& es, T s, T* __restrict__ k) {
  constexpr int DEG = W + 2;
  constexpr int NE = DEG / 2 + 1;    // even coefficients 0,2,..
  constexpr int NO = (DEG + 1) / 2;  // odd coefficients 1,3,..
  const T v = s * s;
#pragma unroll
  for (int i = 0; i < W / 2; ++i) {
    T e = es.c[i][2 * (NE - 1)];
#pragma unroll
    for (int j = NE - 2; j >= 0; --j) e = fma(e, v, es.c[i][2 * j]);
    T o = es.c[i][2 * NO - 1];
#pragma unroll
    for (int j = NO - 2; j >= 0; --j) o = fma(o, v, es.c[i][2 * j + 1]);
    o *= s;
    k[i] = e + o;
    k[W - 1 - i] = e - o;
  }
  if (W & 1) {
    constexpr int i = W / 2;
    T e = es.c[i][2 * (NE - 1)];
#pragma unroll
    for (int j = NE - 2; j >= 0; --j) e = fma(e, v, es.c[i][2 * j]);
    k[i] = e;
  }
}

// 4-word record -> registers with 16-byte loads
template <typename T>
__device__ __forceinline__ void load_record4(const T* __restrict__ p, T* w) {
  if constexpr (sizeof(T) == 8) {
    ld_global_nc_256(reinterpret_cast<const double*>(p), w[0], w[1], w[2], w[3]);
  } else {
    float4 a = __ldg(reinterpret_cast<const float4*>(p));
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
  }
}

__device__ __forceinline__ int wrap_idx(int i, int n) {
  if (i < 0) i += n;
  if (i >= n) i -= n;
  return i;
}

// Slot addressing of the tile-transposed sorted layout (nufft.cuh): thread t of tile `blockIdx.x`
// reads its j-th point at base + j*kTileThreads.
// ------------------------------------------------------------------ 1-D
template <typename T, int W, bool HAS_C>
__global__ void __launch_bounds__(kTileThreads)
spread1d_kernel(const T* __restrict__ rec, const T* __restrict__ cs, int P, int nf0, T* __restrict__ grid,
                const __grid_constant__ ESDev<T> es) {
  const int64_t base = (int64_t)blockIdx.x * kTileThreads * P + threadIdx.x;
  T acc[W];
#pragma unroll
  for (int i = 0; i < W; ++i) acc[i] = 0;
  int cur = -1;
  constexpr int OFF = W / 2 - ((W & 1) ? 0 : 1);
  for (int j = 0; j < P; ++j) {
    const int64_t pos = base + (int64_t)j * kTileThreads;
    T X = rec[pos * 2];
    if (X < T(0)) break;
    T cv = HAS_C ? cs[pos] : T(1);
    int c0 = (int)X;
    if (c0 != cur) {
      if (cur >= 0) {
#pragma unroll
        for (int i = 0; i < W; ++i) atomicAdd(&grid[wrap_idx(cur - OFF + i, nf0)], acc[i]);
      }
#pragma unroll
      for (int i = 0; i < W; ++i) acc[i] = 0;
      cur = c0;
    }
    T k0[W];
    es_eval<T, W>(es, T(2) * (X - (T)c0) - T(1), k0);
#pragma unroll
    for (int i = 0; i < W; ++i) acc[i] = fma(cv, k0[i], acc[i]);
  }
  if (cur >= 0) {
#pragma unroll
    for (int i = 0; i < W; ++i) atomicAdd(&grid[wrap_idx(cur - OFF + i, nf0)], acc[i]);
  }
}

// ------------------------------------------------------------------ 2-D, register footprint
template <typename T, int W>
__device__ __forceinline__ void flush2d(T (&acc)[W][W], int c0, int c1, int nf0, int nf1, T* __restrict__ grid) {
  constexpr int OFF = W / 2 - ((W & 1) ? 0 : 1);
  int cols[W];
#pragma unroll
  for (int j = 0; j < W; ++j) cols[j] = wrap_idx(c1 - OFF + j, nf1);
#pragma unroll
  for (int i = 0; i < W; ++i) {
    T* row = grid + (int64_t)wrap_idx(c0 - OFF + i, nf0) * nf1;
#pragma unroll
    for (int j = 0; j < W; ++j) {
      atomicAdd(row + cols[j], acc[i][j]);
      acc[i][j] = 0;
    }
  }
}

// HAS_C: 0 = all-ones strengths, 1 = strengths in slot order (cs), 2 = strengths in the records' payload word
template <typename T, int W, int HAS_C>
__global__ void __launch_bounds__(kTileThreads, (W <= 5 ? 2 : 1))
spread2d_kernel(const T* __restrict__ rec, const T* __restrict__ cs, int P, int nf0,
                int nf1, T* __restrict__ grid, const __grid_constant__ ESDev<T> es) {
  constexpr int S = kRingStages;
  __shared__ __align__(16) T ring[S][kTileThreads][2];   // (X0, X1) of each thread's next points
  __shared__ T cring[HAS_C ? S : 1][kTileThreads];
  const int t = threadIdx.x;
  const int64_t base = (int64_t)blockIdx.x * kTileThreads * P + t;
  auto issue = [&](int j) {
    if (j < P) {
      const int64_t pos = base + (int64_t)j * kTileThreads;
      const int s = j % S;
      cp_async_word<2 * sizeof(T)>(&ring[s][t][0], rec + pos * 4);   // one 16-byte (8 for f32) copy per point
      if (HAS_C == 1) cp_async_word<sizeof(T)>(&cring[s][t], cs + pos);
      if (HAS_C == 2) cp_async_word<sizeof(T)>(&cring[s][t], rec + pos * 4 + 3);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int j = 0; j < S - 1; ++j) issue(j);
  T acc[W][W];
#pragma unroll
  for (int i = 0; i < W; ++i)
#pragma unroll
    for (int j = 0; j < W; ++j) acc[i][j] = 0;
  int cur0 = -1, cur1 = -1;
  for (int j = 0; j < P; ++j) {
    issue(j + S - 1);
    cp_async_wait<S - 1>();
    const int s = j % S;
    const T X0 = ring[s][t][0];
    if (X0 < T(0)) break;
    const T X1 = ring[s][t][1];
    const T cv = HAS_C ? cring[s][t] : T(1);
    int c0 = (int)X0, c1 = (int)X1;
    if (c0 != cur0 || c1 != cur1) {
      if (cur0 >= 0) flush2d<T, W>(acc, cur0, cur1, nf0, nf1, grid);
      cur0 = c0;
      cur1 = c1;
    }
    T k0[W], k1[W];
    es_eval<T, W>(es, T(2) * (X0 - (T)c0) - T(1), k0);
    es_eval<T, W>(es, T(2) * (X1 - (T)c1) - T(1), k1);
#pragma unroll
    for (int i = 0; i < W; ++i) {
      T a = cv * k0[i];
#pragma unroll
      for (int jj = 0; jj < W; ++jj) acc[i][jj] = fma(a, k1[jj], acc[i][jj]);
    }
  }
  cp_async_wait<0>();
  if (cur0 >= 0) flush2d<T, W>(acc, cur0, cur1, nf0, nf1, grid);
}

// ------------------------------------------------------------------ generic d (2-D wide kernels, 3-D): per-point REDs
// Points are still cell-sorted, so the reductions of neighbouring slots land in the same L2 lines.
template <typename T, int W, int D, int HAS_C>
__global__ void __launch_bounds__(256)
spread_direct_kernel(const T* __restrict__ rec,
                     const T* __restrict__ cs, int64_t npos, int nf0, int nf1, int nf2, T* __restrict__ grid,
                     const __grid_constant__ ESDev<T> es) {
  int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= npos) return;
  constexpr int OFF = W / 2 - ((W & 1) ? 0 : 1);
  T Xr[4];
  load_record4<T>(rec + n * 4, Xr);
  T X0 = Xr[0];
  if (X0 < T(0)) return;
  T cv = HAS_C == 1 ? cs[n] : (HAS_C == 2 ? Xr[3] : T(1));
  T k0[W], k1[W], k2[W];
  int c0 = (int)X0;
  es_eval<T, W>(es, T(2) * (X0 - (T)c0) - T(1), k0);
  T X1 = Xr[1];
  int c1 = (int)X1;
  es_eval<T, W>(es, T(2) * (X1 - (T)c1) - T(1), k1);
  int c2 = 0;
  if (D == 3) {
    T X2 = Xr[2];
    c2 = (int)X2;
    es_eval<T, W>(es, T(2) * (X2 - (T)c2) - T(1), k2);
  }
  if (D == 2) {
#pragma unroll
    for (int i = 0; i < W; ++i) {
      T* row = grid + (int64_t)wrap_idx(c0 - OFF + i, nf0) * nf1;
      T a = cv * k0[i];
#pragma unroll
      for (int j = 0; j < W; ++j) atomicAdd(row + wrap_idx(c1 - OFF + j, nf1), a * k1[j]);
    }
  } else {
    int cols[W];
#pragma unroll
    for (int l = 0; l < W; ++l) cols[l] = wrap_idx(c2 - OFF + l, nf2);
#pragma unroll 1
    for (int i = 0; i < W; ++i) {
      int64_t r0 = (int64_t)wrap_idx(c0 - OFF + i, nf0) * nf1;
      T a = cv * k0[i];
#pragma unroll 1
      for (int j = 0; j < W; ++j) {
        T* row = grid + (r0 + wrap_idx(c1 - OFF + j, nf1)) * nf2;
        T b = a * k1[j];
#pragma unroll
        for (int l = 0; l < W; ++l) atomicAdd(row + cols[l], b * k2[l]);
      }
    }
  }
}

template <typename T, int W>
static void launch_w(const SortedPoints& sp, const ESKernel& k, const T* cs, T* grid, cudaStream_t st, bool from_record) {
  ESDev<T> es;
  es_to_dev<T>(k, es);
  const T* rec = (const T*)sp.rec;
  constexpr int kRegMaxW = sizeof(T) == 8 ? 9 : 10;
  const int nt = (int)sp.ntiles;
  const int mode = from_record ? 2 : (cs ? 1 : 0);
  char nm[48];
  snprintf(nm, sizeof(nm), "spread%dd_w%d_%s%s", sp.dim, W, sizeof(T) == 8 ? "f64" : "f32",
           mode == 0 ? "_ones" : (mode == 2 ? "_rec" : ""));
  ScopedStage stg(nm, st, 1, sp.N);
  const int nf0 = (int)sp.nf[0], nf1 = (int)sp.nf[1], nf2 = (int)sp.nf[2];
  if (sp.dim == 1) {
    if (cs)
      spread1d_kernel<T, W, true><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, grid, es);
    else
      spread1d_kernel<T, W, false><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, grid, es);
  } else if (sp.dim == 2) {
    if constexpr (W <= kRegMaxW) {
      if (mode == 2)
        spread2d_kernel<T, W, 2><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, nf1, grid, es);
      else if (mode == 1)
        spread2d_kernel<T, W, 1><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, nf1, grid, es);
      else
        spread2d_kernel<T, W, 0><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, nf1, grid, es);
    } else {
      int nb = div_up(sp.npos, 256);
      if (mode == 2)
        spread_direct_kernel<T, W, 2, 2><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, 1, grid, es);
      else if (mode == 1)
        spread_direct_kernel<T, W, 2, 1><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, 1, grid, es);
      else
        spread_direct_kernel<T, W, 2, 0><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, 1, grid, es);
    }
  } else {
    int nb = div_up(sp.npos, 256);
    if (cs)
      spread_direct_kernel<T, W, 3, 1><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, nf2, grid, es);
    else
      spread_direct_kernel<T, W, 3, 0><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, nf2, grid, es);
  }
  EFGP_CUDA(cudaGetLastError());
}

template <typename T>
static void launch_t(const SortedPoints& sp, const ESKernel& k, const T* cs, T* grid, cudaStream_t st,
                     bool from_record) {
  switch (k.W) {
#define EFGP_CASE(w) \
  case w:            \
    launch_w<T, w>(sp, k, cs, grid, st, from_record); \
    break;
    EFGP_CASE(2) EFGP_CASE(3) EFGP_CASE(4) EFGP_CASE(5) EFGP_CASE(6) EFGP_CASE(7) EFGP_CASE(8) EFGP_CASE(9)
    EFGP_CASE(10) EFGP_CASE(11) EFGP_CASE(12) EFGP_CASE(13) EFGP_CASE(14) EFGP_CASE(15) EFGP_CASE(16)
#undef EFGP_CASE
    default:
      throw Error(1, "unsupported spreading width");
  }
}

void launch_spread(const SortedPoints& sp, const ESKernel& k, const void* cs, void* grid, cudaStream_t st,
                   bool from_record) {
  if (sp.N == 0) return;
  EFGP_CHECK(!from_record || (sp.has_payload && sp.dim == 2), "sorted point set carries no strength payload");
  EFGP_CHECK((k.W & 1) == sp.odd, "kernel width parity does not match the sorted point set");
  if (sp.dtype == 0)
    launch_t<float>(sp, k, (const float*)cs, (float*)grid, st, from_record);
  else
    launch_t<double>(sp, k, (const double*)cs, (double*)grid, st, from_record);
}

}  // namespace efgp
