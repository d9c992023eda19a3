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
#include "philox.cuh"
#include <optional>

namespace efgp {

template <typename T, int W>
__device__ __forceinline__ void es_eval(const ESDev<T>& es, T s, T* __restrict__ k) {
  constexpr int DEG = es_degree(W);
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

// Deterministic mode (efgp_set_deterministic): every (point, grid cell) contribution is added as a fixed-point integer.
// Integer addition is associative, so the grid is BIT-IDENTICAL from run to run whatever order the atomics land in and
// whatever order the sort left the points of a cell in (the default mode sums runs of points in floating point and
// flushes with fp64 REDs: the last bits, and with them CG iteration counts at tight tolerances, vary).  Only the
// per-point kernels (spread_direct_kernel) implement it; launch_w routes every geometry there in this mode.
// det.scale = 2^k with N * max|c| * 2^k < 2^62: no overflow; resolution 2^-k of the largest possible cell value.
struct DetSpec {
  unsigned long long* ig;   // integer grid (same geometry), nullptr = floating-point mode
  double scale;
};
template <typename T>
__device__ __forceinline__ void grid_add(T* __restrict__ grid, int64_t idx, T v, const DetSpec& det) {
  if (det.ig)
    atomicAdd(det.ig + idx, (unsigned long long)__double2ll_rn((double)v * det.scale));
  else
    atomicAdd(grid + idx, v);
}

__device__ __forceinline__ int wrap_idx(int i, int n) {
  if (i < 0) i += n;
  if (i >= n) i -= n;
  return i;
}

// Slot addressing of the tile-transposed sorted layout (nufft.cuh): thread t of tile `blockIdx.x`
// reads its j-th point at base + j*kTileThreads.
// ------------------------------------------------------------------ 1-D
template <typename T, int W, int HAS_C>
__global__ void __launch_bounds__(kTileThreads)
spread1d_kernel(const T* __restrict__ rec, const T* __restrict__ cs, int P, int nf0, T* __restrict__ grid,
                const __grid_constant__ ESDev<T> es, ProbeSpec pr, DetSpec det) {
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
    T cv = HAS_C == 1 ? cs[pos] : T(1);
    if (HAS_C == 3) {
      uint32_t bits[4];
      philox_bits(pr.seed, probe_counter<T>(pr.use_orig ? rec[pos * 2 + 1] : T(0), pos, pr.use_orig), (uint32_t)(pr.t0 >> 7), bits);
      cv = philox_sign_from_bits<T>(bits, pr.t0);
    }
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

// HAS_C: 0 = all-ones strengths, 1 = strengths in slot order (cs), 2 = strengths in the records' payload word,
// 3 = Rademacher probes regenerated from (seed, probe, slot) (philox.cuh).  NB > 1 (mode 3 only): NB probes in ONE pass
// over the points -- the record is read and the 2 W kernel weights are evaluated once per point, then applied to NB
// register footprints (grid b = grid + b * nf0 * nf1), which is what makes a Hutchinson batch cheaper than NB transforms.
template <int W, int NB>
__host__ __device__ constexpr int spread2d_min_ctas() { return (W * W * NB <= 25) ? 2 : 1; }

template <typename T, int W, int HAS_C, int NB>
__global__ void __launch_bounds__(kTileThreads, (spread2d_min_ctas<W, NB>()))
spread2d_kernel(const T* __restrict__ rec, const T* __restrict__ cs, int P, int nf0,
                int nf1, T* __restrict__ grid, const __grid_constant__ ESDev<T> es, ProbeSpec pr, DetSpec det) {
  static_assert(NB == 1 || HAS_C == 3, "multi-RHS spreading is for regenerated probes");
  constexpr int S = kRingStages;
  __shared__ __align__(16) T ring[S][kTileThreads][2];   // (X0, X1) of each thread's next points
  __shared__ T cring[HAS_C ? S : 1][kTileThreads];      // strengths, or (mode 3) the record word with the original index
  const int t = threadIdx.x;
  const int64_t base = (int64_t)blockIdx.x * kTileThreads * P + t;
  auto issue = [&](int j) {
    if (j < P) {
      const int64_t pos = base + (int64_t)j * kTileThreads;
      const int s = j % S;
      cp_async_word<2 * sizeof(T)>(&ring[s][t][0], rec + pos * 4);   // one 16-byte (8 for f32) copy per point
      if (HAS_C == 1) cp_async_word<sizeof(T)>(&cring[s][t], cs + pos);
      if (HAS_C == 2) cp_async_word<sizeof(T)>(&cring[s][t], rec + pos * 4 + 3);
      if (HAS_C == 3) cp_async_word<sizeof(T)>(&cring[s][t], rec + pos * 4 + 2);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int j = 0; j < S - 1; ++j) issue(j);
  T acc[NB][W][W];
#pragma unroll
  for (int b = 0; b < NB; ++b)
#pragma unroll
    for (int i = 0; i < W; ++i)
#pragma unroll
      for (int j = 0; j < W; ++j) acc[b][i][j] = 0;
  const int64_t cells = (int64_t)nf0 * nf1;
  int cur0 = -1, cur1 = -1;
  for (int j = 0; j < P; ++j) {
    issue(j + S - 1);
    cp_async_wait<S - 1>();
    const int s = j % S;
    const T X0 = ring[s][t][0];
    if (X0 < T(0)) break;
    const T X1 = ring[s][t][1];
    T cv[NB];
#pragma unroll
    for (int b = 0; b < NB; ++b) cv[b] = (HAS_C == 1 || HAS_C == 2) ? cring[s][t] : T(1);
    if (HAS_C == 3) {
      uint32_t bits[4];
      philox_bits(pr.seed, probe_counter<T>(cring[s][t], base + (int64_t)j * kTileThreads, pr.use_orig),
                  (uint32_t)(pr.t0 >> 7), bits);
#pragma unroll
      for (int b = 0; b < NB; ++b) cv[b] = philox_sign_from_bits<T>(bits, pr.t0 + b);
    }
    int c0 = (int)X0, c1 = (int)X1;
    if (c0 != cur0 || c1 != cur1) {
      if (cur0 >= 0) {
#pragma unroll
        for (int b = 0; b < NB; ++b) flush2d<T, W>(acc[b], cur0, cur1, nf0, nf1, grid + b * cells);
      }
      cur0 = c0;
      cur1 = c1;
    }
    T k0[W], k1[W];
    es_eval<T, W>(es, T(2) * (X0 - (T)c0) - T(1), k0);
    es_eval<T, W>(es, T(2) * (X1 - (T)c1) - T(1), k1);
#pragma unroll
    for (int b = 0; b < NB; ++b) {
#pragma unroll
      for (int i = 0; i < W; ++i) {
        T a = cv[b] * k0[i];
#pragma unroll
        for (int jj = 0; jj < W; ++jj) acc[b][i][jj] = fma(a, k1[jj], acc[b][i][jj]);
      }
    }
  }
  cp_async_wait<0>();
  if (cur0 >= 0) {
#pragma unroll
    for (int b = 0; b < NB; ++b) flush2d<T, W>(acc[b], cur0, cur1, nf0, nf1, grid + b * cells);
  }
}

// ------------------------------------------------------------------ 2-D, moment accumulation (dense point sets, fp64)
// The F^*1 spread at W = 9 is bound by the FP64 pipe, not by HBM: per point the register-footprint kernel above spends
// ~80 DFMA on the two piecewise polynomials and W^2 = 81 on the outer product (163 DFMA in the SASS; 1e8 points at the
// measured ~37 TFLOP/s DFMA peak are 0.88 ms, the kernel takes 1.25 ms).  The weights of a point are polynomials of
// its in-cell offsets (s, t) in [-1, 1):  k0_i(s) = sum_a c_i[a] s^a,  k1_j(t) = sum_b c_j[b] t^b,  so the footprint of
// a run of points that share a cell is
//     acc[i][j] = sum_p y_p k0_i(s_p) k1_j(t_p) = sum_{a,b} c_i[a] c_j[b] M[a][b],   M[a][b] = sum_p y_p s_p^a t_p^b.
// The kernel below accumulates the NC x NC MOMENTS M (NC = degree + 1 <= W) instead of the footprint: 2 (NC - 1)
// multiplications for the powers and NC^2 FMAs per point -- 97 instead of 163 FP64 instructions at W = 9 -- and no
// polynomial evaluation at all.  On a cell change the moments are added (RED.ADD.F64, NC^2 consecutive doubles) to a
// per-cell moment grid; moment_to_grid_kernel then applies C (x) C once per CELL (2 NC^2 W FMAs per cell row, ~3e7
// FMAs for a 144^2 grid against 1.6e10 saved) and adds the footprints to the fine grid.  Monomials on [-1, 1] at
// degree <= 8 are well conditioned for this purpose (the coefficients are O(1), the weights are in [0, 1]); the
// transform error stays at the kernel's own aliasing floor (tests/test_ops_gpu.py bars unchanged).
// Used when the points are dense on the grid (launch_w): >= kMomentMinDensity points per cell on average.
constexpr int kMomentMinDensity = 48;
constexpr size_t kMomentMaxBytes = (size_t)192 << 20;

template <int W, int HAS_C, int S>
__global__ void __launch_bounds__(kTileThreads, 1)
spread2d_moment_kernel(const double* __restrict__ rec, const double* __restrict__ cs, int P, int nf0, int nf1,
                       double* __restrict__ mom) {
  using T = double;
  constexpr int NC = es_degree(W) + 1;
  __shared__ __align__(16) T ring[S][kTileThreads][2];
  __shared__ T cring[HAS_C ? S : 1][kTileThreads];
  const int t = threadIdx.x;
  const int64_t base = (int64_t)blockIdx.x * kTileThreads * P + t;
  auto issue = [&](int j, int s) {     // point j into ring slot s (= j mod S, tracked by the caller: no division)
    if (j < P) {
      const int64_t pos = base + (int64_t)j * kTileThreads;
      cp_async_word<2 * sizeof(T)>(&ring[s][t][0], rec + pos * 4);
      if (HAS_C == 1) cp_async_word<sizeof(T)>(&cring[s][t], cs + pos);
      if (HAS_C == 2) cp_async_word<sizeof(T)>(&cring[s][t], rec + pos * 4 + 3);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int j = 0; j < S; ++j) issue(j, j);
  T M[NC][NC];
#pragma unroll
  for (int a = 0; a < NC; ++a)
#pragma unroll
    for (int b = 0; b < NC; ++b) M[a][b] = 0;
  int cur0 = -1, cur1 = -1;
  auto flush = [&]() {
    T* m = mom + ((int64_t)cur0 * nf1 + cur1) * (NC * NC);
#pragma unroll
    for (int a = 0; a < NC; ++a)
#pragma unroll
      for (int b = 0; b < NC; ++b) {
        atomicAdd(m + a * NC + b, M[a][b]);
        M[a][b] = 0;
      }
  };
  // cell and powers of the point in ring slot s (c0 = -1 for the padding sentinel X0 = -1 and past the run's end)
  auto fetch = [&](int j, int s, int& c0, int& c1, T (&pa)[NC], T (&pb)[NC]) {
    const T X0 = j < P ? ring[s][t][0] : T(-1);
    const T X1 = ring[s][t][1];
    c0 = (int)X0;
    c1 = (int)X1;
    const T u = T(2) * (X0 - (T)c0) - T(1), v = T(2) * (X1 - (T)c1) - T(1);
    // powers with a shallow dependency tree: x^k = x^(k/2) * x^(k - k/2)
    T up[NC], vp[NC];
    up[0] = vp[0] = T(1);
    if (NC > 1) { up[1] = u; vp[1] = v; }
#pragma unroll
    for (int k = 2; k < NC; ++k) {
      up[k] = up[k / 2] * up[k - k / 2];
      vp[k] = vp[k / 2] * vp[k - k / 2];
    }
    pa[0] = HAS_C ? cring[s][t] : T(1);
    pb[0] = T(1);
#pragma unroll
    for (int k = 1; k < NC; ++k) {
      pa[k] = HAS_C ? pa[0] * up[k] : up[k];
      pb[k] = vp[k];
    }
  };
  // Software pipeline: the dependent chain of point j + 1 (LDS, conversions, the power tree) is issued in the same
  // basic block as the NC^2 independent FMAs of point j, so that with only two warps per scheduler (200 registers per
  // thread) the FP64 pipe does not idle while both warps walk their chains (ncu before this: FP64 pipe 60 % active,
  // top stall "wait").  Two register sets, the loop is unrolled by hand so that they swap roles without copies.
  T paA[NC], pbA[NC], paB[NC], pbB[NC];
  int a0, a1, b0, b1;
  cp_async_wait<S - 1>();
  fetch(0, 0, a0, a1, paA, pbA);
  int sc = 0;    // ring slot of the current point
  auto step = [&](int j, int c0, int c1, const T (&pa)[NC], const T (&pb)[NC], int& n0, int& n1, T (&npa)[NC],
                  T (&npb)[NC]) {
    issue(j + S, sc);                 // the current point's slot is free: its values are in registers
    cp_async_wait<S - 1>();           // point j + 1 has landed
    if (c0 != cur0 || c1 != cur1) {
      if (cur0 >= 0) flush();
      cur0 = c0;
      cur1 = c1;
    }
    sc = sc + 1 == S ? 0 : sc + 1;
    fetch(j + 1, sc, n0, n1, npa, npb);
#pragma unroll
    for (int a = 0; a < NC; ++a)
#pragma unroll
      for (int b = 0; b < NC; ++b) M[a][b] = fma(pa[a], pb[b], M[a][b]);
  };
  for (int j = 0; j < P; j += 2) {
    if (a0 < 0) break;
    step(j, a0, a1, paA, pbA, b0, b1, paB, pbB);
    if (b0 < 0) break;
    step(j + 1, b0, b1, paB, pbB, a0, a1, paA, pbA);
  }
  cp_async_wait<0>();
  if (cur0 >= 0) flush();
}

// The generator F^*1 (width WG, all-ones strengths) and the right-hand side F^*y (width WR <= WG, strengths in the
// records' payload word) in ONE pass over the records: both kernels are polynomials of the SAME in-cell offsets
// (odd widths share the cell and its centre), so the powers are computed once and feed two moment sets,
// MG[a][b] += u^a v^b (NCG^2) and MR[a][b] += y u^a v^b (NCR^2).  The record stream (3.2 GB at N = 1e8) is read once
// instead of twice, and F^*y costs NCR^2 + NCR extra FP64 instructions per point instead of its own HBM-bound pass
// (0.54 ms).  106 accumulators at (9, 5): the powers of u are produced row by row to stay inside 255 registers.
template <int WG, int WR, int S>
__global__ void __launch_bounds__(kTileThreads, 1)
spread2d_moment_pair_kernel(const double* __restrict__ rec, int P, int nf1, double* __restrict__ momG,
                            double* __restrict__ momR) {
  using T = double;
  constexpr int NG = es_degree(WG) + 1, NR = es_degree(WR) + 1;
  static_assert(NR <= NG, "the right-hand-side kernel is the narrower one");
  __shared__ __align__(16) T ring[S][kTileThreads][2];
  __shared__ T yring[S][kTileThreads];
  const int t = threadIdx.x;
  const int64_t base = (int64_t)blockIdx.x * kTileThreads * P + t;
  auto issue = [&](int j, int s) {
    if (j < P) {
      const int64_t pos = base + (int64_t)j * kTileThreads;
      cp_async_word<2 * sizeof(T)>(&ring[s][t][0], rec + pos * 4);
      cp_async_word<sizeof(T)>(&yring[s][t], rec + pos * 4 + 3);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int j = 0; j < S - 1; ++j) issue(j, j);
  T MG[NG][NG], MR[NR][NR];
#pragma unroll
  for (int a = 0; a < NG; ++a)
#pragma unroll
    for (int b = 0; b < NG; ++b) MG[a][b] = 0;
#pragma unroll
  for (int a = 0; a < NR; ++a)
#pragma unroll
    for (int b = 0; b < NR; ++b) MR[a][b] = 0;
  int cur0 = -1, cur1 = -1;
  auto flush = [&]() {
    const int64_t cell = (int64_t)cur0 * nf1 + cur1;
    T* mg = momG + cell * (NG * NG);
#pragma unroll
    for (int a = 0; a < NG; ++a)
#pragma unroll
      for (int b = 0; b < NG; ++b) {
        atomicAdd(mg + a * NG + b, MG[a][b]);
        MG[a][b] = 0;
      }
    T* mr = momR + cell * (NR * NR);
#pragma unroll
    for (int a = 0; a < NR; ++a)
#pragma unroll
      for (int b = 0; b < NR; ++b) {
        atomicAdd(mr + a * NR + b, MR[a][b]);
        MR[a][b] = 0;
      }
  };
  int s = 0, sn = S - 1;     // ring slots of point j and of point j + S - 1
  for (int j = 0; j < P; ++j) {
    issue(j + S - 1, sn);
    cp_async_wait<S - 1>();
    const T X0 = ring[s][t][0];
    if (X0 < T(0)) break;
    const T X1 = ring[s][t][1];
    const T y = yring[s][t];
    s = s + 1 == S ? 0 : s + 1;
    sn = sn + 1 == S ? 0 : sn + 1;
    const int c0 = (int)X0, c1 = (int)X1;
    if (c0 != cur0 || c1 != cur1) {
      if (cur0 >= 0) flush();
      cur0 = c0;
      cur1 = c1;
    }
    const T u = T(2) * (X0 - (T)c0) - T(1), v = T(2) * (X1 - (T)c1) - T(1);
    T vp[NG];
    vp[0] = T(1);
    vp[1] = v;
#pragma unroll
    for (int k = 2; k < NG; ++k) vp[k] = vp[k / 2] * vp[k - k / 2];
    T ua = T(1);
#pragma unroll
    for (int a = 0; a < NG; ++a) {
      if (a == 1) ua = u;
      if (a > 1) ua *= u;
#pragma unroll
      for (int b = 0; b < NG; ++b) MG[a][b] = fma(ua, vp[b], MG[a][b]);
      if (a < NR) {
        const T ya = y * ua;
#pragma unroll
        for (int b = 0; b < NR; ++b) MR[a][b] = fma(ya, vp[b], MR[a][b]);
      }
    }
  }
  cp_async_wait<0>();
  if (cur0 >= 0) flush();
}

// Per-cell moments -> fine grid.  A CTA takes a segment of kM2GCells consecutive cells of one grid row c0:
//  (1) their moments (NC^2 contiguous doubles per cell) are staged into shared memory with coalesced loads;
//  (2) warp i computes footprint row i of every cell of the segment: T_i[b] = sum_a c_i[a] M[a][b], then
//      F[i][j] = sum_b T_i[b] c_j[b]  (c_{W-1-i}[a] = (-1)^a c_i[a]: es_eval's symmetry, table expanded in shared memory);
//  (3) the footprints of neighbouring cells overlap in all but one column, so they are summed inside the CTA first:
//      out[i][col] = sum_j F_{cell = col - j}[i][j], and ONE RED per (row, column) of the segment goes to the grid --
//      (kM2GCells + W - 1) W REDs per CTA instead of kM2GCells W^2.  (The first version issued every footprint entry as
//      its own RED: 6.7e6 REDs, 81 per grid address, 51 us for a 288^2 grid, whatever the load path.)
constexpr int kM2GCells = 32;

template <int W>
__global__ void __launch_bounds__(kM2GCells * W)
moment_to_grid_kernel(const double* __restrict__ mom, int nf0, int nf1, double* __restrict__ grid,
                      const __grid_constant__ ESDev<double> es) {
  constexpr int NC = es_degree(W) + 1;
  constexpr int NM = NC * NC;
  constexpr int PM = (NM & 1) ? NM : NM + 1;     // odd pitch: the 32 cells of a warp land on different banks
  constexpr int OFF = W / 2 - ((W & 1) ? 0 : 1);
  constexpr int HALF = (W + 1) / 2;
  constexpr int NT = kM2GCells * W;
  __shared__ double cf[W][NC + 1];
  __shared__ double sm[kM2GCells * PM];
  __shared__ double Fs[W * W][kM2GCells + 1];
  const int segs = (nf1 + kM2GCells - 1) / kM2GCells;
  const int c0 = blockIdx.x / segs, c1b = (blockIdx.x - c0 * segs) * kM2GCells;
  const int nc = nf1 - c1b < kM2GCells ? nf1 - c1b : kM2GCells;
  for (int q = threadIdx.x; q < W * NC; q += NT) {
    const int i = q / NC, a = q - i * NC;
    const int ii = i < HALF ? i : W - 1 - i;
    const double c = es.c[ii][a];
    cf[i][a] = (i >= HALF && (a & 1)) ? -c : c;
  }
  const double* src = mom + ((int64_t)c0 * nf1 + c1b) * NM;
  for (int q = threadIdx.x; q < nc * NM; q += NT) {
    const int lc = q / NM;
    sm[lc * PM + (q - lc * NM)] = src[q];
  }
  __syncthreads();
  const int i = threadIdx.x >> 5, lc = threadIdx.x & 31;      // warp i: footprint row i
  if (lc < nc) {
    const double* m = sm + lc * PM;
    double Tb[NC];
#pragma unroll
    for (int b = 0; b < NC; ++b) Tb[b] = 0;
#pragma unroll
    for (int a = 0; a < NC; ++a) {
      const double ca = cf[i][a];
#pragma unroll
      for (int b = 0; b < NC; ++b) Tb[b] = fma(ca, m[a * NC + b], Tb[b]);
    }
#pragma unroll
    for (int j = 0; j < W; ++j) {
      double f = 0;
#pragma unroll
      for (int b = 0; b < NC; ++b) f = fma(Tb[b], cf[j][b], f);
      Fs[i * W + j][lc] = f;
    }
  }
  __syncthreads();
  double* row = grid + (int64_t)wrap_idx(c0 - OFF + i, nf0) * nf1;
  for (int col = lc; col < nc + W - 1; col += 32) {
    double v = 0;
#pragma unroll
    for (int j = 0; j < W; ++j) {
      const int cc = col - j;
      if (cc >= 0 && cc < nc) v += Fs[i * W + j][cc];
    }
    if (v != 0.0) atomicAdd(row + wrap_idx(c1b - OFF + col, nf1), v);
  }
}

// ------------------------------------------------------------------ 3-D, shared-memory tile accumulation
// Points arrive tile-major (points.cu: 8^3-cell tiles).  A CTA takes a chunk of consecutive sorted points and
// accumulates them into a shared-memory copy of the current tile plus its W-1 halo ((8+W-1)^3 values);
// when the tile changes (or the chunk ends) the copy is added to the global grid with one RED per NONZERO
// entry -- ~2e3-5e3 REDs per tile visit instead of W^3 (125..729) REDs per point.
//
// No shared-memory atomics (sm_100a has no native fp64 shared atomic): the 4 warps own disjoint x-planes
// of the shared tile (plane s0 belongs to warp s0 & 3), every warp walks all points of the batch and adds
// only the planes it owns, lanes covering the W x W cells of a plane.  Plain LDS/DFMA/STS, no races, no
// barriers inside a batch.  Kernel weights for a batch of 32 points are evaluated once by 96 threads.
constexpr int kT3Threads = 128;
constexpr int kT3Batch = 32;
constexpr int kT3Chunk = 1024;

template <typename T, int W>
__host__ __device__ constexpr size_t tile3_smem_bytes() {
  constexpr int E = kTile3 + W - 1;
  constexpr int EP = E | 1;   // odd row stride: rows of a plane fall in different shared-memory banks
  return sizeof(T) * (size_t)(E * E * EP + kT3Batch * 3 * W + kT3Batch) + sizeof(int) * kT3Batch * 4;
}

template <typename T, int W, int HAS_C>
__global__ void __launch_bounds__(kT3Threads)
spread3d_tile_kernel(const T* __restrict__ rec, const T* __restrict__ cs, int64_t N, int nf0, int nf1, int nf2,
                     T* __restrict__ grid, const __grid_constant__ ESDev<T> es, ProbeSpec pr, DetSpec det) {
  constexpr int E = kTile3 + W - 1;
  constexpr int EP = E | 1;   // padded (odd) row stride
  constexpr int OFF = W / 2 - ((W & 1) ? 0 : 1);
  extern __shared__ __align__(16) unsigned char smem_raw[];
  int* BI = reinterpret_cast<int*>(smem_raw);   // [batch][4]: tile id, cell0, cell1, cell2 (16-byte aligned)
  T* S = reinterpret_cast<T*>(smem_raw + sizeof(int) * kT3Batch * 4);   // [E][E][EP] tile + halo
  T* KW = S + E * E * EP;                       // [batch][3][W] kernel weights
  T* CV = KW + kT3Batch * 3 * W;                // [batch] strengths
  __shared__ unsigned segmask;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nt1 = (nf1 + kTile3 - 1) / kTile3, nt2 = (nf2 + kTile3 - 1) / kTile3;
  const int64_t p0 = (int64_t)blockIdx.x * kT3Chunk;
  const int64_t p1 = p0 + kT3Chunk < N ? p0 + kT3Chunk : N;
  for (int e = tid; e < E * E * EP; e += kT3Threads) S[e] = T(0);
  int cur_tile = -1, tb0 = 0, tb1 = 0, tb2 = 0;
  // the (j, l) cells of a W x W plane this lane covers: fixed for the whole kernel
  constexpr int R = (W * W + 31) / 32;
  int cj[R], cl[R], cofs[R];
#pragma unroll
  for (int rr = 0; rr < R; ++rr) {
    int r = lane + 32 * rr;
    if (r >= W * W) r = 0;
    cj[rr] = r / W;
    cl[rr] = r - cj[rr] * W;
    cofs[rr] = cj[rr] * EP + cl[rr];
  }

  auto flush = [&]() {   // add the tile copy to the grid and clear it (same thread reads and clears an entry)
    for (int e = tid; e < E * E * EP; e += kT3Threads) {
      T v = S[e];
      if (v != T(0)) {   // (the pad column is never written, so it stays zero)
        int e0 = e / (E * EP), e1 = (e / EP) % E, e2 = e % EP;
        int64_t gi = ((int64_t)wrap_idx(tb0 - OFF + e0, nf0) * nf1 + wrap_idx(tb1 - OFF + e1, nf1)) * nf2 +
                     wrap_idx(tb2 - OFF + e2, nf2);
        atomicAdd(grid + gi, v);
        S[e] = T(0);
      }
    }
  };

  for (int64_t b0 = p0; b0 < p1; b0 += kT3Batch) {
    const int nb = (int)(p1 - b0 < kT3Batch ? p1 - b0 : kT3Batch);
    __syncthreads();   // previous batch fully consumed (KW / BI reuse), S zeroing visible
    if (tid < 3 * nb) {
      const int pt = tid / 3, ax = tid - 3 * pt;
      const T X = rec[(b0 + pt) * 4 + ax];
      const int c = (int)X;
      T kk[W];
      es_eval<T, W>(es, T(2) * (X - (T)c) - T(1), kk);
#pragma unroll
      for (int i = 0; i < W; ++i) KW[(pt * 3 + ax) * W + i] = kk[i];
      BI[pt * 4 + 1 + ax] = c;
    } else if (tid >= 96 && tid < 96 + nb) {
      T cvv = HAS_C == 1 ? cs[b0 + tid - 96] : T(1);
      if (HAS_C == 3) {                      // 3-D sets keep rank order (P = 1): the slot is the sorted rank
        uint32_t bits[4];
        philox_bits(pr.seed, probe_counter<T>(pr.use_orig ? rec[(b0 + tid - 96) * 4 + 3] : T(0), b0 + tid - 96, pr.use_orig),
                    (uint32_t)(pr.t0 >> 7), bits);
        cvv = philox_sign_from_bits<T>(bits, pr.t0);
      }
      CV[tid - 96] = cvv;
    }
    __syncthreads();
    if (tid < 32) {            // tile id per point, and a bit mask of the points that START a new tile run
      int t = -2;
      if (tid < nb) {
        t = ((BI[tid * 4 + 1] / kTile3) * nt1 + BI[tid * 4 + 2] / kTile3) * nt2 + BI[tid * 4 + 3] / kTile3;
        BI[tid * 4] = t;
      }
      int tprev = __shfl_up_sync(0xffffffffu, t, 1);
      unsigned m = __ballot_sync(0xffffffffu, tid == 0 || t != tprev);
      if (tid == 0) segmask = m;
    }
    __syncthreads();
    const unsigned starts = segmask;
    int p = 0;
    while (p < nb) {           // uniform control flow: every thread reads the same shared values
      const int t = BI[p * 4];
      if (t != cur_tile) {
        if (cur_tile >= 0) {
          __syncthreads();     // all warps done with the old tile
          flush();
          __syncthreads();
        }
        cur_tile = t;
        tb2 = (t % nt2) * kTile3;
        tb1 = ((t / nt2) % nt1) * kTile3;
        tb0 = (t / (nt2 * nt1)) * kTile3;
      }
      const unsigned later = p + 1 < 32 ? (starts >> (p + 1)) << (p + 1) : 0u;   // run starts after p
      int q = later ? __ffs(later) - 1 : nb;
      if (q > nb) q = nb;
      for (int pt = p; pt < q; ++pt) {
        const int4 bi = *reinterpret_cast<const int4*>(BI + pt * 4);
        const int l0 = bi.y - tb0;
        const int ifirst = (warp - l0) & 3;       // first x-plane of this footprint owned by this warp
        if (ifirst >= W) continue;
        const T cv = CV[pt];
        const T* k0 = KW + (pt * 3 + 0) * W;
        const T* k1 = KW + (pt * 3 + 1) * W;
        const T* k2 = KW + (pt * 3 + 2) * W;
        T wjl[R];
#pragma unroll
        for (int rr = 0; rr < R; ++rr) wjl[rr] = cv * k1[cj[rr]] * k2[cl[rr]];
        T* base = S + ((l0 * E + (bi.z - tb1)) * EP + (bi.w - tb2));
#pragma unroll 1
        for (int i = ifirst; i < W; i += 4) {
          const T a = k0[i];
          T* plane = base + i * (E * EP);
#pragma unroll
          for (int rr = 0; rr < R; ++rr)
            if (rr + 1 < R || lane + 32 * rr < W * W) plane[cofs[rr]] = fma(a, wjl[rr], plane[cofs[rr]]);
        }
      }
      p = q;
    }
  }
  if (cur_tile >= 0) {
    __syncthreads();
    flush();
  }
}

// ------------------------------------------------------------------ generic d (2-D wide kernels, 3-D): per-point REDs
// Points are still cell-sorted, so the reductions of neighbouring slots land in the same L2 lines.
template <typename T, int W, int D, int HAS_C>
__global__ void __launch_bounds__(256)
spread_direct_kernel(const T* __restrict__ rec,
                     const T* __restrict__ cs, int64_t npos, int nf0, int nf1, int nf2, T* __restrict__ grid,
                     const __grid_constant__ ESDev<T> es, ProbeSpec pr, DetSpec det) {
  int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= npos) return;
  constexpr int OFF = W / 2 - ((W & 1) ? 0 : 1);
  T Xr[4] = {T(0), T(0), T(0), T(0)};
  if constexpr (D == 1) {
    Xr[0] = rec[n * 2];
    Xr[1] = rec[n * 2 + 1];
  } else {
    load_record4<T>(rec + n * 4, Xr);
  }
  T X0 = Xr[0];
  if (X0 < T(0)) return;
  T cv = HAS_C == 1 ? cs[n] : (HAS_C == 2 ? Xr[3] : T(1));
  if (HAS_C == 3) {
    uint32_t bits[4];
    philox_bits(pr.seed, probe_counter<T>(Xr[D], n, pr.use_orig), (uint32_t)(pr.t0 >> 7), bits);
    cv = philox_sign_from_bits<T>(bits, pr.t0);
  }
  T k0[W], k1[W], k2[W];
  int c0 = (int)X0;
  es_eval<T, W>(es, T(2) * (X0 - (T)c0) - T(1), k0);
  if constexpr (D == 1) {
#pragma unroll
    for (int i = 0; i < W; ++i) grid_add<T>(grid, wrap_idx(c0 - OFF + i, nf0), cv * k0[i], det);
    return;
  }
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
      const int64_t row = (int64_t)wrap_idx(c0 - OFF + i, nf0) * nf1;
      T a = cv * k0[i];
#pragma unroll
      for (int j = 0; j < W; ++j) grid_add<T>(grid, row + wrap_idx(c1 - OFF + j, nf1), a * k1[j], det);
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
        const int64_t row = (r0 + wrap_idx(c1 - OFF + j, nf1)) * nf2;
        T b = a * k1[j];
#pragma unroll
        for (int l = 0; l < W; ++l) grid_add<T>(grid, row + cols[l], b * k2[l], det);
      }
    }
  }
}

// Probes one launch may carry for this geometry: the 2-D register kernels hold NB footprints (W^2 NB doubles).
int spread_probe_batch(int dim, int W, int dtype) {
  if (dim != 2) return 1;
  const int reg_max = dtype == 1 ? 9 : 10;
  if (W > reg_max) return 1;
  if (W <= 5) return 4;
  if (W <= 7) return 2;
  return 1;
}

template <typename T, int W>
static void launch_w(const SortedPoints& sp, const ESKernel& k, const T* cs, T* grid, cudaStream_t st, int mode,
                     const ProbeSpec& pr, const DetSpec& det) {
  ESDev<T> es;
  es_to_dev<T>(k, es);
  const T* rec = (const T*)sp.rec;
  constexpr int kRegMaxW = sizeof(T) == 8 ? 9 : 10;
  const int nt = (int)sp.ntiles;
  char nm[56];
  snprintf(nm, sizeof(nm), "spread%dd_w%d_%s%s", sp.dim, W, sizeof(T) == 8 ? "f64" : "f32",
           mode == 0 ? "_ones" : (mode == 2 ? "_rec" : (mode == 3 ? (pr.nb == 4 ? "_probes4" : (pr.nb == 2 ? "_probes2" : "_probes1")) : "")));
  std::optional<ScopedStage> stg_holder;
  stg_holder.emplace(nm, st, 1, sp.N * (mode == 3 ? pr.nb : 1));
  struct { std::optional<ScopedStage>& h; void add_launches(int n) { h->add_launches(n); } void reset() { h.reset(); } }
      stg{stg_holder};
  const int nf0 = (int)sp.nf[0], nf1 = (int)sp.nf[1], nf2 = (int)sp.nf[2];
  // Deterministic mode: one fixed-point RED per (point, grid cell).  Register / shared-memory footprints would sum a
  // run of points in floating point first, in the (run-dependent) order the sort left them inside a cell; per-point
  // integer contributions are order-independent.  Slow by design: a parity / debug mode.
  const bool per_point = det.ig != nullptr;
  if (per_point && sp.dim <= 2) {
    EFGP_CHECK(mode != 3 || pr.nb == 1, "one probe per launch in deterministic mode");
    const int nb = div_up(sp.npos, 256);
#define EFGP_DIRECT(D_)                                                                                          \
    if (mode == 3) spread_direct_kernel<T, W, D_, 3><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, 1, grid, es, pr, det); \
    else if (mode == 2) spread_direct_kernel<T, W, D_, 2><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, 1, grid, es, pr, det); \
    else if (mode == 1) spread_direct_kernel<T, W, D_, 1><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, 1, grid, es, pr, det); \
    else spread_direct_kernel<T, W, D_, 0><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, 1, grid, es, pr, det);
    if (sp.dim == 1) { EFGP_DIRECT(1) } else { EFGP_DIRECT(2) }
#undef EFGP_DIRECT
    EFGP_CUDA(cudaGetLastError());
    return;
  }
  if (sp.dim == 1) {
    if (mode == 3)
      spread1d_kernel<T, W, 3><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, grid, es, pr, det);
    else if (mode == 1)
      spread1d_kernel<T, W, 1><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, grid, es, pr, det);
    else
      spread1d_kernel<T, W, 0><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, grid, es, pr, det);
  } else if (sp.dim == 2) {
    if constexpr (sizeof(T) == 8 && W >= 5 && W <= 9) {
      // dense point sets: accumulate per-cell moments, transform once per cell (see spread2d_moment_kernel)
      static const int mom_env = getenv("EFGP_SPREAD_MOMENTS") ? atoi(getenv("EFGP_SPREAD_MOMENTS")) : -1;
      constexpr int NC = es_degree(W) + 1;
      const int64_t cells = (int64_t)nf0 * nf1;
      const size_t mbytes = sizeof(double) * (size_t)cells * NC * NC;
      const bool dense = sp.N >= (int64_t)kMomentMinDensity * cells && mbytes <= kMomentMaxBytes;
      const bool want = mom_env < 0 ? (W >= 7 && dense) : (mom_env > 0 && mom_env <= W && dense);
      if (mode != 3 && want) {
        double* mom = (double*)pool_alloc(mbytes);
        EFGP_CUDA(cudaMemsetAsync(mom, 0, mbytes, st));
        const double* r8 = (const double*)rec;
        const double* c8 = (const double*)cs;
        if (mode == 2)
          spread2d_moment_kernel<W, 2, kRingStages><<<nt, kTileThreads, 0, st>>>(r8, c8, sp.P, nf0, nf1, mom);
        else if (mode == 1)
          spread2d_moment_kernel<W, 1, kRingStages><<<nt, kTileThreads, 0, st>>>(r8, c8, sp.P, nf0, nf1, mom);
        else
          spread2d_moment_kernel<W, 0, kRingStages><<<nt, kTileThreads, 0, st>>>(r8, c8, sp.P, nf0, nf1, mom);
        stg.add_launches(1);
        stg.reset();      // the per-cell transform is its own line in the stage report
        {
          ScopedStage stg2("type1.moment_to_grid", st, 1, 0);
          ESDev<double> es8;
          es_to_dev<double>(k, es8);
          moment_to_grid_kernel<W><<<nf0 * div_up(nf1, kM2GCells), kM2GCells * W, 0, st>>>(mom, nf0, nf1, (double*)grid, es8);
          EFGP_CUDA(cudaGetLastError());
        }
        pool_free(mom);
        return;
      }
    }
    if constexpr (W <= kRegMaxW) {
      if (mode == 3) {
        if constexpr (W <= 5) {
          if (pr.nb == 4) {
            spread2d_kernel<T, W, 3, 4><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, nf1, grid, es, pr, det);
            EFGP_CUDA(cudaGetLastError());
            return;
          }
        }
        if constexpr (W <= 7) {
          if (pr.nb == 2) {
            spread2d_kernel<T, W, 3, 2><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, nf1, grid, es, pr, det);
            EFGP_CUDA(cudaGetLastError());
            return;
          }
        }
        EFGP_CHECK(pr.nb == 1, "probe batch not supported by this spreading width");
        spread2d_kernel<T, W, 3, 1><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, nf1, grid, es, pr, det);
      } else if (mode == 2)
        spread2d_kernel<T, W, 2, 1><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, nf1, grid, es, pr, det);
      else if (mode == 1)
        spread2d_kernel<T, W, 1, 1><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, nf1, grid, es, pr, det);
      else
        spread2d_kernel<T, W, 0, 1><<<nt, kTileThreads, 0, st>>>(rec, cs, sp.P, nf0, nf1, grid, es, pr, det);
    } else {
      EFGP_CHECK(mode != 3 || pr.nb == 1, "probe batch not supported by this spreading width");
      int nb = div_up(sp.npos, 256);
      if (mode == 3)
        spread_direct_kernel<T, W, 2, 3><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, 1, grid, es, pr, det);
      else if (mode == 2)
        spread_direct_kernel<T, W, 2, 2><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, 1, grid, es, pr, det);
      else if (mode == 1)
        spread_direct_kernel<T, W, 2, 1><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, 1, grid, es, pr, det);
      else
        spread_direct_kernel<T, W, 2, 0><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, 1, grid, es, pr, det);
    }
  } else if (per_point || getenv("EFGP_3D_DIRECT")) {   // per-point REDs (deterministic mode; A/B measurements)
    EFGP_CHECK(mode != 3 || pr.nb == 1, "one probe per launch in 3-D");
    int nb = div_up(sp.npos, 256);
    if (mode == 3)
      spread_direct_kernel<T, W, 3, 3><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, nf2, grid, es, pr, det);
    else if (cs)
      spread_direct_kernel<T, W, 3, 1><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, nf2, grid, es, pr, det);
    else
      spread_direct_kernel<T, W, 3, 0><<<nb, 256, 0, st>>>(rec, cs, sp.npos, nf0, nf1, nf2, grid, es, pr, det);
  } else {
    EFGP_CHECK(mode != 3 || pr.nb == 1, "one probe per launch in 3-D");
    const size_t smem = tile3_smem_bytes<T, W>();
    int nb = div_up(sp.N, kT3Chunk);
    if (mode == 3) {
      EFGP_CUDA(cudaFuncSetAttribute(spread3d_tile_kernel<T, W, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      spread3d_tile_kernel<T, W, 3><<<nb, kT3Threads, smem, st>>>(rec, cs, sp.N, nf0, nf1, nf2, grid, es, pr, det);
    } else if (cs) {
      EFGP_CUDA(cudaFuncSetAttribute(spread3d_tile_kernel<T, W, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      spread3d_tile_kernel<T, W, 1><<<nb, kT3Threads, smem, st>>>(rec, cs, sp.N, nf0, nf1, nf2, grid, es, pr, det);
    } else {
      EFGP_CUDA(cudaFuncSetAttribute(spread3d_tile_kernel<T, W, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      spread3d_tile_kernel<T, W, 0><<<nb, kT3Threads, smem, st>>>(rec, cs, sp.N, nf0, nf1, nf2, grid, es, pr, det);
    }
  }
  EFGP_CUDA(cudaGetLastError());
}

template <typename T>
static void launch_t(const SortedPoints& sp, const ESKernel& k, const T* cs, T* grid, cudaStream_t st, int mode,
                     const ProbeSpec& pr, const DetSpec& det) {
  switch (k.W) {
#define EFGP_CASE(w) \
  case w:            \
    launch_w<T, w>(sp, k, cs, grid, st, mode, pr, det); \
    break;
    EFGP_CASE(2) EFGP_CASE(3) EFGP_CASE(4) EFGP_CASE(5) EFGP_CASE(6) EFGP_CASE(7) EFGP_CASE(8) EFGP_CASE(9)
    EFGP_CASE(10) EFGP_CASE(11) EFGP_CASE(12) EFGP_CASE(13) EFGP_CASE(14) EFGP_CASE(15) EFGP_CASE(16)
#undef EFGP_CASE
    default:
      throw Error(1, "unsupported spreading width");
  }
}

void launch_spread(const SortedPoints& sp, const ESKernel& k, const void* cs, void* grid, cudaStream_t st,
                   bool from_record, void* igrid, double iscale) {
  if (sp.N == 0) return;
  const DetSpec det{(unsigned long long*)igrid, iscale};
  EFGP_CHECK(!from_record || (sp.has_payload && sp.dim == 2), "sorted point set carries no strength payload");
  EFGP_CHECK((k.W & 1) == sp.odd, "kernel width parity does not match the sorted point set");
  const int mode = from_record ? 2 : (cs ? 1 : 0);
  ProbeSpec pr{0, 0, 1, 0};
  if (sp.dtype == 0)
    launch_t<float>(sp, k, (const float*)cs, (float*)grid, st, mode, pr, det);
  else
    launch_t<double>(sp, k, (const double*)cs, (double*)grid, st, mode, pr, det);
}

// ---- fused generator + right-hand side (nufft_type1_pair)
bool spread_pair_supported(const SortedPoints& sp, const ESKernel& kg, const ESKernel& kr) {
  static const bool off = getenv("EFGP_PAIR_FUSED") && atoi(getenv("EFGP_PAIR_FUSED")) == 0;
  if (off || sp.dim != 2 || sp.dtype != 1 || !sp.has_payload || !sp.odd || sp.N == 0) return false;
  if (!((kg.W == 9 || kg.W == 7) && kr.W == 5)) return false;
  const int64_t cells = sp.nf[0] * sp.nf[1];
  const int ng = es_degree(kg.W) + 1;
  return sp.N >= (int64_t)kMomentMinDensity * cells && sizeof(double) * (size_t)cells * ng * ng <= kMomentMaxBytes;
}

template <int WG, int WR>
static void launch_pair_w(const SortedPoints& sp, const ESKernel& kg, const ESKernel& kr, double* gridG, double* gridR,
                          cudaStream_t st) {
  constexpr int NG = es_degree(WG) + 1, NR = es_degree(WR) + 1;
  const int nf0 = (int)sp.nf[0], nf1 = (int)sp.nf[1];
  const int64_t cells = (int64_t)nf0 * nf1;
  const size_t bg = sizeof(double) * (size_t)cells * NG * NG, br = sizeof(double) * (size_t)cells * NR * NR;
  double* mom = (double*)pool_alloc(bg + br);
  double* momR = mom + (size_t)cells * NG * NG;
  {
    char nm[56];
    snprintf(nm, sizeof(nm), "spread2d_pair_w%d_w%d_f64", WG, WR);
    ScopedStage stg(nm, st, 2, sp.N);
    EFGP_CUDA(cudaMemsetAsync(mom, 0, bg + br, st));
    spread2d_moment_pair_kernel<WG, WR, kRingStages><<<(int)sp.ntiles, kTileThreads, 0, st>>>(
        (const double*)sp.rec, sp.P, nf1, mom, momR);
    EFGP_CUDA(cudaGetLastError());
  }
  {
    ScopedStage stg("type1.moment_to_grid", st, 2, 0);
    ESDev<double> eg, er;
    es_to_dev<double>(kg, eg);
    es_to_dev<double>(kr, er);
    moment_to_grid_kernel<WG><<<nf0 * div_up(nf1, kM2GCells), kM2GCells * WG, 0, st>>>(mom, nf0, nf1, gridG, eg);
    moment_to_grid_kernel<WR><<<nf0 * div_up(nf1, kM2GCells), kM2GCells * WR, 0, st>>>(momR, nf0, nf1, gridR, er);
    EFGP_CUDA(cudaGetLastError());
  }
  pool_free(mom);
}

// gridG += F^*1 footprints (kernel kg), gridR += F^*y footprints (kernel kr, y = the records' payload): one pass
void launch_spread_pair(const SortedPoints& sp, const ESKernel& kg, const ESKernel& kr, void* gridG, void* gridR,
                        cudaStream_t st) {
  EFGP_CHECK(spread_pair_supported(sp, kg, kr), "fused pair spreading does not support this geometry");
  if (kg.W == 9)
    launch_pair_w<9, 5>(sp, kg, kr, (double*)gridG, (double*)gridR, st);
  else
    launch_pair_w<7, 5>(sp, kg, kr, (double*)gridG, (double*)gridR, st);
}

// grids b = 0 .. pr.nb-1 (contiguous, nf^d each) += sum over points of Rademacher(seed, pr.t0 + b, slot) psi(l - X)
void launch_spread_probes(const SortedPoints& sp, const ESKernel& k, void* grid, const ProbeSpec& pr, cudaStream_t st,
                          void* igrid, double iscale) {
  if (sp.N == 0) return;
  const DetSpec det{(unsigned long long*)igrid, iscale};
  EFGP_CHECK((k.W & 1) == sp.odd, "kernel width parity does not match the sorted point set");
  EFGP_CHECK(pr.nb >= 1 && (pr.t0 >> 7) == ((pr.t0 + pr.nb - 1) >> 7), "a probe launch must stay inside one 128-probe block");
  if (sp.dtype == 0)
    launch_t<float>(sp, k, (const float*)nullptr, (float*)grid, st, 3, pr, det);
  else
    launch_t<double>(sp, k, (const double*)nullptr, (double*)grid, st, 3, pr, det);
}

}  // namespace efgp
