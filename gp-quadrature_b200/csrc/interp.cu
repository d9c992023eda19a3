// Type-2 interpolation: real fine grid -> values at bin-sorted nonuniform points.
//
// This is the arithmetic FINUFFT performs inside pff.finufft_type2 (reference call sites
// efgpnd.py:1533-1536, 1546-1549, 1679); the algorithm here is this library's own.
//
// Mirror image of spread.cu: a thread walks a contiguous slice of the cell-sorted points and
// keeps the W x W patch of fine-grid values it needs in REGISTERS, reloading it (through the
// read-only path; the whole fine grid lives in L2) only when the cell changes.  Per point the
// work is the kernel evaluation plus W^2 FMAs, and one scattered store of the result.
#include "nufft.cuh"
#include "stream_ring.cuh"
#include "philox.cuh"
#include <cstdlib>

namespace efgp {

template <typename T, int W>
__device__ __forceinline__ void es_eval_i(const ESDev<T>& es, T s, T* __restrict__ k) {
  constexpr int DEG = es_degree(W);
  constexpr int NE = DEG / 2 + 1;
  constexpr int NO = (DEG + 1) / 2;
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

template <typename T>
__device__ __forceinline__ void load_record4i(const T* __restrict__ p, T* w) {
  if constexpr (sizeof(T) == 8) {
    ld_global_nc_256(reinterpret_cast<const double*>(p), w[0], w[1], w[2], w[3]);
  } else {
    float4 a = __ldg(reinterpret_cast<const float4*>(p));
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
  }
}
// 2-D record: (X0, X1, binv, pad)
template <typename T>
__device__ __forceinline__ void load_point2(const T* __restrict__ p, T& X0, T& X1, uint32_t& binv) {
  T w[4];
  load_record4i<T>(p, w);
  X0 = w[0];
  X1 = w[1];
  binv = *reinterpret_cast<const uint32_t*>(&w[2]);
}

__device__ __forceinline__ int wrap_i(int i, int n) {
  if (i < 0) i += n;
  if (i >= n) i -= n;
  return i;
}

template <typename T, int W>
__global__ void __launch_bounds__(kTileThreads)
interp1d_kernel(const T* __restrict__ rec, int P, int nf0,
                const T* __restrict__ grid, T* __restrict__ out, int64_t ostride,
                const __grid_constant__ ESDev<T> es) {
  const int64_t base = (int64_t)blockIdx.x * kTileThreads * P + threadIdx.x;
  constexpr int OFF = W / 2 - ((W & 1) ? 0 : 1);
  T g[W];
  int cur = -1;
  for (int j = 0; j < P; ++j) {
    const int64_t pos = base + (int64_t)j * kTileThreads;
    T X = rec[pos * 2];
    if (X < T(0)) break;
    const int64_t bi = pos;
    int c0 = (int)X;
    if (c0 != cur) {
#pragma unroll
      for (int i = 0; i < W; ++i) g[i] = __ldg(grid + wrap_i(c0 - OFF + i, nf0));
      cur = c0;
    }
    T k0[W];
    es_eval_i<T, W>(es, T(2) * (X - (T)c0) - T(1), k0);
    T acc = 0;
#pragma unroll
    for (int i = 0; i < W; ++i) acc = fma(g[i], k0[i], acc);
    out[bi] = acc;
  }
}

// Narrow kernels (W <= 5) keep their W x W patch in <= 128 registers, so two CTAs share an SM; the point
// records stream through a per-thread cp.async ring (stream_ring.cuh) deep enough to cover DRAM latency.
//
// SCAT: the result of a point goes straight to its level-1 (bucket-order) position, which the record carries in the
// low 32 bits of word 2 (points.cu bin_scatter_kernel), instead of to its slot, and unpermute_output only runs its
// second level.  The idea was to save the slot -> bucket-order pass (0.48 ms at N = 1e8); measured, the scattered
// 8-byte stores are 6x slower than that pass's gathers (see launch_interp), so this is an A/B switch, not the default.
template <typename T, int W, bool SCAT>
__global__ void __launch_bounds__(kTileThreads, (W <= 5 ? 2 : 1))
interp2d_kernel(const T* __restrict__ rec, int P,
                int nf0, int nf1, const T* __restrict__ grid, T* __restrict__ out, int64_t ostride,
                const __grid_constant__ ESDev<T> es) {
  constexpr int S = kRingStages;
  __shared__ __align__(16) T ring[S][kTileThreads][2];      // (X0, X1) of each thread's next points
  __shared__ uint32_t bring[SCAT ? S : 1][kTileThreads];    // level-1 position of the point
  const int t = threadIdx.x;
  const int64_t base = (int64_t)blockIdx.x * kTileThreads * P + t;
  auto issue = [&](int j) {
    if (j < P) {
      const T* src = rec + (base + (int64_t)j * kTileThreads) * 4;
      cp_async_word<2 * sizeof(T)>(&ring[j % S][t][0], src);
      if (SCAT) cp_async_word<4>(&bring[j % S][t], src + 2);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int j = 0; j < S - 1; ++j) issue(j);
  constexpr int OFF = W / 2 - ((W & 1) ? 0 : 1);
  T g[W][W];
  int cur0 = -1, cur1 = -1;
  for (int j = 0; j < P; ++j) {
    issue(j + S - 1);
    cp_async_wait<S - 1>();
    const int s = j % S;
    const T X0 = ring[s][t][0];
    if (X0 < T(0)) break;
    const T X1 = ring[s][t][1];
    const int64_t pn = base + (int64_t)j * kTileThreads;   // results in slot order: coalesced stores
    int c0 = (int)X0, c1 = (int)X1;
    if (c0 != cur0 || c1 != cur1) {
      int cols[W];
#pragma unroll
      for (int jj = 0; jj < W; ++jj) cols[jj] = wrap_i(c1 - OFF + jj, nf1);
#pragma unroll
      for (int i = 0; i < W; ++i) {
        const T* row = grid + (int64_t)wrap_i(c0 - OFF + i, nf0) * nf1;
#pragma unroll
        for (int jj = 0; jj < W; ++jj) g[i][jj] = __ldg(row + cols[jj]);
      }
      cur0 = c0;
      cur1 = c1;
    }
    T k0[W], k1[W];
    es_eval_i<T, W>(es, T(2) * (X0 - (T)c0) - T(1), k0);
    es_eval_i<T, W>(es, T(2) * (X1 - (T)c1) - T(1), k1);
    T acc = 0;
#pragma unroll
    for (int i = 0; i < W; ++i) {
      T r = 0;
#pragma unroll
      for (int jj = 0; jj < W; ++jj) r = fma(g[i][jj], k1[jj], r);
      acc = fma(r, k0[i], acc);
    }
    if (SCAT)
      out[bring[s][t]] = acc;
    else
      out[pn] = acc;
  }
  cp_async_wait<0>();
}

// ------------------------------------------------------------------ 2-D, record slabs staged by bulk async copies (TMA)
// A/B variant of interp2d_kernel (EFGP_INTERP_TMA=1): instead of 256 per-thread 16-byte cp.async per step, ONE elected
// thread issues a 1-D bulk tensor copy (cp.async.bulk.shared::cluster.global, the TMA engine: SASS UBLKCP) of the CTA's
// next kTmaSteps * 256 records = a contiguous 16 KB slab of the tile-transposed layout, completion signalled through an
// mbarrier (complete_tx::bytes); the threads wait on the barrier's phase and read their record from shared memory.
// The slab ring is refilled CTA-wide, so the steps of a slab are followed by one __syncthreads (the per-thread ring of
// interp2d_kernel needs none).  Measured against the cp.async ring in DESIGN.md section 4.
constexpr int kTmaStages = 4;
constexpr int kTmaSteps = 2;       // point steps per slab: 2 * 256 records * 32 bytes = 16 KB (f64)

__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
  unsigned ok;
  do {
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok)
        : "r"(a), "r"(parity)
        : "memory");
  } while (!ok);
}
__device__ __forceinline__ void bulk_copy_g2s(void* smem_dst, const void* gsrc, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   (unsigned)__cvta_generic_to_shared(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"((unsigned)__cvta_generic_to_shared(bar))
               : "memory");
}

template <typename T, int W>
__global__ void __launch_bounds__(kTileThreads, (W <= 5 ? 2 : 1))
interp2d_tma_kernel(const T* __restrict__ rec, int P, int nf0, int nf1, const T* __restrict__ grid, T* __restrict__ out,
                    const __grid_constant__ ESDev<T> es) {
  extern __shared__ __align__(128) unsigned char tma_smem[];
  T* slab = reinterpret_cast<T*>(tma_smem);                          // [stages][steps][256][4]
  __shared__ __align__(8) uint64_t full[kTmaStages];
  const int t = threadIdx.x;
  const int64_t tile0 = (int64_t)blockIdx.x * kTileThreads * P;      // first slot of this CTA's tile
  const int ngroups = (P + kTmaSteps - 1) / kTmaSteps;
  constexpr unsigned kStepBytes = kTileThreads * 4 * sizeof(T);
  if (t == 0) {
#pragma unroll
    for (int s = 0; s < kTmaStages; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  auto issue = [&](int g) {           // thread 0 only
    const int s = g % kTmaStages;
    const int steps = (P - g * kTmaSteps) < kTmaSteps ? (P - g * kTmaSteps) : kTmaSteps;
    const unsigned bytes = (unsigned)steps * kStepBytes;
    mbar_expect_tx(&full[s], bytes);
    bulk_copy_g2s(slab + (size_t)s * kTmaSteps * kTileThreads * 4, rec + (tile0 + (int64_t)g * kTmaSteps * kTileThreads) * 4,
                  bytes, &full[s]);
  };
  if (t == 0)
    for (int g = 0; g < kTmaStages && g < ngroups; ++g) issue(g);
  constexpr int OFF = W / 2 - ((W & 1) ? 0 : 1);
  T g_[W][W];
  int cur0 = -1, cur1 = -1;
  for (int g = 0; g < ngroups; ++g) {
    const int s = g % kTmaStages;
    mbar_wait(&full[s], (unsigned)((g / kTmaStages) & 1));
#pragma unroll
    for (int kk = 0; kk < kTmaSteps; ++kk) {
      const int j = g * kTmaSteps + kk;
      if (j >= P) break;                                   // uniform over the CTA
      const T* r = slab + (((size_t)s * kTmaSteps + kk) * kTileThreads + t) * 4;
      const T X0 = r[0], X1 = r[1];
      if (!(X0 < T(0))) {                                  // (pad slots of the last tile: no early exit, the CTA syncs)
        const int64_t pn = tile0 + (int64_t)j * kTileThreads + t;
        const int c0 = (int)X0, c1 = (int)X1;
        if (c0 != cur0 || c1 != cur1) {
          int cols[W];
#pragma unroll
          for (int jj = 0; jj < W; ++jj) cols[jj] = wrap_i(c1 - OFF + jj, nf1);
#pragma unroll
          for (int i = 0; i < W; ++i) {
            const T* row = grid + (int64_t)wrap_i(c0 - OFF + i, nf0) * nf1;
#pragma unroll
            for (int jj = 0; jj < W; ++jj) g_[i][jj] = __ldg(row + cols[jj]);
          }
          cur0 = c0;
          cur1 = c1;
        }
        T k0[W], k1[W];
        es_eval_i<T, W>(es, T(2) * (X0 - (T)c0) - T(1), k0);
        es_eval_i<T, W>(es, T(2) * (X1 - (T)c1) - T(1), k1);
        T acc = 0;
#pragma unroll
        for (int i = 0; i < W; ++i) {
          T rr = 0;
#pragma unroll
          for (int jj = 0; jj < W; ++jj) rr = fma(g_[i][jj], k1[jj], rr);
          acc = fma(rr, k0[i], acc);
        }
        out[pn] = acc;
      }
    }
    __syncthreads();                                       // every thread is done with slab s
    if (t == 0 && g + kTmaStages < ngroups) issue(g + kTmaStages);
  }
}

// ------------------------------------------------------------------ 3-D, shared-memory staged grid tiles
// Mirror of spread3d_tile_kernel (spread.cu): points arrive tile-major; a CTA stages the current 8^3 tile
// plus halo of the fine grid in shared memory ONCE and every point of the tile gathers its W^3 values from
// there (one warp per point, lanes over the W x W cells of a plane, shuffle reduction) instead of W^3
// loads from L2 per point.  Results are written in slot order (coalesced).
constexpr int kT3ThreadsI = 128;
constexpr int kT3BatchI = 32;
constexpr int kT3ChunkI = 1024;

template <typename T, int W>
__host__ __device__ constexpr size_t tile3i_smem_bytes() {
  constexpr int E = kTile3 + W - 1;
  return sizeof(T) * (size_t)(E * E * E + kT3BatchI * 3 * W) + sizeof(int) * kT3BatchI * 4;
}

template <typename T, int W>
__global__ void __launch_bounds__(kT3ThreadsI)
interp3d_tile_kernel(const T* __restrict__ rec, int64_t N, int nf0, int nf1, int nf2, const T* __restrict__ grid,
                     T* __restrict__ out, const __grid_constant__ ESDev<T> es) {
  constexpr int E = kTile3 + W - 1;
  constexpr int OFF = W / 2 - ((W & 1) ? 0 : 1);
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* S = reinterpret_cast<T*>(smem_raw);        // [E][E][E] tile + halo of the fine grid
  T* KW = S + E * E * E;                        // [batch][3][W]
  int* BI = reinterpret_cast<int*>(KW + kT3BatchI * 3 * W);   // [batch][4]: tile id, cell0, cell1, cell2
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nt1 = (nf1 + kTile3 - 1) / kTile3, nt2 = (nf2 + kTile3 - 1) / kTile3;
  const int64_t p0 = (int64_t)blockIdx.x * kT3ChunkI;
  const int64_t p1 = p0 + kT3ChunkI < N ? p0 + kT3ChunkI : N;
  int cur_tile = -1, tb0 = 0, tb1 = 0, tb2 = 0;
  for (int64_t b0 = p0; b0 < p1; b0 += kT3BatchI) {
    const int nb = (int)(p1 - b0 < kT3BatchI ? p1 - b0 : kT3BatchI);
    __syncthreads();   // previous batch fully consumed
    if (tid < 3 * nb) {
      const int pt = tid / 3, ax = tid - 3 * pt;
      const T X = rec[(b0 + pt) * 4 + ax];
      const int c = (int)X;
      T kk[W];
      es_eval_i<T, W>(es, T(2) * (X - (T)c) - T(1), kk);
#pragma unroll
      for (int i = 0; i < W; ++i) KW[(pt * 3 + ax) * W + i] = kk[i];
      BI[pt * 4 + 1 + ax] = c;
    }
    __syncthreads();
    if (tid < nb)
      BI[tid * 4] = ((BI[tid * 4 + 1] / kTile3) * nt1 + BI[tid * 4 + 2] / kTile3) * nt2 + BI[tid * 4 + 3] / kTile3;
    __syncthreads();
    int p = 0;
    while (p < nb) {           // uniform control flow
      const int t = BI[p * 4];
      if (t != cur_tile) {
        cur_tile = t;
        tb2 = (t % nt2) * kTile3;
        tb1 = ((t / nt2) % nt1) * kTile3;
        tb0 = (t / (nt2 * nt1)) * kTile3;
        __syncthreads();       // everyone finished reading the old tile
        for (int e = tid; e < E * E * E; e += kT3ThreadsI) {
          int e0 = e / (E * E), e1 = (e / E) % E, e2 = e % E;
          int64_t gi = ((int64_t)wrap_i(tb0 - OFF + e0, nf0) * nf1 + wrap_i(tb1 - OFF + e1, nf1)) * nf2 +
                       wrap_i(tb2 - OFF + e2, nf2);
          S[e] = __ldg(grid + gi);
        }
        __syncthreads();
      }
      int q = p + 1;
      while (q < nb && BI[q * 4] == t) ++q;
      for (int pt = p + warp; pt < q; pt += kT3ThreadsI / 32) {     // one warp per point
        const int l0 = BI[pt * 4 + 1] - tb0, l1 = BI[pt * 4 + 2] - tb1, l2 = BI[pt * 4 + 3] - tb2;
        const T* k0 = KW + (pt * 3 + 0) * W;
        const T* k1 = KW + (pt * 3 + 1) * W;
        const T* k2 = KW + (pt * 3 + 2) * W;
        T acc = T(0);
        for (int r = lane; r < W * W; r += 32) {
          const int j = r / W, l = r - j * W;
          const T* col = S + ((l0 * E + l1 + j) * E + l2 + l);
          T ssum = T(0);
#pragma unroll
          for (int i = 0; i < W; ++i) ssum = fma(col[i * E * E], k0[i], ssum);
          acc = fma(ssum, k1[j] * k2[l], acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
        if (lane == 0) out[b0 + pt] = acc;
      }
      p = q;
    }
  }
}

// generic: wide 2-D kernels and 3-D, one slot per thread, gathers through L1/L2
template <typename T, int W, int D>
__global__ void __launch_bounds__(256)
interp_direct_kernel(const T* __restrict__ rec, int64_t npos, int nf0, int nf1, int nf2,
                     const T* __restrict__ grid, T* __restrict__ out, int64_t ostride,
                     const __grid_constant__ ESDev<T> es) {
  int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= npos) return;
  constexpr int OFF = W / 2 - ((W & 1) ? 0 : 1);
  T k0[W], k1[W], k2[W];
  T Xr[4];
  load_record4i<T>(rec + n * 4, Xr);
  T X0 = Xr[0];
  if (X0 < T(0)) return;
  const int64_t bi = n;
  int c0 = (int)X0;
  es_eval_i<T, W>(es, T(2) * (X0 - (T)c0) - T(1), k0);
  T X1 = Xr[1];
  int c1 = (int)X1;
  es_eval_i<T, W>(es, T(2) * (X1 - (T)c1) - T(1), k1);
  int c2 = 0;
  if (D == 3) {
    T X2 = Xr[2];
    c2 = (int)X2;
    es_eval_i<T, W>(es, T(2) * (X2 - (T)c2) - T(1), k2);
  }
  T acc = 0;
  if (D == 2) {
#pragma unroll
    for (int i = 0; i < W; ++i) {
      const T* row = grid + (int64_t)wrap_i(c0 - OFF + i, nf0) * nf1;
      T r = 0;
#pragma unroll
      for (int j = 0; j < W; ++j) r = fma(__ldg(row + wrap_i(c1 - OFF + j, nf1)), k1[j], r);
      acc = fma(r, k0[i], acc);
    }
  } else {
    int cols[W];
#pragma unroll
    for (int l = 0; l < W; ++l) cols[l] = wrap_i(c2 - OFF + l, nf2);
#pragma unroll 1
    for (int i = 0; i < W; ++i) {
      int64_t r0 = (int64_t)wrap_i(c0 - OFF + i, nf0) * nf1;
      T ri = 0;
#pragma unroll 1
      for (int j = 0; j < W; ++j) {
        const T* row = grid + (r0 + wrap_i(c1 - OFF + j, nf1)) * nf2;
        T r = 0;
#pragma unroll
        for (int l = 0; l < W; ++l) r = fma(__ldg(row + cols[l]), k2[l], r);
        ri = fma(r, k1[j], ri);
      }
      acc = fma(ri, k0[i], acc);
    }
  }
  out[bi] = acc;
}

// ------------------------------------------------------------------ fused Hutchinson epilogue
// sum_n Z_t[n] * (F f_t)[n] without ever writing F f_t: the interpolation kernels below multiply each point's result by
// the regenerated Rademacher sign (philox.cuh: the SAME (seed, probe, slot) stream the spread used) and reduce it --
// per thread over its points, per CTA in shared memory, then a fixed-order sum over the CTA partials (deterministic).
// This is term1 of the gradient (efgpnd.py:235-241: mean_t sum_n Z * Alpha); no (T, N) array exists.
__device__ __forceinline__ double cta_sum(double v, double* sh) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) sh[w] = v;
  __syncthreads();
  double s = 0;
  if (threadIdx.x == 0)
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += sh[i];
  return s;   // valid in thread 0
}

template <typename T, int W, int NB>
__global__ void __launch_bounds__(kTileThreads, ((W * W * NB <= 25) ? 2 : 1))
interp2d_dot_kernel(const T* __restrict__ rec, int P, int nf0, int nf1, const T* __restrict__ grid,
                    double* __restrict__ partial, const __grid_constant__ ESDev<T> es, ProbeSpec pr) {
  constexpr int S = kRingStages;
  __shared__ __align__(16) T ring[S][kTileThreads][2];
  __shared__ T oring[S][kTileThreads];       // record word 2: level-1 position | original index
  __shared__ double sh[kTileThreads / 32];
  const int t = threadIdx.x;
  const int64_t base = (int64_t)blockIdx.x * kTileThreads * P + t;
  auto issue = [&](int j) {
    if (j < P) {
      const T* src = rec + (base + (int64_t)j * kTileThreads) * 4;
      cp_async_word<2 * sizeof(T)>(&ring[j % S][t][0], src);
      cp_async_word<sizeof(T)>(&oring[j % S][t], src + 2);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int j = 0; j < S - 1; ++j) issue(j);
  constexpr int OFF = W / 2 - ((W & 1) ? 0 : 1);
  const int64_t cells = (int64_t)nf0 * nf1;
  T g[NB][W][W];
  double sum[NB];
#pragma unroll
  for (int b = 0; b < NB; ++b) sum[b] = 0;
  int cur0 = -1, cur1 = -1;
  for (int j = 0; j < P; ++j) {
    issue(j + S - 1);
    cp_async_wait<S - 1>();
    const int s = j % S;
    const T X0 = ring[s][t][0];
    if (X0 < T(0)) break;
    const T X1 = ring[s][t][1];
    int c0 = (int)X0, c1 = (int)X1;
    if (c0 != cur0 || c1 != cur1) {
      int cols[W];
#pragma unroll
      for (int jj = 0; jj < W; ++jj) cols[jj] = wrap_i(c1 - OFF + jj, nf1);
#pragma unroll
      for (int i = 0; i < W; ++i) {
        const T* row = grid + (int64_t)wrap_i(c0 - OFF + i, nf0) * nf1;
#pragma unroll
        for (int b = 0; b < NB; ++b)
#pragma unroll
          for (int jj = 0; jj < W; ++jj) g[b][i][jj] = __ldg(row + b * cells + cols[jj]);
      }
      cur0 = c0;
      cur1 = c1;
    }
    T k0[W], k1[W];
    es_eval_i<T, W>(es, T(2) * (X0 - (T)c0) - T(1), k0);
    es_eval_i<T, W>(es, T(2) * (X1 - (T)c1) - T(1), k1);
    uint32_t bits[4];
    philox_bits(pr.seed, probe_counter<T>(oring[s][t], base + (int64_t)j * kTileThreads, pr.use_orig),
                (uint32_t)(pr.t0 >> 7), bits);
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      T acc = 0;
#pragma unroll
      for (int i = 0; i < W; ++i) {
        T r = 0;
#pragma unroll
        for (int jj = 0; jj < W; ++jj) r = fma(g[b][i][jj], k1[jj], r);
        acc = fma(r, k0[i], acc);
      }
      sum[b] += (double)(philox_sign_from_bits<T>(bits, pr.t0 + b) * acc);
    }
  }
  cp_async_wait<0>();
#pragma unroll
  for (int b = 0; b < NB; ++b) {
    const double tot = cta_sum(sum[b], sh);
    if (t == 0) partial[(int64_t)blockIdx.x * NB + b] = tot;
  }
}

// 1-D, wide 2-D and 3-D: one slot per thread (as interp_direct_kernel), one probe per launch
template <typename T, int W, int D>
__global__ void __launch_bounds__(256)
interp_direct_dot_kernel(const T* __restrict__ rec, int64_t npos, int nf0, int nf1, int nf2, const T* __restrict__ grid,
                         double* __restrict__ partial, const __grid_constant__ ESDev<T> es, ProbeSpec pr) {
  __shared__ double sh[8];
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  constexpr int OFF = W / 2 - ((W & 1) ? 0 : 1);
  constexpr int RWD = D == 1 ? 2 : 4;
  double val = 0;
  if (n < npos) {
    T Xr[4] = {T(0), T(0), T(0), T(0)};
    if constexpr (D == 1) {
      Xr[0] = rec[n * 2];
      Xr[1] = rec[n * 2 + 1];
    } else {
      load_record4i<T>(rec + n * RWD, Xr);
    }
    const T X0 = Xr[0];
    if (!(X0 < T(0))) {
      T k0[W], k1[W], k2[W];
      const int c0 = (int)X0;
      es_eval_i<T, W>(es, T(2) * (X0 - (T)c0) - T(1), k0);
      T acc = 0;
      if constexpr (D == 1) {
#pragma unroll
        for (int i = 0; i < W; ++i) acc = fma(__ldg(grid + wrap_i(c0 - OFF + i, nf0)), k0[i], acc);
      } else {
        const T X1 = Xr[1];
        const int c1 = (int)X1;
        es_eval_i<T, W>(es, T(2) * (X1 - (T)c1) - T(1), k1);
        if constexpr (D == 2) {
#pragma unroll
          for (int i = 0; i < W; ++i) {
            const T* row = grid + (int64_t)wrap_i(c0 - OFF + i, nf0) * nf1;
            T r = 0;
#pragma unroll
            for (int j = 0; j < W; ++j) r = fma(__ldg(row + wrap_i(c1 - OFF + j, nf1)), k1[j], r);
            acc = fma(r, k0[i], acc);
          }
        } else {
          const T X2 = Xr[2];
          const int c2 = (int)X2;
          es_eval_i<T, W>(es, T(2) * (X2 - (T)c2) - T(1), k2);
          int cols[W];
#pragma unroll
          for (int l = 0; l < W; ++l) cols[l] = wrap_i(c2 - OFF + l, nf2);
#pragma unroll 1
          for (int i = 0; i < W; ++i) {
            const int64_t r0 = (int64_t)wrap_i(c0 - OFF + i, nf0) * nf1;
            T ri = 0;
#pragma unroll 1
            for (int j = 0; j < W; ++j) {
              const T* row = grid + (r0 + wrap_i(c1 - OFF + j, nf1)) * nf2;
              T r = 0;
#pragma unroll
              for (int l = 0; l < W; ++l) r = fma(__ldg(row + cols[l]), k2[l], r);
              ri = fma(r, k1[j], ri);
            }
            acc = fma(ri, k0[i], acc);
          }
        }
      }
      uint32_t bits[4];
      philox_bits(pr.seed, probe_counter<T>(Xr[D], n, pr.use_orig), (uint32_t)(pr.t0 >> 7), bits);
      val = (double)(philox_sign_from_bits<T>(bits, pr.t0) * acc);
    }
  }
  const double tot = cta_sum(val, sh);
  if (threadIdx.x == 0) partial[blockIdx.x] = tot;
}

// out[b] = sum over CTAs of partial[cta * nb + b], fixed order: one CTA per probe
__global__ void __launch_bounds__(256)
probe_dot_reduce_kernel(const double* __restrict__ partial, int64_t nctas, int nb, double* __restrict__ out) {
  __shared__ double sh[8];
  const int b = blockIdx.x;
  double v = 0;
  for (int64_t c = threadIdx.x; c < nctas; c += blockDim.x) v += partial[c * nb + b];
  const double tot = cta_sum(v, sh);
  if (threadIdx.x == 0) out[b] = tot;
}

// Z itself, slot order (tests / the parity path compare the regenerated probes with explicit ones)
template <typename T>
__global__ void __launch_bounds__(256)
probe_fill_kernel(const T* __restrict__ rec, int RW, int64_t npos, T* __restrict__ z, ProbeSpec pr) {
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= npos) return;
  uint32_t bits[4];
  philox_bits(pr.seed, (uint64_t)n, (uint32_t)(pr.t0 >> 7), bits);
  z[n] = rec[n * RW] < T(0) ? T(0) : philox_sign_from_bits<T>(bits, pr.t0);
}

// counter = original index: Z in original order needs no sorted set at all
template <typename T>
__global__ void __launch_bounds__(256) probe_fill_orig_kernel(int64_t N, T* __restrict__ z, int64_t ostride, ProbeSpec pr) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  uint32_t bits[4];
  philox_bits(pr.seed, (uint64_t)i, (uint32_t)(pr.t0 >> 7), bits);
  z[i * ostride] = philox_sign_from_bits<T>(bits, pr.t0);
}

template <typename T, int W>
static void launch_w(const SortedPoints& sp, const ESKernel& k, const T* grid, T* out, int64_t ostride,
                     cudaStream_t st, bool scatter = false) {
  ESDev<T> es;
  es_to_dev<T>(k, es);
  const T* rec = (const T*)sp.rec;
  constexpr int kRegMaxW = sizeof(T) == 8 ? 9 : 10;
  const int nt = (int)sp.ntiles;
  char nm[48];
  snprintf(nm, sizeof(nm), "interp%dd_w%d_%s", sp.dim, W, sizeof(T) == 8 ? "f64" : "f32");
  ScopedStage stg(nm, st, 1, sp.N);
  if (sp.dim == 1) {
    interp1d_kernel<T, W><<<nt, kTileThreads, 0, st>>>(rec, sp.P, (int)sp.nf[0], grid, out, ostride, es);
  } else if (sp.dim == 2) {
    if constexpr (W <= kRegMaxW) {
      static const bool use_tma = getenv("EFGP_INTERP_TMA") != nullptr && atoi(getenv("EFGP_INTERP_TMA")) != 0;
      if (use_tma) {
        const size_t smem = (size_t)kTmaStages * kTmaSteps * kTileThreads * 4 * sizeof(T);
        EFGP_CUDA(cudaFuncSetAttribute(interp2d_tma_kernel<T, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        interp2d_tma_kernel<T, W><<<nt, kTileThreads, smem, st>>>(rec, sp.P, (int)sp.nf[0], (int)sp.nf[1], grid, out, es);
      } else if (scatter) {
        interp2d_kernel<T, W, true><<<nt, kTileThreads, 0, st>>>(rec, sp.P, (int)sp.nf[0], (int)sp.nf[1], grid,
                                                                 out, ostride, es);
      } else {
        interp2d_kernel<T, W, false><<<nt, kTileThreads, 0, st>>>(rec, sp.P, (int)sp.nf[0], (int)sp.nf[1], grid,
                                                                  out, ostride, es);
      }
    } else {
      interp_direct_kernel<T, W, 2><<<div_up(sp.npos, 256), 256, 0, st>>>(
          rec, sp.npos, (int)sp.nf[0], (int)sp.nf[1], 1, grid, out, ostride, es);
    }
  } else if (getenv("EFGP_3D_TILE_INTERP")) {
    // measured slower than the per-point gather below (one thread per point gathering through L1/L2 is
    // instruction-lean; the tile-major sort gives it the locality): kept for A/B runs only
    const size_t smem = tile3i_smem_bytes<T, W>();
    EFGP_CUDA(cudaFuncSetAttribute(interp3d_tile_kernel<T, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    interp3d_tile_kernel<T, W><<<div_up(sp.N, kT3ChunkI), kT3ThreadsI, smem, st>>>(
        rec, sp.N, (int)sp.nf[0], (int)sp.nf[1], (int)sp.nf[2], grid, out, es);
  } else {
    interp_direct_kernel<T, W, 3><<<div_up(sp.npos, 256), 256, 0, st>>>(
        rec, sp.npos, (int)sp.nf[0], (int)sp.nf[1], (int)sp.nf[2], grid, out, ostride, es);
  }
  EFGP_CUDA(cudaGetLastError());
}

template <typename T>
static void launch_t(const SortedPoints& sp, const ESKernel& k, const T* grid, T* out, int64_t ostride,
                     cudaStream_t st, bool scatter = false) {
  switch (k.W) {
#define EFGP_CASE(w) \
  case w:            \
    launch_w<T, w>(sp, k, grid, out, ostride, st, scatter); \
    break;
    EFGP_CASE(2) EFGP_CASE(3) EFGP_CASE(4) EFGP_CASE(5) EFGP_CASE(6) EFGP_CASE(7) EFGP_CASE(8) EFGP_CASE(9)
    EFGP_CASE(10) EFGP_CASE(11) EFGP_CASE(12) EFGP_CASE(13) EFGP_CASE(14) EFGP_CASE(15) EFGP_CASE(16)
#undef EFGP_CASE
    default:
      throw Error(1, "unsupported interpolation width");
  }
}

void launch_interp(const SortedPoints& sp, const ESKernel& k, const void* grid, void* out, int64_t ostride,
                   cudaStream_t st) {
  if (sp.N == 0) return;
  EFGP_CHECK((k.W & 1) == sp.odd, "kernel width parity does not match the sorted point set");
  // the kernels write slot-order results (coalesced); unpermute_output gathers them back to original order
  const size_t esz = sp.dtype == 0 ? 4 : 8;
  void* tmp = pool_alloc(esz * (size_t)sp.npos);
  // 2-D register-patch kernels: results go straight to bucket order (interp2d_kernel<.., SCAT>), one un-permute level
  // A/B only (EFGP_INTERP_SCATTER=1).  Measured at N = 1e8: 3.81 ms against 0.61 (interp) + 0.48 (slot -> bucket order):
  // 1e8 scattered 8-byte partial-sector stores run at ~2.6e10/s, the gathers of the separate pass at ~2e11/s.
  static const bool scat_env = getenv("EFGP_INTERP_SCATTER") && atoi(getenv("EFGP_INTERP_SCATTER")) != 0;
  static const bool tma_env = getenv("EFGP_INTERP_TMA") != nullptr && atoi(getenv("EFGP_INTERP_TMA")) != 0;
  const bool scatter = scat_env && !tma_env && sp.dim == 2 && k.W <= (sp.dtype == 0 ? 10 : 9);
  if (sp.dtype == 0)
    launch_t<float>(sp, k, (const float*)grid, (float*)tmp, 1, st, scatter);
  else
    launch_t<double>(sp, k, (const double*)grid, (double*)tmp, 1, st, scatter);
  unpermute_output(sp, tmp, out, ostride, st, scatter);
  pool_free(tmp);  // stream-ordered reuse
}

template <typename T, int W>
static void launch_dot_w(const SortedPoints& sp, const ESKernel& k, const T* grid, double* out, const ProbeSpec& pr,
                         cudaStream_t st) {
  ESDev<T> es;
  es_to_dev<T>(k, es);
  const T* rec = (const T*)sp.rec;
  constexpr int kRegMaxW = sizeof(T) == 8 ? 9 : 10;
  char nm[56];
  snprintf(nm, sizeof(nm), "interp%dd_w%d_%s_dot%d", sp.dim, W, sizeof(T) == 8 ? "f64" : "f32", pr.nb);
  ScopedStage stg(nm, st, 2, sp.N * pr.nb);
  int64_t nctas = 0;
  double* partial = nullptr;
  if (sp.dim == 2 && W <= kRegMaxW) {
    nctas = sp.ntiles;
    partial = (double*)pool_alloc(sizeof(double) * (size_t)nctas * pr.nb);
    if constexpr (W <= kRegMaxW) {
      bool done = false;
      if constexpr (W <= 5) {
        if (pr.nb == 4) {
          interp2d_dot_kernel<T, W, 4><<<(int)nctas, kTileThreads, 0, st>>>(rec, sp.P, (int)sp.nf[0], (int)sp.nf[1], grid,
                                                                            partial, es, pr);
          done = true;
        }
      }
      if constexpr (W <= 7) {
        if (!done && pr.nb == 2) {
          interp2d_dot_kernel<T, W, 2><<<(int)nctas, kTileThreads, 0, st>>>(rec, sp.P, (int)sp.nf[0], (int)sp.nf[1], grid,
                                                                            partial, es, pr);
          done = true;
        }
      }
      if (!done) {
        EFGP_CHECK(pr.nb == 1, "probe batch not supported by this interpolation width");
        interp2d_dot_kernel<T, W, 1><<<(int)nctas, kTileThreads, 0, st>>>(rec, sp.P, (int)sp.nf[0], (int)sp.nf[1], grid,
                                                                          partial, es, pr);
      }
    }
  } else {
    EFGP_CHECK(pr.nb == 1, "one probe per launch for this geometry");
    nctas = div_up(sp.npos, 256);
    partial = (double*)pool_alloc(sizeof(double) * (size_t)nctas);
    if (sp.dim == 1)
      interp_direct_dot_kernel<T, W, 1><<<(int)nctas, 256, 0, st>>>(rec, sp.npos, (int)sp.nf[0], 1, 1, grid, partial, es, pr);
    else if (sp.dim == 2)
      interp_direct_dot_kernel<T, W, 2><<<(int)nctas, 256, 0, st>>>(rec, sp.npos, (int)sp.nf[0], (int)sp.nf[1], 1, grid,
                                                                    partial, es, pr);
    else
      interp_direct_dot_kernel<T, W, 3><<<(int)nctas, 256, 0, st>>>(rec, sp.npos, (int)sp.nf[0], (int)sp.nf[1],
                                                                    (int)sp.nf[2], grid, partial, es, pr);
  }
  EFGP_CUDA(cudaGetLastError());
  probe_dot_reduce_kernel<<<pr.nb, 256, 0, st>>>(partial, nctas, pr.nb, out);
  EFGP_CUDA(cudaGetLastError());
  pool_free(partial);   // stream-ordered reuse
}

template <typename T>
static void launch_dot_t(const SortedPoints& sp, const ESKernel& k, const T* grid, double* out, const ProbeSpec& pr,
                         cudaStream_t st) {
  switch (k.W) {
#define EFGP_CASE(w) \
  case w:            \
    launch_dot_w<T, w>(sp, k, grid, out, pr, st); \
    break;
    EFGP_CASE(2) EFGP_CASE(3) EFGP_CASE(4) EFGP_CASE(5) EFGP_CASE(6) EFGP_CASE(7) EFGP_CASE(8) EFGP_CASE(9)
    EFGP_CASE(10) EFGP_CASE(11) EFGP_CASE(12) EFGP_CASE(13) EFGP_CASE(14) EFGP_CASE(15) EFGP_CASE(16)
#undef EFGP_CASE
    default:
      throw Error(1, "unsupported interpolation width");
  }
}

// out[b] = sum_n Rademacher(seed, pr.t0 + b, slot(n)) * (interpolation of grid b at point n),  b < pr.nb
void launch_interp_probe_dots(const SortedPoints& sp, const ESKernel& k, const void* grid, double* out,
                              const ProbeSpec& pr, cudaStream_t st) {
  EFGP_CHECK((k.W & 1) == sp.odd, "kernel width parity does not match the sorted point set");
  EFGP_CHECK(pr.nb >= 1 && (pr.t0 >> 7) == ((pr.t0 + pr.nb - 1) >> 7), "a probe launch must stay inside one 128-probe block");
  if (sp.N == 0) {
    EFGP_CUDA(cudaMemsetAsync(out, 0, sizeof(double) * pr.nb, st));
    return;
  }
  if (sp.dtype == 0)
    launch_dot_t<float>(sp, k, (const float*)grid, out, pr, st);
  else
    launch_dot_t<double>(sp, k, (const double*)grid, out, pr, st);
}

// z (N,) in ORIGINAL point order = the probe pr.t0 the kernels regenerate for this sorted set
void launch_probe_export(const SortedPoints& sp, void* z_out, int64_t ostride, const ProbeSpec& pr, cudaStream_t st) {
  if (sp.N == 0) return;
  if (pr.use_orig) {
    if (sp.dtype == 0)
      probe_fill_orig_kernel<float><<<div_up(sp.N, 256), 256, 0, st>>>(sp.N, (float*)z_out, ostride, pr);
    else
      probe_fill_orig_kernel<double><<<div_up(sp.N, 256), 256, 0, st>>>(sp.N, (double*)z_out, ostride, pr);
    EFGP_CUDA(cudaGetLastError());
    return;
  }
  const size_t esz = sp.dtype == 0 ? 4 : 8;
  void* tmp = pool_alloc(esz * (size_t)sp.npos);
  if (sp.dtype == 0)
    probe_fill_kernel<float><<<div_up(sp.npos, 256), 256, 0, st>>>((const float*)sp.rec, sp.RW, sp.npos, (float*)tmp, pr);
  else
    probe_fill_kernel<double><<<div_up(sp.npos, 256), 256, 0, st>>>((const double*)sp.rec, sp.RW, sp.npos, (double*)tmp, pr);
  EFGP_CUDA(cudaGetLastError());
  unpermute_output(sp, tmp, z_out, ostride, st);
  pool_free(tmp);
}

}  // namespace efgp
