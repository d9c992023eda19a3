// Lockstep batched preconditioned conjugate gradients on the fused Toeplitz operator.
//
// Semantics follow cg.py of the reference exactly:
//   * single right-hand side  (_solve_single, cg.py:86-153): stop when ||r|| / (||b|| + 1e-16) < tol,
//     tested after the residual update; iters_completed = (index of the stopping iteration) + 1.
//   * batched (_solve_batched, cg.py:155-244): per-row early stop (rel < tol OR ||r|| < 1e-12), rows
//     frozen once converged, loop ends at the top of the next iteration -> iters_completed is one
//     larger than in the single case (capped by max_iter).
//   * alpha = rz / (pAp + 1e-16), beta = rz_new / (rz + 1e-16) (cg.py:57, 121, 145, 205, 222);
//     convergence is relative to ||b|| (1 if b == 0), not to the initial residual (cg.py:108-111).
//
// Per iteration there are four small kernels around the cuFFT pair:
//   A: beta, p = M^-1 r + beta p, write ws*p into the padded FFT input        (fused axpy + pad)
//   M: pointwise multiply by the cached generator spectrum
//   B: Ap = c1*ws*crop(.) + c2*p and the partial sums of Re<p,Ap>             (fused crop + dot)
//   C: alpha, x += alpha p, r -= alpha Ap, partial sums of ||r||^2 and <r,M^-1 r>  (fused axpy + dots)
// Reductions are two-stage with a fixed order (deterministic); every block re-reduces the <=64
// partials of its row, so no separate scalar kernels and no atomics on the critical path.
// Converged rows are skipped by all kernels via a double-buffered per-row flag.
#include "toeplitz.cuh"
#include <cstdlib>
#include <cooperative_groups.h>

namespace efgp {

template <typename T>
void fft_pair_begin(Toeplitz& t, int rows, typename cplx_of<T>::type* P, typename cplx_of<T>::type* Q,
                    cudaStream_t st);
template <typename T>
void fft_pair_end(Toeplitz& t, int rows, typename cplx_of<T>::type* Q, cudaStream_t st);

struct TzGeomC {
  int dim;
  int n[kMaxDim];
  int F[kMaxDim];
  int64_t M, Ftot;
  int mc[kMaxDim];     // centre of the block per axis, (n - 1) / 2   (Hermitian path)
  int Fh;              // F[last] / 2 + 1: last extent of the half spectrum
  int64_t Htot;        // elements of the half spectrum per row
};

__device__ __forceinline__ void tzc_offsets(const TzGeomC& g, int64_t k64, int64_t& pad_off, int64_t& crop_off) {
  int id[kMaxDim] = {0, 0, 0};
  uint32_t k = (uint32_t)k64;          // M < 2^31 (checked on the host): 32-bit divisions
  for (int a = g.dim - 1; a >= 0; --a) {
    const uint32_t q = k / (uint32_t)g.n[a];
    id[a] = (int)(k - q * (uint32_t)g.n[a]);
    k = q;
  }
  pad_off = 0;
  crop_off = 0;
  for (int a = 0; a < g.dim; ++a) {
    pad_off = pad_off * g.F[a] + id[a];
    crop_off = crop_off * g.F[a] + (id[a] + g.n[a] - 1);
  }
}

// Hermitian path: position of mode k (neg = false) or of -k (neg = true, taken when k lies in the half that the
// half spectrum does not store: centred last index < 0) inside the (F_0, .., F_last/2+1) half-spectrum array.
__device__ __forceinline__ void tzc_sym_offset(const TzGeomC& g, int64_t k64, int64_t& hoff, bool& neg) {
  int id[kMaxDim] = {0, 0, 0};
  uint32_t k = (uint32_t)k64;
  for (int a = g.dim - 1; a >= 0; --a) {
    const uint32_t q = k / (uint32_t)g.n[a];
    id[a] = (int)(k - q * (uint32_t)g.n[a]);
    k = q;
  }
  const int last = g.dim - 1;
  neg = id[last] < g.mc[last];
  hoff = 0;
  for (int a = 0; a < g.dim; ++a) {
    int ka = id[a] - g.mc[a];
    if (neg) ka = -ka;
    if (a < last)
      hoff = hoff * g.F[a] + (ka < 0 ? ka + g.F[a] : ka);
    else
      hoff = hoff * g.Fh + ka;
  }
}

constexpr int kMaxChunks = 256;    // <= blockDim: a block re-reduces a row's partials with one value per thread
constexpr int kChunkLen = 512;     // two elements per thread: the small-M kernels are latency chains otherwise

struct CgScal {
  double* rz;        // [2][B]
  double* denom;     // [B]
  double* p_pAp;     // [B][NC]
  double* p_rr;      // [B][NC]
  double* p_rz;      // [B][NC]
  int* act;          // [2][B]
  int* conv_iter;    // [B]
  int* counters;     // [0] = n_active, [1] = max conv iter
  int B, NC;
  int64_t len;       // elements per chunk
};

__device__ __forceinline__ double block_sum(double v, double* sh) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) sh[w] = v;
  __syncthreads();
  double s = 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += sh[i];
  }
  return s;  // valid in thread 0
}

// Sum of a row's <= 256 chunk partials, same value (and summation order) in every thread of every block.
__device__ __forceinline__ double sum_partials(const double* p, int nc) {
  __shared__ double shp[9];
  double v = (int)threadIdx.x < nc ? p[threadIdx.x] : 0.0;
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();                       // protects shp against the previous call's readers
  if ((threadIdx.x & 31) == 0) shp[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0;
  for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += shp[i];
  return s;
}
// r = b - Ax0 (Ax0 in Ap), p = minv * r, partials of <r, minv r> and ||b||^2
template <typename T>
__global__ void __launch_bounds__(256)
cg_init_kernel(const typename cplx_of<T>::type* __restrict__ b, const typename cplx_of<T>::type* __restrict__ Ap,
               const T* __restrict__ minv, typename cplx_of<T>::type* __restrict__ r,
               typename cplx_of<T>::type* __restrict__ p, int64_t M, CgScal s) {
  using C = typename cplx_of<T>::type;
  __shared__ double sh[8];
  int row = blockIdx.y, c = blockIdx.x;
  int64_t k0 = c * s.len, k1 = k0 + s.len < M ? k0 + s.len : M;
  double a_rz = 0, a_bb = 0;
  for (int64_t k = k0 + threadIdx.x; k < k1; k += blockDim.x) {
    int64_t i = (int64_t)row * M + k;
    C bv = b[i], av;
    if (Ap) {
      av = Ap[i];
    } else {            // x0 == 0: A x0 = 0 was not computed
      av.x = 0;
      av.y = 0;
    }
    C rv;
    rv.x = bv.x - av.x;
    rv.y = bv.y - av.y;
    T mi = minv ? minv[k] : T(1);
    C pv;
    pv.x = rv.x * mi;
    pv.y = rv.y * mi;
    r[i] = rv;
    p[i] = pv;
    a_rz += (double)mi * ((double)rv.x * rv.x + (double)rv.y * rv.y);
    a_bb += (double)bv.x * bv.x + (double)bv.y * bv.y;
  }
  double t1 = block_sum(a_rz, sh);
  double t2 = block_sum(a_bb, sh);
  if (threadIdx.x == 0) {
    s.p_rz[row * s.NC + c] = t1;
    s.p_rr[row * s.NC + c] = t2;  // holds ||b||^2 partials during init
  }
}

__global__ void __launch_bounds__(256) cg_init_scal_kernel(CgScal s) {   // one block per row
  const int row = blockIdx.x;
  double rz = sum_partials(s.p_rz + row * s.NC, s.NC);
  double bb = sqrt(sum_partials(s.p_rr + row * s.NC, s.NC));
  if (threadIdx.x != 0) return;
  s.rz[s.B + row] = rz;  // parity 1 == "previous" for iteration 0
  s.denom[row] = bb > 0 ? bb : 1.0;
  s.act[s.B + row] = 1;
  s.conv_iter[row] = -1;
  if (row == 0) {
    s.counters[0] = s.B;
    s.counters[1] = -1;
  }
}

// Convergence decision shared by kernel A and the finalize kernel (reads partials of iteration it-1).
__device__ __forceinline__ bool row_converged(const CgScal& s, int row, double tol, int early, int batched,
                                              double& rz_new) {
  double rr = sum_partials(s.p_rr + row * s.NC, s.NC);
  rz_new = sum_partials(s.p_rz + row * s.NC, s.NC);
  if (!early) return false;
  double rn = sqrt(rr);
  bool conv = rn / (s.denom[row] + 1e-16) < tol;
  if (batched) conv = conv || rn < 1e-12;
  return conv;
}

template <typename T, bool SYM>
__global__ void __launch_bounds__(256)
cg_A_kernel(typename cplx_of<T>::type* __restrict__ p, const typename cplx_of<T>::type* __restrict__ r,
            const T* __restrict__ minv, const T* __restrict__ ws, TzGeomC g, int it, double tol, int early,
            int batched, typename cplx_of<T>::type* __restrict__ P, CgScal s) {
  using C = typename cplx_of<T>::type;
  int row = blockIdx.y, c = blockIdx.x;
  const int cur = it & 1, prev = cur ^ 1;
  if (!s.act[prev * s.B + row]) {
    if (c == 0 && threadIdx.x == 0) s.act[cur * s.B + row] = 0;
    return;
  }
  double beta = 0;
  if (it > 0) {
    double rz_new;
    bool conv = row_converged(s, row, tol, early, batched, rz_new);
    if (conv) {
      if (c == 0 && threadIdx.x == 0) {
        s.act[cur * s.B + row] = 0;
        s.conv_iter[row] = it - 1;
        atomicSub(&s.counters[0], 1);
        atomicMax(&s.counters[1], it - 1);
      }
      return;
    }
    beta = rz_new / (s.rz[prev * s.B + row] + 1e-16);
    if (c == 0 && threadIdx.x == 0) {
      s.rz[cur * s.B + row] = rz_new;
      s.act[cur * s.B + row] = 1;
    }
  } else if (c == 0 && threadIdx.x == 0) {
    s.rz[cur * s.B + row] = s.rz[prev * s.B + row];
    s.act[cur * s.B + row] = 1;
  }
  int64_t k0 = c * s.len, k1 = k0 + s.len < g.M ? k0 + s.len : g.M;
  const T bt = (T)beta;
  for (int64_t k = k0 + threadIdx.x; k < k1; k += blockDim.x) {
    int64_t i = (int64_t)row * g.M + k;
    C pv = p[i];
    if (it > 0) {
      C rv = r[i];
      T mi = minv ? minv[k] : T(1);
      pv.x = rv.x * mi + bt * pv.x;
      pv.y = rv.y * mi + bt * pv.y;
      p[i] = pv;
    }
    T w = ws ? ws[k] : T(1);
    int64_t po, co;
    tzc_offsets(g, k, po, co);
    C o;
    o.x = pv.x * w;
    o.y = pv.y * w;
    if (SYM) {        // half spectrum of the real array X(f): conj(ws p) at the modes with centred last index >= 0
      int64_t ho;
      bool neg;
      tzc_sym_offset(g, k, ho, neg);
      if (!neg) {
        o.y = -o.y;
        P[(int64_t)row * g.Htot + ho] = o;
      }
    } else {
      P[(int64_t)row * g.Ftot + po] = o;
    }
  }
}

// Hermitian path: v <- (v + conj(mirror v)) / 2 per row.  On an odd centred grid the mirror of flat index k is M-1-k.
// The real-FFT matvec reads only half of p and returns an exactly symmetric Ap, so an asymmetric component of the
// start (rounding-level in a warm start) would never be reduced and would leak into x through p: project it out once.
template <typename T>
__global__ void __launch_bounds__(256)
cg_symmetrize_kernel(typename cplx_of<T>::type* __restrict__ v, int64_t M) {
  using C = typename cplx_of<T>::type;
  const int row = blockIdx.y;
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t km = M - 1 - k;
  if (k > km) return;
  C* base = v + (int64_t)row * M;
  C a = base[k], b = base[km];
  C sm;
  sm.x = T(0.5) * (a.x + b.x);
  sm.y = T(0.5) * (a.y - b.y);
  if (k == km) sm.y = 0;
  base[k] = sm;
  sm.y = -sm.y;
  base[km] = sm;
}

// Hermitian path: X(f) *= V(f), both real
template <typename T>
__global__ void __launch_bounds__(256)
cg_Mr_kernel(T* __restrict__ R, const T* __restrict__ vr, int64_t Ftot, const int* __restrict__ act) {
  using V2 = typename cplx_of<T>::type;      // two reals per access (rows start 16-byte aligned when Ftot is even)
  int row = blockIdx.y;
  if (!act[row]) return;
  int64_t f = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 2;
  if (f >= Ftot) return;
  T* base = R + (int64_t)row * Ftot;
  if ((Ftot & 1) == 0) {
    V2 a = *reinterpret_cast<V2*>(base + f);
    const V2 h = *reinterpret_cast<const V2*>(vr + f);
    a.x *= h.x;
    a.y *= h.y;
    *reinterpret_cast<V2*>(base + f) = a;
  } else {
    base[f] *= vr[f];
    if (f + 1 < Ftot) base[f + 1] *= vr[f + 1];
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
cg_M_kernel(typename cplx_of<T>::type* __restrict__ Q, const typename cplx_of<T>::type* __restrict__ vhat,
            int64_t Ftot, const int* __restrict__ act) {
  using C = typename cplx_of<T>::type;
  int row = blockIdx.y;
  if (!act[row]) return;
  int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= Ftot) return;
  C a = Q[(int64_t)row * Ftot + f];
  C h = vhat[f];
  C o;
  o.x = a.x * h.x - a.y * h.y;
  o.y = a.x * h.y + a.y * h.x;
  Q[(int64_t)row * Ftot + f] = o;
}

template <typename T, bool SYM>
__global__ void __launch_bounds__(256)
cg_B_kernel(const typename cplx_of<T>::type* __restrict__ Q, const typename cplx_of<T>::type* __restrict__ p,
            const T* __restrict__ ws, TzGeomC g, T c1, T c2, int it,
            typename cplx_of<T>::type* __restrict__ Ap, CgScal s) {
  using C = typename cplx_of<T>::type;
  __shared__ double sh[8];
  int row = blockIdx.y, c = blockIdx.x;
  if (!s.act[(it & 1) * s.B + row]) return;
  int64_t k0 = c * s.len, k1 = k0 + s.len < g.M ? k0 + s.len : g.M;
  double acc = 0;
  for (int64_t k = k0 + threadIdx.x; k < k1; k += blockDim.x) {
    int64_t i = (int64_t)row * g.M + k;
    int64_t po, co;
    tzc_offsets(g, k, po, co);
    C q;
    if (SYM) {        // y[k] = conj(out[k]) for stored modes, y[k] = conj(y[-k]) = out[-k] for the other half
      int64_t ho;
      bool neg;
      tzc_sym_offset(g, k, ho, neg);
      q = Q[(int64_t)row * g.Htot + ho];
      if (!neg) q.y = -q.y;
    } else {
      q = Q[(int64_t)row * g.Ftot + co];
    }
    C pv = p[i];
    T w = ws ? ws[k] * c1 : c1;
    C o;
    o.x = q.x * w + c2 * pv.x;
    o.y = q.y * w + c2 * pv.y;
    Ap[i] = o;
    acc += (double)pv.x * o.x + (double)pv.y * o.y;
  }
  double t = block_sum(acc, sh);
  if (threadIdx.x == 0) s.p_pAp[row * s.NC + c] = t;
}

template <typename T>
__device__ __forceinline__ void cg_C_body(typename cplx_of<T>::type* __restrict__ x, typename cplx_of<T>::type* __restrict__ r,
                                          const typename cplx_of<T>::type* __restrict__ p,
                                          const typename cplx_of<T>::type* __restrict__ Ap, const T* __restrict__ minv,
                                          int64_t M, int it, const CgScal& s, int row, int c) {
  using C = typename cplx_of<T>::type;
  __shared__ double sh[8];
  if (!s.act[(it & 1) * s.B + row]) return;
  double pAp = sum_partials(s.p_pAp + row * s.NC, s.NC) + 1e-16;
  const T alpha = (T)(s.rz[(it & 1) * s.B + row] / pAp);
  int64_t k0 = c * s.len, k1 = k0 + s.len < M ? k0 + s.len : M;
  double a_rr = 0, a_rz = 0;
  for (int64_t k = k0 + threadIdx.x; k < k1; k += blockDim.x) {
    int64_t i = (int64_t)row * M + k;
    C xv = x[i], rv = r[i], pv = p[i], av = Ap[i];
    xv.x += alpha * pv.x;
    xv.y += alpha * pv.y;
    rv.x -= alpha * av.x;
    rv.y -= alpha * av.y;
    x[i] = xv;
    r[i] = rv;
    double n2 = (double)rv.x * rv.x + (double)rv.y * rv.y;
    a_rr += n2;
    a_rz += (double)(minv ? minv[k] : T(1)) * n2;
  }
  double t1 = block_sum(a_rr, sh);
  double t2 = block_sum(a_rz, sh);
  if (threadIdx.x == 0) {
    s.p_rr[row * s.NC + c] = t1;
    s.p_rz[row * s.NC + c] = t2;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
cg_C_kernel(typename cplx_of<T>::type* __restrict__ x, typename cplx_of<T>::type* __restrict__ r,
            const typename cplx_of<T>::type* __restrict__ p, const typename cplx_of<T>::type* __restrict__ Ap,
            const T* __restrict__ minv, int64_t M, int it, CgScal s) {
  cg_C_body<T>(x, r, p, Ap, minv, M, it, s, blockIdx.y, blockIdx.x);
}

// ------------------------------------------------------------------ tiny 2-D blocks: matvec by direct DFTs
// For the posterior-mean system the block is tiny (71 x 71 unknowns, 144 x 144 embedding at the headline size):
// the cuFFT pair is 4 launches of ~3 us each and the iteration is bound by launch latency, not by work.  A
// zero-padded, cropped 2-D transform written as plain DFT sums costs 2 n0 n1 F1 + 2 F0 F1 n0 complex MACs
// (4.4e6 here: about a microsecond of FP64), needs no padded buffer, and fuses with the vector updates, so an
// iteration is 4 launches (D1, D2, D3, C) instead of 8:
//   D1  rows   : beta, p = M^-1 r + beta p,   U[j0, k1] = sum_j1 ws p[j0, j1] w1^(k1 j1)
//   D2  columns: V[k0] = sum_j0 U[j0, k1] w0^(k0 j0);  V *= vhat[k0, k1];  X[i0, k1] = sum_k0 V[k0] w0^-(k0 (i0 + n0 - 1))
//   D3  rows   : Y[i0, i1] = sum_k1 X[i0, k1] w1^-(k1 (i1 + n1 - 1));  Ap = c1 ws Y + c2 p,  partial <p, Ap>
// with w = exp(-2 pi i / F): the same unnormalised transforms cuFFT computes, so the cached vhat (pre-scaled by
// 1 / Ftot) is shared with the FFT path.  Chunks of the reductions are the block's rows (NC = n0, len = n1).
template <typename C>
__device__ __forceinline__ C cmul(C a, C b) {
  C o;
  o.x = a.x * b.x - a.y * b.y;
  o.y = a.x * b.y + a.y * b.x;
  return o;
}
template <typename C>
__device__ __forceinline__ void cmac(C& acc, C a, C b) {   // acc += a * b
  acc.x = fma(a.x, b.x, acc.x);
  acc.x = fma(-a.y, b.y, acc.x);
  acc.y = fma(a.x, b.y, acc.y);
  acc.y = fma(a.y, b.x, acc.y);
}
template <typename C>
__device__ __forceinline__ void cmac_conj(C& acc, C a, C b) {   // acc += a * conj(b)
  acc.x = fma(a.x, b.x, acc.x);
  acc.x = fma(a.y, b.y, acc.x);
  acc.y = fma(a.y, b.x, acc.y);
  acc.y = fma(-a.x, b.y, acc.y);
}

// One output of a DFT sum, split over G = 4 adjacent lanes (terms j = g, g + 4, ...) with two independent
// accumulators per lane, then combined with two shuffles: sum_j data[j] * tw[(j * step) mod F] (or its conjugate
// twiddle).  Every lane of the warp must call it (the shuffles are full-mask); lanes of one group get the sum.
constexpr int kDftG = 4;
template <typename C, bool CONJ>
__device__ __forceinline__ C dft_sum4(const C* __restrict__ data, int nterms, const C* __restrict__ tw, int F, int step) {
  const int g = threadIdx.x & (kDftG - 1);
  C a0, a1;
  a0.x = a0.y = a1.x = a1.y = 0;
  int idx = (g * step) % F;
  const int d1 = (kDftG * step) % F;
  int j = g;
  for (; j + kDftG < nterms; j += 2 * kDftG) {
    int idx2 = idx + d1;
    if (idx2 >= F) idx2 -= F;
    if (CONJ) {
      cmac_conj(a0, data[j], tw[idx]);
      cmac_conj(a1, data[j + kDftG], tw[idx2]);
    } else {
      cmac(a0, data[j], tw[idx]);
      cmac(a1, data[j + kDftG], tw[idx2]);
    }
    idx = idx2 + d1;
    if (idx >= F) idx -= F;
  }
  if (j < nterms) {
    if (CONJ) cmac_conj(a0, data[j], tw[idx]); else cmac(a0, data[j], tw[idx]);
  }
  a0.x += a1.x;
  a0.y += a1.y;
#pragma unroll
  for (int o = 1; o < kDftG; o <<= 1) {
    a0.x += __shfl_xor_sync(0xffffffffu, a0.x, o);
    a0.y += __shfl_xor_sync(0xffffffffu, a0.y, o);
  }
  return a0;
}

template <typename T>
__global__ void dft_twiddle_kernel(typename cplx_of<T>::type* tw, int F) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= F) return;
  double sn, cs;
  sincospi(-2.0 * (double)k / (double)F, &sn, &cs);
  tw[k].x = (T)cs;
  tw[k].y = (T)sn;
}

constexpr int kDftMaxF = 256;

// Returns whether `row` is still active in iteration `it` (the same answer in every CTA that asks: the decision
// only reads the reduced scalars of iteration it - 1).  CTA c = 0 records the decision; CTAs with c >= n0 (the
// persistent kernel has more CTAs than block rows) only decide.
template <typename T>
__device__ __forceinline__ bool cg_D1_body(typename cplx_of<T>::type* __restrict__ p,
                                           const typename cplx_of<T>::type* __restrict__ r, const T* __restrict__ minv,
                                           const T* __restrict__ ws, int n0, int n1, int F1, int it, double tol, int early,
                                           int batched, const typename cplx_of<T>::type* __restrict__ tw1g,
                                           typename cplx_of<T>::type* __restrict__ U, const CgScal& s, int row, int c) {
  using C = typename cplx_of<T>::type;
  __shared__ C zrow[kDftMaxF];
  __shared__ C tw1[kDftMaxF];
  const int cur = it & 1, prev = cur ^ 1;
  if (!s.act[prev * s.B + row]) {
    if (c == 0 && threadIdx.x == 0) s.act[cur * s.B + row] = 0;
    return false;
  }
  double beta = 0;
  if (it > 0) {
    double rz_new;
    bool conv = row_converged(s, row, tol, early, batched, rz_new);
    if (conv) {
      if (c == 0 && threadIdx.x == 0) {
        s.act[cur * s.B + row] = 0;
        s.conv_iter[row] = it - 1;
        atomicSub(&s.counters[0], 1);
        atomicMax(&s.counters[1], it - 1);
      }
      return false;
    }
    beta = rz_new / (s.rz[prev * s.B + row] + 1e-16);
    if (c == 0 && threadIdx.x == 0) {
      s.rz[cur * s.B + row] = rz_new;
      s.act[cur * s.B + row] = 1;
    }
  } else if (c == 0 && threadIdx.x == 0) {
    s.rz[cur * s.B + row] = s.rz[prev * s.B + row];
    s.act[cur * s.B + row] = 1;
  }
  if (c >= n0) return true;
  const int64_t M = (int64_t)n0 * n1;
  const T bt = (T)beta;
  __syncthreads();            // (persistent kernel) the previous use of zrow / tw1 by this CTA is over
  for (int j = threadIdx.x; j < F1; j += blockDim.x) tw1[j] = tw1g[j];
  for (int j1 = threadIdx.x; j1 < n1; j1 += blockDim.x) {
    const int64_t k = (int64_t)c * n1 + j1, i = (int64_t)row * M + k;
    C pv = p[i];
    if (it > 0) {
      C rv = r[i];
      T mi = minv ? minv[k] : T(1);
      pv.x = rv.x * mi + bt * pv.x;
      pv.y = rv.y * mi + bt * pv.y;
      p[i] = pv;
    }
    T w = ws ? ws[k] : T(1);
    C o;
    o.x = pv.x * w;
    o.y = pv.y * w;
    zrow[j1] = o;
  }
  __syncthreads();
  C* Urow = U + ((int64_t)row * n0 + c) * F1;
  const int per_pass = blockDim.x / kDftG;
  for (int base = 0; base < F1; base += per_pass) {          // uniform trip count: full-mask shuffles inside
    const int k1 = base + (int)(threadIdx.x / kDftG);
    const C acc = dft_sum4<C, false>(zrow, k1 < F1 ? n1 : 0, tw1, F1, k1 < F1 ? k1 : 0);
    if (k1 < F1 && (threadIdx.x & (kDftG - 1)) == 0) Urow[k1] = acc;
  }
  return true;
}

template <typename T>
__global__ void __launch_bounds__(256)
cg_D1_kernel(typename cplx_of<T>::type* __restrict__ p, const typename cplx_of<T>::type* __restrict__ r,
             const T* __restrict__ minv, const T* __restrict__ ws, int n1, int F1, int it, double tol, int early,
             int batched, const typename cplx_of<T>::type* __restrict__ tw1g, typename cplx_of<T>::type* __restrict__ U,
             CgScal s) {
  cg_D1_body<T>(p, r, minv, ws, gridDim.x, n1, F1, it, tol, early, batched, tw1g, U, s, blockIdx.y, blockIdx.x);
}

template <typename T>
__device__ __forceinline__ void cg_D2_body(const typename cplx_of<T>::type* __restrict__ U,
                                           const typename cplx_of<T>::type* __restrict__ vhat, int n0, int F0, int F1,
                                           const typename cplx_of<T>::type* __restrict__ tw0g,
                                           typename cplx_of<T>::type* __restrict__ X, const int* __restrict__ act,
                                           int row, int k1) {
  using C = typename cplx_of<T>::type;
  __shared__ C ucol[kDftMaxF];
  __shared__ C vcol[kDftMaxF];
  __shared__ C tw0[kDftMaxF];
  if (!act[row]) return;
  __syncthreads();            // (persistent kernel) the previous use of the shared columns by this CTA is over
  for (int j = threadIdx.x; j < F0; j += blockDim.x) tw0[j] = tw0g[j];
  for (int j0 = threadIdx.x; j0 < n0; j0 += blockDim.x) ucol[j0] = U[((int64_t)row * n0 + j0) * F1 + k1];
  __syncthreads();
  const int per_pass = blockDim.x / kDftG, og = threadIdx.x / kDftG;
  const bool lead = (threadIdx.x & (kDftG - 1)) == 0;
  for (int base = 0; base < F0; base += per_pass) {
    const int k0 = base + og;
    C h;
    h.x = h.y = 0;
    if (k0 < F0 && lead) h = vhat[(int64_t)k0 * F1 + k1];     // in flight during the sum
    const C acc = dft_sum4<C, false>(ucol, k0 < F0 ? n0 : 0, tw0, F0, k0 < F0 ? k0 : 0);
    if (k0 < F0 && lead) vcol[k0] = cmul(acc, h);
  }
  __syncthreads();
  for (int base = 0; base < n0; base += per_pass) {
    const int i0 = base + og;
    const C acc = dft_sum4<C, true>(vcol, i0 < n0 ? F0 : 0, tw0, F0, i0 < n0 ? (i0 + n0 - 1) % F0 : 0);
    if (i0 < n0 && lead) X[((int64_t)row * n0 + i0) * F1 + k1] = acc;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
cg_D2_kernel(const typename cplx_of<T>::type* __restrict__ U, const typename cplx_of<T>::type* __restrict__ vhat,
             int n0, int F0, int F1, const typename cplx_of<T>::type* __restrict__ tw0g,
             typename cplx_of<T>::type* __restrict__ X, const int* __restrict__ act) {
  cg_D2_body<T>(U, vhat, n0, F0, F1, tw0g, X, act, blockIdx.y, blockIdx.x);
}

template <typename T>
__device__ __forceinline__ void cg_D3_body(const typename cplx_of<T>::type* __restrict__ X,
                                           const typename cplx_of<T>::type* __restrict__ p, const T* __restrict__ ws,
                                           int n0, int n1, int F1, T c1, T c2, int it,
                                           const typename cplx_of<T>::type* __restrict__ tw1g,
                                           typename cplx_of<T>::type* __restrict__ Ap, const CgScal& s, int row, int c) {
  using C = typename cplx_of<T>::type;
  __shared__ C xrow[kDftMaxF];
  __shared__ C tw1[kDftMaxF];
  __shared__ double sh[8];
  if (!s.act[(it & 1) * s.B + row]) return;
  const int64_t M = (int64_t)n0 * n1;
  __syncthreads();            // (persistent kernel) the previous use of xrow / tw1 by this CTA is over
  for (int j = threadIdx.x; j < F1; j += blockDim.x) {
    tw1[j] = tw1g[j];
    xrow[j] = X[((int64_t)row * n0 + c) * F1 + j];
  }
  __syncthreads();
  double dot = 0;
  const int per_pass = blockDim.x / kDftG, og = threadIdx.x / kDftG;
  const bool lead = (threadIdx.x & (kDftG - 1)) == 0;
  for (int base = 0; base < n1; base += per_pass) {
    const int i1 = base + og;
    const bool on = i1 < n1 && lead;
    const int64_t k = (int64_t)c * n1 + (i1 < n1 ? i1 : 0), i = (int64_t)row * M + k;
    C pv;
    pv.x = pv.y = 0;
    T w = c1;
    if (on) {                                                  // in flight during the sum
      pv = p[i];
      if (ws) w = ws[k] * c1;
    }
    const C acc = dft_sum4<C, true>(xrow, i1 < n1 ? F1 : 0, tw1, F1, i1 < n1 ? (i1 + n1 - 1) % F1 : 0);
    if (on) {
      C o;
      o.x = acc.x * w + c2 * pv.x;
      o.y = acc.y * w + c2 * pv.y;
      Ap[i] = o;
      dot += (double)pv.x * o.x + (double)pv.y * o.y;
    }
  }
  double tsum = block_sum(dot, sh);
  if (threadIdx.x == 0) s.p_pAp[row * s.NC + c] = tsum;
}

template <typename T>
__global__ void __launch_bounds__(256)
cg_D3_kernel(const typename cplx_of<T>::type* __restrict__ X, const typename cplx_of<T>::type* __restrict__ p,
             const T* __restrict__ ws, int n1, int F1, T c1, T c2, int it,
             const typename cplx_of<T>::type* __restrict__ tw1g, typename cplx_of<T>::type* __restrict__ Ap, CgScal s) {
  cg_D3_body<T>(X, p, ws, gridDim.x, n1, F1, c1, c2, it, tw1g, Ap, s, blockIdx.y, blockIdx.x);
}

// The whole solve of a tiny block as ONE cooperative launch: the four phases of an iteration are separated by
// grid-wide barriers instead of kernel boundaries, so an iteration costs four barriers (~1.5 us each on 144 CTAs)
// instead of four launches with their dependent-launch latency, and the host neither launches per iteration nor
// reads convergence counters in between.  Every CTA takes the same decisions from the same reduced scalars, so
// all of them leave the loop in the same iteration (a barrier is never skipped by part of the grid).
template <typename T>
struct CgPersistArgs {
  typename cplx_of<T>::type *x, *r, *p, *Ap, *U, *X;
  const typename cplx_of<T>::type *vhat, *tw0, *tw1;
  const T *minv, *ws;
  int n0, n1, F0, F1;
  T c1, c2;
  double tol;
  int early, batched, max_iter;
  CgScal s;
};

template <typename T>
__global__ void __launch_bounds__(256) cg_persist_kernel(const CgPersistArgs<T> a) {
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  const int c = blockIdx.x, B = a.s.B;
  const int64_t M = (int64_t)a.n0 * a.n1;
  int it = 0;
  for (; it < a.max_iter; ++it) {
    bool any = false;
    for (int row = 0; row < B; ++row)
      any |= cg_D1_body<T>(a.p, a.r, a.minv, a.ws, a.n0, a.n1, a.F1, it, a.tol, a.early, a.batched, a.tw1, a.U, a.s,
                           row, c);
    if (!any) break;                     // identical in every CTA
    grid.sync();
    if (c < a.F1)
      for (int row = 0; row < B; ++row)
        cg_D2_body<T>(a.U, a.vhat, a.n0, a.F0, a.F1, a.tw0, a.X, a.s.act + (it & 1) * B, row, c);
    grid.sync();
    if (c < a.n0)
      for (int row = 0; row < B; ++row)
        cg_D3_body<T>(a.X, a.p, a.ws, a.n0, a.n1, a.F1, a.c1, a.c2, it, a.tw1, a.Ap, a.s, row, c);
    grid.sync();
    if (c < a.n0)
      for (int row = 0; row < B; ++row) cg_C_body<T>(a.x, a.r, a.p, a.Ap, a.minv, M, it, a.s, row, c);
    grid.sync();
  }
  if (c == 0 && it == a.max_iter && it > 0) {      // what cg_final_kernel does after the last iteration
    const int prev = (it & 1) ^ 1;
    for (int row = 0; row < B; ++row) {
      if (!a.s.act[prev * B + row]) continue;
      double rz_new;
      if (row_converged(a.s, row, a.tol, a.early, a.batched, rz_new) && threadIdx.x == 0) {
        a.s.act[prev * B + row] = 0;
        a.s.conv_iter[row] = it - 1;
        atomicSub(&a.s.counters[0], 1);
        atomicMax(&a.s.counters[1], it - 1);
      }
    }
  }
}

// After `it` iterations: register rows that converged in iteration it-1.
__global__ void __launch_bounds__(256) cg_final_kernel(int it, double tol, int early, int batched, CgScal s) {
  const int row = blockIdx.x;           // one block per row
  if (it <= 0) return;
  const int prev = (it & 1) ^ 1;
  if (!s.act[prev * s.B + row]) return;
  double rz_new;
  if (row_converged(s, row, tol, early, batched, rz_new) && threadIdx.x == 0) {
    s.act[prev * s.B + row] = 0;
    s.conv_iter[row] = it - 1;
    atomicSub(&s.counters[0], 1);
    atomicMax(&s.counters[1], it - 1);
  }
}

template <typename T>
static void solve_group(Toeplitz& t, const T* ws, double sigmasq, int mode, const T* minv, const void* b_, void* x_,
                        int B, double tol, int max_iter, int early, int batched, int x0_zero, int& all_conv, int& jmax,
                        cudaStream_t st, int hermitian, int* async_counters) {
  using C = typename cplx_of<T>::type;
  const int64_t M = t.M;
  TzGeomC g;
  g.dim = t.dim;
  for (int a = 0; a < kMaxDim; ++a) {
    g.n[a] = (int)t.n[a];
    g.F[a] = (int)t.F[a];
  }
  g.M = M;
  g.Ftot = t.Ftot;
  bool odd = true;
  g.Htot = 1;
  for (int a = 0; a < kMaxDim; ++a) {
    g.mc[a] = (g.n[a] - 1) / 2;
    if (a < t.dim) {
      odd = odd && (g.n[a] & 1);
      g.Htot *= (a == t.dim - 1) ? g.F[a] / 2 + 1 : g.F[a];
    }
  }
  g.Fh = g.F[t.dim - 1] / 2 + 1;
  const C* b = (const C*)b_;
  C* x = (C*)x_;
  T c1 = 1, c2 = 0;
  const T* w = ws;
  if (mode == kModeT) w = nullptr;
  if (mode == kModeAmean) c2 = (T)sigmasq;
  if (mode == kModeAvar) {
    c1 = (T)(1.0 / sigmasq);
    c2 = 1;
  }
  // scratch
  C* vec = (C*)t.cg_vec.get((size_t)3 * B * M * sizeof(C));
  C *r = vec, *p = vec + (int64_t)B * M, *Ap = vec + (int64_t)2 * B * M;
  CgScal s;
  s.B = B;
  s.NC = (int)std::min<int64_t>(kMaxChunks, std::max<int64_t>(1, (M + kChunkLen - 1) / kChunkLen));
  s.len = (M + s.NC - 1) / s.NC;
  // tiny 2-D blocks with few right-hand sides: direct-DFT matvec, 4 launches per iteration (see cg_D1_kernel)
  static const bool dft_off = std::getenv("EFGP_CG_FFT") != nullptr;     // A/B switch: force the cuFFT path
  const bool direct = !dft_off && t.dim == 2 && B <= 4 && t.F[0] <= kDftMaxF && t.F[1] <= kDftMaxF &&
                      t.n[0] <= kMaxChunks;
  const int n0 = (int)t.n[0], n1 = (int)t.n[1], F0 = (int)t.F[0], F1 = (int)t.F[1];
  C *tw0 = nullptr, *tw1 = nullptr, *U = nullptr, *X = nullptr;
  if (direct) {
    s.NC = n0;            // reduction chunks = rows of the block
    s.len = n1;
    const bool fresh = t.dft_tw.p == nullptr;
    tw0 = (C*)t.dft_tw.get(((size_t)F0 + F1) * sizeof(C));
    tw1 = tw0 + F0;
    U = (C*)t.dft.get((size_t)2 * B * n0 * F1 * sizeof(C));
    X = U + (size_t)B * n0 * F1;
    if (fresh || !t.dft_ready) {
      dft_twiddle_kernel<T><<<div_up(F0, 128), 128, 0, st>>>(tw0, F0);
      dft_twiddle_kernel<T><<<div_up(F1, 128), 128, 0, st>>>(tw1, F1);
      t.dft_ready = true;
    }
  }
  size_t nd = (size_t)2 * B + B + (size_t)3 * B * s.NC;
  size_t ni = (size_t)2 * B + B + 4;
  char* sc = (char*)t.cg_scal.get(nd * sizeof(double) + ni * sizeof(int));
  double* dp = (double*)sc;
  s.rz = dp;
  s.denom = dp + 2 * B;
  s.p_pAp = dp + 3 * B;
  s.p_rr = s.p_pAp + (size_t)B * s.NC;
  s.p_rz = s.p_rr + (size_t)B * s.NC;
  int* ip = (int*)(dp + nd);
  s.act = ip;
  s.conv_iter = ip + 2 * B;
  s.counters = ip + 3 * B;

  // FFT buffers (P zero outside the block)
  size_t need = (size_t)B * t.Ftot * sizeof(C);
  bool grow = need > t.P.cap;
  C* P = (C*)t.P.get(need);
  if (grow || t.P_rows < B) {
    EFGP_CUDA(cudaMemsetAsync(P, 0, need, st));
    t.P_rows = B;
  }
  C* Q = (C*)t.Q.get(need);
  // Hermitian right-hand sides (caller's promise) on odd blocks: real-FFT matvec in the cuFFT loop below
  static const bool sym_off = std::getenv("EFGP_CG_NO_HERMITIAN") != nullptr;     // A/B switch
  const bool sym = hermitian && odd && !direct && !sym_off;
  const T* vreal = sym ? (const T*)toeplitz_real_spectrum(t, st) : nullptr;

  // single Hermitian right-hand side on a small odd 2-D block: direct-convolution solve (cg_small.cu)
  static const bool conv_off = std::getenv("EFGP_CG_CONV") && std::atoi(std::getenv("EFGP_CG_CONV")) == 0;
  const bool try_conv = direct && !conv_off && hermitian && odd && B == 1 && M <= 6400 && t.v_raw != nullptr && w != nullptr;
  // Ap = A x0  (generic apply writes through P/Q as well); skipped for a cold start
  const dim3 gsym(div_up((M + 1) / 2, 256), B);
  if ((sym || try_conv) && !x0_zero) cg_symmetrize_kernel<T><<<gsym, 256, 0, st>>>(x, M);
  if (!x0_zero) toeplitz_apply(t, x, B, ws, sigmasq, mode, Ap, st);
  dim3 grid(s.NC, B), gf(div_up(t.Ftot, 256), B);
  {
    ScopedStage stg_init("cg.init", st, 2);
    cg_init_kernel<T><<<grid, 256, 0, st>>>(b, x0_zero ? (const C*)nullptr : (const C*)Ap, minv, r, p, M, s);
    cg_init_scal_kernel<<<B, 256, 0, st>>>(s);
    if (sym || try_conv) {
      cg_symmetrize_kernel<T><<<gsym, 256, 0, st>>>(r, M);
      cg_symmetrize_kernel<T><<<gsym, 256, 0, st>>>(p, M);
    }
  }
  EFGP_CUDA(cudaGetLastError());

  static int* h_counters = nullptr;   // pinned, process lifetime (cudaMallocHost costs ~0.5 ms per call)
  if (!h_counters) EFGP_CUDA(cudaMallocHost(&h_counters, 2 * sizeof(int)));
  h_counters[0] = B;
  h_counters[1] = -1;
  // When to look at the convergence counters first: every read-back stalls the host behind the stream, after
  // which the remaining iterations are launch-bound.  Solves of the same kind (operator mode, single/batched)
  // repeat with nearly the same iteration count in the EFGP loop, so the first check is scheduled where the
  // previous such solve stopped; iterations past convergence are no-ops (rows are frozen on the device).
  static int hint[4][2][2] = {};         // per operator mode, single/batched, general/Hermitian matvec
  int& my_hint = hint[mode][batched ? 1 : 0][sym ? 1 : 0];
  int it = 0;
  int check = my_hint > 0 ? std::min(my_hint, 64) : 4;
  bool first = true;
  ScopedStage stg("cg.iterations", st, 0);
  if (try_conv && cg_conv_launch<T>(t, w, minv, x, r, p, c1, c2, tol, early, max_iter, s.rz + s.B, s.denom, s.counters,
                                    s.conv_iter, st)) {
    stg.add_launches(1);
    if (async_counters) {
      EFGP_CUDA(cudaMemcpyAsync(async_counters, s.counters, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
      all_conv = -1;
      return;
    }
    EFGP_CUDA(cudaMemcpyAsync(h_counters, s.counters, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
    EFGP_CUDA(cudaStreamSynchronize(st));
    all_conv = early && h_counters[0] <= 0;
    jmax = h_counters[1];
    return;
  }
  if (direct) {
    static const bool persist_off = std::getenv("EFGP_CG_PERSIST") && std::atoi(std::getenv("EFGP_CG_PERSIST")) == 0;
    static int sms = 0, coop = 0;
    if (!sms) {
      int dev = 0;
      EFGP_CUDA(cudaGetDevice(&dev));
      EFGP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      EFGP_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    }
    int per_sm = 0;
    EFGP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, cg_persist_kernel<T>, 256, 0));
    const int ctas = std::max(n0, F1);
    if (!persist_off && coop && ctas <= per_sm * sms) {
      CgPersistArgs<T> a;
      a.x = x; a.r = r; a.p = p; a.Ap = Ap; a.U = U; a.X = X;
      a.vhat = (const C*)t.vhat; a.tw0 = tw0; a.tw1 = tw1;
      a.minv = minv; a.ws = w;
      a.n0 = n0; a.n1 = n1; a.F0 = F0; a.F1 = F1;
      a.c1 = c1; a.c2 = c2;
      a.tol = tol; a.early = early; a.batched = batched; a.max_iter = max_iter;
      a.s = s;
      void* args[] = {(void*)&a};
      stg.add_launches(1);
      EFGP_CUDA(cudaLaunchCooperativeKernel((void*)cg_persist_kernel<T>, dim3(ctas), dim3(256), args, 0, st));
      if (async_counters) {      // the whole solve is on the stream: hand the counters over without waiting for it
        EFGP_CUDA(cudaMemcpyAsync(async_counters, s.counters, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
        all_conv = -1;
        return;
      }
      EFGP_CUDA(cudaMemcpyAsync(h_counters, s.counters, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
      EFGP_CUDA(cudaStreamSynchronize(st));
      all_conv = early && h_counters[0] <= 0;
      jmax = h_counters[1];
      return;
    }
  }
  while (it < max_iter) {
    int until = std::min(max_iter, it + check);
    stg.add_launches(4 * (until - it));
    for (; direct && it < until; ++it) {
      cg_D1_kernel<T><<<grid, 256, 0, st>>>(p, r, minv, w, n1, F1, it, tol, early, batched, tw1, U, s);
      cg_D2_kernel<T><<<dim3(F1, B), 256, 0, st>>>(U, (const C*)t.vhat, n0, F0, F1, tw0, X, s.act + (it & 1) * B);
      cg_D3_kernel<T><<<grid, 256, 0, st>>>(X, p, w, n1, F1, c1, c2, it, tw1, Ap, s);
      cg_C_kernel<T><<<grid, 256, 0, st>>>(x, r, p, Ap, minv, M, it, s);
    }
    for (; sym && it < until; ++it) {
      // half-size matvec: H = conj(ws p) on the stored half of the modes (C2R input, clobbered by cuFFT -> zeroed
      // every iteration), X = Z2D(H) real, X *= V, H = D2Z(X), Ap from conj(H)
      C* H = P;
      T* R = (T*)Q;
      EFGP_CUDA(cudaMemsetAsync(H, 0, (size_t)B * g.Htot * sizeof(C), st));
      cg_A_kernel<T, true><<<grid, 256, 0, st>>>(p, r, minv, w, g, it, tol, early, batched, H, s);
      {
        cufftHandle pl = global_fft().get(t.dim, t.F, sizeof(T) == 4 ? CUFFT_C2R : CUFFT_Z2D, B, st);
        if (sizeof(T) == 4)
          EFGP_CUFFT(cufftExecC2R(pl, (cufftComplex*)H, (cufftReal*)R));
        else
          EFGP_CUFFT(cufftExecZ2D(pl, (cufftDoubleComplex*)H, (cufftDoubleReal*)R));
      }
      cg_Mr_kernel<T><<<dim3(div_up((t.Ftot + 1) / 2, 256), B), 256, 0, st>>>(R, vreal, t.Ftot, s.act + (it & 1) * B);
      {
        cufftHandle pl = global_fft().get(t.dim, t.F, sizeof(T) == 4 ? CUFFT_R2C : CUFFT_D2Z, B, st);
        if (sizeof(T) == 4)
          EFGP_CUFFT(cufftExecR2C(pl, (cufftReal*)R, (cufftComplex*)H));
        else
          EFGP_CUFFT(cufftExecD2Z(pl, (cufftDoubleReal*)R, (cufftDoubleComplex*)H));
      }
      cg_B_kernel<T, true><<<grid, 256, 0, st>>>(H, p, w, g, c1, c2, it, Ap, s);
      cg_C_kernel<T><<<grid, 256, 0, st>>>(x, r, p, Ap, minv, M, it, s);
    }
    for (; it < until; ++it) {
      cg_A_kernel<T, false><<<grid, 256, 0, st>>>(p, r, minv, w, g, it, tol, early, batched, P, s);
      fft_pair_begin<T>(t, B, P, Q, st);
      cg_M_kernel<T><<<gf, 256, 0, st>>>(Q, (const C*)t.vhat, t.Ftot, s.act + (it & 1) * B);
      fft_pair_end<T>(t, B, Q, st);
      cg_B_kernel<T, false><<<grid, 256, 0, st>>>(Q, p, w, g, c1, c2, it, Ap, s);
      cg_C_kernel<T><<<grid, 256, 0, st>>>(x, r, p, Ap, minv, M, it, s);
    }
    EFGP_CUDA(cudaGetLastError());
    if (early) {
      cg_final_kernel<<<B, 256, 0, st>>>(it, tol, early, batched, s);
      EFGP_CUDA(cudaMemcpyAsync(h_counters, s.counters, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
      EFGP_CUDA(cudaStreamSynchronize(st));
      if (h_counters[0] <= 0) break;
    }
    if (first) {
      first = false;
      check = 4;
    } else if (check < 32) {
      check *= 2;
    }
  }
  if (!early) EFGP_CUDA(cudaStreamSynchronize(st));
  if (sym) t.P_rows = 0;                 // P was used as the half spectrum: its zero padding is gone
  all_conv = early && h_counters[0] <= 0;
  jmax = h_counters[1];
  if (all_conv) my_hint = jmax + 1;      // iterations 0..jmax: the final kernel then sees every row converged
}

void cg_solve(Toeplitz& t, const void* ws, double sigmasq, int mode, const void* minv, const void* b, void* x,
              int batch, double tol, int max_iter, int early_stopping, int batched_semantics, int x0_is_zero,
              int* iters_completed, cudaStream_t st, int hermitian, int* async_counters) {
  EFGP_CHECK(batch >= 1, "batch must be >= 1");
  EFGP_CHECK(mode >= 0 && mode <= 3, "unknown operator mode");
  EFGP_CHECK(mode == kModeT || ws != nullptr, "ws is required for Gv / A_mean / A_var");
  EFGP_CHECK(t.M < (int64_t)1 << 31, "Toeplitz block too large for the 32-bit index arithmetic of the CG kernels");
  if (max_iter < 0) max_iter = (int)std::min<int64_t>(2 * t.M, 2147483647);
  size_t esz = t.dtype == 0 ? 8 : 16;
  int group = (int)std::max<int64_t>(1, std::min<int64_t>(batch, (int64_t)((12ull << 30) / ((size_t)t.Ftot * esz))));
  int iters = 0;
  bool all_conv_all = true;
  int jmax_all = -1;
  int* async_one = (async_counters && batch <= group) ? async_counters : nullptr;   // one group: may stay asynchronous
  for (int b0 = 0; b0 < batch; b0 += group) {
    int nb = std::min(group, batch - b0);
    int all_conv = 0, jmax = -1;
    const char* bb = (const char*)b + (size_t)b0 * t.M * esz;
    char* xx = (char*)x + (size_t)b0 * t.M * esz;
    if (t.dtype == 0)
      solve_group<float>(t, (const float*)ws, sigmasq, mode, (const float*)minv, bb, xx, nb, tol, max_iter,
                         early_stopping, batched_semantics, x0_is_zero, all_conv, jmax, st, hermitian, async_one);
    else
      solve_group<double>(t, (const double*)ws, sigmasq, mode, (const double*)minv, bb, xx, nb, tol, max_iter,
                          early_stopping, batched_semantics, x0_is_zero, all_conv, jmax, st, hermitian, async_one);
    if (all_conv < 0) {                  // enqueued only: the counters arrive in async_counters with the stream
      if (iters_completed) *iters_completed = -1;
      return;
    }
    all_conv_all = all_conv_all && all_conv;
    jmax_all = std::max(jmax_all, jmax);
    int it_g = max_iter;
    if (all_conv) it_g = batched_semantics ? std::min(jmax + 2, max_iter) : jmax + 1;
    iters = std::max(iters, it_g);
  }
  if (iters_completed) *iters_completed = iters;
  if (async_counters) {                  // the solve ran synchronously: same hand-over format
    async_counters[0] = all_conv_all ? 0 : 1;
    async_counters[1] = jmax_all;
  }
}

}  // namespace efgp
