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
    C bv = b[i], av = Ap[i];
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

template <typename T>
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
    P[(int64_t)row * g.Ftot + po] = o;
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

template <typename T>
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
    C q = Q[(int64_t)row * g.Ftot + co];
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
__global__ void __launch_bounds__(256)
cg_C_kernel(typename cplx_of<T>::type* __restrict__ x, typename cplx_of<T>::type* __restrict__ r,
            const typename cplx_of<T>::type* __restrict__ p, const typename cplx_of<T>::type* __restrict__ Ap,
            const T* __restrict__ minv, int64_t M, int it, CgScal s) {
  using C = typename cplx_of<T>::type;
  __shared__ double sh[8];
  int row = blockIdx.y, c = blockIdx.x;
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
                        int B, double tol, int max_iter, int early, int batched, int& all_conv, int& jmax,
                        cudaStream_t st) {
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

  // Ap = A x0  (generic apply writes through P/Q as well)
  toeplitz_apply(t, x, B, ws, sigmasq, mode, Ap, st);
  dim3 grid(s.NC, B), gf(div_up(t.Ftot, 256), B);
  {
    ScopedStage stg_init("cg.init", st, 2);
    cg_init_kernel<T><<<grid, 256, 0, st>>>(b, Ap, minv, r, p, M, s);
    cg_init_scal_kernel<<<B, 256, 0, st>>>(s);
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
  static int hint[4][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};
  int& my_hint = hint[mode][batched ? 1 : 0];
  int it = 0;
  int check = my_hint > 0 ? std::min(my_hint, 64) : 4;
  bool first = true;
  ScopedStage stg("cg.iterations", st, 0);
  while (it < max_iter) {
    int until = std::min(max_iter, it + check);
    stg.add_launches(4 * (until - it));
    for (; it < until; ++it) {
      cg_A_kernel<T><<<grid, 256, 0, st>>>(p, r, minv, w, g, it, tol, early, batched, P, s);
      fft_pair_begin<T>(t, B, P, Q, st);
      cg_M_kernel<T><<<gf, 256, 0, st>>>(Q, (const C*)t.vhat, t.Ftot, s.act + (it & 1) * B);
      fft_pair_end<T>(t, B, Q, st);
      cg_B_kernel<T><<<grid, 256, 0, st>>>(Q, p, w, g, c1, c2, it, Ap, s);
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
  all_conv = early && h_counters[0] <= 0;
  jmax = h_counters[1];
  if (all_conv) my_hint = jmax + 1;      // iterations 0..jmax: the final kernel then sees every row converged
}

void cg_solve(Toeplitz& t, const void* ws, double sigmasq, int mode, const void* minv, const void* b, void* x,
              int batch, double tol, int max_iter, int early_stopping, int batched_semantics,
              int* iters_completed, cudaStream_t st) {
  EFGP_CHECK(batch >= 1, "batch must be >= 1");
  EFGP_CHECK(mode >= 0 && mode <= 3, "unknown operator mode");
  EFGP_CHECK(mode == kModeT || ws != nullptr, "ws is required for Gv / A_mean / A_var");
  EFGP_CHECK(t.M < (int64_t)1 << 31, "Toeplitz block too large for the 32-bit index arithmetic of the CG kernels");
  if (max_iter < 0) max_iter = (int)std::min<int64_t>(2 * t.M, 2147483647);
  size_t esz = t.dtype == 0 ? 8 : 16;
  int group = (int)std::max<int64_t>(1, std::min<int64_t>(batch, (int64_t)((12ull << 30) / ((size_t)t.Ftot * esz))));
  int iters = 0;
  for (int b0 = 0; b0 < batch; b0 += group) {
    int nb = std::min(group, batch - b0);
    int all_conv = 0, jmax = -1;
    const char* bb = (const char*)b + (size_t)b0 * t.M * esz;
    char* xx = (char*)x + (size_t)b0 * t.M * esz;
    if (t.dtype == 0)
      solve_group<float>(t, (const float*)ws, sigmasq, mode, (const float*)minv, bb, xx, nb, tol, max_iter,
                         early_stopping, batched_semantics, all_conv, jmax, st);
    else
      solve_group<double>(t, (const double*)ws, sigmasq, mode, (const double*)minv, bb, xx, nb, tol, max_iter,
                          early_stopping, batched_semantics, all_conv, jmax, st);
    int it_g = max_iter;
    if (all_conv) it_g = batched_semantics ? std::min(jmax + 2, max_iter) : jmax + 1;
    iters = std::max(iters, it_g);
  }
  if (iters_completed) *iters_completed = iters;
}

}  // namespace efgp
