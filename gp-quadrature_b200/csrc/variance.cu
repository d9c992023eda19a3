// Device stages of the two posterior-variance estimators around the batched PCG.
//
//  * stochastic (diag_sums_nd, efgpnd.py:1634-1664): lag sums
//        c[r] = mean_j ifftn( fftn(gamma_j, s) * conj(fftn(eta_j, s)), s ),   s = (2 mtot - 1)^d,
//    the zero-padded cross-correlation of gamma_j = ws * A_var^-1 (ws * eta_j) with the probe eta_j.  The inverse
//    transform is linear, so the J products are summed in the spectral domain and ONE inverse FFT is taken (the
//    reference takes J); any FFT size >= 2 mtot - 1 gives the same linear correlation, so a 2-3-5-smooth size is used
//    and the lags are mapped back to the reference's (2 mtot - 1)^d FFT ordering.
//  * regular (compute_prediction_variance 'regular', efgpnd.py:1805-1820): per test point x_i
//        rhs_i = ws * conj(fx_i),  var_i = max(0, Re sum_k fx_i[k] ws[k] gamma_i[k]),  fx_i[k] = exp(2 pi i x_i . xi_k);
//    the reference stores the dense (b, M) matrix fx; here both kernels rebuild the separable phases
//    exp(2 pi i h x_ia (k_a - hm)) per axis in shared memory, so fx is never stored.
#include "toeplitz.cuh"
#include <algorithm>

namespace efgp {

struct XcGeom {
  int dim;
  int n[kMaxDim];    // block extents (mtot)
  int F[kMaxDim];    // FFT extents
  int s[kMaxDim];    // output extents 2 n - 1
  int64_t M, Ftot, Stot;
};

// P[j, pad(k)] = gamma[j, k], Q[j, pad(k)] = eta[j, k] (imaginary part 0); the rest of P / Q is zero (memset)
template <typename T>
__global__ void __launch_bounds__(256)
xc_pad_kernel(const typename cplx_of<T>::type* __restrict__ gam, const T* __restrict__ eta, const T* __restrict__ ws,
              XcGeom g, typename cplx_of<T>::type* __restrict__ P, typename cplx_of<T>::type* __restrict__ Q) {
  using C = typename cplx_of<T>::type;
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y;
  if (k >= g.M) return;
  int64_t r = k, off = 0, mul = 1;
  for (int a = g.dim - 1; a >= 0; --a) {
    off += (r % g.n[a]) * mul;
    r /= g.n[a];
    mul *= g.F[a];
  }
  C gv = gam[(int64_t)j * g.M + k];
  if (ws) {                        // gamma = ws * u (efgpnd.py:1657)
    const T w = ws[k];
    gv.x *= w;
    gv.y *= w;
  }
  P[(int64_t)j * g.Ftot + off] = gv;
  C e;
  e.x = eta[(int64_t)j * g.M + k];
  e.y = 0;
  Q[(int64_t)j * g.Ftot + off] = e;
}

// acc[f] += sum_j P[j, f] conj(Q[j, f])
template <typename T>
__global__ void __launch_bounds__(256)
xc_accumulate_kernel(const typename cplx_of<T>::type* __restrict__ P, const typename cplx_of<T>::type* __restrict__ Q,
                     int rows, int64_t Ftot, int first, double2* __restrict__ acc) {
  const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= Ftot) return;
  double ar = 0, ai = 0;
  for (int j = 0; j < rows; ++j) {
    const auto p = P[(int64_t)j * Ftot + f];
    const auto q = Q[(int64_t)j * Ftot + f];
    ar += (double)p.x * q.x + (double)p.y * q.y;
    ai += (double)p.y * q.x - (double)p.x * q.y;
  }
  double2 o;
  o.x = ar;
  o.y = ai;
  if (!first) {
    const double2 old = acc[f];
    o.x += old.x;
    o.y += old.y;
  }
  acc[f] = o;
}

// out[r] (FFT ordering on (2n-1)^d) = R[lag(r) mod F] * scale
template <typename T>
__global__ void __launch_bounds__(256)
xc_crop_kernel(const double2* __restrict__ R, XcGeom g, double scale, typename cplx_of<T>::type* __restrict__ out) {
  using C = typename cplx_of<T>::type;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= g.Stot) return;
  int64_t r = t, off = 0, mul = 1;
  for (int a = g.dim - 1; a >= 0; --a) {
    const int ra = (int)(r % g.s[a]);
    r /= g.s[a];
    const int lag = ra < g.n[a] ? ra : ra - g.s[a];
    off += (int64_t)(lag < 0 ? lag + g.F[a] : lag) * mul;
    mul *= g.F[a];
  }
  const double2 v = R[off];
  C o;
  o.x = (T)(v.x * scale);
  o.y = (T)(v.y * scale);
  out[t] = o;
}

template <typename T>
static void xcorr_mean_t(const void* gam_, const void* eta_, const void* ws_, int J, int dim, int mtot, void* out_,
                         cudaStream_t st) {
  using C = typename cplx_of<T>::type;
  XcGeom g;
  g.dim = dim;
  g.M = g.Ftot = g.Stot = 1;
  int64_t Fs[kMaxDim] = {1, 1, 1};
  for (int a = 0; a < kMaxDim; ++a) {
    g.n[a] = a < dim ? mtot : 1;
    g.s[a] = a < dim ? 2 * mtot - 1 : 1;
    g.F[a] = a < dim ? (int)next_smooth(2 * mtot - 1) : 1;
    Fs[a] = g.F[a];
    g.M *= g.n[a];
    g.Ftot *= g.F[a];
    g.Stot *= g.s[a];
  }
  const C* gam = (const C*)gam_;
  const T* eta = (const T*)eta_;
  // probe chunks: P and Q stay under ~1 GB each
  int rows = (int)std::max<int64_t>(1, std::min<int64_t>(J, ((int64_t)1 << 30) / (g.Ftot * (int64_t)sizeof(C))));
  rows = std::min(rows, 65535);
  C* P = (C*)pool_alloc((size_t)rows * g.Ftot * sizeof(C));
  C* Q = (C*)pool_alloc((size_t)rows * g.Ftot * sizeof(C));
  double2* acc = (double2*)pool_alloc((size_t)g.Ftot * sizeof(double2));
  ScopedStage stg("variance.xcorr_mean", st, 0, (long long)J * g.M);
  for (int j0 = 0; j0 < J; j0 += rows) {
    const int nb = std::min(rows, J - j0);
    EFGP_CUDA(cudaMemsetAsync(P, 0, (size_t)nb * g.Ftot * sizeof(C), st));
    EFGP_CUDA(cudaMemsetAsync(Q, 0, (size_t)nb * g.Ftot * sizeof(C), st));
    xc_pad_kernel<T><<<dim3(div_up(g.M, 256), nb), 256, 0, st>>>(gam + (int64_t)j0 * g.M, eta + (int64_t)j0 * g.M,
                                                                 (const T*)ws_, g, P, Q);
    cufftHandle plan = global_fft().get(dim, Fs, sizeof(T) == 4 ? CUFFT_C2C : CUFFT_Z2Z, nb, st);
    if (sizeof(T) == 4) {
      EFGP_CUFFT(cufftExecC2C(plan, (cufftComplex*)P, (cufftComplex*)P, CUFFT_FORWARD));
      EFGP_CUFFT(cufftExecC2C(plan, (cufftComplex*)Q, (cufftComplex*)Q, CUFFT_FORWARD));
    } else {
      EFGP_CUFFT(cufftExecZ2Z(plan, (cufftDoubleComplex*)P, (cufftDoubleComplex*)P, CUFFT_FORWARD));
      EFGP_CUFFT(cufftExecZ2Z(plan, (cufftDoubleComplex*)Q, (cufftDoubleComplex*)Q, CUFFT_FORWARD));
    }
    xc_accumulate_kernel<T><<<div_up(g.Ftot, 256), 256, 0, st>>>(P, Q, nb, g.Ftot, j0 == 0 ? 1 : 0, acc);
    stg.add_launches(2);
    EFGP_CUDA(cudaGetLastError());
  }
  // one inverse transform of the summed spectrum (always in double: it carries the mean over all probes)
  cufftHandle inv = global_fft().get(dim, Fs, CUFFT_Z2Z, 1, st);
  EFGP_CUFFT(cufftExecZ2Z(inv, (cufftDoubleComplex*)acc, (cufftDoubleComplex*)acc, CUFFT_INVERSE));
  xc_crop_kernel<T><<<div_up(g.Stot, 256), 256, 0, st>>>(acc, g, 1.0 / ((double)g.Ftot * (double)J), (C*)out_);
  stg.add_launches(1);
  EFGP_CUDA(cudaGetLastError());
  pool_free(acc);
  pool_free(Q);
  pool_free(P);
}

void xcorr_mean(const void* us, const void* etas, const void* ws, int J, int dim, int mtot, int dtype, void* out,
                cudaStream_t st) {
  EFGP_CHECK(J >= 1 && mtot >= 1 && dim >= 1 && dim <= kMaxDim, "bad cross-correlation geometry");
  if (dtype == 0)
    xcorr_mean_t<float>(us, etas, ws, J, dim, mtot, out, st);
  else
    xcorr_mean_t<double>(us, etas, ws, J, dim, mtot, out, st);
}

// ------------------------------------------------------------------ 'regular' variance rows
constexpr int kVarMaxM = 768;   // modes per axis a CTA's phase tables hold

// E[a][k] = exp(2 pi i h x_a (k - hm)), k = 0..mtot-1, for the CTA's point; phases through sincospi of the product
// reduced in double (exact period handling for large |k|)
template <typename T>
__device__ __forceinline__ void var_phase_tables(const T* __restrict__ x, int64_t i, int dim, int mtot, double h,
                                                 double2 (*E)[kVarMaxM]) {
  const int hm = (mtot - 1) / 2;
  for (int q = threadIdx.x; q < dim * mtot; q += blockDim.x) {
    const int a = q / mtot, k = q - a * mtot;
    // xi_k in the model's dtype, like xis built by the reference (efgpnd.py:759-766)
    const T xi = (T)((double)(k - hm) * h);
    const double ph = 2.0 * (double)((T)(x[i * dim + a] * xi));
    double s, c;
    sincospi(ph, &s, &c);
    E[a][k].x = c;
    E[a][k].y = s;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
var_rhs_kernel(const T* __restrict__ x, int dim, int mtot, double h, const T* __restrict__ ws, int64_t M,
               typename cplx_of<T>::type* __restrict__ rhs) {
  using C = typename cplx_of<T>::type;
  __shared__ double2 E[kMaxDim][kVarMaxM];
  const int64_t i = blockIdx.x;
  var_phase_tables<T>(x, i, dim, mtot, h, E);
  __syncthreads();
  for (int64_t k = threadIdx.x; k < M; k += blockDim.x) {
    int64_t r = k;
    double pr = 1.0, pi = 0.0;
    for (int a = dim - 1; a >= 0; --a) {
      const double2 e = E[a][r % mtot];
      r /= mtot;
      const double t = pr * e.x - pi * e.y;
      pi = pr * e.y + pi * e.x;
      pr = t;
    }
    const double w = (double)ws[k];
    C o;
    o.x = (T)(w * pr);
    o.y = (T)(-w * pi);       // ws * conj(fx)
    rhs[i * M + k] = o;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
var_reduce_kernel(const T* __restrict__ x, int dim, int mtot, double h, const T* __restrict__ ws, int64_t M,
                  const typename cplx_of<T>::type* __restrict__ gamma, T* __restrict__ var) {
  __shared__ double2 E[kMaxDim][kVarMaxM];
  __shared__ double sh[8];
  const int64_t i = blockIdx.x;
  var_phase_tables<T>(x, i, dim, mtot, h, E);
  __syncthreads();
  double acc = 0;
  for (int64_t k = threadIdx.x; k < M; k += blockDim.x) {
    int64_t r = k;
    double pr = 1.0, pi = 0.0;
    for (int a = dim - 1; a >= 0; --a) {
      const double2 e = E[a][r % mtot];
      r /= mtot;
      const double t = pr * e.x - pi * e.y;
      pi = pr * e.y + pi * e.x;
      pr = t;
    }
    const auto gq = gamma[i * M + k];
    acc += (double)ws[k] * (pr * (double)gq.x - pi * (double)gq.y);      // Re(fx ws gamma)
  }
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0;
    for (int q = 0; q < (int)(blockDim.x >> 5); ++q) s += sh[q];
    var[i] = (T)(s > 0 ? s : 0);
  }
}

void variance_rows(const void* x, int64_t b, int dim, int mtot, double h, const void* ws, int dtype, void* rhs,
                   cudaStream_t st) {
  EFGP_CHECK(mtot <= kVarMaxM, "mtot too large for the regular-variance kernels");
  if (b == 0) return;
  int64_t M = 1;
  for (int a = 0; a < dim; ++a) M *= mtot;
  ScopedStage stg("variance.rhs_kernel", st, 1, b * M);
  if (dtype == 0)
    var_rhs_kernel<float><<<(unsigned)b, 256, 0, st>>>((const float*)x, dim, mtot, h, (const float*)ws, M, (float2*)rhs);
  else
    var_rhs_kernel<double><<<(unsigned)b, 256, 0, st>>>((const double*)x, dim, mtot, h, (const double*)ws, M, (double2*)rhs);
  EFGP_CUDA(cudaGetLastError());
}

void variance_reduce(const void* x, int64_t b, int dim, int mtot, double h, const void* ws, int dtype, const void* gamma,
                     void* var, cudaStream_t st) {
  EFGP_CHECK(mtot <= kVarMaxM, "mtot too large for the regular-variance kernels");
  if (b == 0) return;
  int64_t M = 1;
  for (int a = 0; a < dim; ++a) M *= mtot;
  ScopedStage stg("variance.reduce_kernel", st, 1, b * M);
  if (dtype == 0)
    var_reduce_kernel<float><<<(unsigned)b, 256, 0, st>>>((const float*)x, dim, mtot, h, (const float*)ws, M,
                                                         (const float2*)gamma, (float*)var);
  else
    var_reduce_kernel<double><<<(unsigned)b, 256, 0, st>>>((const double*)x, dim, mtot, h, (const double*)ws, M,
                                                           (const double2*)gamma, (double*)var);
  EFGP_CUDA(cudaGetLastError());
}

}  // namespace efgp
