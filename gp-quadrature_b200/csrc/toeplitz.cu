// F^H F Toeplitz embedding and matvec.
//
// Reference: ToeplitzND (efgpnd.py:1244-1393): zero-pad the (4m+1)^d generator to the FFT
// shape, cache its FFT, and apply  T x = crop(ifftn(fftn(pad x) * vhat))  with the crop window
// [n-1, 2n-1) per axis; operator factories create_Gv / create_A_mean / create_A_var
// (efgpnd.py:1572-1609) wrap it in the spectral-weight diagonal:  Gv = ws * T(ws * v),
// A_mean = Gv + sigma^2 I,  A_var = Gv / sigma^2 + I.
//
// Here the diagonal scalings, the zero padding, the 1/F^d normalisation, the pointwise multiply
// and the crop are fused into three kernels around a cuFFT forward/inverse pair.  The padded
// input buffer is written only inside the block (its zero region persists across applies because
// the forward FFT is out of place).
#include "toeplitz.cuh"

namespace efgp {

Toeplitz::~Toeplitz() {
  if (vhat) pool_free(vhat);
  if (vhat_real) pool_free(vhat_real);
}

struct TzGeom {
  int dim;
  int n[kMaxDim];
  int F[kMaxDim];
  int64_t M, Ftot;
};

__device__ __forceinline__ void tz_offsets(const TzGeom& g, int64_t k, int64_t& pad_off, int64_t& crop_off) {
  // k: flat C-order index into the n^d block
  int id[kMaxDim] = {0, 0, 0};
  for (int a = g.dim - 1; a >= 0; --a) {
    id[a] = (int)(k % g.n[a]);
    k /= g.n[a];
  }
  pad_off = 0;
  crop_off = 0;
  for (int a = 0; a < g.dim; ++a) {
    pad_off = pad_off * g.F[a] + id[a];
    crop_off = crop_off * g.F[a] + (id[a] + g.n[a] - 1);
  }
}

template <typename C>
__global__ void __launch_bounds__(256)
tz_pad_v_kernel(const C* __restrict__ v, TzGeom g /* n = L here */, C* __restrict__ vp) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= g.Ftot) return;
  int id[kMaxDim] = {0, 0, 0};
  int64_t r = t;
  for (int a = g.dim - 1; a >= 0; --a) {
    id[a] = (int)(r % g.F[a]);
    r /= g.F[a];
  }
  bool in = true;
  int64_t src = 0;
  for (int a = 0; a < g.dim; ++a) {
    in = in && id[a] < g.n[a];
    src = src * g.n[a] + id[a];
  }
  C o;
  o.x = 0;
  o.y = 0;
  if (in) o = v[src];
  vp[t] = o;
}

template <typename C, typename T>
__global__ void __launch_bounds__(256) tz_scale_kernel(C* __restrict__ a, int64_t n, T s) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  C v = a[t];
  v.x *= s;
  v.y *= s;
  a[t] = v;
}

template <typename T>
__global__ void __launch_bounds__(256)
tz_pad_x_kernel(const typename cplx_of<T>::type* __restrict__ x, const T* __restrict__ ws, TzGeom g, int batch,
                const int* __restrict__ act, typename cplx_of<T>::type* __restrict__ P) {
  using C = typename cplx_of<T>::type;
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int b = blockIdx.y;
  if (k >= g.M) return;
  if (act && !act[b]) return;
  int64_t po, co;
  tz_offsets(g, k, po, co);
  C v = x[(int64_t)b * g.M + k];
  if (ws) {
    T w = ws[k];
    v.x *= w;
    v.y *= w;
  }
  P[(int64_t)b * g.Ftot + po] = v;
}

template <typename T>
__global__ void __launch_bounds__(256)
tz_mul_kernel(typename cplx_of<T>::type* __restrict__ Q, const typename cplx_of<T>::type* __restrict__ vhat,
              int64_t Ftot, const int* __restrict__ act) {
  using C = typename cplx_of<T>::type;
  int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int b = blockIdx.y;
  if (f >= Ftot) return;
  if (act && !act[b]) return;
  C a = Q[(int64_t)b * Ftot + f];
  C h = vhat[f];
  C o;
  o.x = a.x * h.x - a.y * h.y;
  o.y = a.x * h.y + a.y * h.x;
  Q[(int64_t)b * Ftot + f] = o;
}

template <typename T>
__global__ void __launch_bounds__(256)
tz_crop_kernel(const typename cplx_of<T>::type* __restrict__ Q, const typename cplx_of<T>::type* __restrict__ x,
               const T* __restrict__ ws, TzGeom g, T c1, T c2, typename cplx_of<T>::type* __restrict__ y) {
  using C = typename cplx_of<T>::type;
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int b = blockIdx.y;
  if (k >= g.M) return;
  int64_t po, co;
  tz_offsets(g, k, po, co);
  C q = Q[(int64_t)b * g.Ftot + co];
  T w = ws ? ws[k] * c1 : c1;
  C o;
  o.x = q.x * w;
  o.y = q.y * w;
  if (c2 != T(0)) {
    C xv = x[(int64_t)b * g.M + k];
    o.x += c2 * xv.x;
    o.y += c2 * xv.y;
  }
  y[(int64_t)b * g.M + k] = o;
}

static TzGeom make_tz_geom(const Toeplitz& t) {
  TzGeom g;
  g.dim = t.dim;
  for (int a = 0; a < kMaxDim; ++a) {
    g.n[a] = (int)t.n[a];
    g.F[a] = (int)t.F[a];
  }
  g.M = t.M;
  g.Ftot = t.Ftot;
  return g;
}

template <typename T>
static void create_t(Toeplitz& t, const void* v, cudaStream_t st) {
  using C = typename cplx_of<T>::type;
  t.vhat = pool_alloc(sizeof(C) * (size_t)t.Ftot);
  TzGeom g = make_tz_geom(t);
  for (int a = 0; a < kMaxDim; ++a) g.n[a] = (int)t.L[a];
  ScopedStage stg("toeplitz.create", st, 2);
  tz_pad_v_kernel<C><<<div_up(t.Ftot, 256), 256, 0, st>>>((const C*)v, g, (C*)t.vhat);
  EFGP_CUDA(cudaGetLastError());
  cufftHandle plan = global_fft().get(t.dim, t.F, sizeof(T) == 4 ? CUFFT_C2C : CUFFT_Z2Z, 1, st);
  if (sizeof(T) == 4)
    EFGP_CUFFT(cufftExecC2C(plan, (cufftComplex*)t.vhat, (cufftComplex*)t.vhat, CUFFT_FORWARD));
  else
    EFGP_CUFFT(cufftExecZ2Z(plan, (cufftDoubleComplex*)t.vhat, (cufftDoubleComplex*)t.vhat, CUFFT_FORWARD));
  tz_scale_kernel<C, T><<<div_up(t.Ftot, 256), 256, 0, st>>>((C*)t.vhat, t.Ftot, (T)(1.0 / (double)t.Ftot));
  EFGP_CUDA(cudaGetLastError());
}

// vhat holds the spectrum of the generator stored from index 0 (centre at c_a = n_a - 1); multiplying by
// exp(+2 pi i f.c / F) moves the centre to the origin, where a Hermitian generator has a real spectrum.
template <typename T>
__global__ void tz_real_spectrum_kernel(const typename cplx_of<T>::type* __restrict__ vhat, TzGeom g,
                                        T* __restrict__ out) {
  int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= g.Ftot) return;
  int64_t rem = f;
  double re = 1.0, im = 0.0;
  for (int a = g.dim - 1; a >= 0; --a) {
    const int fa = (int)(rem % g.F[a]);
    rem /= g.F[a];
    const int64_t num = ((int64_t)fa * (g.n[a] - 1)) % g.F[a];      // exact phase index
    double s, c;
    sincospi(2.0 * (double)num / (double)g.F[a], &s, &c);
    const double r2 = re * c - im * s;
    im = re * s + im * c;
    re = r2;
  }
  const typename cplx_of<T>::type h = vhat[f];
  out[f] = (T)((double)h.x * re - (double)h.y * im);
}

const void* toeplitz_real_spectrum(Toeplitz& t, cudaStream_t st) {
  if (t.vhat_real) return t.vhat_real;
  TzGeom g = make_tz_geom(t);
  const size_t esz = t.dtype == 0 ? 4 : 8;
  t.vhat_real = pool_alloc(esz * (size_t)t.Ftot);
  if (t.dtype == 0)
    tz_real_spectrum_kernel<float><<<div_up(t.Ftot, 256), 256, 0, st>>>((const float2*)t.vhat, g, (float*)t.vhat_real);
  else
    tz_real_spectrum_kernel<double><<<div_up(t.Ftot, 256), 256, 0, st>>>((const double2*)t.vhat, g,
                                                                         (double*)t.vhat_real);
  EFGP_CUDA(cudaGetLastError());
  return t.vhat_real;
}

void toeplitz_create(Toeplitz& t, const void* v, int dim, const int64_t* Ls, int dtype, int force_pow2,
                     cudaStream_t st) {
  EFGP_CHECK(dim >= 1 && dim <= kMaxDim, "dim must be 1..3");
  t.dim = dim;
  t.v_raw = v;
  t.dtype = dtype;
  t.M = 1;
  t.Ftot = 1;
  for (int a = 0; a < kMaxDim; ++a) {
    t.L[a] = a < dim ? Ls[a] : 1;
    EFGP_CHECK(t.L[a] >= 1, "Toeplitz generator extents must be >= 1");
    t.n[a] = (t.L[a] + 1) / 2;
    t.F[a] = a < dim ? (force_pow2 ? next_pow2(t.L[a]) : next_smooth(t.L[a])) : 1;
    t.M *= t.n[a];
    t.Ftot *= t.F[a];
  }
  if (dtype == 0)
    create_t<float>(t, v, st);
  else
    create_t<double>(t, v, st);
}

// Ensure P has `rows` rows whose out-of-block region is zero.
template <typename T>
static typename cplx_of<T>::type* ensure_P(Toeplitz& t, int rows, cudaStream_t st) {
  using C = typename cplx_of<T>::type;
  size_t need = (size_t)rows * t.Ftot * sizeof(C);
  bool grow = need > t.P.cap;
  C* P = (C*)t.P.get(need);
  if (grow || t.P_rows < rows) {
    EFGP_CUDA(cudaMemsetAsync(P, 0, need, st));
    t.P_rows = rows;
  }
  return P;
}

template <typename T>
void fft_pair_begin(Toeplitz& t, int rows, typename cplx_of<T>::type* P, typename cplx_of<T>::type* Q,
                    cudaStream_t st) {
  cufftHandle plan = global_fft().get(t.dim, t.F, sizeof(T) == 4 ? CUFFT_C2C : CUFFT_Z2Z, rows, st);
  if (sizeof(T) == 4)
    EFGP_CUFFT(cufftExecC2C(plan, (cufftComplex*)P, (cufftComplex*)Q, CUFFT_FORWARD));
  else
    EFGP_CUFFT(cufftExecZ2Z(plan, (cufftDoubleComplex*)P, (cufftDoubleComplex*)Q, CUFFT_FORWARD));
}
template <typename T>
void fft_pair_end(Toeplitz& t, int rows, typename cplx_of<T>::type* Q, cudaStream_t st) {
  cufftHandle plan = global_fft().get(t.dim, t.F, sizeof(T) == 4 ? CUFFT_C2C : CUFFT_Z2Z, rows, st);
  if (sizeof(T) == 4)
    EFGP_CUFFT(cufftExecC2C(plan, (cufftComplex*)Q, (cufftComplex*)Q, CUFFT_INVERSE));
  else
    EFGP_CUFFT(cufftExecZ2Z(plan, (cufftDoubleComplex*)Q, (cufftDoubleComplex*)Q, CUFFT_INVERSE));
}
template void fft_pair_begin<float>(Toeplitz&, int, float2*, float2*, cudaStream_t);
template void fft_pair_begin<double>(Toeplitz&, int, double2*, double2*, cudaStream_t);
template void fft_pair_end<float>(Toeplitz&, int, float2*, cudaStream_t);
template void fft_pair_end<double>(Toeplitz&, int, double2*, cudaStream_t);

static int max_rows_per_pass(const Toeplitz& t, size_t elem) {
  // keep P and Q under ~6 GB each
  int64_t r = (int64_t)((6ull << 30) / ((size_t)t.Ftot * elem));
  return (int)std::max<int64_t>(1, r);
}

template <typename T>
static void apply_t(Toeplitz& t, const void* x, int batch, const void* ws, double sigmasq, int mode, void* y,
                    cudaStream_t st) {
  using C = typename cplx_of<T>::type;
  TzGeom g = make_tz_geom(t);
  const T* w = (mode == kModeT) ? nullptr : (const T*)ws;
  EFGP_CHECK(mode == kModeT || ws != nullptr, "ws is required for Gv / A_mean / A_var");
  T c1 = 1, c2 = 0;
  if (mode == kModeAmean) c2 = (T)sigmasq;
  if (mode == kModeAvar) {
    c1 = (T)(1.0 / sigmasq);
    c2 = 1;
  }
  int chunk = std::min(batch, max_rows_per_pass(t, sizeof(C)));
  C* P = ensure_P<T>(t, chunk, st);
  C* Q = (C*)t.Q.get((size_t)chunk * t.Ftot * sizeof(C));
  for (int b0 = 0; b0 < batch; b0 += chunk) {
    int nb = std::min(chunk, batch - b0);
    const C* xb = (const C*)x + (int64_t)b0 * t.M;
    C* yb = (C*)y + (int64_t)b0 * t.M;
    dim3 gm(div_up(t.M, 256), nb), gf(div_up(t.Ftot, 256), nb);
    ScopedStage stg("toeplitz.apply", st, 3);
    tz_pad_x_kernel<T><<<gm, 256, 0, st>>>(xb, w, g, nb, nullptr, P);
    fft_pair_begin<T>(t, nb, P, Q, st);
    tz_mul_kernel<T><<<gf, 256, 0, st>>>(Q, (const C*)t.vhat, t.Ftot, nullptr);
    fft_pair_end<T>(t, nb, Q, st);
    tz_crop_kernel<T><<<gm, 256, 0, st>>>(Q, xb, w, g, c1, c2, yb);
    EFGP_CUDA(cudaGetLastError());
  }
}

void toeplitz_apply(Toeplitz& t, const void* x, int batch, const void* ws, double sigmasq, int mode, void* y,
                    cudaStream_t st) {
  EFGP_CHECK(batch >= 1, "batch must be >= 1");
  EFGP_CHECK(mode >= 0 && mode <= 3, "unknown apply mode");
  if (t.dtype == 0)
    apply_t<float>(t, x, batch, ws, sigmasq, mode, y, st);
  else
    apply_t<double>(t, x, batch, ws, sigmasq, mode, y, st);
}

}  // namespace efgp
