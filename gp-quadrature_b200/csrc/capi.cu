// extern "C" boundary of libefgp_b200 (see include/efgp_b200.h).
#include "../../include/efgp_b200.h"
#include "toeplitz.cuh"
#include "fit.cuh"
#include <memory>

namespace efgp {
static thread_local std::string g_last_error;
void set_last_error(const std::string& msg) { g_last_error = msg; }
}  // namespace efgp

struct efgp_nufft {
  efgp::Nufft nu;
  cudaStream_t last_stream = nullptr;     // blocks released by the destructor are tagged with it (pool.cpp)
};
struct efgp_toeplitz {
  efgp::Toeplitz t;
  cudaStream_t last_stream = nullptr;
};

// Every entry point that touches the device: refuse a second GPU in this process and publish the stream to the pool.
#define EFGP_ENTER(stream)                               \
  efgp::check_device();                                  \
  efgp::set_current_stream((cudaStream_t)(stream))

#define EFGP_TRY(body)                                  \
  try {                                                 \
    body;                                               \
    return 0;                                           \
  } catch (const efgp::Error& e) {                      \
    efgp::set_last_error(e.what());                     \
    return e.code;                                      \
  } catch (const std::exception& e) {                   \
    efgp::set_last_error(e.what());                     \
    return 99;                                          \
  }

extern "C" {

int efgp_version(void) { return 100; }

const char* efgp_last_error(void) { return efgp::g_last_error.c_str(); }

/* per-stage device timing (bench / diagnostics) */
int efgp_timing_enable(int on) {
  EFGP_TRY({ efgp::timing_enable(on); })
}

int efgp_timing_report(char* buf, int64_t cap) {
  EFGP_TRY({
    EFGP_CHECK(buf != nullptr && cap > 0, "null report buffer");
    std::string r = efgp::timing_report();
    if ((int64_t)r.size() + 1 > cap) r.resize((size_t)cap - 1);
    memcpy(buf, r.c_str(), r.size() + 1);
  })
}

int efgp_set_deterministic(int on) {
  EFGP_TRY({ efgp::set_deterministic(on); })
}

int efgp_pool_trim(void) {
  EFGP_TRY({ efgp::pool_trim(); })
}

// FP64 FMA throughput of this device, measured: 16 independent DFMA chains per thread, 8 warps x 8 CTAs per SM.
// The roofline denominator of the FP64-pipe-bound spreading kernels (bench.py); not used on the product path.
namespace {
__global__ void __launch_bounds__(256) fp64_peak_kernel(double* out, int iters, double seed) {
  double a[16];
#pragma unroll
  for (int q = 0; q < 16; ++q) a[q] = seed + q + threadIdx.x;
  const double m = 1.0000001, c = 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int q = 0; q < 16; ++q) a[q] = fma(a[q], m, c);
  }
  double s = 0;
#pragma unroll
  for (int q = 0; q < 16; ++q) s += a[q];
  if (s == 12345.678) out[0] = s;   // never true: keeps the chains alive
}
}  // namespace

static double measure_fp64_peak(cudaStream_t st) {
  int dev = 0;
  int sms = 0;
  EFGP_CUDA(cudaGetDevice(&dev));
  EFGP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  double* d = (double*)efgp::pool_alloc(sizeof(double));
  cudaEvent_t e0;
  cudaEvent_t e1;
  EFGP_CUDA(cudaEventCreate(&e0));
  EFGP_CUDA(cudaEventCreate(&e1));
  const int iters = 8192;
  const int blocks = sms * 8;
  double best = 0;
  for (int r = 0; r < 4; ++r) {
    EFGP_CUDA(cudaEventRecord(e0, st));
    fp64_peak_kernel<<<blocks, 256, 0, st>>>(d, iters, 1.0 + r);
    EFGP_CUDA(cudaEventRecord(e1, st));
    EFGP_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    EFGP_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    double tf = 2.0 * 16.0 * iters * 256.0 * blocks / (ms * 1e-3) * 1e-12;
    if (r > 0 && tf > best) best = tf;
  }
  EFGP_CUDA(cudaEventDestroy(e0));
  EFGP_CUDA(cudaEventDestroy(e1));
  efgp::pool_free(d);
  return best;
}

int efgp_measure_fp64_peak(double* tflops, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    EFGP_CHECK(tflops != nullptr, "null output");
    *tflops = measure_fp64_peak((cudaStream_t)stream);
  })
}

int efgp_nufft_create(efgp_nufft** out, int dim, int dtype) {
  EFGP_TRY({
    EFGP_CHECK(out != nullptr, "null output handle");
    EFGP_CHECK(dim >= 1 && dim <= efgp::kMaxDim, "dim must be 1, 2 or 3");
    EFGP_CHECK(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    efgp_nufft* h = new efgp_nufft();
    h->nu.dim = dim;
    h->nu.dtype = dtype;
    *out = h;
  })
}

int efgp_nufft_set_points(efgp_nufft* nu, const void* x, int64_t npts, double h, const double* xcen,
                          void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    if (nu) nu->last_stream = (cudaStream_t)stream;
    EFGP_CHECK(nu != nullptr, "null handle");
    EFGP_CHECK(npts >= 0, "npts must be >= 0");
    EFGP_CHECK(npts == 0 || x != nullptr, "null points");
    nu->nu.x = x;
    nu->nu.N = npts;
    nu->nu.h = h;
    for (int a = 0; a < efgp::kMaxDim; ++a) nu->nu.xcen[a] = (xcen && a < nu->nu.dim) ? xcen[a] : 0.0;
    nu->nu.sorted.clear();
  })
}

int efgp_nufft_rescale(efgp_nufft* nu, double h, const double* xcen) {
  EFGP_TRY({
    EFGP_CHECK(nu != nullptr, "null handle");
    bool same = h == nu->nu.h;
    for (int a = 0; a < nu->nu.dim; ++a) same = same && (xcen ? xcen[a] : 0.0) == nu->nu.xcen[a];
    if (!same) {
      nu->nu.h = h;
      for (int a = 0; a < efgp::kMaxDim; ++a) nu->nu.xcen[a] = (xcen && a < nu->nu.dim) ? xcen[a] : 0.0;
      efgp::set_current_stream(nu->last_stream);
      nu->nu.sorted.clear();      // the cell order depends on h: rebuilt by the next transform
    }
  })
}

int efgp_nufft_carry_index(efgp_nufft* nu, int on) {
  EFGP_TRY({
    EFGP_CHECK(nu != nullptr, "null handle");
    if (nu->nu.carry_index != (on != 0)) {
      nu->nu.carry_index = on != 0;
      efgp::set_current_stream(nu->last_stream);
      nu->nu.sorted.clear();
    }
  })
}

void efgp_nufft_destroy(efgp_nufft* nu) {
  if (!nu) return;
  efgp::set_current_stream(nu->last_stream);
  delete nu;
}

int efgp_nufft_type1(efgp_nufft* nu, const void* strengths, int is_complex, int batch,
                     const int64_t* out_shape, double eps, int isign, int modeord, const void* scale,
                     void* out, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    if (nu) nu->last_stream = (cudaStream_t)stream;
    EFGP_CHECK(nu != nullptr && out != nullptr && out_shape != nullptr, "null argument");
    efgp::nufft_type1(nu->nu, strengths, is_complex, batch, out_shape, eps, isign, modeord, scale, out,
                      (cudaStream_t)stream);
  })
}

int efgp_nufft_type1_pair(efgp_nufft* nu, const void* y, const int64_t* shape_gen, double eps_gen,
                          const int64_t* shape_rhs, double eps_rhs, const void* scale, void* out_gen, void* out_rhs,
                          void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    if (nu) nu->last_stream = (cudaStream_t)stream;
    EFGP_CHECK(nu != nullptr && y != nullptr && shape_gen != nullptr && shape_rhs != nullptr && out_gen != nullptr &&
                   out_rhs != nullptr,
               "null argument");
    efgp::nufft_type1_pair(nu->nu, y, shape_gen, eps_gen, shape_rhs, eps_rhs, scale, out_gen, out_rhs,
                           (cudaStream_t)stream);
  })
}

int efgp_nufft_type2(efgp_nufft* nu, const void* modes, int batch, const int64_t* shape, double eps,
                     int isign, int modeord, const void* scale, int real_out, void* out, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    if (nu) nu->last_stream = (cudaStream_t)stream;
    EFGP_CHECK(nu != nullptr && modes != nullptr && shape != nullptr, "null argument");
    EFGP_CHECK(nu->nu.N == 0 || out != nullptr, "null output");
    efgp::nufft_type2(nu->nu, modes, batch, shape, eps, isign, modeord, scale, real_out, out,
                      (cudaStream_t)stream);
  })
}

int efgp_nufft_plan_info(efgp_nufft* nu, const int64_t* shape, double eps, int64_t* nf_out, int* width_out) {
  EFGP_TRY({
    EFGP_CHECK(nu != nullptr && shape != nullptr, "null argument");
    if (nu->nu.dtype == 0 && eps < 1e-7) eps = 1e-7;
    int w0 = efgp::es_width_for(eps, 2.0, efgp::kMaxW);
    for (int a = 0; a < nu->nu.dim; ++a)
      if (nf_out) nf_out[a] = efgp::next_smooth_even(std::max<int64_t>(2 * shape[a], 2 * w0));
    if (width_out) *width_out = w0;
  })
}

int efgp_toeplitz_create(efgp_toeplitz** out, const void* v, int dim, const int64_t* Ls, int dtype,
                         int force_pow2, void* stream) {
  EFGP_TRY({
    EFGP_CHECK(out != nullptr && v != nullptr && Ls != nullptr, "null argument");
    EFGP_CHECK(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    EFGP_ENTER(stream);
    std::unique_ptr<efgp_toeplitz> h(new efgp_toeplitz());
    h->last_stream = (cudaStream_t)stream;
    efgp::toeplitz_create(h->t, v, dim, Ls, dtype, force_pow2, (cudaStream_t)stream);
    *out = h.release();
  })
}

void efgp_toeplitz_destroy(efgp_toeplitz* t) {
  if (!t) return;
  efgp::set_current_stream(t->last_stream);
  delete t;
}

int efgp_toeplitz_fft_shape(efgp_toeplitz* t, int64_t* shape_out) {
  EFGP_TRY({
    EFGP_CHECK(t != nullptr && shape_out != nullptr, "null argument");
    for (int a = 0; a < t->t.dim; ++a) shape_out[a] = t->t.F[a];
  })
}

int efgp_toeplitz_apply(efgp_toeplitz* t, const void* x, int batch, const void* ws, double sigmasq,
                        int mode, void* y, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    if (t) t->last_stream = (cudaStream_t)stream;
    EFGP_CHECK(t != nullptr && x != nullptr && y != nullptr, "null argument");
    efgp::toeplitz_apply(t->t, x, batch, ws, sigmasq, mode, y, (cudaStream_t)stream);
  })
}

int efgp_cg_solve(efgp_toeplitz* t, const void* ws, double sigmasq, int mode, const void* minv,
                  const void* b, void* x, int batch, double tol, int max_iter, int early_stopping,
                  int batched_semantics, int x0_is_zero, int* iters_completed, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    if (t) t->last_stream = (cudaStream_t)stream;
    EFGP_CHECK(t != nullptr && b != nullptr && x != nullptr, "null argument");
    efgp::cg_solve(t->t, ws, sigmasq, mode, minv, b, x, batch, tol, max_iter, early_stopping,
                   batched_semantics, x0_is_zero, iters_completed, (cudaStream_t)stream);
  })
}

int efgp_cg_solve_hermitian(efgp_toeplitz* t, const void* ws, double sigmasq, int mode, const void* minv,
                            const void* b, void* x, int batch, double tol, int max_iter, int early_stopping,
                            int batched_semantics, int x0_is_zero, int* iters_completed, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    if (t) t->last_stream = (cudaStream_t)stream;
    EFGP_CHECK(t != nullptr && b != nullptr && x != nullptr, "null argument");
    efgp::cg_solve(t->t, ws, sigmasq, mode, minv, b, x, batch, tol, max_iter, early_stopping,
                   batched_semantics, x0_is_zero, iters_completed, (cudaStream_t)stream, 1);
  })
}

int efgp_cg_solve_async(efgp_toeplitz* t, const void* ws, double sigmasq, int mode, const void* minv, const void* b,
                        void* x, int batch, double tol, int max_iter, int early_stopping, int batched_semantics,
                        int x0_is_zero, int hermitian, int* counters_pinned, int* iters_completed, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    if (t) t->last_stream = (cudaStream_t)stream;
    EFGP_CHECK(t != nullptr && b != nullptr && x != nullptr && counters_pinned != nullptr, "null argument");
    efgp::cg_solve(t->t, ws, sigmasq, mode, minv, b, x, batch, tol, max_iter, early_stopping,
                   batched_semantics, x0_is_zero, iters_completed, (cudaStream_t)stream, hermitian, counters_pinned);
  })
}

int efgp_lanczos(efgp_toeplitz* t, const void* ws, double sigmasq, int mode, void* q, int batch, int steps,
                 double breakdown_tol, double* alphas, double* betas, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    if (t) t->last_stream = (cudaStream_t)stream;
    EFGP_CHECK(t != nullptr && q != nullptr && alphas != nullptr && betas != nullptr, "null argument");
    efgp::lanczos_tridiag(t->t, ws, sigmasq, mode, q, batch, steps, breakdown_tol, alphas, betas,
                          (cudaStream_t)stream);
  })
}

/* ---- library communicator */
int efgp_comm_unique_id(void* id128) {
  EFGP_TRY({
    EFGP_CHECK(id128 != nullptr, "null argument");
    efgp::comm_unique_id(id128);
  })
}

int efgp_comm_init(int rank, int world, const void* id128) {
  EFGP_TRY({
    EFGP_CHECK(id128 != nullptr, "null argument");
    efgp::check_device();
    efgp::comm_init(rank, world, id128);
  })
}

int efgp_comm_destroy(void) {
  EFGP_TRY({ efgp::comm_destroy(); })
}

int efgp_comm_world(void) { return efgp::comm_world(); }

int efgp_allreduce(void* buf, int64_t count, int dtype, int op, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    EFGP_CHECK(buf != nullptr && count >= 0, "bad argument");
    efgp::comm_allreduce(buf, count, dtype, op, (cudaStream_t)stream);
  })
}

/* ---- fused fit */
int efgp_bbox_async(const void* x, int64_t npts, int dim, int dtype, int sharded, double* ext_pinned, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    EFGP_CHECK(x != nullptr && ext_pinned != nullptr, "null argument");
    EFGP_CHECK(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    efgp::bbox_async(x, npts, dim, dtype, sharded, ext_pinned, (cudaStream_t)stream);
  })
}

int efgp_fit(efgp_nufft* nu, const void* y, const efgp_fit_args* args, void* ws, void* vr, void* minv, void* beta,
             efgp_toeplitz** toep_out, int* counters_pinned, int* iters_completed, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    EFGP_CHECK(nu != nullptr && y != nullptr && args != nullptr && toep_out != nullptr && counters_pinned != nullptr,
               "null argument");
    EFGP_CHECK(nu->nu.N >= 1, "fit needs at least one point");
    nu->last_stream = (cudaStream_t)stream;
    efgp::FitArgs a;
    a.family = args->family;
    a.c = args->c;
    a.a = args->a;
    a.b = args->b;
    a.p = args->p;
    a.h = args->h;
    a.hd = args->hd;
    a.mtot = args->mtot;
    a.sigmasq = args->sigmasq;
    a.nufft_eps = args->nufft_eps;
    a.gen_eps = args->gen_eps;
    a.cg_tol = args->cg_tol;
    a.max_iter = args->max_iter;
    a.precondition = args->precondition;
    a.force_pow2 = args->force_pow2;
    a.sharded = args->sharded;
    a.warm_start = args->warm_start;
    EFGP_CHECK(a.family == 0 || a.family == 1, "unknown spectral-density family");
    if (a.h != nu->nu.h) {           // same points, new spacing: the cell order of the sorted sets is stale
      nu->nu.h = a.h;
      nu->nu.sorted.clear();
    }
    std::unique_ptr<efgp_toeplitz> h(new efgp_toeplitz());
    h->last_stream = (cudaStream_t)stream;
    efgp::fit_run(nu->nu, h->t, y, a, ws, vr, minv, beta, counters_pinned, iters_completed, (cudaStream_t)stream);
    *toep_out = h.release();
  })
}

/* ---- posterior variance: device stages around the batched solves */
int efgp_xcorr_mean(const void* us, const void* etas, const void* ws, int nprobes, int dim, int mtot, int dtype, void* out,
                    void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    EFGP_CHECK(us != nullptr && etas != nullptr && out != nullptr, "null argument");
    EFGP_CHECK(dtype == 0 || dtype == 1, "dtype must be 0 (f32) or 1 (f64)");
    efgp::xcorr_mean(us, etas, ws, nprobes, dim, mtot, dtype, out, (cudaStream_t)stream);
  })
}

int efgp_variance_rows(const void* x, int64_t npts, int dim, int mtot, double h, const void* ws, int dtype, void* rhs,
                       void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    EFGP_CHECK(x != nullptr && ws != nullptr && rhs != nullptr && npts >= 0, "bad argument");
    EFGP_CHECK(dim >= 1 && dim <= efgp::kMaxDim && mtot >= 1, "bad geometry");
    efgp::variance_rows(x, npts, dim, mtot, h, ws, dtype, rhs, (cudaStream_t)stream);
  })
}

int efgp_variance_reduce(const void* x, int64_t npts, int dim, int mtot, double h, const void* ws, int dtype,
                         const void* gamma, void* var, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    EFGP_CHECK(x != nullptr && ws != nullptr && gamma != nullptr && var != nullptr && npts >= 0, "bad argument");
    EFGP_CHECK(dim >= 1 && dim <= efgp::kMaxDim && mtot >= 1, "bad geometry");
    efgp::variance_reduce(x, npts, dim, mtot, h, ws, dtype, gamma, var, (cudaStream_t)stream);
  })
}

/* ---- Hutchinson probes streamed through the transforms */
int efgp_nufft_type1_probes(efgp_nufft* nu, uint64_t seed, int t0, int nprobes, const int64_t* out_shape, double eps,
                            const void* scale, void* out, uint64_t* sort_id, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    EFGP_CHECK(nu != nullptr && out_shape != nullptr && out != nullptr, "null argument");
    nu->last_stream = (cudaStream_t)stream;
    efgp::nufft_type1_probes(nu->nu, seed, t0, nprobes, out_shape, eps, scale, out, sort_id, (cudaStream_t)stream);
  })
}

int efgp_nufft_type2_probe_dots(efgp_nufft* nu, uint64_t seed, int t0, int nprobes, const void* modes, const int64_t* shape,
                                double eps, const void* scale, double* sums, uint64_t* sort_id, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    EFGP_CHECK(nu != nullptr && modes != nullptr && shape != nullptr && sums != nullptr, "null argument");
    nu->last_stream = (cudaStream_t)stream;
    efgp::nufft_type2_probe_dots(nu->nu, seed, t0, nprobes, modes, shape, eps, scale, sums, sort_id, (cudaStream_t)stream);
  })
}

int efgp_nufft_probes_export(efgp_nufft* nu, uint64_t seed, int t0, int nprobes, const int64_t* shape, double eps, void* z,
                             uint64_t* sort_id, void* stream) {
  EFGP_TRY({
    EFGP_ENTER(stream);
    EFGP_CHECK(nu != nullptr && shape != nullptr && z != nullptr, "null argument");
    nu->last_stream = (cudaStream_t)stream;
    efgp::nufft_probes_export(nu->nu, seed, t0, nprobes, shape, eps, z, sort_id, (cudaStream_t)stream);
  })
}

}  // extern "C"
