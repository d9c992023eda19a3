// Shared declarations for libefgp_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cufft.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <stdexcept>

namespace efgp {

// ---------------------------------------------------------------- errors
void set_last_error(const std::string& msg);

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

#define EFGP_CUDA(call)                                                                       \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess)                                                                    \
      throw ::efgp::Error(2, std::string(#call) + ": " + cudaGetErrorString(e_) + " @" +      \
                                 __FILE__ + ":" + std::to_string(__LINE__));                  \
  } while (0)

#define EFGP_CUFFT(call)                                                                      \
  do {                                                                                        \
    cufftResult r_ = (call);                                                                  \
    if (r_ != CUFFT_SUCCESS)                                                                  \
      throw ::efgp::Error(3, std::string(#call) + ": cufft error " + std::to_string((int)r_) + \
                                 " @" + __FILE__ + ":" + std::to_string(__LINE__));           \
  } while (0)

#define EFGP_CHECK(cond, msg)                                                                 \
  do {                                                                                        \
    if (!(cond)) throw ::efgp::Error(1, std::string(msg));                                    \
  } while (0)

// ---------------------------------------------------------------- ES kernel (host side)
constexpr int kMaxW = 16;      // widest spreading kernel (eps ~ 1e-14)
constexpr int kMaxDeg = 20;    // highest Horner degree
constexpr int kMaxDim = 3;
constexpr double kMaxSigma = 16.0;   // beta keeps growing with the upsampling factor up to here
// Degree of the piecewise polynomials, a function of the width only (the device Horner loops are unrolled at compile
// time).  Measured with scripts/es_study.py against the exact NUDFT: for every width 2..16 and sigma in [2, 8] the
// transform error at degree W - 1 is within 10 % of the error at degree W + 2 (W >= 5; W for the narrow kernels),
// because the aliasing error exp(-pi W sqrt(1 - 1/sigma)) of the kernel itself is the floor.
__host__ __device__ constexpr int es_degree(int W) { return W <= 4 ? W : W - 1; }

// Piecewise-polynomial "exponential of semicircle" kernel phi(z)=exp(beta(sqrt(1-z^2)-1)).
// For a point with in-cell offset u in [-1/2,1/2) the W weights are
//   P_i(u) = phi((i-(W-1)/2-u)/(W/2)),  i=0..W-1,   and   P_{W-1-i}(u) = P_i(-u),
// so only the first ceil(W/2) polynomials are stored; coefficients are in s=2u (monomial).
struct ESKernel {
  int W = 0;
  int deg = 0;           // polynomial degree (deg+1 coefficients per interval)
  double beta = 0;
  double coef[(kMaxW + 1) / 2][kMaxDeg + 1];  // coef[i][j] multiplies s^j
  double fit_err = 0;    // max |poly - phi| over the support (diagnostic)
};

// Width and beta for a requested tolerance and effective upsampling factor sigma = nf/nmodes.
int es_width_for(double eps, double sigma, int max_w);
double es_error_model(int W, double sigma);   // modelled relative l2 error of a transform with this kernel
void es_build(ESKernel& k, int W, double sigma, double target_err);
double es_eval_exact(const ESKernel& k, double z);  // z in [-1,1]
// psi_hat(k/nf) for k = 0..kmax, psi(z)=phi(2z/W): deconvolution factors.
void es_fourier_series(const ESKernel& k, int64_t nf, int64_t kmax, std::vector<double>& out);
// next even 2,3,5-smooth integer >= n
int64_t next_smooth_even(int64_t n);
int64_t next_pow2(int64_t n);
int64_t next_smooth(int64_t n);

// device-side copy of the kernel table (passed by value to kernels -> constant bank)
template <typename T>
struct ESDev {
  T c[(kMaxW + 1) / 2][kMaxDeg + 1];
};
template <typename T>
inline void es_to_dev(const ESKernel& k, ESDev<T>& d) {
  for (int i = 0; i < (kMaxW + 1) / 2; ++i)
    for (int j = 0; j <= kMaxDeg; ++j) d.c[i][j] = (T)k.coef[i][j];
}

// sm_100a 256-bit global accesses (SASS LDG.E.ENL2.256 / STG.E.ENL2.256): a 32-byte point record moves
// as ONE L2 request covering one full sector.
#ifdef __CUDACC__
__device__ __forceinline__ void ld_global_nc_256(const double* p, double& a, double& b, double& c, double& d) {
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
}
__device__ __forceinline__ void st_global_256(double* p, double a, double b, double c, double d) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}
#endif

template <typename T> struct cplx_of;
template <> struct cplx_of<float> { using type = float2; };
template <> struct cplx_of<double> { using type = double2; };

inline int div_up(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// Process-wide caching pool for large device scratch (sort temporaries, sorted point sets):
// cudaMalloc/cudaFree of GB-sized blocks costs milliseconds and synchronises the device, and the
// EFGP loop rebuilds its plans every hyperparameter step.
// Optional per-stage device timing (timing.cpp): a ScopedStage records CUDA events on `st` around the
// launches issued during its lifetime when timing is enabled, and costs one branch otherwise.
bool timing_on();
void timing_enable(int on);
std::string timing_report();
struct ScopedStage {
  ScopedStage(const char* name, cudaStream_t st, int launches = 1, long long units = 0);
  ~ScopedStage();
  void add_launches(int n);
  cudaStream_t st_;
  long long idx_;
};

void* pool_alloc(size_t bytes);
void pool_free(void* p);
void pool_trim();
// stream of the C-ABI call in progress on this thread (pool.cpp: blocks released on one stream are not handed to
// another without waiting), and the one-GPU-per-process check
void set_current_stream(cudaStream_t st);
cudaStream_t current_stream();
void check_device();

}  // namespace efgp
