// Toeplitz (F^H F) operator and lockstep batched CG -- internal declarations.
#pragma once
#include "nufft.cuh"

namespace efgp {

struct Toeplitz {
  int dim = 0;
  int dtype = 1;
  int64_t L[kMaxDim] = {1, 1, 1};   // generator extents (4m+1)
  int64_t n[kMaxDim] = {1, 1, 1};   // block extents (L+1)/2
  int64_t F[kMaxDim] = {1, 1, 1};   // FFT extents
  int64_t M = 1, Ftot = 1;
  const void* v_raw = nullptr;      // the generator itself (borrowed: the caller keeps it alive with the handle)
  void* vhat = nullptr;             // FFT of the zero-padded generator, pre-scaled by 1/Ftot
  void* vhat_real = nullptr;        // real spectrum of the CENTRED generator (Hermitian v), built on first use
  DevBuf P, Q;                      // padded input (zeros persist), spectrum / output
  int64_t P_rows = 0;               // rows of P known to be zero outside the block
  DevBuf cg_vec, cg_scal;
  DevBuf dft_tw, dft;               // direct-DFT CG path (cg.cu): twiddle tables, U / X scratch
  bool dft_ready = false;           // twiddles written
  ~Toeplitz();
};

enum ApplyMode { kModeT = 0, kModeGv = 1, kModeAmean = 2, kModeAvar = 3 };

void toeplitz_create(Toeplitz& t, const void* v, int dim, const int64_t* Ls, int dtype, int force_pow2,
                     cudaStream_t st);
void toeplitz_apply(Toeplitz& t, const void* x, int batch, const void* ws, double sigmasq, int mode, void* y,
                    cudaStream_t st);
// hermitian != 0: the caller promises b[-k] = conj(b[k]) in the centred mode index for every row of b and x0 and
// a Hermitian generator v (both hold for D F^* c with real c and an even real diagonal D): the cuFFT matvec then
// runs as Z2D -> real multiply -> D2Z on half-size arrays.
void cg_solve(Toeplitz& t, const void* ws, double sigmasq, int mode, const void* minv, const void* b, void* x,
              int batch, double tol, int max_iter, int early_stopping, int batched_semantics, int x0_is_zero,
              int* iters_completed, cudaStream_t st, int hermitian = 0, int* async_counters = nullptr);
// async_counters (pinned host int[2], optional): when the solve runs as ONE cooperative kernel the call returns without
// synchronising, *iters_completed = -1, and the stream delivers {rows still active, last converging iteration}
// there; otherwise the call is synchronous and fills the same two ints itself.
// V[f] = (1 / Ftot) sum_j v[j] exp(-2 pi i f.j / F) over the centred generator index j: real when v[-j] = conj v[j].
const void* toeplitz_real_spectrum(Toeplitz& t, cudaStream_t st);
// cg_small.cu: the single-RHS Hermitian solve of a small 2-D block as one cooperative kernel with one barrier per
// iteration (direct block-Toeplitz convolution); false = geometry does not qualify
template <typename T>
bool cg_conv_launch(Toeplitz& t, const T* ws, const T* minv, void* x, const void* r0, const void* p0, T c1, T c2,
                    double tol, int early, int max_iter, const double* rz0, const double* denom, int* counters,
                    int* conv_iter, cudaStream_t st);
// lanczos.cu: `steps` lockstep Lanczos steps from the normalised start block q0 (overwritten); alphas / betas are
// (batch, steps) doubles on the device.  Enqueues only (no synchronisation).
void lanczos_tridiag(Toeplitz& t, const void* ws, double sigmasq, int mode, void* q0, int batch, int steps,
                     double breakdown, double* alphas, double* betas, cudaStream_t st);

// variance.cu: device stages of diag_sums_nd (efgpnd.py:1657-1664) and of the 'regular' variance (efgpnd.py:1805-1820)
void xcorr_mean(const void* us, const void* etas, const void* ws, int J, int dim, int mtot, int dtype, void* out,
                cudaStream_t st);
void variance_rows(const void* x, int64_t b, int dim, int mtot, double h, const void* ws, int dtype, void* rhs,
                   cudaStream_t st);
void variance_reduce(const void* x, int64_t b, int dim, int mtot, double h, const void* ws, int dtype, const void* gamma,
                     void* var, cudaStream_t st);

}  // namespace efgp
