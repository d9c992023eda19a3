/* libefgp_b200 -- C ABI of the B200-native EFGP hot path.
 *
 * The reference (danbider/gp-quadrature) is pure Python and has no FFI layer; its hot path
 * bottoms out in four external calls.  Each entry point below replaces one
 * reference interface, cited as file:line relative to the reference tree.  INTEGRATION.md shows
 * the ctypes binding a reference maintainer would add.
 *
 * Conventions
 *   - every data pointer is a caller-owned DEVICE pointer (e.g. torch.Tensor.data_ptr()); the
 *     caller keeps the allocation alive for the duration of the call (and, for set_points, for
 *     the lifetime of the handle);
 *   - dtype: 0 = float32 / complex64, 1 = float64 / complex128 (interleaved re,im);
 *   - `stream` is a cudaStream_t passed as void* (0 = default stream); calls enqueue on it and
 *     do not synchronise unless stated;
 *   - return value 0 = ok; otherwise efgp_last_error() holds a thread-local message;
 *   - handles are not thread-safe.
 */
#ifndef EFGP_B200_H
#define EFGP_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct efgp_nufft efgp_nufft;
typedef struct efgp_toeplitz efgp_toeplitz;

int efgp_version(void);
const char* efgp_last_error(void);

/* ---- diagnostics (no reference counterpart; the reference's torch.profiler ranges, efgpnd.py:46-59).
 * efgp_timing_enable(1) starts recording CUDA events around every launch group the library issues;
 * efgp_timing_report writes "stage,scopes,launches,units,total_ms\n" lines (synchronises the device).
 * efgp_pool_trim releases the cached device scratch blocks.
 * efgp_measure_fp64_peak times a pure DFMA kernel: the FP64 roofline denominator bench.py reports for the
 * FP64-pipe-bound spreading kernels (MEASURED_PEAKS.json only holds HBM and bf16 peaks). */
int efgp_timing_enable(int on);
int efgp_timing_report(char* buf, int64_t cap);
int efgp_pool_trim(void);
/* Deterministic spreading (also EFGP_DETERMINISTIC=1): the type-1 kernels accumulate 64-bit FIXED-POINT integers, whose
 * addition is associative, so every transform -- and with it every CG iterate and iteration count -- is bit-identical
 * from run to run (the default fp64 atomics land in a varying order: last-bit differences).  Resolution: 2^-62 of
 * N * max|strength| per addend; one host synchronisation per transform with explicit strengths.  A parity / debug
 * mode ("deterministic-reduction mode for parity tests"). */
int efgp_set_deterministic(int on);
int efgp_measure_fp64_peak(double* tflops, void* stream);

/* ---- NUFFT: replaces class NUFFT (efgpnd.py:1423-1549) and the pytorch_finufft calls under it */

/* NUFFT.__init__ (efgpnd.py:1432-1452): phi = 2*pi*h*(x - xcen).  x is (N, dim) row-major. */
int efgp_nufft_create(efgp_nufft** out, int dim, int dtype);
int efgp_nufft_set_points(efgp_nufft* nu, const void* x, int64_t npts, double h, const double* xcen,
                          void* stream);
void efgp_nufft_destroy(efgp_nufft* nu);
/* Same points, new spacing / centre: what the reference's training loop does at every optimiser step (a new
 * NUFFT(x, xcen, h, eps) object on the same x, efgpnd.py:103-107 inside efgpnd.py:1138-1195).  Keeps the handle and its
 * borrowed point pointer; the bin-sorted sets are rebuilt by the next transform if (h, xcen) changed. */
int efgp_nufft_rescale(efgp_nufft* nu, double h, const double* xcen);
/* on != 0: the sorted sets of this handle carry every point's ORIGINAL index (float64 handles: it rides in the unused
 * half of a record word, +8 bytes per point of sort traffic), which makes the regenerated Hutchinson probes below a
 * function of (seed, probe, original index): independent of the run, of the order inside a cell and of how probes are
 * split over ranks.  Off (default, and always for float32): the counter is the point's slot in the sorted set. */
int efgp_nufft_carry_index(efgp_nufft* nu, int on);

/* NUFFT.type1 -> pff.finufft_type1(phi, vals, out_shape, eps, isign=-1, modeord=False)
 * (efgpnd.py:1454-1508).  strengths: (batch, npts) real (is_complex=0) or complex (1), NULL = all
 * ones (the Toeplitz generator, efgpnd.py:1395-1421).  out: (batch, *out_shape) complex, C order,
 * modes -(n/2)..(n-1)/2 per axis (FFT order if modeord != 0).  scale: optional real (prod out_shape,)
 * diagonal multiplied into the output (the spectral weights ws, efgpnd.py:786). */
int efgp_nufft_type1(efgp_nufft* nu, const void* strengths, int is_complex, int batch,
                     const int64_t* out_shape, double eps, int isign, int modeord, const void* scale,
                     void* out, void* stream);

/* The two type-1 transforms of a fit in one call (EFGPND._compute_common_parameters, efgpnd.py:786-790:
 * rhs = ws * F^* y and compute_convolution_vector_vectorized_dD = F^* 1 at eps 6e-8): out_gen is
 * (*shape_gen) = F^* 1 at eps_gen, out_rhs is (*shape_rhs) = scale * F^* y at eps_rhs, y real (npts,).
 * One sort of the points serves both and carries y as payload (no strength permutation pass). */
int efgp_nufft_type1_pair(efgp_nufft* nu, const void* y, const int64_t* shape_gen, double eps_gen,
                          const int64_t* shape_rhs, double eps_rhs, const void* scale, void* out_gen, void* out_rhs,
                          void* stream);

/* NUFFT.type2 -> pff.finufft_type2(phi, fk, eps, isign=+1, modeord) (efgpnd.py:1510-1549, 1679).
 * modes: (batch, *shape) complex; scale: optional real diagonal applied to the modes first
 * (ws * beta, efgpnd.py:922).  out: (batch, npts) complex, or real when real_out != 0 (exactly the
 * real part of the complex result: every reference caller takes .real). */
int efgp_nufft_type2(efgp_nufft* nu, const void* modes, int batch, const int64_t* shape, double eps,
                     int isign, int modeord, const void* scale, int real_out, void* out, void* stream);

/* diagnostics: spreading width / fine grid the plan would use for (shape, eps) */
int efgp_nufft_plan_info(efgp_nufft* nu, const int64_t* shape, double eps, int64_t* nf_out, int* width_out);

/* ---- Toeplitz: replaces class ToeplitzND (efgpnd.py:1239-1393) */
int efgp_toeplitz_create(efgp_toeplitz** out, const void* v, int dim, const int64_t* Ls, int dtype,
                         int force_pow2, void* stream);
void efgp_toeplitz_destroy(efgp_toeplitz* t);
int efgp_toeplitz_fft_shape(efgp_toeplitz* t, int64_t* shape_out);

/* mode 0: y = T x                          ToeplitzND.__call__ (efgpnd.py:1331-1393)
 * mode 1: y = ws * T(ws * x)               create_Gv      (efgpnd.py:1572-1591)
 * mode 2: y = Gv(x) + sigmasq * x          create_A_mean  (efgpnd.py:1593-1600)
 * mode 3: y = Gv(x) / sigmasq + x          create_A_var   (efgpnd.py:1602-1609)
 * x, y: (batch, M) complex; ws: (M,) real. */
int efgp_toeplitz_apply(efgp_toeplitz* t, const void* x, int batch, const void* ws, double sigmasq,
                        int mode, void* y, void* stream);

/* ---- CG: replaces ConjugateGradients.solve (cg.py:73-244) for operators of the four modes above.
 * b: (batch, M) complex; x: (batch, M) complex, holds x0 on entry and the solution on exit;
 * minv: optional (M,) real Jacobi diagonal M^-1 (create_jacobi_precond, efgpnd.py:1619-1631);
 * max_iter < 0 means 2*M (cg.py:60-65); batched_semantics selects _solve_batched (1) or
 * _solve_single (0) stopping rules and iteration counting; x0_is_zero = 1 is the caller's promise that x holds
 * zeros on entry (the cold start of every fit, efgpnd.py:805), which saves the matvec A x0.  Synchronises the stream. */
int efgp_cg_solve(efgp_toeplitz* t, const void* ws, double sigmasq, int mode, const void* minv,
                  const void* b, void* x, int batch, double tol, int max_iter, int early_stopping,
                  int batched_semantics, int x0_is_zero, int* iters_completed, void* stream);

/* efgp_cg_solve for right-hand sides that are Hermitian-symmetric in the centred mode index, b[-k] = conj(b[k]) for
 * every row of b (and of x when x0_is_zero = 0), with a Hermitian generator v.  That is the case for every system the
 * reference builds as D F^* c with real point values c and an even real diagonal D: the posterior-mean solve
 * (efgpnd.py:786-812), the kernel-derivative probe rows of the gradient (efgpnd.py:183-197), the 'regular' variance rows
 * (efgpnd.py:1805-1820).  The operator preserves the symmetry, the transform of the padded iterate is then real, and the
 * matvec runs as C2R -> real multiply -> R2C on half-size arrays (blocks with odd extents; otherwise, and for the tiny
 * blocks solved without cuFFT, this is efgp_cg_solve).  Same arguments, semantics and iteration counting. */
int efgp_cg_solve_hermitian(efgp_toeplitz* t, const void* ws, double sigmasq, int mode, const void* minv,
                            const void* b, void* x, int batch, double tol, int max_iter, int early_stopping,
                            int batched_semantics, int x0_is_zero, int* iters_completed, void* stream);

/* efgp_cg_solve without the host synchronisation where the solve does not need the host: small 2-D blocks with <= 4
 * right-hand sides (the posterior-mean system of a fit) run as ONE cooperative kernel that takes every convergence
 * decision on the device; the call then returns at once with *iters_completed = -1 and the stream later writes
 * counters_pinned[0] = rows still active (<= 0: all converged) and counters_pinned[1] = index j of the last converging
 * iteration into the caller's PINNED host ints, from which iters_completed = all converged ? (batched_semantics ?
 * min(j + 2, max_iter) : j + 1) : max_iter (cg.py:132-153, 198-244).  x is valid in stream order.  Every other solve
 * runs synchronously exactly like efgp_cg_solve / efgp_cg_solve_hermitian (hermitian != 0), sets *iters_completed
 * and fills the two ints itself.  This lets the caller enqueue the type-2 transform of predict() behind the solve
 * instead of idling the GPU while the host learns an iteration count it only reports (efgpnd.py:805-812, 922). */
int efgp_cg_solve_async(efgp_toeplitz* t, const void* ws, double sigmasq, int mode, const void* minv, const void* b,
                        void* x, int batch, double tol, int max_iter, int early_stopping, int batched_semantics,
                        int x0_is_zero, int hermitian, int* counters_pinned, int* iters_completed, void* stream);

/* ---- Lanczos: replaces the probe loop of logdet_slq (efgpnd.py:1719-1746), all probes in lockstep, for
 * operators of the four modes above (the reference uses A = I + Gv / sigma^2, mode 3).
 * q: (batch, M) complex start vectors z / ||z|| (efgpnd.py:1723), OVERWRITTEN (used as work space);
 * alphas, betas: (batch, steps) float64 device arrays = the diagonal and the off-diagonal (betas[:, j] couples
 * rows j and j+1; the last column is unused, as in the reference's T, efgpnd.py:1748-1753).  A row whose beta falls
 * below breakdown_tol (1e-12, efgpnd.py:1742) stops: its remaining entries are alpha = 1, beta = 0, decoupled unit
 * blocks that add log 1 = 0 to the quadrature.  Enqueues on the stream, does not synchronise. */
int efgp_lanczos(efgp_toeplitz* t, const void* ws, double sigmasq, int mode, void* q, int batch, int steps,
                 double breakdown_tol, double* alphas, double* betas, void* stream);

/* ---- library communicator: SURVEY.md 8b's efgp_init(nccl) / efgp_allreduce_grid.  The reference has no
 * multi-device code; these serve the point-sharded fit (one rank per GPU).  Rank 0 calls
 * efgp_comm_unique_id, the host program broadcasts the 128 bytes (e.g. torch.distributed), every rank calls
 * efgp_comm_init.  NCCL is loaded with dlopen only here.  efgp_allreduce: in place on the stream, dtype 0 = f32,
 * 1 = f64, op 0 = sum, 1 = max; a no-op without a communicator. */
int efgp_comm_unique_id(void* id128);
int efgp_comm_init(int rank, int world, const void* id128);
int efgp_comm_destroy(void);
int efgp_comm_world(void);
int efgp_allreduce(void* buf, int64_t count, int dtype, int op, void* stream);

/* ---- the posterior-mean fit in one call: EFGPND._compute_common_parameters (efgpnd.py:745-822).
 * efgp_bbox_async: [max x_0.., -min x_0..] of the points ((npts, dim) row-major) -> ext_pinned (2 dim doubles of PINNED
 * host memory, valid once the stream reaches that point; MAX-reduced over the communicator when sharded != 0):
 * L = max_a (max - min) of efgpnd.py:751.  The host derives (h, mtot) from L (get_xis, utils/kernels.py:72-143). */
int efgp_bbox_async(const void* x, int64_t npts, int dim, int dtype, int sharded, double* ext_pinned, void* stream);

/* efgp_bbox_static: NO COUNTERPART IN THE REFERENCE -- added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same
 * reduction as efgp_bbox_async ([max x_0.., -min x_0..], bit-identical: max/min are exact) with its per-CTA partials in
 * a file-scope static __device__ array (csrc/noalloc.cu; 786,688 bytes held by every GPU that loads the library until
 * exit) instead of the scratch pool: it enqueues one cudaMemsetAsync (the finishing-CTA ticket must start at zero) and
 * one kernel, and calls no cudaMalloc, cudaFree or allocating library routine.  ext_dev: 2 dim doubles of DEVICE
 * memory supplied by the caller (no communicator reduction, no host copy).  npts > EFGP_BBOX_STATIC_MAX_POINTS (one
 * 8192-point tile per slot of the array) is refused with error code 1 and a message naming the limit.  Every call
 * shares the one array: calls that OVERLAP on the GPU (two streams, two graphs replayed at once) race -- serialise them. */
#define EFGP_BBOX_STATIC_MAX_POINTS 134217728LL
int efgp_bbox_static(const void* x, int64_t npts, int dim, int dtype, double* ext_dev, void* stream);

typedef struct efgp_fit_args {
  int family;            /* spectral density on xi = h * (k - (mtot-1)/2):  0: S = c exp(-a |xi|^2)  (squared_exponential.py:64-91) */
  double c, a, b, p;     /*                                                 1: S = c (b + 4 pi^2 |xi|^2)^(-p)  (matern.py:100-122) */
  double h, hd;          /* grid spacing and h^dim (efgpnd.py:756-780: ws = sqrt(S h^d)) */
  int mtot;              /* modes per axis (odd) */
  double sigmasq, nufft_eps, gen_eps, cg_tol;   /* gen_eps: tolerance of the F^*1 transform (6e-8, efgpnd.py:1418) */
  int max_iter;          /* < 0: 2 M (cg.py:60-65) */
  int precondition;      /* Jacobi preconditioner of efgpnd.py:795-799 */
  int force_pow2;        /* Toeplitz FFT size: next power of two (efgpnd.py:791) or next 2-3-5-smooth size */
  int sharded;           /* points sharded over the communicator: [F^*1 ; ws F^*y] is sum-reduced before the solve */
  int warm_start;        /* beta holds the starting guess (efgpnd.py:803-806); otherwise it must hold zeros */
} efgp_fit_args;

/* nu: handle of the training points (efgp_nufft_set_points); y: (npts,) real.  Caller-owned device outputs:
 * ws (M,) real; vr ((4m+1)^d + M,) complex = [F^*1 ; ws F^*y] (kept alive by the caller while *toep_out is used: the
 * generator is read once, the right-hand side by the solve); minv (M,) real Jacobi diagonal (may be NULL when
 * precondition = 0); beta (M,) complex in/out.  *toep_out is a new Toeplitz handle (efgp_toeplitz_destroy).  The solve
 * is enqueued like efgp_cg_solve_async (counters_pinned, iters_completed have the same meaning). */
int efgp_fit(efgp_nufft* nu, const void* y, const efgp_fit_args* args, void* ws, void* vr, void* minv, void* beta,
             efgp_toeplitz** toep_out, int* counters_pinned, int* iters_completed, void* stream);

/* ---- posterior variance: the device stages around the batched solves.
 * efgp_xcorr_mean: diag_sums_nd's lag sums (efgpnd.py:1657-1664),
 *   out[r] = mean_j ifftn( fftn(ws * us_j, s) * conj(fftn(etas_j, s)), s ),  s = (2 mtot - 1)^dim, FFT ordering,
 * us (nprobes, mtot^dim) complex (the CG solutions), etas (nprobes, mtot^dim) REAL probes, ws (mtot^dim,) real or NULL;
 * out ((2 mtot - 1)^dim) complex.
 * efgp_variance_rows / efgp_variance_reduce: compute_prediction_variance 'regular' (efgpnd.py:1805-1820) for test points
 * x (npts, dim) on the grid xi_k = h (k - (mtot-1)/2):  rhs[i] = ws * conj(fx_i)  ((npts, mtot^dim) complex) and
 * var[i] = max(0, Re sum_k fx_i[k] ws[k] gamma[i, k])  ((npts,) real); the dense fx of the reference is never stored. */
int efgp_xcorr_mean(const void* us, const void* etas, const void* ws, int nprobes, int dim, int mtot, int dtype, void* out,
                    void* stream);
int efgp_variance_rows(const void* x, int64_t npts, int dim, int mtot, double h, const void* ws, int dtype, void* rhs,
                       void* stream);
int efgp_variance_reduce(const void* x, int64_t npts, int dim, int mtot, double h, const void* ws, int dtype,
                         const void* gamma, void* var, void* stream);

/* ---- Hutchinson probes streamed through the transforms: the Monte-Carlo trace of efgpnd_gradient_batched
 * (efgpnd.py:176-241) without its (T, N) arrays Z, F(D'F^*Z) and F(ws beta).  Probe t is the Rademacher vector
 * Z_t[n] = sign bit of Philox-4x32-10(key = seed, counter = (index of point n, t / 128)) -- the ORIGINAL index with
 * efgp_nufft_carry_index(nu, 1), else the slot of the point in the sorted set --, regenerated
 * inside the spread kernel and inside the interpolation epilogue, never stored.
 *   efgp_nufft_type1_probes:     out[b] = scale * F^* Z_{t0+b}, b < nprobes: (nprobes, *out_shape) complex (efgpnd.py:183)
 *   efgp_nufft_type2_probe_dots: sums[b] = sum_n Z_{t0+b}[n] * Re(F (scale * modes[b]))[n]: modes (nprobes, *shape)
 *                                complex -> sums (nprobes,) FLOAT64 on the device (efgpnd.py:235-241, any dtype)
 *   efgp_nufft_probes_export:    z (nprobes, npts) real, original point order: the probes themselves (tests)
 * The signs are tied to the point ORDER of one sorted set: *sort_id (in/out, start with 0) carries its identity from
 * the first call to the later ones, which fail loudly if the handle re-sorted in between.  2-D kernels take up to 4
 * probes per pass over the points (kernel weights evaluated once per point). */
int efgp_nufft_type1_probes(efgp_nufft* nu, uint64_t seed, int t0, int nprobes, const int64_t* out_shape, double eps,
                            const void* scale, void* out, uint64_t* sort_id, void* stream);
int efgp_nufft_type2_probe_dots(efgp_nufft* nu, uint64_t seed, int t0, int nprobes, const void* modes, const int64_t* shape,
                                double eps, const void* scale, double* sums, uint64_t* sort_id, void* stream);
int efgp_nufft_probes_export(efgp_nufft* nu, uint64_t seed, int t0, int nprobes, const int64_t* shape, double eps, void* z,
                             uint64_t* sort_id, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* EFGP_B200_H */
