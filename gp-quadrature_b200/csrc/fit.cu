// The posterior-mean fit as ONE library call: everything EFGPND._compute_common_parameters (efgpnd.py:745-822 of
// the reference) enqueues after the frequency grid (h, mtot) is known --
//   ws = sqrt(S(xi) h^d) on the mode grid                       (efgpnd.py:767-780, kernels/*.spectral_density)
//   v = F^*1 at eps 6e-8 and rhs = ws * F^*y, one sort          (efgpnd.py:786-790)
//   [points sharded over GPUs: ONE NCCL all-reduce of [v ; rhs], issued right behind the deconvolution kernels]
//   Toeplitz spectrum, Jacobi diagonal 1 / (c0 ws^2 + sigma^2)   (efgpnd.py:791-799)
//   PCG for beta, enqueued without a host synchronisation        (efgpnd.py:807-814)
// -- with no Python between the launches: the host part of a step was ~1.3 ms of interpreter and ctypes time, which
// bounded the 8-GPU strong scaling of round 1 at 3.96x.  The bounding box that fixes h (efgpnd.py:751) is a separate
// asynchronous call (bbox_async) so that the host's L-independent work overlaps it.
#include "toeplitz.cuh"
#include "fit.cuh"
#include <algorithm>

namespace efgp {

// ------------------------------------------------------------------ bounding box
constexpr int kBoxBlocks = 592;     // 4 CTAs per SM
constexpr int kBoxThreads = 256;

template <typename T, int D>
__global__ void __launch_bounds__(kBoxThreads)
bbox_partial_kernel(const T* __restrict__ x, int64_t N, double* __restrict__ part /* [blocks][2 D] */) {
  double hi[D], nlo[D];
#pragma unroll
  for (int a = 0; a < D; ++a) hi[a] = nlo[a] = -1.0e308;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x) {
    T xv[D];
    if constexpr (D == 2) {
      using V2 = typename cplx_of<T>::type;
      const V2 p = reinterpret_cast<const V2*>(x)[i];
      xv[0] = p.x;
      xv[1] = p.y;
    } else {
#pragma unroll
      for (int a = 0; a < D; ++a) xv[a] = x[i * D + a];
    }
#pragma unroll
    for (int a = 0; a < D; ++a) {
      const double v = (double)xv[a];
      hi[a] = v > hi[a] ? v : hi[a];
      nlo[a] = -v > nlo[a] ? -v : nlo[a];
    }
  }
  __shared__ double sh[2 * D][kBoxThreads / 32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int a = 0; a < D; ++a) {
    double u = hi[a], l = nlo[a];
    for (int o = 16; o > 0; o >>= 1) {
      u = fmax(u, __shfl_xor_sync(0xffffffffu, u, o));
      l = fmax(l, __shfl_xor_sync(0xffffffffu, l, o));
    }
    if (lane == 0) {
      sh[a][w] = u;
      sh[D + a][w] = l;
    }
  }
  __syncthreads();
  if (threadIdx.x < 2 * D) {
    double m = -1.0e308;
    for (int q = 0; q < kBoxThreads / 32; ++q) m = fmax(m, sh[threadIdx.x][q]);
    part[(int64_t)blockIdx.x * 2 * D + threadIdx.x] = m;
  }
}

// one WARP per output [hi_0..hi_{d-1}, -lo_0..-lo_{d-1}]: lanes stride over the CTA partials (a single thread per
// output walked 592 dependent loads: 44 us in the ncu launch list of round 2, replicated on every rank)
__global__ void bbox_final_kernel(const double* __restrict__ part, int blocks, int dim, double* __restrict__ ext) {
  const int c = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (c >= 2 * dim) return;
  double m = -1.0e308;
  for (int b = lane; b < blocks; b += 32) m = fmax(m, part[(int64_t)b * 2 * dim + c]);
  for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (lane == 0) ext[c] = m;
}

void bbox_async(const void* x, int64_t N, int dim, int dtype, int sharded, double* ext_pinned, cudaStream_t st) {
  EFGP_CHECK(dim >= 1 && dim <= kMaxDim, "dim must be 1..3");
  EFGP_CHECK(N >= 1, "bounding box of an empty point set");
  double* part = (double*)pool_alloc(sizeof(double) * (size_t)(kBoxBlocks + 1) * 2 * kMaxDim);
  double* ext = part + (size_t)kBoxBlocks * 2 * kMaxDim;
  const int blocks = (int)std::min<int64_t>(kBoxBlocks, (N + kBoxThreads - 1) / kBoxThreads);
  {
    ScopedStage stg("fit.bbox", st, 2, N);
#define EFGP_BOX(T, D) bbox_partial_kernel<T, D><<<blocks, kBoxThreads, 0, st>>>((const T*)x, N, part)
    if (dtype == 0) {
      if (dim == 1) EFGP_BOX(float, 1); else if (dim == 2) EFGP_BOX(float, 2); else EFGP_BOX(float, 3);
    } else {
      if (dim == 1) EFGP_BOX(double, 1); else if (dim == 2) EFGP_BOX(double, 2); else EFGP_BOX(double, 3);
    }
#undef EFGP_BOX
    bbox_final_kernel<<<1, 32 * 2 * dim, 0, st>>>(part, blocks, dim, ext);
    EFGP_CUDA(cudaGetLastError());
  }
  if (sharded) comm_allreduce(ext, 2 * dim, 1, 1, st);
  EFGP_CUDA(cudaMemcpyAsync(ext_pinned, ext, sizeof(double) * 2 * dim, cudaMemcpyDeviceToHost, st));
  pool_free(part);   // stream-ordered reuse
}

// ------------------------------------------------------------------ spectral weights and Jacobi diagonal
template <typename T>
__global__ void __launch_bounds__(256)
ws_kernel(T* __restrict__ ws, int dim, int mtot, int64_t M, FitArgs a) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= M) return;
  const int hm = (mtot - 1) / 2;
  int64_t r = k;
  double q = 0;
  for (int ax = 0; ax < dim; ++ax) {
    const int i = (int)(r % mtot);
    r /= mtot;
    double xi = (double)(i - hm) * a.h;
    if (sizeof(T) == 4) xi = (double)(float)xi;       // the float32 model holds its frequency grid in float32
    q += xi * xi;
  }
  const double S = a.family == 0 ? a.c * exp(-a.a * q) : a.c * pow(a.b + 39.47841760435743447533796 * q, -a.p);
  ws[k] = (T)sqrt(S * a.hd);
}

// minv[k] = 1 / (c0 ws[k]^2 + sigma^2), c0 = Re v[centre] (efgpnd.py:795-799, 1619-1631)
template <typename T>
__global__ void __launch_bounds__(256)
jacobi_kernel(const T* __restrict__ ws, const typename cplx_of<T>::type* __restrict__ v, int64_t centre, T sigmasq,
              int64_t M, T* __restrict__ minv) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= M) return;
  const T c0 = v[centre].x;
  const T w = ws[k];
  minv[k] = T(1) / (c0 * (w * w) + sigmasq);
}

void fit_run(Nufft& nu, Toeplitz& t, const void* y, const FitArgs& a, void* ws, void* vr, void* minv, void* beta,
             int* counters_pinned, int* iters_completed, cudaStream_t st) {
  const int dim = nu.dim;
  EFGP_CHECK(a.mtot >= 1 && (a.mtot & 1), "mtot must be odd");
  EFGP_CHECK(ws && vr && beta, "null output buffer");
  int64_t M = 1, Lg = 1, centre = 0;
  int64_t shape_rhs[kMaxDim] = {1, 1, 1}, shape_gen[kMaxDim] = {1, 1, 1};
  const int64_t lgen = 2 * (int64_t)(a.mtot - 1) + 1;          // 4 m + 1 with m = (mtot - 1) / 2
  for (int ax = 0; ax < dim; ++ax) {
    shape_rhs[ax] = a.mtot;
    shape_gen[ax] = lgen;
    M *= a.mtot;
    Lg *= lgen;
    centre = centre * lgen + (lgen - 1) / 2;
  }
  const size_t csz = nu.dtype == 0 ? 8 : 16;
  char* v = (char*)vr;
  char* rhs = v + (size_t)Lg * csz;
  {
    ScopedStage stg("fit.ws_kernel", st, 1, M);
    if (nu.dtype == 0)
      ws_kernel<float><<<div_up(M, 256), 256, 0, st>>>((float*)ws, dim, a.mtot, M, a);
    else
      ws_kernel<double><<<div_up(M, 256), 256, 0, st>>>((double*)ws, dim, a.mtot, M, a);
    EFGP_CUDA(cudaGetLastError());
  }
  // sharded: all ranks choose the fine grid as one rank holding every point would (the rank counts differ by at most one point)
  nu.cost_npts = a.sharded ? nu.N * (int64_t)std::max(1, comm_world()) : 0;
  nufft_type1_pair(nu, y, shape_gen, a.gen_eps, shape_rhs, a.nufft_eps, ws, v, rhs, st);
  if (a.sharded) {
    ScopedStage stg("fit.allreduce_grids", st, 0, Lg + M);
    comm_allreduce(vr, 2 * (Lg + M), nu.dtype, 0, st);
  }
  toeplitz_create(t, v, dim, shape_gen, nu.dtype, a.force_pow2, st);
  const void* mi = nullptr;
  if (a.precondition) {
    EFGP_CHECK(minv != nullptr, "null Jacobi buffer");
    ScopedStage stg("fit.jacobi_kernel", st, 1, M);
    if (nu.dtype == 0)
      jacobi_kernel<float><<<div_up(M, 256), 256, 0, st>>>((const float*)ws, (const float2*)v, centre, (float)a.sigmasq, M,
                                                          (float*)minv);
    else
      jacobi_kernel<double><<<div_up(M, 256), 256, 0, st>>>((const double*)ws, (const double2*)v, centre, a.sigmasq, M,
                                                            (double*)minv);
    EFGP_CUDA(cudaGetLastError());
    mi = minv;
  }
  cg_solve(t, ws, a.sigmasq, kModeAmean, mi, rhs, beta, 1, a.cg_tol, a.max_iter, 1, 0, a.warm_start ? 0 : 1,
           iters_completed, st, 1, counters_pinned);
}

}  // namespace efgp
