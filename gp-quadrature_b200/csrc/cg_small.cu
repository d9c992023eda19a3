// Posterior-mean PCG for SMALL 2-D blocks by direct block-Toeplitz convolution: one cooperative kernel, ONE grid barrier
// per iteration.
//
// The mean solve of a fit (EFGPND._compute_common_parameters, efgpnd.py:807-814 -> cg.py:86-153) on M = mtot^2 <= ~6000
// unknowns is pure latency: cg_persist_kernel (cg.cu) needs four grid barriers per iteration, because its zero-padded
// 2-D DFT matvec transposes twice (rows -> columns -> rows) and its two reductions are global.  Here
//   * the matvec is the DEFINITION of the operator, (T q)[k] = sum_j v[k - j + n - 1] q[j]  (ToeplitzND,
//     efgpnd.py:1331-1393: the crop window [n-1, 2n-1) of the linear convolution with the generator), evaluated
//     directly: M^2 complex MACs (2.5e7 for 71^2), halved by the Hermitian symmetry of the iterates
//     (b[-k] = conj b[k] for D F^* y with real y: only the modes k <= mid are computed), spread over <= 148 CTAs that
//     each own a strip of outputs -- no transposes;
//   * every CTA keeps the (half) vectors p and r in shared memory and performs the vector updates and BOTH reductions of
//     the iteration redundantly, in the same order, so all CTAs hold bit-identical scalars and take the same decisions
//     without a second or third barrier.  The only exchange is A p (40 KB for 71^2), published to global memory and read
//     back (L1 bypassed) after the one grid.sync().
// Arithmetic of an iteration is cg.py's (alpha = rz / (pAp + 1e-16), beta = rz' / (rz + 1e-16), stop when
// ||r|| / (||b|| + 1e-16) < tol, tested after the residual update); only the summation order of the dot products and of
// the matvec differs from the FFT path (rounding level).
#include "toeplitz.cuh"
#include <cooperative_groups.h>
#include <cstdlib>

namespace efgp {

constexpr int kConvThreads = 256;
// outputs per thread (sliding window of generator values): template parameter kConvR in {3, 4, 6}, see cg_conv_launch
constexpr int kConvMaxM = 6400;      // shared memory: 32 M bytes for q, p, r (+ partial sums)

template <typename T>
struct ConvArgs {
  using C = typename cplx_of<T>::type;
  const C* v;            // generator (2 n0 - 1, 2 n1 - 1), C order
  const T* ws;           // (M)
  const T* minv;         // (M) or nullptr
  C* x;                  // (M) in/out
  const C* r0;           // (M) initial residual b - A x0 (symmetric)
  const C* p0;           // (M) initial direction M^-1 r0
  C* Ap;                 // (2, H) exchange buffer, double-buffered by iteration parity
  const double* rz0;     // <r0, M^-1 r0>
  const double* denom;   // ||b|| (1 if b = 0)
  int* counters;         // [0] rows still active, [1] index of the converging iteration
  int* conv_iter;
  int n0, n1, M, H;      // H = (M + 1) / 2: modes 0 .. H-1, mode H-1 is its own mirror
  int chunks, cper;      // column chunks per output row, columns per chunk
  int L, G;              // lanes per group (ceil(cper / R)), groups per CTA
  T c1, c2;
  double tol;
  int early, max_iter;
};

template <typename C>
__device__ __forceinline__ void cmacc(C& acc, C a, C b) {
  acc.x = fma(a.x, b.x, acc.x);
  acc.x = fma(-a.y, b.y, acc.x);
  acc.y = fma(a.x, b.y, acc.y);
  acc.y = fma(a.y, b.x, acc.y);
}

// sum over the CTA, result broadcast to every thread (fixed order: identical in every CTA)
__device__ __forceinline__ double cta_allsum(double v, double* sh) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0;
#pragma unroll
  for (int i = 0; i < kConvThreads / 32; ++i) s += sh[i];
  return s;
}

template <typename T, int kConvR>
__global__ void __launch_bounds__(kConvThreads) cg_conv_kernel(const ConvArgs<T> a) {
  using C = typename cplx_of<T>::type;
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  extern __shared__ __align__(16) unsigned char conv_smem[];
  C* q = reinterpret_cast<C*>(conv_smem);      // [M]  ws * p, both halves
  C* ph = q + a.M;                             // [H]
  C* rh = ph + a.H;                            // [H]
  C* part = rh + a.H;                          // [G][cper]
  __shared__ double sh[kConvThreads / 32];
  const int tid = threadIdx.x;
  const int M = a.M, H = a.H, n0 = a.n0, n1 = a.n1, L1 = 2 * n1 - 1;
  const int k0 = blockIdx.x / a.chunks, col0 = (blockIdx.x % a.chunks) * a.cper;
  const int ncol = (n1 - col0) < a.cper ? (n1 - col0) : a.cper;
  for (int k = tid; k < H; k += kConvThreads) {
    ph[k] = a.p0[k];
    rh[k] = a.r0[k];
  }
  double rz = a.rz0[0];
  const double denom = a.denom[0];
  double rr = 0, rz_new = 0;
  // the output this thread owns after the partial sums are combined (tid < ncol)
  const int kown = k0 * n1 + col0 + tid;
  const bool owner = tid < ncol && kown < H;
  const int l = tid % a.L, g = tid / a.L;
  int it = 0, conv_at = -1;
  for (; it < a.max_iter; ++it) {
    if (it > 0) {
      if (a.early && sqrt(rr) / (denom + 1e-16) < a.tol) {
        conv_at = it - 1;
        break;
      }
      const T bt = (T)(rz_new / (rz + 1e-16));
      rz = rz_new;
      for (int k = tid; k < H; k += kConvThreads) {
        C pv = ph[k];
        const C rv = rh[k];
        const T mi = a.minv ? a.minv[k] : T(1);
        pv.x = rv.x * mi + bt * pv.x;
        pv.y = rv.y * mi + bt * pv.y;
        ph[k] = pv;
      }
    }
    __syncthreads();
    for (int k = tid; k < H; k += kConvThreads) {      // q = ws p on both halves (ws is even)
      const C pv = ph[k];
      const T w = a.ws[k];
      C qv;
      qv.x = pv.x * w;
      qv.y = (k == H - 1) ? T(0) : pv.y * w;            // the self-mirrored mode is real
      q[k] = qv;
      qv.y = -qv.y;
      q[M - 1 - k] = qv;
    }
    __syncthreads();
    // ---- matvec strip: outputs (k0, col0 .. col0 + ncol): work units = half rows of q, dealt to the lane groups
    if (g < a.G) {
      C acc[kConvR];
#pragma unroll
      for (int r = 0; r < kConvR; ++r) acc[r].x = acc[r].y = 0;
      const int hl = (n1 + 1) / 2;                       // first half: j1 in [0, hl), second: [hl, n1)
      const int o0 = l * kConvR;                         // first output column (relative) of this thread
      for (int u = g; u < 2 * n0; u += a.G) {
        const int j0 = u >> 1;
        const int ja = (u & 1) ? hl : 0, jb = (u & 1) ? n1 : hl;
        const C* vrow = a.v + (size_t)(k0 - j0 + n0 - 1) * L1;
        const C* qrow = q + j0 * n1;
        // window w[r] = vrow[col0 + o0 + r + n1 - 1 - j1]: the index falls by one per step of j1
        int base = col0 + o0 + n1 - 1 - ja;
        C w[kConvR];
#pragma unroll
        for (int r = 0; r < kConvR; ++r) {
          int idx = base + r;
          idx = idx < L1 ? idx : L1 - 1;
          w[r] = __ldg(vrow + idx);
        }
        for (int j1 = ja; j1 < jb; ++j1) {
          const C qv = qrow[j1];
#pragma unroll
          for (int r = 0; r < kConvR; ++r) cmacc(acc[r], w[r], qv);
#pragma unroll
          for (int r = kConvR - 1; r > 0; --r) w[r] = w[r - 1];
          --base;
          w[0] = __ldg(vrow + (base >= 0 ? base : 0));
        }
      }
#pragma unroll
      for (int r = 0; r < kConvR; ++r)
        if (o0 + r < a.cper) part[g * a.cper + o0 + r] = acc[r];
    }
    __syncthreads();
    // A p is double-buffered by iteration parity: a CTA that is already publishing iteration it + 1 cannot disturb a CTA
    // still reading iteration it, and nobody can run two iterations ahead of the barrier
    C* Apb = a.Ap + (size_t)(it & 1) * H;
    C pown, apown;
    pown.x = pown.y = apown.x = apown.y = 0;
    if (owner) {
      C s;
      s.x = s.y = 0;
      for (int gg = 0; gg < a.G; ++gg) {
        const C v2 = part[gg * a.cper + tid];
        s.x += v2.x;
        s.y += v2.y;
      }
      pown = ph[kown];
      const T w = a.ws[kown] * a.c1;
      apown.x = s.x * w + a.c2 * pown.x;
      apown.y = s.y * w + a.c2 * pown.y;
      if (kown == H - 1) apown.y = 0;                    // the self-mirrored mode is real
      Apb[kown] = apown;
    }
    grid.sync();                                         // the ONE exchange (and barrier) of the iteration
    // ---- redundant vector phase: identical in every CTA
    double s_pap = 0;
    for (int k = tid; k < H; k += kConvThreads) {
      const C av = __ldcg(Apb + k);
      const C pv = ph[k];
      const double wgt = (k == H - 1) ? 1.0 : 2.0;
      s_pap += wgt * ((double)pv.x * av.x + (double)pv.y * av.y);
    }
    const double pAp = cta_allsum(s_pap, sh) + 1e-16;
    const T alpha = (T)(rz / pAp);
    double s_rr = 0, s_rz = 0;
    for (int k = tid; k < H; k += kConvThreads) {
      const C av = __ldcg(Apb + k);
      C rv = rh[k];
      rv.x -= alpha * av.x;
      rv.y -= alpha * av.y;
      rh[k] = rv;
      const double wgt = (k == H - 1) ? 1.0 : 2.0;
      const double n2 = (double)rv.x * rv.x + (double)rv.y * rv.y;
      s_rr += wgt * n2;
      s_rz += wgt * (double)(a.minv ? a.minv[k] : T(1)) * n2;
    }
    if (owner) {                                         // x += alpha p on the owned mode and on its mirror
      C xv = a.x[kown];
      xv.x += alpha * pown.x;
      xv.y += alpha * pown.y;
      if (kown == H - 1) xv.y = 0;
      a.x[kown] = xv;
      if (kown != M - 1 - kown) {
        xv.y = -xv.y;
        a.x[M - 1 - kown] = xv;
      }
    }
    rr = cta_allsum(s_rr, sh);
    rz_new = cta_allsum(s_rz, sh);
  }
  if (conv_at < 0 && it == a.max_iter && it > 0 && a.early && sqrt(rr) / (denom + 1e-16) < a.tol) conv_at = it - 1;
  if (blockIdx.x == 0 && tid == 0) {
    a.counters[0] = conv_at >= 0 ? 0 : 1;
    a.counters[1] = conv_at;
    a.conv_iter[0] = conv_at;
  }
}

// Returns false when the geometry does not qualify (the caller falls back to cg_persist_kernel).
template <typename T>
bool cg_conv_launch(Toeplitz& t, const T* ws, const T* minv, void* x, const void* r0, const void* p0, T c1, T c2,
                    double tol, int early, int max_iter, const double* rz0, const double* denom, int* counters,
                    int* conv_iter, cudaStream_t st) {
  using C = typename cplx_of<T>::type;
  static const bool off = std::getenv("EFGP_CG_CONV") && std::atoi(std::getenv("EFGP_CG_CONV")) == 0;
  if (off || t.dim != 2 || !t.v_raw) return false;
  const int n0 = (int)t.n[0], n1 = (int)t.n[1];
  const int64_t M = t.M;
  if (!(n0 & 1) || !(n1 & 1) || M > kConvMaxM || M < 9) return false;
  static int sms = 0, coop = 0;
  if (!sms) {
    int dev = 0;
    EFGP_CUDA(cudaGetDevice(&dev));
    EFGP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    EFGP_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
  }
  if (!coop) return false;
  ConvArgs<T> a;
  a.n0 = n0;
  a.n1 = n1;
  a.M = (int)M;
  a.H = (int)((M + 1) / 2);
  const int rows = (n0 + 1) / 2;                       // output rows 0 .. (n0-1)/2 hold every mode k <= mid
  a.chunks = std::max(1, std::min(n1, sms / rows));
  a.cper = (n1 + a.chunks - 1) / a.chunks;
  a.chunks = (n1 + a.cper - 1) / a.cper;
  // outputs per thread: fatter threads amortise the q loads and the window shifts over more MACs (measured at 71^2,
  // cper = 18: R = 3 / 4 / 6 / 9 / 12 / 18 -> 0.206 / 0.203 / 0.178 / 0.206 / 0.244 / 0.214 ms for 10 iterations); among
  // {3, 4, 6} take the one that wastes the fewest lanes on the column chunk, the larger one on a tie
  static const int r_env = std::getenv("EFGP_CG_CONV_R") ? std::atoi(std::getenv("EFGP_CG_CONV_R")) : 0;
  int kConvR = 3;
  if (r_env == 3 || r_env == 4 || r_env == 6) {
    kConvR = r_env;
  } else {
    int best_waste = 1 << 30;
    for (int r : {3, 4, 6}) {
      const int waste = (a.cper + r - 1) / r * r - a.cper;
      if (waste <= best_waste) {
        best_waste = waste;
        kConvR = r;
      }
    }
  }
  if (a.cper > kConvThreads) return false;
  const int ctas = rows * a.chunks;
  size_t smem = 0;
  for (int attempt = 0; attempt < 2; ++attempt) {        // more lane groups mean more partial sums: fall back to R = 3
    a.L = (a.cper + kConvR - 1) / kConvR;
    a.G = std::min(kConvThreads / a.L, 2 * n0);
    smem = sizeof(C) * ((size_t)M + 2 * (size_t)a.H + (size_t)a.G * a.cper);
    if (smem <= 220 * 1024 || kConvR == 3) break;
    kConvR = 3;
  }
  void* fn = kConvR == 6 ? (void*)cg_conv_kernel<T, 6> : (kConvR == 4 ? (void*)cg_conv_kernel<T, 4> : (void*)cg_conv_kernel<T, 3>);
  if (smem > 220 * 1024) return false;
  EFGP_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  EFGP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kConvThreads, smem));
  if (per_sm < 1 || ctas > per_sm * sms) return false;
  a.v = (const C*)t.v_raw;
  a.ws = ws;
  a.minv = minv;
  a.x = (C*)x;
  a.r0 = (const C*)r0;
  a.p0 = (const C*)p0;
  a.Ap = (C*)t.dft.get(sizeof(C) * 2 * (size_t)a.H);
  a.rz0 = rz0;
  a.denom = denom;
  a.counters = counters;
  a.conv_iter = conv_iter;
  a.c1 = c1;
  a.c2 = c2;
  a.tol = tol;
  a.early = early;
  a.max_iter = max_iter;
  void* args[] = {(void*)&a};
  EFGP_CUDA(cudaLaunchCooperativeKernel(fn, dim3(ctas), dim3(kConvThreads), args, smem, st));
  return true;
}

template bool cg_conv_launch<float>(Toeplitz&, const float*, const float*, void*, const void*, const void*, float, float,
                                    double, int, int, const double*, const double*, int*, int*, cudaStream_t);
template bool cg_conv_launch<double>(Toeplitz&, const double*, const double*, void*, const void*, const void*, double,
                                     double, double, int, int, const double*, const double*, int*, int*, cudaStream_t);

}  // namespace efgp
