// Lockstep batched Lanczos tridiagonalisation on the fused Toeplitz operator: the inner loop of the
// stochastic Lanczos quadrature log-determinant (logdet_slq, reference efgpnd.py:1722-1746), all probes at once.
//
// The recurrence is the reference's, in its order of operations, per probe row:
//     v = A q - beta_prev q_prev;  alpha = Re<q, v>;  v -= alpha q;  beta = ||v||;
//     record (alpha, beta);  stop the row if beta < breakdown;  q_prev = q;  q = v / beta
// A row that broke down keeps emitting (alpha, beta) = (1, 0): a decoupled unit block of the tridiagonal
// matrix whose Ritz value 1 adds log 1 = 0 to the quadrature, which is what ending that row's loop does.
// Nothing is read back during the loop (no host synchronisation): per step there is one batched matvec
// (toeplitz.cu) and three fused vector kernels with deterministic two-stage reductions.
#include "toeplitz.cuh"

namespace efgp {
namespace {

constexpr int kLzMaxChunks = 256;   // <= blockDim
constexpr int kLzChunkLen = 512;

struct LzScal {
  double* p_a;     // [B][NC] partials of Re<q, v>
  double* p_b;     // [B][NC] partials of ||v||^2
  int* alive;      // [2][B]
  int B, NC;
  int64_t len;
};

__device__ __forceinline__ double lz_block_sum(double v, double* sh) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) sh[w] = v;
  __syncthreads();
  double s = 0;
  if (threadIdx.x == 0)
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += sh[i];
  return s;  // valid in thread 0
}

// same value and summation order in every thread of every block of the row
__device__ __forceinline__ double lz_sum_partials(const double* p, int nc) {
  __shared__ double shp[9];
  double v = (int)threadIdx.x < nc ? p[threadIdx.x] : 0.0;
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) shp[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0;
  for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += shp[i];
  return s;
}

__global__ void lz_init_kernel(int* alive, int B) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B) alive[i] = 1;
}

// v = Aq - beta_prev * q_prev (in place in Aq), partials of Re<q, v>
template <typename T>
__global__ void __launch_bounds__(256)
lz_v_kernel(typename cplx_of<T>::type* __restrict__ v, const typename cplx_of<T>::type* __restrict__ q,
            const typename cplx_of<T>::type* __restrict__ qprev, int64_t M, int step, LzScal s) {
  using C = typename cplx_of<T>::type;
  __shared__ double sh[8];
  const int row = blockIdx.y, c = blockIdx.x;
  if (!s.alive[(step & 1) * s.B + row]) return;
  T bp = 0;
  if (step > 0) bp = (T)sqrt(lz_sum_partials(s.p_b + (size_t)row * s.NC, s.NC));
  const int64_t k0 = c * s.len, k1 = k0 + s.len < M ? k0 + s.len : M;
  double acc = 0;
  for (int64_t k = k0 + threadIdx.x; k < k1; k += blockDim.x) {
    const int64_t i = (int64_t)row * M + k;
    C a = v[i];
    const C qq = q[i];
    if (step > 0) {
      const C qp = qprev[i];
      a.x -= bp * qp.x;
      a.y -= bp * qp.y;
      v[i] = a;
    }
    acc += (double)qq.x * a.x + (double)qq.y * a.y;
  }
  const double t = lz_block_sum(acc, sh);
  if (threadIdx.x == 0) s.p_a[(size_t)row * s.NC + c] = t;
}

// alpha = sum of the partials; v -= alpha q; partials of ||v||^2; alphas[row, step] = alpha (1 for a stopped row)
template <typename T>
__global__ void __launch_bounds__(256)
lz_alpha_kernel(typename cplx_of<T>::type* __restrict__ v, const typename cplx_of<T>::type* __restrict__ q,
                int64_t M, int step, int steps, double* __restrict__ alphas, LzScal s) {
  using C = typename cplx_of<T>::type;
  __shared__ double sh[8];
  const int row = blockIdx.y, c = blockIdx.x;
  if (!s.alive[(step & 1) * s.B + row]) {
    if (c == 0 && threadIdx.x == 0) alphas[(size_t)row * steps + step] = 1.0;
    return;
  }
  const double alpha_d = lz_sum_partials(s.p_a + (size_t)row * s.NC, s.NC);
  const T alpha = (T)alpha_d;
  const int64_t k0 = c * s.len, k1 = k0 + s.len < M ? k0 + s.len : M;
  double acc = 0;
  for (int64_t k = k0 + threadIdx.x; k < k1; k += blockDim.x) {
    const int64_t i = (int64_t)row * M + k;
    C a = v[i];
    const C qq = q[i];
    a.x -= alpha * qq.x;
    a.y -= alpha * qq.y;
    v[i] = a;
    acc += (double)a.x * a.x + (double)a.y * a.y;
  }
  const double t = lz_block_sum(acc, sh);
  if (threadIdx.x == 0) {
    s.p_b[(size_t)row * s.NC + c] = t;
    if (c == 0) alphas[(size_t)row * steps + step] = (double)alpha;
  }
}

// beta = ||v||; betas[row, step] = beta (0 for a stopped row); the row stops if beta < breakdown; q_next = v / beta
template <typename T>
__global__ void __launch_bounds__(256)
lz_beta_kernel(typename cplx_of<T>::type* __restrict__ v, int64_t M, int step, int steps, double breakdown,
               double* __restrict__ betas, LzScal s) {
  using C = typename cplx_of<T>::type;
  const int row = blockIdx.y, c = blockIdx.x;
  int* alive_next = s.alive + ((step + 1) & 1) * s.B + row;
  if (!s.alive[(step & 1) * s.B + row]) {
    if (c == 0 && threadIdx.x == 0) {
      betas[(size_t)row * steps + step] = 0.0;
      *alive_next = 0;
    }
    return;
  }
  const T beta = (T)sqrt(lz_sum_partials(s.p_b + (size_t)row * s.NC, s.NC));
  const bool go = (double)beta >= breakdown;
  if (c == 0 && threadIdx.x == 0) {
    betas[(size_t)row * steps + step] = (double)beta;
    *alive_next = go ? 1 : 0;
  }
  if (!go) return;
  const int64_t k0 = c * s.len, k1 = k0 + s.len < M ? k0 + s.len : M;
  for (int64_t k = k0 + threadIdx.x; k < k1; k += blockDim.x) {
    const int64_t i = (int64_t)row * M + k;
    C a = v[i];
    a.x /= beta;
    a.y /= beta;
    v[i] = a;
  }
}

template <typename T>
void lanczos_typed(Toeplitz& t, const void* ws, double sigmasq, int mode, void* q0, int B, int steps, double breakdown,
                   double* alphas, double* betas, cudaStream_t st) {
  using C = typename cplx_of<T>::type;
  const int64_t M = t.M;
  LzScal s;
  s.B = B;
  s.NC = (int)std::min<int64_t>(kLzMaxChunks, std::max<int64_t>(1, (M + kLzChunkLen - 1) / kLzChunkLen));
  s.len = (M + s.NC - 1) / s.NC;
  C* vec = (C*)t.cg_vec.get((size_t)2 * B * M * sizeof(C));
  char* sc = (char*)t.cg_scal.get((size_t)2 * B * s.NC * sizeof(double) + (size_t)2 * B * sizeof(int));
  s.p_a = (double*)sc;
  s.p_b = s.p_a + (size_t)B * s.NC;
  s.alive = (int*)(s.p_b + (size_t)B * s.NC);
  lz_init_kernel<<<div_up(B, 256), 256, 0, st>>>(s.alive, B);
  // three rotating (B, M) buffers: q_prev, q, v.  The caller's start block is the first q (it is overwritten).
  C* qprev = vec;
  C* q = (C*)q0;
  C* v = vec + (int64_t)B * M;
  const dim3 grid(s.NC, B);
  for (int j = 0; j < steps; ++j) {
    toeplitz_apply(t, q, B, ws, sigmasq, mode, v, st);
    ScopedStage stg("lanczos.update", st, 3);
    lz_v_kernel<T><<<grid, 256, 0, st>>>(v, q, qprev, M, j, s);
    lz_alpha_kernel<T><<<grid, 256, 0, st>>>(v, q, M, j, steps, alphas, s);
    lz_beta_kernel<T><<<grid, 256, 0, st>>>(v, M, j, steps, breakdown, betas, s);
    C* old = qprev;
    qprev = q;
    q = v;
    v = old;
  }
  EFGP_CUDA(cudaGetLastError());
}

}  // namespace

void lanczos_tridiag(Toeplitz& t, const void* ws, double sigmasq, int mode, void* q0, int batch, int steps,
                     double breakdown, double* alphas, double* betas, cudaStream_t st) {
  EFGP_CHECK(batch >= 1 && steps >= 1, "batch and steps must be positive");
  EFGP_CHECK(batch <= 65535, "at most 65535 probes per call");
  EFGP_CHECK(mode >= kModeT && mode <= kModeAvar, "bad operator mode");
  EFGP_CHECK(mode == kModeT || ws != nullptr, "ws is required for the weighted operators");
  if (t.dtype == 0)
    lanczos_typed<float>(t, ws, sigmasq, mode, q0, batch, steps, breakdown, alphas, betas, st);
  else
    lanczos_typed<double>(t, ws, sigmasq, mode, q0, batch, steps, breakdown, alphas, betas, st);
}

}  // namespace efgp
