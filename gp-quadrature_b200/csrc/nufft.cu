// NUFFT plans: type-1 (nonuniform -> modes) and type-2 (modes -> nonuniform) built from the
// spread / interp kernels, cuFFT real<->complex transforms on the fine grid, and fused
// deconvolve+crop(+spectral-weight) / pre-correct+pad(+spectral-weight) kernels.
//
// Replaces pff.finufft_type1 / finufft_type2 as used by efgpnd.py:1496-1499 (isign=-1,
// centred modes), :1533-1549 (isign=+1) and :1679 (modeord=True).  Mode conventions follow
// those call sites: per axis k = -(n//2) .. (n-1)//2, C-order, or FFT order when modeord != 0.
//
// All spreading is done with REAL strengths on a REAL fine grid (complex strengths are two real
// transforms; complex type-2 outputs are the transforms of the Hermitian and anti-Hermitian parts
// of the mode array), so every FFT is R2C/C2R on half the data of a complex transform.
#include "nufft.cuh"
#include "philox.cuh"
#include <mutex>
#include <map>
#include <utility>
#include <cmath>
#include <algorithm>
#include <cstdlib>

namespace efgp {

// ------------------------------------------------------------------ cuFFT plan cache
cufftHandle FftCache::get(int dim, const int64_t* n, cufftType type, int batch, cudaStream_t st) {
  FftPlanKey key;
  key.dim = dim;
  key.type = (int)type;
  key.batch = batch;
  for (int a = 0; a < kMaxDim; ++a) key.n[a] = a < dim ? n[a] : 1;
  auto it = plans.find(key);
  cufftHandle h;
  if (it == plans.end()) {
    if (plans.size() > 24) {  // bound the cache
      for (auto& kv : plans) cufftDestroy(kv.second);
      plans.clear();
    }
    EFGP_CUFFT(cufftCreate(&h));
    EFGP_CUFFT(cufftSetAutoAllocation(h, 0));
    long long nn[kMaxDim];
    for (int a = 0; a < dim; ++a) nn[a] = n[a];
    size_t ws = 0;
    EFGP_CUFFT(cufftMakePlanMany64(h, dim, nn, nullptr, 1, 0, nullptr, 1, 0, type, batch, &ws));
    plans[key] = h;
    work_bytes = std::max(work_bytes, ws);
  } else {
    h = it->second;
  }
  size_t ws = 0;
  EFGP_CUFFT(cufftGetSize(h, &ws));
  EFGP_CUFFT(cufftSetWorkArea(h, work.get(std::max<size_t>(ws, 256))));
  EFGP_CUFFT(cufftSetStream(h, st));
  return h;
}
FftCache::~FftCache() {
  // process teardown: the CUDA context may already be gone, leave the plans to the driver
}
FftCache& global_fft() {
  static FftCache* g = new FftCache();   // intentionally leaked (see ~FftCache)
  return *g;
}
static std::map<std::string, std::unique_ptr<DevBuf>>& global_dec_tables() {
  static auto* g = new std::map<std::string, std::unique_ptr<DevBuf>>();
  return *g;
}

// ------------------------------------------------------------------ deterministic (fixed-point) spreading
static int g_deterministic = -1;
void set_deterministic(int on) { g_deterministic = on ? 1 : 0; }
bool deterministic() {
  if (g_deterministic < 0) {
    const char* e = std::getenv("EFGP_DETERMINISTIC");
    g_deterministic = (e && std::atoi(e) != 0) ? 1 : 0;
  }
  return g_deterministic == 1;
}

template <typename T>
__global__ void __launch_bounds__(256)
absmax_kernel(const T* __restrict__ c, int64_t n, int64_t stride, unsigned long long* __restrict__ out) {
  double m = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double v = fabs((double)c[i * stride]);
    m = v > m ? v : m;
  }
  for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) atomicMax(out, (unsigned long long)__double_as_longlong(m));   // order-preserving for m >= 0
}

// 2^k with (sum of |strengths|, bounded by N * cmax) * 2^k < 2^62: the integer grid cannot overflow
static double det_scale_for(double n_times_cmax) {
  if (!(n_times_cmax > 0)) n_times_cmax = 1.0;
  int e = 0;
  std::frexp(n_times_cmax, &e);            // n_times_cmax < 2^e
  return std::ldexp(1.0, 62 - e);
}

// largest |strength| of (n, stride) values on the device (synchronises: deterministic mode is a parity / debug mode)
static double device_absmax(const void* c, int64_t n, int64_t stride, int dtype, cudaStream_t st) {
  unsigned long long* d = (unsigned long long*)pool_alloc(sizeof(unsigned long long));
  EFGP_CUDA(cudaMemsetAsync(d, 0, sizeof(unsigned long long), st));
  const int blocks = (int)std::min<int64_t>(1024, (n + 255) / 256);
  if (dtype == 0)
    absmax_kernel<float><<<blocks, 256, 0, st>>>((const float*)c, n, stride, d);
  else
    absmax_kernel<double><<<blocks, 256, 0, st>>>((const double*)c, n, stride, d);
  unsigned long long h = 0;
  EFGP_CUDA(cudaMemcpyAsync(&h, d, sizeof(h), cudaMemcpyDeviceToHost, st));
  EFGP_CUDA(cudaStreamSynchronize(st));
  pool_free(d);
  double v;
  std::memcpy(&v, &h, sizeof(v));
  return v;
}

template <typename T>
__global__ void __launch_bounds__(256)
det_convert_kernel(const unsigned long long* __restrict__ ig, int64_t n, double inv_scale, T* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = (T)((double)(long long)ig[i] * inv_scale);      // (in place when T = double: same element size)
}

// ------------------------------------------------------------------ mode bookkeeping
struct ModeGeom {
  int dim;
  int n[kMaxDim];       // modes per axis
  int nf[kMaxDim];      // fine grid per axis
  int nh;               // nf[last]/2+1
  int modeord;
  int flip;             // 1: look up / place at -k
  int64_t M;            // prod n
};

__device__ __forceinline__ int signed_mode(int idx, int n, int modeord) {
  if (modeord) return idx <= (n - 1) / 2 ? idx : idx - n;
  return idx - n / 2;
}
__device__ __forceinline__ int mode_index(int k, int n, int modeord) {  // inverse of signed_mode, -1 if absent
  if (k < -(n / 2) || k > (n - 1) / 2) return -1;
  if (modeord) return k >= 0 ? k : k + n;
  return k + n / 2;
}

// out[b, idx] = (spec_re[b] (+ i spec_im[b]))[k mod nf] / prod_a psihat_a(k_a) * scale[idx]
template <typename T>
__global__ void __launch_bounds__(256)
deconv_kernel(const typename cplx_of<T>::type* __restrict__ spec, int ncomp, int batch, ModeGeom g,
              const T* __restrict__ dinv0, const T* __restrict__ dinv1, const T* __restrict__ dinv2,
              const T* __restrict__ scale, typename cplx_of<T>::type* __restrict__ out) {
  using C = typename cplx_of<T>::type;
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= g.M * batch) return;
  int64_t idx = t % g.M;
  int b = (int)(t / g.M);
  int id[kMaxDim] = {0, 0, 0};
  {
    int64_t r = idx;
    for (int a = g.dim - 1; a >= 0; --a) {
      id[a] = (int)(r % g.n[a]);
      r /= g.n[a];
    }
  }
  const T* dinv[kMaxDim] = {dinv0, dinv1, dinv2};
  int q[kMaxDim] = {0, 0, 0};
  T f = T(1);
  for (int a = 0; a < g.dim; ++a) {
    int k = signed_mode(id[a], g.n[a], g.modeord);
    f *= dinv[a][k < 0 ? -k : k];
    if (g.flip) k = -k;
    q[a] = k < 0 ? k + g.nf[a] : k;
  }
  if (scale) f *= scale[idx];
  const int last = g.dim - 1;
  bool conj = q[last] >= g.nh;
  if (conj)
    for (int a = 0; a < g.dim; ++a) q[a] = q[a] == 0 ? 0 : g.nf[a] - q[a];
  int64_t off = 0;
  int64_t slab = 1;
  for (int a = 0; a < g.dim; ++a) {
    int ext = a == last ? g.nh : g.nf[a];
    off = off * ext + q[a];
    slab *= ext;
  }
  C v = spec[(int64_t)(b * ncomp) * slab + off];
  T re = v.x, im = conj ? -v.y : v.y;
  if (ncomp == 2) {
    C w = spec[(int64_t)(b * ncomp + 1) * slab + off];
    T wr = w.x, wi = conj ? -w.y : w.y;
    re -= wi;  // (vr + i vi) + i (wr + i wi)
    im += wr;
  }
  C o;
  o.x = re * f;
  o.y = im * f;
  out[t] = o;
}

// Half-spectrum of the fine grid for a real-output type-2: comp 0 -> Hermitian part of the scaled,
// pre-corrected modes; comp 1 -> (f[k] - conj f[-k]) / (2i).  Writes every entry (zero padding).
template <typename T>
__global__ void __launch_bounds__(256)
precorrect_kernel(const typename cplx_of<T>::type* __restrict__ modes, int ncomp, int batch, ModeGeom g,
                  const T* __restrict__ dinv0, const T* __restrict__ dinv1, const T* __restrict__ dinv2,
                  const T* __restrict__ scale, typename cplx_of<T>::type* __restrict__ spec) {
  using C = typename cplx_of<T>::type;
  int64_t slab = 1;
  const int last = g.dim - 1;
  for (int a = 0; a < g.dim; ++a) slab *= (a == last ? g.nh : g.nf[a]);
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= slab * batch * ncomp) return;
  int64_t off = t % slab;
  int bc = (int)(t / slab);
  int b = bc / ncomp, comp = bc % ncomp;
  int q[kMaxDim] = {0, 0, 0};
  {
    int64_t r = off;
    for (int a = g.dim - 1; a >= 0; --a) {
      int ext = a == last ? g.nh : g.nf[a];
      q[a] = (int)(r % ext);
      r /= ext;
    }
  }
  const T* dinv[kMaxDim] = {dinv0, dinv1, dinv2};
  T f = T(1);
  int64_t ip = 0, im = 0;  // flat indices of +k and -k in the mode array
  bool has_p = true, has_m = true;
  bool in_range = true;
  for (int a = 0; a < g.dim; ++a) {
    int k = q[a] <= g.nf[a] / 2 ? q[a] : q[a] - g.nf[a];
    if (2 * q[a] == g.nf[a] && g.n[a] > 0) {
      // Nyquist bin of the fine grid: only reachable if n == nf, which never happens (nf >= 2n)
    }
    if (g.flip) k = -k;
    int ka = k < 0 ? -k : k;
    if (ka > g.n[a] / 2) {
      in_range = false;
      break;
    }
    f *= dinv[a][ka];
    int jp = mode_index(k, g.n[a], g.modeord);
    int jm = mode_index(-k, g.n[a], g.modeord);
    if (jp < 0) has_p = false;
    if (jm < 0) has_m = false;
    ip = ip * g.n[a] + (jp < 0 ? 0 : jp);
    im = im * g.n[a] + (jm < 0 ? 0 : jm);
  }
  C o;
  o.x = 0;
  o.y = 0;
  if (in_range && (has_p || has_m)) {
    T pr = 0, pi = 0, mr = 0, mi = 0;
    if (has_p) {
      C v = modes[(int64_t)b * g.M + ip];
      T s = scale ? scale[ip] : T(1);
      pr = v.x * s;
      pi = v.y * s;
    }
    if (has_m) {
      C v = modes[(int64_t)b * g.M + im];
      T s = scale ? scale[im] : T(1);
      mr = v.x * s;
      mi = v.y * s;
    }
    if (comp == 0) {  // (p + conj m)/2
      o.x = T(0.5) * (pr + mr) * f;
      o.y = T(0.5) * (pi - mi) * f;
    } else {  // (p - conj m)/(2i) = ((pi + mi) - i (pr - mr))/2
      o.x = T(0.5) * (pi + mi) * f;
      o.y = -T(0.5) * (pr - mr) * f;
    }
  }
  spec[t] = o;
}

// ------------------------------------------------------------------ plan selection
struct GridChoice {
  int64_t nf[kMaxDim];
  ESKernel ker;
  SortedPoints* sp;
};

static void upload_dinv(Nufft& nu, const ESKernel& ker, const int64_t* nf, const int64_t* n, int dim, int dtype,
                        void* dptr[kMaxDim], cudaStream_t st) {
  // dinv_a[k] = 1/psihat(k/nf_a), k = 0..n_a/2 ; all axes packed into nu.dec
  size_t esz = dtype == 0 ? 4 : 8;
  size_t total = 0;
  for (int a = 0; a < dim; ++a) total += (size_t)(n[a] / 2 + 1);
  // tables are cached per geometry: repeated transforms (probes, CG) skip the host work and the sync
  char keybuf[160];
  snprintf(keybuf, sizeof(keybuf), "%d|%d|%.17g|%lld,%lld,%lld|%lld,%lld,%lld", dtype, ker.W, ker.beta, (long long)nf[0],
           (long long)(dim > 1 ? nf[1] : 0), (long long)(dim > 2 ? nf[2] : 0), (long long)n[0],
           (long long)(dim > 1 ? n[1] : 0), (long long)(dim > 2 ? n[2] : 0));
  std::string key(keybuf);
  auto& tables = global_dec_tables();
  (void)nu;
  auto hit = tables.find(key);
  if (hit != tables.end()) {
    char* b0 = (char*)hit->second->p;
    size_t pos0 = 0;
    for (int a = 0; a < dim; ++a) {
      dptr[a] = b0 + pos0 * esz;
      pos0 += (size_t)(n[a] / 2 + 1);
    }
    for (int a = dim; a < kMaxDim; ++a) dptr[a] = b0;
    return;
  }
  if (tables.size() > 64) tables.clear();
  std::unique_ptr<DevBuf> holder(new DevBuf());
  char* base = (char*)holder->get(total * esz);
  tables[key] = std::move(holder);
  std::vector<double> host;
  std::vector<float> hostf;
  size_t pos = 0;
  std::vector<double> all(total);
  for (int a = 0; a < dim; ++a) {
    es_fourier_series(ker, nf[a], n[a] / 2, host);
    for (size_t k = 0; k < host.size(); ++k) all[pos + k] = 1.0 / host[k];
    dptr[a] = base + pos * esz;
    pos += host.size();
  }
  for (int a = dim; a < kMaxDim; ++a) dptr[a] = base;
  if (dtype == 0) {
    hostf.assign(all.begin(), all.end());
    EFGP_CUDA(cudaMemcpyAsync(base, hostf.data(), total * 4, cudaMemcpyHostToDevice, st));
  } else {
    EFGP_CUDA(cudaMemcpyAsync(base, all.data(), total * 8, cudaMemcpyHostToDevice, st));
  }
  // (pageable-memory cudaMemcpyAsync stages the host buffer before returning: no sync needed)
}

static SortedPoints* get_sorted(Nufft& nu, const int64_t* nf, int odd, cudaStream_t st) {
  for (auto& sp : nu.sorted) {
    bool same = sp->odd == odd;
    for (int a = 0; a < nu.dim; ++a) same = same && sp->nf[a] == nf[a];
    if (same) {
      sp->last_use = ++nu.clock;
      return sp.get();
    }
  }
  if (nu.sorted.size() >= 3) {  // evict least recently used
    size_t j = 0;
    for (size_t i = 1; i < nu.sorted.size(); ++i)
      if (nu.sorted[i]->last_use < nu.sorted[j]->last_use) j = i;
    nu.sorted.erase(nu.sorted.begin() + j);
  }
  std::unique_ptr<SortedPoints> sp(new SortedPoints());
  static uint64_t next_id = 0;
  sp->id = ++next_id;
  build_sorted_points(*sp, nu.x, nu.pending_payload, nu.N, nu.dim, nu.dtype, nu.h, nu.xcen, nf, odd, st, nu.carry_index);
  if (sp->has_payload) nu.fresh_payload_sp = sp.get();
  sp->last_use = ++nu.clock;
  nu.sorted.push_back(std::move(sp));
  return nu.sorted.back().get();
}

// Pick the fine grid, kernel and sorted point set for `n` modes at tolerance eps.  An existing
// sorted set on a finer grid is reused when possible (larger effective upsampling -> narrower
// kernel, and no second sort): this is how the F*y transform shares the Toeplitz-vector grid.
// The polynomial fit of the ES kernel depends on (W, sigma) only and costs tens of microseconds of host time
// right in front of the first kernel launch of every transform: keep the fitted tables for the process.
static void es_build_cached(ESKernel& k, int W, double sigma) {
  static auto* cache = new std::map<std::pair<int, double>, ESKernel>();
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  auto key = std::make_pair(W, sigma);
  auto it = cache->find(key);
  if (it == cache->end()) {
    ESKernel fresh;
    es_build(fresh, W, sigma, 0);
    if (cache->size() > 256) cache->clear();
    it = cache->emplace(key, fresh).first;
  }
  k = it->second;
}

// Modelled cost of one transform in units of one FP64 FMA on the points' side: per point W^d footprint FMAs (the 3-D
// shared-memory tile kernels run ~3x below the FP64 rate of the 2-D register kernels) plus d even/odd Horner
// evaluations; per fine-grid cell the memset, the flushes, the cuFFT passes and the deconvolution (~200 FMA-times of
// HBM traffic, profiles/ncu_r01_traffic.json).  Only used to rank candidate upsampling factors.
static double transform_cost(int dim, double npts, int W, double cells) {
  const double foot = std::pow((double)W, dim) * (dim == 3 ? 3.0 : 1.0);
  const double horner = dim * 0.5 * W * (es_degree(W) + 3);
  return npts * (foot + horner) + 200.0 * cells;
}

// Point count the grid-choice cost model sees.  Point-sharded fits set Nufft::cost_npts to the GLOBAL count: every rank
// (and a one-rank run on all the points) then picks the same fine grid and kernel, so the all-reduced sum of the shard
// transforms is the one-rank transform to rounding, not merely to the NUFFT tolerance.
static double cost_points(const Nufft& nu) { return (double)(nu.cost_npts > 0 ? nu.cost_npts : nu.N); }

// Fine grid of a transform that builds its OWN sorted set: EFGP grids are tiny next to N, so a larger upsampling
// factor (narrower kernel: fewer FMAs per point) is nearly free for SMALL mode counts.  For the Toeplitz generator of
// the headline solve (141^2 modes, eps 6e-8, 1e8 points) it is not: W = 7 on a 720^2 grid saves 0.36 ms of spreading
// and costs 0.72 ms of level-2 sort (profiles/ab_sigma_r02.md), which is what the sort term below encodes; the model
// keeps sigma = 2 / W = 9 there.  EFGP_SIGMA pins the factor (A/B measurements).
static double pick_sigma(const Nufft& nu, const int64_t* n, double eps, int64_t* nf_out, int* w_out) {
  static const double sigma_env = [] {
    const char* e = std::getenv("EFGP_SIGMA");
    double v = e ? std::atof(e) : 0.0;
    return v >= 1.25 ? v : 0.0;
  }();
  // The choice is a pure function of (dim, dtype, mode counts, eps, point count): remembered, because this runs on the
  // host between the bounding-box read-back and the first sort kernel of every fit (~15 us of exp / pow calls).
  struct Memo { int dim, dtype; int64_t n[kMaxDim]; double eps, npts; int64_t nf[kMaxDim]; int w; double sigma; };
  static std::mutex memo_mu;
  static std::vector<Memo> memo;
  {
    std::lock_guard<std::mutex> lk(memo_mu);
    for (const Memo& m : memo) {
      bool same = m.dim == nu.dim && m.dtype == nu.dtype && m.eps == eps && m.npts == cost_points(nu);
      for (int a = 0; a < nu.dim; ++a) same = same && m.n[a] == n[a];
      if (same) {
        for (int a = 0; a < kMaxDim; ++a) nf_out[a] = m.nf[a];
        *w_out = m.w;
        return m.sigma;
      }
    }
  }
  const double cands[] = {2.0, 2.5, 3.0, 4.0, 5.0, 6.0, 8.0};
  double best_cost = 0, best_sigma = 0;
  for (double sg : cands) {
    if (sigma_env > 0) sg = sigma_env;
    int64_t nf[kMaxDim] = {1, 1, 1};
    // the width depends on the achieved factor, which depends on the rounded-up grid: two passes settle it
    int w = es_width_for(eps, sg, kMaxW);
    double cells = 1, sig_eff = 1e30;
    for (int pass = 0; pass < 2; ++pass) {
      cells = 1;
      sig_eff = 1e30;
      for (int a = 0; a < nu.dim; ++a) {
        nf[a] = next_smooth_even(std::max<int64_t>((int64_t)std::ceil(sg * (double)n[a]), 2 * w));
        cells *= (double)nf[a];
        sig_eff = std::min(sig_eff, (double)nf[a] / (double)n[a]);
      }
      w = es_width_for(eps, sig_eff, kMaxW);
    }
    const size_t esz = nu.dtype == 0 ? 4 : 8;
    const bool fits = cells < 3.9e9 && cells * esz <= (double)((size_t)2 << 30);
    if (!fits && best_sigma > 0) break;
    double cost = transform_cost(nu.dim, cost_points(nu), w, cells);
    if (nu.dim <= 2) {
      // this grid also keys the cell sort: the level-2 counting sort pays per (2048-point tile, bin) pair, measured at
      // N = 1e8 as 1.2 ms + 1.35e-6 ms per cell (288^2 .. 864^2 cells) = 1.6e-4 FMA-times per point and cell
      cost += 1.6e-4 * cells * cost_points(nu);
      // an even width fixes the parity of every later transform on these sorted points: the narrow F*y / type-2
      // kernels (odd W = 5 at the reference's default eps) would be forced one step wider
      if (nu.dim == 2 && (w % 2) == 0) cost += 40.0 * cost_points(nu);
    }
    if (best_sigma == 0 || cost < best_cost) {
      best_cost = cost;
      best_sigma = sg;
      for (int a = 0; a < kMaxDim; ++a) nf_out[a] = a < nu.dim ? nf[a] : 1;
      *w_out = w;
    }
    if (sigma_env > 0) break;
  }
  {
    std::lock_guard<std::mutex> lk(memo_mu);
    if (memo.size() >= 64) memo.erase(memo.begin());
    Memo m;
    m.dim = nu.dim;
    m.dtype = nu.dtype;
    for (int a = 0; a < kMaxDim; ++a) {
      m.n[a] = a < nu.dim ? n[a] : 1;
      m.nf[a] = nf_out[a];
    }
    m.eps = eps;
    m.npts = cost_points(nu);
    m.w = *w_out;
    m.sigma = best_sigma;
    memo.push_back(m);
  }
  return best_sigma;
}

static GridChoice choose_grid(Nufft& nu, const int64_t* n, double eps, cudaStream_t st) {
  GridChoice gc;
  const int max_w = kMaxW;
  if (nu.dtype == 0 && eps < 1e-7) eps = 1e-7;  // single precision floor
  int64_t own[kMaxDim] = {1, 1, 1};
  int w_own = 2;
  pick_sigma(nu, n, eps, own, &w_own);
  double own_cells = 1;
  for (int a = 0; a < nu.dim; ++a) own_cells *= (double)own[a];
  const double own_cost = transform_cost(nu.dim, cost_points(nu), w_own, own_cells);
  // An existing sorted set on a grid at least 2x the mode count can serve this transform without another sort
  // (larger effective upsampling -> narrower kernel): this is how F*y and the type-2 share the generator's grid.
  SortedPoints* best = nullptr;
  int best_w = 0;
  double best_cost = 0;
  for (auto& sp : nu.sorted) {
    bool ok = true;
    double cells = 1, sigma = 1e30;
    for (int a = 0; a < nu.dim; ++a) {
      ok = ok && sp->nf[a] >= 2 * n[a];
      cells *= (double)sp->nf[a];
      sigma = std::min(sigma, (double)sp->nf[a] / (double)n[a]);
    }
    if (!ok) continue;
    int w = es_width_for(eps, sigma, max_w);
    if ((w & 1) != sp->odd) ++w;       // the half-cell shift of the stored coordinates fixes the parity
    if (w > max_w) continue;
    for (int a = 0; a < nu.dim; ++a) ok = ok && sp->nf[a] >= 2 * w;
    if (!ok) continue;
    // a second sort costs about as much as three narrow transforms and holds another copy of the points
    const double cost = transform_cost(nu.dim, cost_points(nu), w, cells);
    if (cost > 4.0 * own_cost + 100.0 * cost_points(nu)) continue;
    if (!best || cost < best_cost) {
      best = sp.get();
      best_w = w;
      best_cost = cost;
    }
  }
  if (best) {
    double sigma = 1e30;
    for (int a = 0; a < nu.dim; ++a) sigma = std::min(sigma, (double)best->nf[a] / (double)n[a]);
    for (int a = 0; a < kMaxDim; ++a) gc.nf[a] = best->nf[a];
    es_build_cached(gc.ker, best_w, sigma);
    best->last_use = ++nu.clock;
    gc.sp = best;
    return gc;
  }
  double sigma = 1e30;
  for (int a = 0; a < nu.dim; ++a) sigma = std::min(sigma, (double)own[a] / (double)n[a]);
  for (int a = 0; a < kMaxDim; ++a) gc.nf[a] = own[a];
  es_build_cached(gc.ker, w_own, sigma);
  gc.sp = get_sorted(nu, gc.nf, w_own & 1, st);
  return gc;
}

static ModeGeom make_geom(int dim, const int64_t* n, const int64_t* nf, int modeord, int flip) {
  ModeGeom g;
  g.dim = dim;
  g.M = 1;
  for (int a = 0; a < kMaxDim; ++a) {
    g.n[a] = a < dim ? (int)n[a] : 1;
    g.nf[a] = a < dim ? (int)nf[a] : 1;
    if (a < dim) g.M *= n[a];
  }
  g.nh = g.nf[dim - 1] / 2 + 1;
  g.modeord = modeord;
  g.flip = flip;
  return g;
}

template <typename T>
static void type1_t(Nufft& nu, const void* strengths, int is_complex, int batch, const int64_t* n, double eps,
                    int isign, int modeord, const void* scale, void* out, cudaStream_t st, bool payload_ok = false) {
  using C = typename cplx_of<T>::type;
  const int dim = nu.dim;
  GridChoice gc = choose_grid(nu, n, eps, st);
  ModeGeom g = make_geom(dim, n, gc.nf, modeord, isign > 0 ? 1 : 0);
  void* dinv[kMaxDim];
  upload_dinv(nu, gc.ker, gc.nf, n, dim, nu.dtype, dinv, st);
  int64_t cells = 1, slab = 1;
  for (int a = 0; a < dim; ++a) {
    cells *= gc.nf[a];
    slab *= (a == dim - 1 ? gc.nf[a] / 2 + 1 : gc.nf[a]);
  }
  const int ncomp = is_complex ? 2 : 1;
  // chunk the batch so the fine grids stay under ~8 GB
  int64_t per = (int64_t)ncomp * (cells * sizeof(T) + slab * sizeof(C));
  int chunk = (int)std::max<int64_t>(1, std::min<int64_t>(batch, (8ll << 30) / std::max<int64_t>(per, 1)));
  T* fine = (T*)nu.fine.get((size_t)chunk * ncomp * cells * sizeof(T));
  C* spec = (C*)nu.spec.get((size_t)chunk * ncomp * slab * sizeof(C));
  for (int b0 = 0; b0 < batch; b0 += chunk) {
    int nb = std::min(chunk, batch - b0);
    // deterministic mode: integer accumulation (for T = double the integer grid IS the fine-grid buffer)
    const bool det = deterministic() && nu.N > 0;
    unsigned long long* ig = nullptr;
    double iscale = 0;
    if (det) {
      double cmax = 1.0;
      if (strengths) cmax = device_absmax((const T*)strengths + (int64_t)b0 * nu.N * ncomp, (int64_t)nb * nu.N * ncomp, 1, nu.dtype, st);
      iscale = det_scale_for((double)nu.N * cmax);
      ig = sizeof(T) == 8 ? (unsigned long long*)fine
                          : (unsigned long long*)nu.tmp.get((size_t)nb * ncomp * cells * sizeof(unsigned long long));
      EFGP_CUDA(cudaMemsetAsync(ig, 0, (size_t)nb * ncomp * cells * sizeof(unsigned long long), st));
    } else {
      ScopedStage stg("type1.memset_fine", st, 0);
      EFGP_CUDA(cudaMemsetAsync(fine, 0, (size_t)nb * ncomp * cells * sizeof(T), st));
    }
    // strengths already ride in the records of a set sorted during this very call (nufft_type1_pair)
    const bool from_rec = payload_ok && strengths && !is_complex && batch == 1 && gc.sp == nu.fresh_payload_sp;
    T* cs = (strengths && !from_rec) ? (T*)nu.cperm.get((size_t)gc.sp->npos * sizeof(T)) : nullptr;
    for (int b = 0; b < nb; ++b) {
      for (int comp = 0; comp < ncomp; ++comp) {
        const int64_t goff = (int64_t)(b * ncomp + comp) * cells;
        if (from_rec) {
          launch_spread(*gc.sp, gc.ker, nullptr, fine + goff, st, true, ig ? ig + goff : nullptr, iscale);
          continue;
        }
        if (strengths) {
          const T* c = (const T*)strengths + (int64_t)(b0 + b) * nu.N * ncomp + comp;
          permute_strengths(*gc.sp, c, ncomp, cs, st);
        }
        launch_spread(*gc.sp, gc.ker, cs, fine + goff, st, false, ig ? ig + goff : nullptr, iscale);
      }
    }
    if (det) {
      const int64_t tot = (int64_t)nb * ncomp * cells;
      det_convert_kernel<T><<<div_up(tot, 256), 256, 0, st>>>(ig, tot, 1.0 / iscale, fine);
      EFGP_CUDA(cudaGetLastError());
    }
    cufftHandle plan = global_fft().get(dim, gc.nf, sizeof(T) == 4 ? CUFFT_R2C : CUFFT_D2Z, nb * ncomp, st);
    {
      ScopedStage stg("type1.cufft_r2c", st, 0);
      if (sizeof(T) == 4)
        EFGP_CUFFT(cufftExecR2C(plan, (cufftReal*)fine, (cufftComplex*)spec));
      else
        EFGP_CUFFT(cufftExecD2Z(plan, (cufftDoubleReal*)fine, (cufftDoubleComplex*)spec));
    }
    int64_t tot = g.M * nb;
    ScopedStage stg("type1.deconv_kernel", st);
    deconv_kernel<T><<<div_up(tot, 256), 256, 0, st>>>(spec, ncomp, nb, g, (const T*)dinv[0], (const T*)dinv[1],
                                                       (const T*)dinv[2], (const T*)scale,
                                                       (C*)out + (int64_t)b0 * g.M);
    EFGP_CUDA(cudaGetLastError());
  }
}

template <typename T>
static void type2_t(Nufft& nu, const void* modes, int batch, const int64_t* n, double eps, int isign,
                    int modeord, const void* scale, int real_out, void* out, cudaStream_t st) {
  using C = typename cplx_of<T>::type;
  const int dim = nu.dim;
  GridChoice gc = choose_grid(nu, n, eps, st);
  // cuFFT's inverse is e^{+i}: isign > 0 places mode k at +k, isign < 0 at -k
  ModeGeom g = make_geom(dim, n, gc.nf, modeord, isign > 0 ? 0 : 1);
  void* dinv[kMaxDim];
  upload_dinv(nu, gc.ker, gc.nf, n, dim, nu.dtype, dinv, st);
  int64_t cells = 1, slab = 1;
  for (int a = 0; a < dim; ++a) {
    cells *= gc.nf[a];
    slab *= (a == dim - 1 ? gc.nf[a] / 2 + 1 : gc.nf[a]);
  }
  const int ncomp = real_out ? 1 : 2;
  int64_t per = (int64_t)ncomp * (cells * sizeof(T) + slab * sizeof(C));
  int chunk = (int)std::max<int64_t>(1, std::min<int64_t>(batch, (8ll << 30) / std::max<int64_t>(per, 1)));
  T* fine = (T*)nu.fine.get((size_t)chunk * ncomp * cells * sizeof(T));
  C* spec = (C*)nu.spec.get((size_t)chunk * ncomp * slab * sizeof(C));
  for (int b0 = 0; b0 < batch; b0 += chunk) {
    int nb = std::min(chunk, batch - b0);
    int64_t tot = slab * nb * ncomp;
    {
    ScopedStage stg("type2.precorrect_kernel", st);
    precorrect_kernel<T><<<div_up(tot, 256), 256, 0, st>>>((const C*)modes + (int64_t)b0 * g.M, ncomp, nb, g,
                                                           (const T*)dinv[0], (const T*)dinv[1],
                                                           (const T*)dinv[2], (const T*)scale, spec);
    }
    EFGP_CUDA(cudaGetLastError());
    cufftHandle plan = global_fft().get(dim, gc.nf, sizeof(T) == 4 ? CUFFT_C2R : CUFFT_Z2D, nb * ncomp, st);
    {
      ScopedStage stg("type2.cufft_c2r", st, 0);
      if (sizeof(T) == 4)
        EFGP_CUFFT(cufftExecC2R(plan, (cufftComplex*)spec, (cufftReal*)fine));
      else
        EFGP_CUFFT(cufftExecZ2D(plan, (cufftDoubleComplex*)spec, (cufftDoubleReal*)fine));
    }
    for (int b = 0; b < nb; ++b)
      for (int comp = 0; comp < ncomp; ++comp)
        launch_interp(*gc.sp, gc.ker, fine + (int64_t)(b * ncomp + comp) * cells,
                      (T*)out + (int64_t)(b0 + b) * nu.N * ncomp + comp, ncomp, st);
  }
}

void nufft_type1(Nufft& nu, const void* strengths, int is_complex, int batch, const int64_t* n, double eps,
                 int isign, int modeord, const void* scale, void* out, cudaStream_t st) {
  EFGP_CHECK(batch >= 1, "batch must be >= 1");
  for (int a = 0; a < nu.dim; ++a) EFGP_CHECK(n[a] >= 1, "mode counts must be >= 1");
  if (nu.dtype == 0)
    type1_t<float>(nu, strengths, is_complex, batch, n, eps, isign, modeord, scale, out, st);
  else
    type1_t<double>(nu, strengths, is_complex, batch, n, eps, isign, modeord, scale, out, st);
}

// Generator and right-hand side in ONE pass over the records (spread.cu spread2d_moment_pair_kernel), one batched
// R2C transform of the two fine grids and two deconvolutions.  Returns false (nothing enqueued besides, possibly, the
// sort both paths need) when the geometry is not the dense 2-D fp64 case the fused kernel covers.
static bool type1_pair_fused(Nufft& nu, const int64_t* shape_gen, double eps_gen, const int64_t* shape_rhs,
                             double eps_rhs, const void* scale, void* out_gen, void* out_rhs, cudaStream_t st) {
  using T = double;
  using C = double2;
  if (nu.dtype != 1 || nu.dim != 2 || nu.N == 0 || deterministic()) return false;
  GridChoice gg = choose_grid(nu, shape_gen, eps_gen, st);     // sorts with the payload (pending_payload is set)
  nu.pending_payload = nullptr;
  GridChoice gr = choose_grid(nu, shape_rhs, eps_rhs, st);
  if (gg.sp != gr.sp || gg.sp != nu.fresh_payload_sp || !spread_pair_supported(*gg.sp, gg.ker, gr.ker)) return false;
  const int64_t* nf = gg.nf;
  const int64_t cells = nf[0] * nf[1], slab = nf[0] * (nf[1] / 2 + 1);
  ModeGeom geg = make_geom(2, shape_gen, nf, 0, 0), ger = make_geom(2, shape_rhs, nf, 0, 0);
  void* dg[kMaxDim];
  void* dr[kMaxDim];
  upload_dinv(nu, gg.ker, nf, shape_gen, 2, 1, dg, st);
  upload_dinv(nu, gr.ker, nf, shape_rhs, 2, 1, dr, st);
  T* fine = (T*)nu.fine.get((size_t)2 * cells * sizeof(T));
  C* spec = (C*)nu.spec.get((size_t)2 * slab * sizeof(C));
  {
    ScopedStage stg("type1.memset_fine", st, 0);
    EFGP_CUDA(cudaMemsetAsync(fine, 0, (size_t)2 * cells * sizeof(T), st));
  }
  launch_spread_pair(*gg.sp, gg.ker, gr.ker, fine, fine + cells, st);
  cufftHandle plan = global_fft().get(2, nf, CUFFT_D2Z, 2, st);
  {
    ScopedStage stg("type1.cufft_r2c", st, 0);
    EFGP_CUFFT(cufftExecD2Z(plan, (cufftDoubleReal*)fine, (cufftDoubleComplex*)spec));
  }
  ScopedStage stg("type1.deconv_kernel", st, 2);
  deconv_kernel<T><<<div_up(geg.M, 256), 256, 0, st>>>(spec, 1, 1, geg, (const T*)dg[0], (const T*)dg[1], (const T*)dg[2],
                                                       (const T*)nullptr, (C*)out_gen);
  deconv_kernel<T><<<div_up(ger.M, 256), 256, 0, st>>>(spec + slab, 1, 1, ger, (const T*)dr[0], (const T*)dr[1],
                                                       (const T*)dr[2], (const T*)scale, (C*)out_rhs);
  EFGP_CUDA(cudaGetLastError());
  return true;
}

void nufft_type1_pair(Nufft& nu, const void* y, const int64_t* shape_gen, double eps_gen, const int64_t* shape_rhs,
                      double eps_rhs, const void* scale, void* out_gen, void* out_rhs, cudaStream_t st) {
  for (int a = 0; a < nu.dim; ++a) EFGP_CHECK(shape_gen[a] >= 1 && shape_rhs[a] >= 1, "mode counts must be >= 1");
  nu.pending_payload = nu.dim == 2 ? y : nullptr;
  nu.fresh_payload_sp = nullptr;
  try {
    if (type1_pair_fused(nu, shape_gen, eps_gen, shape_rhs, eps_rhs, scale, out_gen, out_rhs, st)) {
      nu.pending_payload = nullptr;
      nu.fresh_payload_sp = nullptr;
      return;
    }
    // (a declined fused attempt may already have built the payload-carrying sort: the two passes below find it)
    if (nu.dtype == 0)
      type1_t<float>(nu, nullptr, 0, 1, shape_gen, eps_gen, -1, 0, nullptr, out_gen, st);
    else
      type1_t<double>(nu, nullptr, 0, 1, shape_gen, eps_gen, -1, 0, nullptr, out_gen, st);
    nu.pending_payload = nullptr;
    if (nu.dtype == 0)
      type1_t<float>(nu, y, 0, 1, shape_rhs, eps_rhs, -1, 0, scale, out_rhs, st, true);
    else
      type1_t<double>(nu, y, 0, 1, shape_rhs, eps_rhs, -1, 0, scale, out_rhs, st, true);
  } catch (...) {
    nu.pending_payload = nullptr;
    nu.fresh_payload_sp = nullptr;
    throw;
  }
  nu.fresh_payload_sp = nullptr;
}

void nufft_type2(Nufft& nu, const void* modes, int batch, const int64_t* n, double eps, int isign, int modeord,
                 const void* scale, int real_out, void* out, cudaStream_t st) {
  EFGP_CHECK(batch >= 1, "batch must be >= 1");
  for (int a = 0; a < nu.dim; ++a) EFGP_CHECK(n[a] >= 1, "mode counts must be >= 1");
  if (nu.N == 0) return;
  if (nu.dtype == 0)
    type2_t<float>(nu, modes, batch, n, eps, isign, modeord, scale, real_out, out, st);
  else
    type2_t<double>(nu, modes, batch, n, eps, isign, modeord, scale, real_out, out, st);
}

// ------------------------------------------------------------------ streamed Hutchinson probes
static int probe_group(int t, int remaining, int nb_max) {
  int g = 1;
  while (g * 2 <= nb_max && g * 2 <= remaining) g *= 2;
  while (g > 1 && (t >> 7) != ((t + g - 1) >> 7)) g /= 2;
  return g;
}

static void check_sort_id(const SortedPoints& sp, uint64_t* sort_id) {
  if (!sort_id) return;
  if (sp.has_orig) {            // the probes are a function of the ORIGINAL index: any sorted set serves
    *sort_id = ~0ull;
    return;
  }
  if (*sort_id == 0)
    *sort_id = sp.id;
  else
    EFGP_CHECK(*sort_id == sp.id,
               "the sorted point set changed between two transforms of the same probes: the regenerated Rademacher "
               "signs are tied to the point order (re-run the type-1 stage)");
}

template <typename T>
static void type1_probes_t(Nufft& nu, uint64_t seed, int t0, int nb_total, const int64_t* n, double eps, const void* scale,
                           void* out, uint64_t* sort_id, cudaStream_t st) {
  using C = typename cplx_of<T>::type;
  const int dim = nu.dim;
  GridChoice gc = choose_grid(nu, n, eps, st);
  check_sort_id(*gc.sp, sort_id);
  ModeGeom g = make_geom(dim, n, gc.nf, 0, 0);
  void* dinv[kMaxDim];
  upload_dinv(nu, gc.ker, gc.nf, n, dim, nu.dtype, dinv, st);
  int64_t cells = 1, slab = 1;
  for (int a = 0; a < dim; ++a) {
    cells *= gc.nf[a];
    slab *= (a == dim - 1 ? gc.nf[a] / 2 + 1 : gc.nf[a]);
  }
  const int64_t per = cells * (int64_t)sizeof(T) + slab * (int64_t)sizeof(C);
  const int chunk = (int)std::max<int64_t>(1, std::min<int64_t>(nb_total, (8ll << 30) / std::max<int64_t>(per, 1)));
  T* fine = (T*)nu.fine.get((size_t)chunk * cells * sizeof(T));
  C* spec = (C*)nu.spec.get((size_t)chunk * slab * sizeof(C));
  const int nb_max = deterministic() ? 1 : spread_probe_batch(dim, gc.ker.W, nu.dtype);
  for (int b0 = 0; b0 < nb_total; b0 += chunk) {
    const int nb = std::min(chunk, nb_total - b0);
    const bool det = deterministic() && nu.N > 0;
    unsigned long long* ig = nullptr;
    double iscale = 0;
    if (det) {
      iscale = det_scale_for((double)nu.N);
      ig = sizeof(T) == 8 ? (unsigned long long*)fine
                          : (unsigned long long*)nu.tmp.get((size_t)nb * cells * sizeof(unsigned long long));
      EFGP_CUDA(cudaMemsetAsync(ig, 0, (size_t)nb * cells * sizeof(unsigned long long), st));
    } else {
      ScopedStage stg("type1.memset_fine", st, 0);
      EFGP_CUDA(cudaMemsetAsync(fine, 0, (size_t)nb * cells * sizeof(T), st));
    }
    for (int b = 0; b < nb;) {
      const int gsz = probe_group(t0 + b0 + b, nb - b, nb_max);
      launch_spread_probes(*gc.sp, gc.ker, fine + (int64_t)b * cells, ProbeSpec{seed, t0 + b0 + b, gsz, gc.sp->has_orig ? 1 : 0}, st,
                           ig ? ig + (int64_t)b * cells : nullptr, iscale);
      b += gsz;
    }
    if (det) {
      const int64_t tot = (int64_t)nb * cells;
      det_convert_kernel<T><<<div_up(tot, 256), 256, 0, st>>>(ig, tot, 1.0 / iscale, fine);
      EFGP_CUDA(cudaGetLastError());
    }
    cufftHandle plan = global_fft().get(dim, gc.nf, sizeof(T) == 4 ? CUFFT_R2C : CUFFT_D2Z, nb, st);
    {
      ScopedStage stg("type1.cufft_r2c", st, 0);
      if (sizeof(T) == 4)
        EFGP_CUFFT(cufftExecR2C(plan, (cufftReal*)fine, (cufftComplex*)spec));
      else
        EFGP_CUFFT(cufftExecD2Z(plan, (cufftDoubleReal*)fine, (cufftDoubleComplex*)spec));
    }
    const int64_t tot = g.M * nb;
    ScopedStage stg("type1.deconv_kernel", st);
    deconv_kernel<T><<<div_up(tot, 256), 256, 0, st>>>(spec, 1, nb, g, (const T*)dinv[0], (const T*)dinv[1],
                                                       (const T*)dinv[2], (const T*)scale, (C*)out + (int64_t)b0 * g.M);
    EFGP_CUDA(cudaGetLastError());
  }
}

template <typename T>
static void type2_probe_dots_t(Nufft& nu, uint64_t seed, int t0, int nb_total, const void* modes, const int64_t* n,
                               double eps, const void* scale, double* sums, uint64_t* sort_id, cudaStream_t st) {
  using C = typename cplx_of<T>::type;
  const int dim = nu.dim;
  GridChoice gc = choose_grid(nu, n, eps, st);
  check_sort_id(*gc.sp, sort_id);
  ModeGeom g = make_geom(dim, n, gc.nf, 0, 0);       // isign = +1, centred modes
  void* dinv[kMaxDim];
  upload_dinv(nu, gc.ker, gc.nf, n, dim, nu.dtype, dinv, st);
  int64_t cells = 1, slab = 1;
  for (int a = 0; a < dim; ++a) {
    cells *= gc.nf[a];
    slab *= (a == dim - 1 ? gc.nf[a] / 2 + 1 : gc.nf[a]);
  }
  const int64_t per = cells * (int64_t)sizeof(T) + slab * (int64_t)sizeof(C);
  const int chunk = (int)std::max<int64_t>(1, std::min<int64_t>(nb_total, (8ll << 30) / std::max<int64_t>(per, 1)));
  T* fine = (T*)nu.fine.get((size_t)chunk * cells * sizeof(T));
  C* spec = (C*)nu.spec.get((size_t)chunk * slab * sizeof(C));
  const int nb_max = spread_probe_batch(dim, gc.ker.W, nu.dtype);
  for (int b0 = 0; b0 < nb_total; b0 += chunk) {
    const int nb = std::min(chunk, nb_total - b0);
    const int64_t tot = slab * nb;
    {
      ScopedStage stg("type2.precorrect_kernel", st);
      precorrect_kernel<T><<<div_up(tot, 256), 256, 0, st>>>((const C*)modes + (int64_t)b0 * g.M, 1, nb, g, (const T*)dinv[0],
                                                             (const T*)dinv[1], (const T*)dinv[2], (const T*)scale, spec);
      EFGP_CUDA(cudaGetLastError());
    }
    cufftHandle plan = global_fft().get(dim, gc.nf, sizeof(T) == 4 ? CUFFT_C2R : CUFFT_Z2D, nb, st);
    {
      ScopedStage stg("type2.cufft_c2r", st, 0);
      if (sizeof(T) == 4)
        EFGP_CUFFT(cufftExecC2R(plan, (cufftComplex*)spec, (cufftReal*)fine));
      else
        EFGP_CUFFT(cufftExecZ2D(plan, (cufftDoubleComplex*)spec, (cufftDoubleReal*)fine));
    }
    for (int b = 0; b < nb;) {
      const int gsz = probe_group(t0 + b0 + b, nb - b, nb_max);
      launch_interp_probe_dots(*gc.sp, gc.ker, fine + (int64_t)b * cells, sums + b0 + b,
                               ProbeSpec{seed, t0 + b0 + b, gsz, gc.sp->has_orig ? 1 : 0}, st);
      b += gsz;
    }
  }
}

void nufft_type1_probes(Nufft& nu, uint64_t seed, int t0, int nb, const int64_t* n, double eps, const void* scale, void* out,
                        uint64_t* sort_id, cudaStream_t st) {
  EFGP_CHECK(nb >= 1 && t0 >= 0, "bad probe range");
  for (int a = 0; a < nu.dim; ++a) EFGP_CHECK(n[a] >= 1, "mode counts must be >= 1");
  if (nu.dtype == 0)
    type1_probes_t<float>(nu, seed, t0, nb, n, eps, scale, out, sort_id, st);
  else
    type1_probes_t<double>(nu, seed, t0, nb, n, eps, scale, out, sort_id, st);
}

void nufft_type2_probe_dots(Nufft& nu, uint64_t seed, int t0, int nb, const void* modes, const int64_t* n, double eps,
                            const void* scale, double* sums, uint64_t* sort_id, cudaStream_t st) {
  EFGP_CHECK(nb >= 1 && t0 >= 0, "bad probe range");
  for (int a = 0; a < nu.dim; ++a) EFGP_CHECK(n[a] >= 1, "mode counts must be >= 1");
  if (nu.dtype == 0)
    type2_probe_dots_t<float>(nu, seed, t0, nb, modes, n, eps, scale, sums, sort_id, st);
  else
    type2_probe_dots_t<double>(nu, seed, t0, nb, modes, n, eps, scale, sums, sort_id, st);
}

void nufft_probes_export(Nufft& nu, uint64_t seed, int t0, int nb, const int64_t* n, double eps, void* z, uint64_t* sort_id,
                         cudaStream_t st) {
  EFGP_CHECK(nb >= 1 && t0 >= 0, "bad probe range");
  GridChoice gc = choose_grid(nu, n, eps, st);
  check_sort_id(*gc.sp, sort_id);
  const size_t esz = nu.dtype == 0 ? 4 : 8;
  for (int b = 0; b < nb; ++b)
    launch_probe_export(*gc.sp, (char*)z + (size_t)b * nu.N * esz, 1, ProbeSpec{seed, t0 + b, 1, gc.sp->has_orig ? 1 : 0}, st);
}

}  // namespace efgp
