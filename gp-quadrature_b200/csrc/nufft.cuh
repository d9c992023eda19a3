// Internal NUFFT structures: bin-sorted point set, fine-grid plan, launch helpers.
#pragma once
#include "common.cuh"
#include <map>
#include <memory>

namespace efgp {

// Points sorted by fine-grid cell for one fine grid geometry.
// xs[a][i] = Xc = X + shift (mod nf), X = nf * frac(h (x - xcen)); shift = 1/2 for odd W, 0 for even
// (so cell = floor(Xc) identifies the W^d footprint for either parity, see spread.cu).
struct SortedPoints {
  int dim = 0;
  int dtype = 1;  // 0 f32, 1 f64
  int64_t N = 0;
  int64_t nf[kMaxDim] = {1, 1, 1};
  int odd = 0;
  void* xs[kMaxDim] = {nullptr, nullptr, nullptr};
  uint32_t* perm = nullptr;
  uint64_t last_use = 0;
  ~SortedPoints();
};

struct FftPlanKey {
  int dim, type, batch;
  int64_t n[kMaxDim];
  bool operator<(const FftPlanKey& o) const {
    if (dim != o.dim) return dim < o.dim;
    if (type != o.type) return type < o.type;
    if (batch != o.batch) return batch < o.batch;
    for (int a = 0; a < kMaxDim; ++a)
      if (n[a] != o.n[a]) return n[a] < o.n[a];
    return false;
  }
};

// growable device scratch
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  void* get(size_t bytes);
  ~DevBuf();
};

struct FftCache {
  std::map<FftPlanKey, cufftHandle> plans;
  DevBuf work;
  size_t work_bytes = 0;
  cufftHandle get(int dim, const int64_t* n, cufftType type, int batch, cudaStream_t st);
  ~FftCache();
};

struct Nufft {
  int dim = 0;
  int dtype = 1;
  int64_t N = 0;
  const void* x = nullptr;  // borrowed device pointer, (N, dim) row-major
  double h = 0;
  double xcen[kMaxDim] = {0, 0, 0};
  uint64_t clock = 0;
  std::vector<std::unique_ptr<SortedPoints>> sorted;  // small LRU cache
  FftCache fft;
  DevBuf fine, spec, tmp, dec;
  std::map<std::string, std::vector<double>> dec_cache_host;
};

// ---- points.cu
void build_sorted_points(SortedPoints& sp, const void* x, int64_t N, int dim, int dtype, double h,
                         const double* xcen, const int64_t* nf, int odd, cudaStream_t st);

// ---- spread.cu : grid (real, nf^d) += sum_n c_n psi(l - X_n).  c == nullptr means all ones.
// c is indexed in ORIGINAL point order through perm (c[perm[i] * cstride]).
void launch_spread(const SortedPoints& sp, const ESKernel& k, const void* c, int64_t cstride,
                   void* grid, cudaStream_t st);

// ---- interp.cu : out[perm[i] * ostride] = sum_l grid[l] psi(l - X_i)
void launch_interp(const SortedPoints& sp, const ESKernel& k, const void* grid, void* out,
                   int64_t ostride, cudaStream_t st);

// ---- nufft.cu
void nufft_type1(Nufft& nu, const void* strengths, int is_complex, int batch, const int64_t* n, double eps,
                 int isign, int modeord, const void* scale, void* out, cudaStream_t st);
void nufft_type2(Nufft& nu, const void* modes, int batch, const int64_t* n, double eps, int isign, int modeord,
                 const void* scale, int real_out, void* out, cudaStream_t st);

}  // namespace efgp
