// Internal NUFFT structures: bin-sorted point set, fine-grid plan, launch helpers.
#pragma once
#include "common.cuh"
#include <map>
#include <memory>

namespace efgp {

// Points sorted by fine-grid cell for one fine grid geometry.
// Per point a RECORD of RW words of T is stored:  X[0..dim)  then the 32-bit level-1 position `binv`
// (raw bits in word `dim`), padded to RW = 2 words (dim 1) or 4 words (dim 2, 3) so that a record is
// one aligned 8/16/32-byte unit: the counting sort places a point with ONE full-sector store, and a
// spread / interp thread fetches it with one 8- or 16-byte cp.async / vector load.
// X[a] = Xc = X + shift (mod nf), X = nf * frac(h (x - xcen)); shift = 1/2 for odd W, 0 for even
// (so cell = floor(Xc) identifies the W^d footprint for either parity, see spread.cu).
//
// Storage is TILE-TRANSPOSED at record granularity: sorted ranks are cut into tiles of kTileThreads*P
// consecutive points; inside a tile thread t owns the P consecutive ranks [t*P, (t+1)*P) (one long register
// run), and its j-th point is stored at slot  tile*kTileThreads*P + j*kTileThreads + t  so that at every
// step j a warp reads 32 consecutive records (fully coalesced, prefetchable).  Slots past N hold X[0] = -1.
constexpr int kTileThreads = 256;
constexpr int kTile3 = 8;   // 3-D: edge (in cells) of the fine-grid tiles the sort groups points by
struct SortedPoints {
  int dim = 0;
  int dtype = 1;  // 0 f32, 1 f64
  int64_t N = 0;
  int64_t nf[kMaxDim] = {1, 1, 1};
  int odd = 0;
  int P = 1;            // points per thread
  int RW = 2;           // words of T per record
  bool has_payload = false;   // 2-D only: word 3 of every record holds a real per-point strength
  bool has_orig = false;      // 64-bit records: the upper half of the position word holds the point's ORIGINAL index
  int64_t ntiles = 0;
  int64_t npos = 0;     // ntiles * kTileThreads * P
  void* rec = nullptr;        // npos records
  uint16_t* A = nullptr;      // A[i] = position of original point i inside its 2048-point tile, bucket-major (N entries)
  uint32_t* B = nullptr;      // B[a] = slot of the point at level-1 position a (inverse of the records' binv)
  uint32_t* tileStart = nullptr;  // [nb1][ntiles1]: first level-1 position of (bucket, 2048-point tile)
  int ntiles1 = 0, nb1 = 0;
  uint64_t last_use = 0;
  uint64_t id = 0;            // unique per build: the regenerated probes are a function of (seed, probe, slot)
  ~SortedPoints();
};
__host__ __device__ constexpr int rec_words(int dim) { return dim == 1 ? 2 : 4; }

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

// cuFFT plans are cached process-wide (one process drives one GPU): the EFGP loop rebuilds its NUFFT
// and Toeplitz handles at every hyperparameter step, and plan creation costs far more than the
// transforms on these small grids.  All plans share one work area (calls are stream-ordered).
struct FftCache {
  std::map<FftPlanKey, cufftHandle> plans;
  DevBuf work;
  size_t work_bytes = 0;
  cufftHandle get(int dim, const int64_t* n, cufftType type, int batch, cudaStream_t st);
  ~FftCache();
};
FftCache& global_fft();

struct Nufft {
  int dim = 0;
  int dtype = 1;
  int64_t N = 0;
  const void* x = nullptr;  // borrowed device pointer, (N, dim) row-major
  double h = 0;
  double xcen[kMaxDim] = {0, 0, 0};
  uint64_t clock = 0;
  std::vector<std::unique_ptr<SortedPoints>> sorted;  // small LRU cache
  DevBuf fine, spec, tmp, dec, cperm;
  bool carry_index = false;                    // sorted sets carry the original point index (probe streaming)
  int64_t cost_npts = 0;                       // > 0: point count used by the grid-choice cost model (sharded fits)
  const void* pending_payload = nullptr;       // set by nufft_type1_pair: next sort carries these strengths
  const SortedPoints* fresh_payload_sp = nullptr;  // the sorted set built with them during this call
};

// ---- points.cu
// payload: optional real (N,) strengths carried through the sort into the records (2-D only)
void build_sorted_points(SortedPoints& sp, const void* x, const void* payload, int64_t N, int dim, int dtype,
                         double h, const double* xcen, const int64_t* nf, int odd, cudaStream_t st,
                         bool carry_index = false);

// ---- points.cu : cs[slot] = c[i * cstride] for the point i stored in that slot (npos entries), in two
// sector-efficient stages (original -> bucket order -> slot order)
void permute_strengths(const SortedPoints& sp, const void* c, int64_t cstride, void* cs, cudaStream_t st);
// out[i * ostride] = r[B[A[i]]]: slot-order results (written coalesced by interp) back to original order
// r holds slot-order results, or (bucket_order) results already at their level-1 positions
void unpermute_output(const SortedPoints& sp, const void* r, void* out, int64_t ostride, cudaStream_t st,
                      bool bucket_order = false);

// ---- spread.cu : grid (real, nf^d) += sum_n c_n psi(l - X_n).  cs == nullptr means all ones;
// otherwise cs holds the strengths in SLOT order (permute_strengths).
// from_record: take the strengths from the records' payload word instead of cs (needs sp.has_payload)
// igrid != nullptr: deterministic mode -- the flushes add round(v * iscale) to the 64-bit INTEGER grid igrid instead of
// v to grid (see DetSpec in spread.cu); the caller converts afterwards (det_convert in nufft.cu)
void launch_spread(const SortedPoints& sp, const ESKernel& k, const void* cs, void* grid, cudaStream_t st,
                   bool from_record = false, void* igrid = nullptr, double iscale = 0.0);
void set_deterministic(int on);
bool deterministic();

// ---- Rademacher probes regenerated inside the kernels (philox.cuh): spread of pr.nb probes into pr.nb contiguous
// grids, interpolation with the fused sum_n Z * (.) epilogue, and Z itself in original order (tests)
struct ProbeSpec;
int spread_probe_batch(int dim, int W, int dtype);
bool spread_pair_supported(const SortedPoints& sp, const ESKernel& kg, const ESKernel& kr);
void launch_spread_pair(const SortedPoints& sp, const ESKernel& kg, const ESKernel& kr, void* gridG, void* gridR,
                        cudaStream_t st);
void launch_spread_probes(const SortedPoints& sp, const ESKernel& k, void* grid, const ProbeSpec& pr, cudaStream_t st,
                          void* igrid = nullptr, double iscale = 0.0);
void launch_interp_probe_dots(const SortedPoints& sp, const ESKernel& k, const void* grid, double* out,
                              const ProbeSpec& pr, cudaStream_t st);
void launch_probe_export(const SortedPoints& sp, void* z_out, int64_t ostride, const ProbeSpec& pr, cudaStream_t st);

// ---- interp.cu : out[i * ostride] = sum_l grid[l] psi(l - X_i)  (i = original point index)
void launch_interp(const SortedPoints& sp, const ESKernel& k, const void* grid, void* out,
                   int64_t ostride, cudaStream_t st);

// ---- nufft.cu
void nufft_type1(Nufft& nu, const void* strengths, int is_complex, int batch, const int64_t* n, double eps,
                 int isign, int modeord, const void* scale, void* out, cudaStream_t st);
// F^*1 on shape_gen at eps_gen and scale * F^*y on shape_rhs at eps_rhs in one call: one sort, y carried
// through it as record payload (no strength permutation)
void nufft_type1_pair(Nufft& nu, const void* y, const int64_t* shape_gen, double eps_gen, const int64_t* shape_rhs,
                      double eps_rhs, const void* scale, void* out_gen, void* out_rhs, cudaStream_t st);
void nufft_type2(Nufft& nu, const void* modes, int batch, const int64_t* n, double eps, int isign, int modeord,
                 const void* scale, int real_out, void* out, cudaStream_t st);
// Hutchinson probes streamed through the transforms (efgpnd.py:176-241 without the (T, N) arrays):
//   type1_probes:     out[b] = scale * F^* Z_{t0+b}            (nb, *n) complex, Z regenerated in the spread kernel
//   type2_probe_dots: sums[b] = sum_n Z_{t0+b}[n] Re(F f_b)[n]  modes (nb, *n) complex -> (nb,) float64
//   probes_export:    Z_{t0..t0+nb-1} in original point order   (nb, N) real
// All three must see the same sorted point set: `sort_id` returns / checks its identity.
void nufft_type1_probes(Nufft& nu, uint64_t seed, int t0, int nb, const int64_t* n, double eps, const void* scale, void* out,
                        uint64_t* sort_id, cudaStream_t st);
void nufft_type2_probe_dots(Nufft& nu, uint64_t seed, int t0, int nb, const void* modes, const int64_t* n, double eps,
                            const void* scale, double* sums, uint64_t* sort_id, cudaStream_t st);
void nufft_probes_export(Nufft& nu, uint64_t seed, int t0, int nb, const int64_t* n, double eps, void* z, uint64_t* sort_id,
                         cudaStream_t st);

}  // namespace efgp
