// Bin-sorting of nonuniform points by fine-grid cell: a two-level counting sort written for B200.
//
// Replaces the point handling FINUFFT does inside every pff.finufft_type1/type2 call
// (reference call sites efgpnd.py:1496-1499, 1533-1549; phi = 2*pi*h*(x - xcen), efgpnd.py:1451).
// Points are folded to one period, scaled to fine-grid units and ordered by cell once per
// (point set, fine grid); every later spread / interp streams them in cell order.
//
// Why not a library radix sort + gather: moving an 8- or 16-byte payload through a RANDOM permutation
// of 1e8 elements costs one 32-byte DRAM sector per element (measured: 2.0-2.3 ms per payload array,
// 4.2 ms for the scattered type-2 result), and an LSD radix sort on (key, index) needs three more passes.
// Here the permutation is factored into two stages that are each sector-efficient:
//
//   level 1  original order -> 128 (<= 256) coarse buckets (contiguous cell ranges = bands of grid rows).
//            One CTA bucket-sorts a tile of 2048 points in shared memory and writes each bucket's
//            run contiguously, so DRAM sees full sectors.  A[i] = level-1 position of point i.
//   level 2  inside a bucket (N/128 points: a few MB, L2 resident) a counting sort by cell places the
//            points in their final slots.  Binv[slot] = level-1 position of the point in that slot.
//
// Later transforms move per-point data through the same two stages: strengths c[i] -> tmp[A[i]]
// (coalesced reads, bucket runs) -> cs[slot] = tmp[Binv[slot]] (gather inside an L2-sized window);
// type-2 results are written in slot order (coalesced), gathered into bucket order (tmp[a] = r[B[a]],
// B = Binv^-1, inside the bucket window) and streamed back as out[i] = tmp[A[i]].
#include <cstdlib>
#include "nufft.cuh"
#include <cub/device/device_scan.cuh>
#include <algorithm>

namespace efgp {

SortedPoints::~SortedPoints() {
  if (rec) pool_free(rec);
  if (A) pool_free(A);
  if (B) pool_free(B);
  if (tileStart) pool_free(tileStart);
}

// Scratch comes from the caching pool (pool.cpp): handles are rebuilt every hyperparameter step and a
// cudaMalloc/cudaFree pair per buffer would synchronise the device each time.
void* DevBuf::get(size_t bytes) {
  if (bytes > cap) {
    if (p) pool_free(p);
    p = nullptr;
    cap = 0;
    p = pool_alloc(bytes);
    cap = bytes;
  }
  return p;
}
DevBuf::~DevBuf() {
  if (p) pool_free(p);
}

template <typename T>
struct FoldParams {
  T h;
  T xcen[kMaxDim];
  T nf[kMaxDim];
  int inf[kMaxDim];
  T shift;
  int ntile[kMaxDim];   // 3-D: fine-grid tiles of kTile3 cells per axis (tile-major sort keys), else unused
};

struct SortGeom {
  uint32_t cpb;     // cells per level-1 bucket (contiguous range of flattened cell indices)
  uint32_t cpb_inv; // floor(2^32 / cpb) (0 for cpb == 1), g_inv likewise: the per-point divisions by these run-time
  uint32_t g_inv;   // constants are a multiply-high and one fix-up instead of ~20-instruction 32-bit divisions
  uint32_t P_inv, tile_inv;   // the same for P and for the tile length kTileThreads * P (slot_of_rank)
  uint32_t g;       // cells per level-2 bin (1 = exact cell sort)
  int nb1;          // level-1 buckets (<= kMaxBuckets)
  int nbins2;       // level-2 bins per bucket (<= kMaxBins2)
  int P;            // points per thread of the tile-transposed layout (any value in 1..kMaxP)
};

// n / d for a run-time d with inv = floor(2^32 / d) (inv == 0 means d == 1): the multiply-high underestimates the
// quotient by at most one for every 32-bit n, one comparison fixes it.  Exact.
__device__ __forceinline__ uint32_t fast_div(uint32_t n, uint32_t d, uint32_t inv) {
  if (inv == 0) return n;
  uint32_t q = __umulhi(n, inv);
  if (n - q * d >= d) ++q;
  return q;
}

// slot_of_rank with the divisions by the tile length and by P done through their reciprocals (SortGeom)
__device__ __forceinline__ int64_t slot_of_rank_fast(uint32_t r32, uint32_t P, uint32_t P_inv, uint32_t tile_inv) {
  const uint32_t tile_len = (uint32_t)kTileThreads * P;
  const uint32_t tile = fast_div(r32, tile_len, tile_inv);
  const uint32_t rr = r32 - tile * tile_len;
  const uint32_t t = fast_div(rr, P, P_inv), j = rr - t * P;
  return (int64_t)tile * tile_len + (int64_t)j * kTileThreads + t;
}

constexpr int kMaxBuckets = 256;
constexpr int kDefaultBuckets = 128;
constexpr int kMaxBins2 = 8192;
constexpr int kSortThreads = 256;
constexpr int kSortPer = 8;
static_assert(kSortThreads == kMaxBuckets, "the permutation kernels use one thread per bucket");
constexpr int kL1Threads = 512;   // level-1 kernels: 512 threads x 4 points (fewer live registers per thread,
constexpr int kL1Per = 4;         // 3 CTAs = 1536 threads per SM) over the same 2048-point tile
constexpr int kSortTile = kSortThreads * kSortPer;  // 2048 points per CTA (43 KB of staging: 5 CTAs per SM)
static_assert(kL1Threads * kL1Per == kSortTile, "level-1 and permutation kernels share one tile size");

// fine-grid coordinate of one coordinate value, shifted by 1/2 for odd widths, in [0, nf)
template <typename T>
__device__ __forceinline__ T fold_coord(T xv, T h, T xc, T nf, T shift, int inf) {
  T t = h * (xv - xc);
  t -= floor(t);
  T X = t * nf + shift;
  if (X >= nf) X -= nf;
  if (!(X >= T(0)) || X >= nf) X = T(0);  // rounding at the period edge (and NaN guard)
  if ((int)X >= inf) X = T(0);
  return X;
}

// Sort key of a point.  1-D / 2-D: the flattened (row-major) cell index.  3-D: TILE-major -- cells are grouped
// in kTile3^3 blocks and the key is (tile index) * kTile3^3 + (cell within the tile), so that all points of a
// tile are consecutive after the sort: the 3-D spread / interp kernels stage one tile (+ halo) of the fine
// grid in shared memory at a time (spread.cu, interp.cu).
template <typename T, int D>
__device__ __forceinline__ uint32_t fold_values(const T* xv, const FoldParams<T>& p, T* X);

template <typename T, int D>
__device__ __forceinline__ uint32_t fold_point(const T* __restrict__ x, int64_t i, const FoldParams<T>& p, T* X) {
  T xv[D];
#pragma unroll
  for (int a = 0; a < D; ++a) xv[a] = x[i * D + a];
  return fold_values<T, D>(xv, p, X);
}

// xv may alias X
template <typename T, int D>
__device__ __forceinline__ uint32_t fold_values(const T* xv, const FoldParams<T>& p, T* X) {
  uint32_t key = 0;
  int c[D];
#pragma unroll
  for (int a = 0; a < D; ++a) {
    X[a] = fold_coord(xv[a], p.h, p.xcen[a], p.nf[a], p.shift, p.inf[a]);
    c[a] = (int)X[a];
    key = key * (uint32_t)p.inf[a] + (uint32_t)c[a];
  }
  if (D == 3) {
    const uint32_t tile = ((uint32_t)(c[0] / kTile3) * p.ntile[1] + (uint32_t)(c[1] / kTile3)) * p.ntile[2] +
                          (uint32_t)(c[2] / kTile3);
    const uint32_t within = ((uint32_t)(c[0] % kTile3) * kTile3 + (uint32_t)(c[1] % kTile3)) * kTile3 +
                            (uint32_t)(c[2] % kTile3);
    key = tile * (kTile3 * kTile3 * kTile3) + within;
  }
  return key;
}

// slot of sorted rank r in the tile-transposed layout (nufft.cuh); r < 2^31 (checked on the host): 32-bit divisions
__device__ __forceinline__ int64_t slot_of_rank(int64_t r, int P) {
  const uint32_t tile_len = (uint32_t)kTileThreads * (uint32_t)P;
  const uint32_t r32 = (uint32_t)r;
  const uint32_t tile = r32 / tile_len;
  const uint32_t rr = r32 - tile * tile_len;
  const uint32_t t = rr / (uint32_t)P, j = rr - t * (uint32_t)P;
  return (int64_t)tile * tile_len + (int64_t)j * kTileThreads + t;
}

// ------------------------------------------------------------------ level 1
template <typename T, int D>
__global__ void __launch_bounds__(kL1Threads)
bucket_hist_kernel(const T* __restrict__ x, int64_t N, FoldParams<T> p, SortGeom g, int ntiles1,
                   uint32_t* __restrict__ tileHist) {
  __shared__ uint32_t cnt[kMaxBuckets];
  const int tid = threadIdx.x;
  const int64_t tile = blockIdx.x;
  if (tid < kMaxBuckets) cnt[tid] = 0;
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kL1Per; ++k) {
    int64_t i = tile * kSortTile + k * kL1Threads + tid;
    if (i < N) {
      T X[D];
      uint32_t key = fold_point<T, D>(x, i, p, X);
      atomicAdd(&cnt[fast_div(key, g.cpb, g.cpb_inv)], 1u);
    }
  }
  __syncthreads();
  for (int b = tid; b < g.nb1; b += kL1Threads) tileHist[(int64_t)b * ntiles1 + tile] = cnt[b];
}

// One CTA: bucket-sort its 2048 points in shared memory, then write each bucket's run contiguously.
// NS = D staged words per point, or D + 1 when a per-point payload (real strengths y) rides along.
template <typename T, int D, int NS>
__global__ void __launch_bounds__(kL1Threads, 3)
bucket_scatter_kernel(const T* __restrict__ x, const T* __restrict__ ypay, int64_t N, FoldParams<T> p, SortGeom g,
                      int ntiles1, const uint32_t* __restrict__ tileStart, T* __restrict__ R0, T* __restrict__ R1,
                      T* __restrict__ R2, T* __restrict__ R3, uint16_t* __restrict__ K2, uint16_t* __restrict__ A,
                      uint32_t* __restrict__ RI) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* S = (T*)smem_raw;                                   // [NS][kSortTile]
  uint32_t* SKB = (uint32_t*)(S + NS * kSortTile);       // [kSortTile] level-1 bucket | level-2 bin << 16 (one word:
                                                         // one scattered store and one load instead of two 16-bit ones)
  uint32_t* cnt = SKB + kSortTile;                       // [kMaxBuckets]
  uint32_t* lstart = cnt + kMaxBuckets;
  uint32_t* gstart = lstart + kMaxBuckets;
  uint32_t* SI = gstart + kMaxBuckets;                    // [kSortTile] original indices (only when RI != nullptr)
  // per input slot (k * kL1Threads + tid): bucket | level-2 bin << 16.  Held in shared memory across the scan instead
  // of registers: at 3 CTAs of 512 threads per SM (42 registers) they were spilled to local memory, whose reloads
  // missed L1 (ncu round 2: 4.3e6 LDL, 1.35 GB of local-memory sectors through L2, "long scoreboard" the top stall)
  __shared__ uint32_t slotkey[kSortTile];
  __shared__ uint16_t slotrank[kSortTile];   // rank of the point inside its bucket within this tile
  __shared__ uint32_t wtot[kMaxBuckets / 32];
  const int tid = threadIdx.x;
  const int64_t tile = blockIdx.x;
  // this tile's first output position in every bucket: issued now, consumed after the local scan
  const uint32_t gs = tid < g.nb1 ? tileStart[(int64_t)tid * ntiles1 + tile] : 0u;
  if (tid < kMaxBuckets) cnt[tid] = 0;
  __syncthreads();
  T X[kL1Per][NS];
#pragma unroll
  for (int k = 0; k < kL1Per; ++k) {
    int64_t i = tile * kSortTile + k * kL1Threads + tid;
    uint32_t sk = 0xffffffffu;
    if (i < N) {
      uint32_t key = fold_point<T, D>(x, i, p, X[k]);
      if (NS > D) X[k][D] = ypay[i];
      uint32_t b = fast_div(key, g.cpb, g.cpb_inv);
      sk = b | (fast_div(key - b * g.cpb, g.g, g.g_inv) << 16);
      slotrank[k * kL1Threads + tid] = (uint16_t)atomicAdd(&cnt[b], 1u);
    }
    slotkey[k * kL1Threads + tid] = sk;
  }
  __syncthreads();
  uint32_t c = 0, v = 0;
  if (tid < kMaxBuckets) {   // exclusive scan of the 256 bucket counts: warp scans + 8 warp totals
    c = cnt[tid];
    v = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
      if ((tid & 31) >= o) v += u;
    }
    if ((tid & 31) == 31) wtot[tid >> 5] = v;
  }
  __syncthreads();
  if (tid < kMaxBuckets) {
    uint32_t woff = 0;
    for (int q = 0; q < (tid >> 5); ++q) woff += wtot[q];
    lstart[tid] = v - c + woff;
    gstart[tid] = gs;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kL1Per; ++k) {
    const uint32_t sk = slotkey[k * kL1Threads + tid];
    if (sk != 0xffffffffu) {
      int64_t i = tile * kSortTile + k * kL1Threads + tid;
      const uint32_t bkk = sk & 0xffffu;
      uint32_t lp = lstart[bkk] + slotrank[k * kL1Threads + tid];
#pragma unroll
      for (int a = 0; a < NS; ++a) S[a * kSortTile + lp] = X[k][a];
      SKB[lp] = sk;
      if (RI) SI[lp] = (uint32_t)i;
      A[i] = (uint16_t)lp;      // tile-local position in bucket-major order (< kSortTile)
    }
  }
  __syncthreads();
  const int64_t rem = N - tile * kSortTile;
  const int n_tile = rem < kSortTile ? (int)rem : kSortTile;
  T* R[4] = {R0, R1, R2, R3};
  for (int lp = tid; lp < n_tile; lp += kL1Threads) {
    const uint32_t skb = SKB[lp];
    const uint32_t b = skb & 0xffffu;
    uint32_t a = gstart[b] + ((uint32_t)lp - lstart[b]);
#pragma unroll
    for (int d = 0; d < NS; ++d) R[d][a] = S[d * kSortTile + lp];
    K2[a] = (uint16_t)(skb >> 16);
    if (RI) RI[a] = SI[lp];
  }
}

// one record = RW words, written with one aligned vector store (a 32-byte record is one 256-bit store)
template <typename T, int RW>
__device__ __forceinline__ void store_record(T* dst, const T (&w)[RW]) {
  if constexpr (sizeof(T) * RW == 32) {
    const double* wd = reinterpret_cast<const double*>(w);
    st_global_256(reinterpret_cast<double*>(dst), wd[0], wd[1], wd[2], wd[3]);
  } else if constexpr (sizeof(T) == 8) {            // 1-D double: (X, index) = 16 bytes
    reinterpret_cast<double2*>(dst)[0] = make_double2((double)w[0], (double)w[1]);
  } else if constexpr (RW == 4) {                   // float records: 16 bytes, values stay in registers
    reinterpret_cast<float4*>(dst)[0] = make_float4((float)w[0], (float)w[1], (float)w[2], (float)w[3]);
  } else {
    reinterpret_cast<float2*>(dst)[0] = make_float2((float)w[0], (float)w[1]);
  }
}

// ------------------------------------------------------------------ level 2
// Level-2 tiles never straddle buckets: bucket b owns tiles [tile2Base[b], tile2Base[b+1]).
__global__ void __launch_bounds__(kMaxBuckets)
level2_tiles_kernel(const uint32_t* __restrict__ tileStart, int ntiles1, int nb1, int64_t N,
                    uint32_t* __restrict__ bucketStart /* nb1+1 */, uint32_t* __restrict__ tile2Base /* nb1+1 */) {
  __shared__ uint32_t tl[kMaxBuckets + 1];
  const int b = threadIdx.x;
  uint32_t s = 0, e = 0;
  if (b < nb1) {
    s = tileStart[(int64_t)b * ntiles1];
    e = b + 1 < nb1 ? tileStart[(int64_t)(b + 1) * ntiles1] : (uint32_t)N;
    bucketStart[b] = s;
    if (b == nb1 - 1) bucketStart[nb1] = (uint32_t)N;
  }
  tl[b] = b < nb1 ? (e - s + kSortTile - 1) / kSortTile : 0u;
  __syncthreads();
  if (b == 0) {
    uint32_t acc = 0;
    for (int q = 0; q < nb1; ++q) {
      uint32_t c = tl[q];
      tile2Base[q] = acc;
      acc += c;
    }
    tile2Base[nb1] = acc;
  }
}

__device__ __forceinline__ bool level2_tile_range_of(uint32_t tile, const uint32_t* __restrict__ bucketStart,
                                                     const uint32_t* __restrict__ tile2Base, int nb1, int& b,
                                                     uint32_t& s, uint32_t& e) {
  if (tile >= tile2Base[nb1]) return false;
  int lo = 0, hi = nb1;                 // last b with tile2Base[b] <= tile
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (tile2Base[mid] <= tile) lo = mid; else hi = mid;
  }
  b = lo;
  s = bucketStart[b] + (tile - tile2Base[b]) * kSortTile;
  uint32_t be = bucketStart[b + 1];
  e = s + kSortTile < be ? s + kSortTile : be;
  return true;
}
__device__ __forceinline__ bool level2_tile_range(const uint32_t* __restrict__ bucketStart,
                                                  const uint32_t* __restrict__ tile2Base, int nb1, int& b,
                                                  uint32_t& s, uint32_t& e) {
  return level2_tile_range_of(blockIdx.x, bucketStart, tile2Base, nb1, b, s, e);
}

// Per-(bucket, bin) point counts.  A CTA walks kHistTiles CONSECUTIVE level-2 tiles (tiles are ordered by bucket, a
// bucket has hundreds of them) and adds its shared-memory counts to the global ones only when the bucket changes or
// at the end: one global atomic per (CTA, non-empty bin) instead of one per (tile, non-empty bin) -- 3e7 -> 4e6 at
// N = 1e8, which is what the one-tile version spent its 0.17 ms on (it reads 2 bytes per point).
constexpr int kHistTiles = 16;

__global__ void __launch_bounds__(kSortThreads)
bin_hist_kernel(const uint16_t* __restrict__ K2, const uint32_t* __restrict__ bucketStart,
                const uint32_t* __restrict__ tile2Base, SortGeom g, uint32_t* __restrict__ binCount) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint32_t* cnt = (uint32_t*)smem_raw;  // [nbins2]
  for (int q = threadIdx.x; q < g.nbins2; q += kSortThreads) cnt[q] = 0;
  int cur = -1;
  for (int k = 0; k < kHistTiles; ++k) {
    int b;
    uint32_t s, e;
    const bool ok = level2_tile_range_of(blockIdx.x * kHistTiles + k, bucketStart, tile2Base, g.nb1, b, s, e);
    if (!ok || b != cur) {                     // (uniform over the CTA) publish and clear the finished bucket
      __syncthreads();
      if (cur >= 0) {
        for (int q = threadIdx.x; q < g.nbins2; q += kSortThreads) {
          const uint32_t c = cnt[q];
          if (c) {
            atomicAdd(&binCount[(int64_t)cur * g.nbins2 + q], c);
            cnt[q] = 0;
          }
        }
      }
      __syncthreads();
      if (!ok) return;
      cur = b;
    }
    for (uint32_t a = s + threadIdx.x; a < e; a += kSortThreads) atomicAdd(&cnt[K2[a]], 1u);
  }
  __syncthreads();
  for (int q = threadIdx.x; q < g.nbins2; q += kSortThreads) {
    const uint32_t c = cnt[q];
    if (c) atomicAdd(&binCount[(int64_t)cur * g.nbins2 + q], c);
  }
}

// One level-2 tile (<= 2048 consecutive points of one bucket): all of the tile's words are loaded into registers
// FIRST (8 points x 4 loads per thread in flight, their latency hidden behind the histogram phase), K2 is read
// once, and every point leaves as ONE 256-bit record store.
template <typename T, int D>
__global__ void __launch_bounds__(kSortThreads, 3)
bin_scatter_kernel(const T* __restrict__ R0, const T* __restrict__ R1, const T* __restrict__ R2,
                   const T* __restrict__ RY, const uint16_t* __restrict__ K2, const uint32_t* __restrict__ bucketStart,
                   const uint32_t* __restrict__ tile2Base, SortGeom g, const uint32_t* __restrict__ binStart,
                   uint32_t* __restrict__ binCursor, T* __restrict__ rec, uint32_t* __restrict__ B,
                   const uint32_t* __restrict__ RI) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint32_t* cnt = (uint32_t*)smem_raw;   // [nbins2] counts, then running ranks
  uint32_t* base = cnt + g.nbins2;       // [nbins2] first sorted rank of this tile's points in the bin
  int b;
  uint32_t s, e;
  if (!level2_tile_range(bucketStart, tile2Base, g.nb1, b, s, e)) return;
  const T* R[3] = {R0, R1, R2};
  constexpr int RW = rec_words(D);
  constexpr uint32_t kNone = 0xffffffffu;
  // level-2 bin and original index of every input slot (k * kSortThreads + tid): in shared memory, not registers --
  // with the 8 records (64 registers) they exceeded the 85-register budget of 3 CTAs per SM and were spilled
  __shared__ uint32_t sq[kSortTile];
  __shared__ uint32_t soi[kSortTile];
  T w[kSortPer][RW];
#pragma unroll
  for (int k = 0; k < kSortPer; ++k) {
    const uint32_t a = s + k * kSortThreads + threadIdx.x;
    uint32_t qk = kNone;
#pragma unroll
    for (int d = 0; d < RW; ++d) w[k][d] = T(0);
    if (a < e) {
      qk = K2[a];
      if (RI) soi[k * kSortThreads + threadIdx.x] = RI[a];
#pragma unroll
      for (int d = 0; d < D; ++d) w[k][d] = R[d][a];
      if (D == 2 && RY) w[k][3] = RY[a];     // payload rides in the record's spare word
    }
    sq[k * kSortThreads + threadIdx.x] = qk;
  }
  for (int i = threadIdx.x; i < g.nbins2; i += kSortThreads) cnt[i] = 0;
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kSortPer; ++k) {
    const uint32_t qk = sq[k * kSortThreads + threadIdx.x];
    if (qk != kNone) atomicAdd(&cnt[qk], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < g.nbins2; i += kSortThreads) {
    uint32_t c = cnt[i];
    if (c) {
      int64_t gb = (int64_t)b * g.nbins2 + i;
      base[i] = binStart[gb] + atomicAdd(&binCursor[gb], c);
    }
    cnt[i] = 0;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kSortPer; ++k) {
    const uint32_t qk = sq[k * kSortThreads + threadIdx.x];
    if (qk == kNone) continue;
    const uint32_t a = s + k * kSortThreads + threadIdx.x;
    uint32_t r = base[qk] + atomicAdd(&cnt[qk], 1u);
    int64_t slot = slot_of_rank_fast(r, (uint32_t)g.P, g.P_inv, g.tile_inv);
    *reinterpret_cast<uint32_t*>(&w[k][D]) = a;
    // 64-bit words have room for the ORIGINAL index next to the level-1 position: the regenerated Hutchinson probes
    // are a function of (seed, probe, original index), independent of the (run-dependent) order inside a cell
    if constexpr (sizeof(T) == 8) reinterpret_cast<uint32_t*>(&w[k][D])[1] = RI ? soi[k * kSortThreads + threadIdx.x] : 0u;
    store_record<T, RW>(rec + slot * RW, w[k]);
    B[a] = (uint32_t)slot;
  }
}

template <typename T, int D>
__global__ void __launch_bounds__(256)
pad_slots_kernel(int64_t N, int64_t npos, int P, T* __restrict__ rec) {
  int64_t r = N + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= npos) return;
  int64_t slot = slot_of_rank(r, P);
  constexpr int RW = rec_words(D);
  T w[RW];
#pragma unroll
  for (int d = 0; d < RW; ++d) w[d] = T(0);
  w[0] = T(-1);
  *reinterpret_cast<uint32_t*>(&w[D]) = 0xffffffffu;
  store_record<T, RW>(rec + slot * RW, w);
}

// ------------------------------------------------------------------ per-point data through the two stages
// Per-tile view of the level-1 permutation, rebuilt from the persistent tileStart table: bucket b's
// points of this tile sit at global positions [gstart[b], gstart[b] + cnt[b]) and, bucket-major, at
// tile-local positions [lstart[b], lstart[b] + cnt[b]).  Needs blockDim.x == kMaxBuckets.
__device__ __forceinline__ void tile_bucket_ranges(const uint32_t* __restrict__ tileStart, int ntiles1, int nb1,
                                                   int64_t N, int64_t tile, uint32_t* gstart, uint32_t* lstart,
                                                   uint32_t* wsum) {
  const int tid = threadIdx.x;
  uint32_t gs = 0xffffffffu, c = 0;
  if (tid < nb1) {
    const int64_t idx = (int64_t)tid * ntiles1 + tile;
    gs = tileStart[idx];
    const uint32_t ge = idx + 1 < (int64_t)nb1 * ntiles1 ? tileStart[idx + 1] : (uint32_t)N;
    c = ge - gs;
  }
  uint32_t v = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
    if ((tid & 31) >= o) v += u;
  }
  if ((tid & 31) == 31) wsum[tid >> 5] = v;
  __syncthreads();
  uint32_t woff = 0;
  for (int q = 0; q < (tid >> 5); ++q) woff += wsum[q];
  gstart[tid] = gs;
  lstart[tid] = woff + v - c;
  __syncthreads();
}
// Lanes-per-bucket of the run copies below: the power of two nearest the mean run length kSortTile / nb1.
__host__ __device__ inline int run_group_shift(int nb1) {
  int per = kSortTile / (nb1 > 0 ? nb1 : 1);
  int sh = 3;
  while (sh < 5 && (1 << sh) < per) ++sh;
  return sh;   // 8, 16 or 32 lanes
}

// tmp[a(i)] = c[i]: the tile's values are bucket-ordered in shared memory through A (tile-local position of
// point i in bucket-major order), then each bucket's run is written contiguously by a group of lanes (full
// sectors, few L2 requests) instead of one scattered 8-byte store per point.
template <typename T>
__global__ void __launch_bounds__(kMaxBuckets)
to_bucket_order_kernel(const T* __restrict__ c, int64_t cstride, const uint16_t* __restrict__ A, int64_t N,
                       const uint32_t* __restrict__ tileStart, int ntiles1, int nb1, int gshift,
                       T* __restrict__ tmp) {
  __shared__ T S[kSortTile];
  __shared__ uint32_t gstart[kMaxBuckets], lstart[kMaxBuckets], wsum[8];
  __shared__ uint16_t longlist[kMaxBuckets];
  __shared__ int nlong;
  const int tid = threadIdx.x;
  const int64_t tile = blockIdx.x;
  if (tid == 0) nlong = 0;
  T v[kSortPer];
  uint32_t lp[kSortPer];
#pragma unroll
  for (int k = 0; k < kSortPer; ++k) {
    int64_t i = tile * kSortTile + k * kSortThreads + tid;
    lp[k] = 0xffffffffu;
    if (i < N) {
      lp[k] = A[i];
      v[k] = c[i * cstride];
    }
  }
  tile_bucket_ranges(tileStart, ntiles1, nb1, N, tile, gstart, lstart, wsum);
#pragma unroll
  for (int k = 0; k < kSortPer; ++k)
    if (lp[k] != 0xffffffffu) S[lp[k]] = v[k];
  __syncthreads();
  const int64_t rem = N - tile * kSortTile;
  const uint32_t n_tile = rem < kSortTile ? (uint32_t)rem : (uint32_t)kSortTile;
  const int G = 1 << gshift, lane = tid & (G - 1);
  const uint32_t kLong = 4u << gshift;       // runs up to here are copied by one lane group
  for (int b = tid >> gshift; b < nb1; b += (kMaxBuckets >> gshift)) {
    const uint32_t ls = lstart[b], gs = gstart[b];
    uint32_t cnt = (b + 1 < nb1 ? lstart[b + 1] : n_tile) - ls;
    cnt = cnt < kLong ? cnt : kLong;
    for (uint32_t j = lane; j < cnt; j += G) tmp[gs + j] = S[ls + j];
  }
  // spatially coherent input (gridded data sets, presorted points, clusters): a tile falls into a few buckets with runs
  // of hundreds of points -- the rest of such a run is copied by the whole CTA
  if (tid < nb1) {
    const uint32_t cb = (tid + 1 < nb1 ? lstart[tid + 1] : n_tile) - lstart[tid];
    if (cb > kLong) longlist[atomicAdd(&nlong, 1)] = (uint16_t)tid;
  }
  __syncthreads();
  for (int q = 0; q < nlong; ++q) {
    const int b = longlist[q];
    const uint32_t ls = lstart[b];
    const uint32_t cnt = (b + 1 < nb1 ? lstart[b + 1] : n_tile) - ls;
    const uint32_t gs = gstart[b];
    for (uint32_t j = kLong + tid; j < cnt; j += kMaxBuckets) tmp[gs + j] = S[ls + j];
  }
}

// tmp[a] = r[B[a]]: slot-order results into bucket order.  a runs sequentially, so B and tmp stream and
// the gathers from r stay inside the current bucket's L2-sized window.  Four positions per thread: one
// 16-byte load of B, four independent gathers in flight.
template <typename T>
__global__ void __launch_bounds__(256)
slot_to_bucket_order_kernel(const T* __restrict__ r, const uint32_t* __restrict__ B, int64_t N,
                            T* __restrict__ tmp) {
  const int64_t a0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (a0 + 3 < N) {
    const uint4 bv = __ldg(reinterpret_cast<const uint4*>(B + a0));
    const T v0 = r[bv.x], v1 = r[bv.y], v2 = r[bv.z], v3 = r[bv.w];
    if constexpr (sizeof(T) == 8) {
      st_global_256(reinterpret_cast<double*>(tmp + a0), v0, v1, v2, v3);
    } else {
      reinterpret_cast<float4*>(tmp + a0)[0] = make_float4(v0, v1, v2, v3);
    }
  } else {
    for (int64_t a = a0; a < N; ++a) tmp[a] = r[B[a]];
  }
}

// out[i] = tmp[a(i)]: each bucket's run of this tile is read contiguously into shared memory by a group of
// lanes, then every point picks its value at its tile-local position A[i] and the output is written coalesced.
template <typename T>
__global__ void __launch_bounds__(kMaxBuckets, 5)
from_bucket_order_kernel(const T* __restrict__ tmp, const uint16_t* __restrict__ A,
                         int64_t N, const uint32_t* __restrict__ tileStart, int ntiles1, int nb1, int gshift,
                         T* __restrict__ out, int64_t ostride) {
  __shared__ T S[kSortTile];
  __shared__ uint32_t gstart[kMaxBuckets], lstart[kMaxBuckets], wsum[8];
  __shared__ uint16_t longlist[kMaxBuckets];
  __shared__ int nlong;
  const int tid = threadIdx.x;
  const int64_t tile = blockIdx.x;
  if (tid == 0) nlong = 0;                  // (tile_bucket_ranges synchronises before anybody adds to it)
  uint32_t lp[kSortPer];
#pragma unroll
  for (int k = 0; k < kSortPer; ++k) {      // issued first: consumed after the staging phase
    int64_t i = tile * kSortTile + k * kSortThreads + tid;
    lp[k] = i < N ? (uint32_t)A[i] : 0xffffffffu;
  }
  tile_bucket_ranges(tileStart, ntiles1, nb1, N, tile, gstart, lstart, wsum);
  const int64_t rem = N - tile * kSortTile;
  const uint32_t n_tile = rem < kSortTile ? (uint32_t)rem : (uint32_t)kSortTile;
  const int G = 1 << gshift, lane = tid & (G - 1), ngroups = kMaxBuckets >> gshift;
  const uint32_t kLong = 4u << gshift;
  // first G elements of 4 runs per lane group at a time: their global loads are in flight together
  for (int b0 = tid >> gshift; b0 < nb1; b0 += 4 * ngroups) {
    uint32_t ls[4], gs[4], cnt[4];
    T v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int b = b0 + q * ngroups;
      cnt[q] = 0;
      if (b < nb1) {
        ls[q] = lstart[b];
        gs[q] = gstart[b];
        cnt[q] = (b + 1 < nb1 ? lstart[b + 1] : n_tile) - ls[q];
        if ((uint32_t)lane < cnt[q]) v[q] = tmp[gs[q] + lane];
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if ((uint32_t)lane < cnt[q]) S[ls[q] + lane] = v[q];
      const uint32_t lim = cnt[q] < kLong ? cnt[q] : kLong;
      for (uint32_t j = lane + G; j < lim; j += G) S[ls[q] + j] = tmp[gs[q] + j];
    }
  }
  // long runs (spatially coherent input: gridded data, presorted points, clusters) are finished by the whole CTA
  // instead of one lane group (measured on Morton-presorted points: 3.4 -> 1.0 ms); the buckets with long runs are
  // listed first so that the usual case (none) costs one shared-memory flag
  if (tid < nb1) {
    const uint32_t cb = (tid + 1 < nb1 ? lstart[tid + 1] : n_tile) - lstart[tid];
    if (cb > kLong) longlist[atomicAdd(&nlong, 1)] = (uint16_t)tid;
  }
  __syncthreads();
  for (int q = 0; q < nlong; ++q) {
    const int b = longlist[q];
    const uint32_t lsb = lstart[b];
    const uint32_t cb = (b + 1 < nb1 ? lstart[b + 1] : n_tile) - lsb;
    const uint32_t gsb = gstart[b];
    for (uint32_t j = kLong + tid; j < cb; j += kMaxBuckets) S[lsb + j] = tmp[gsb + j];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kSortPer; ++k) {
    int64_t i = tile * kSortTile + k * kSortThreads + tid;
    if (lp[k] != 0xffffffffu) out[i * ostride] = S[lp[k]];
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
to_slot_order_kernel(const T* __restrict__ tmp, const T* __restrict__ rec, int RW, int dim, int64_t npos,
                     T* __restrict__ cs) {
  int64_t pos = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= npos) return;
  uint32_t a = *reinterpret_cast<const uint32_t*>(rec + pos * RW + dim);
  cs[pos] = a == 0xffffffffu ? T(0) : tmp[a];
}

void permute_strengths(const SortedPoints& sp, const void* c, int64_t cstride, void* cs, cudaStream_t st) {
  if (sp.npos == 0) return;
  const size_t esz = sp.dtype == 0 ? 4 : 8;
  void* tmp = pool_alloc(esz * (size_t)sp.N);
  const int gsh = run_group_shift(sp.nb1);
  {
    ScopedStage stg("permute.to_bucket_order", st, 1, sp.N);
    if (sp.dtype == 0)
      to_bucket_order_kernel<float><<<sp.ntiles1, kMaxBuckets, 0, st>>>((const float*)c, cstride, sp.A, sp.N,
                                                                        sp.tileStart, sp.ntiles1, sp.nb1, gsh, (float*)tmp);
    else
      to_bucket_order_kernel<double><<<sp.ntiles1, kMaxBuckets, 0, st>>>((const double*)c, cstride, sp.A, sp.N,
                                                                         sp.tileStart, sp.ntiles1, sp.nb1, gsh, (double*)tmp);
  }
  {
    ScopedStage stg("permute.to_slot_order", st, 1, sp.N);
    if (sp.dtype == 0)
      to_slot_order_kernel<float><<<div_up(sp.npos, 256), 256, 0, st>>>((const float*)tmp, (const float*)sp.rec, sp.RW,
                                                                        sp.dim, sp.npos, (float*)cs);
    else
      to_slot_order_kernel<double><<<div_up(sp.npos, 256), 256, 0, st>>>((const double*)tmp, (const double*)sp.rec,
                                                                         sp.RW, sp.dim, sp.npos, (double*)cs);
  }
  EFGP_CUDA(cudaGetLastError());
  pool_free(tmp);  // stream-ordered reuse
}

void unpermute_output(const SortedPoints& sp, const void* r, void* out, int64_t ostride, cudaStream_t st,
                      bool bucket_order) {
  if (sp.N == 0) return;
  const size_t esz = sp.dtype == 0 ? 4 : 8;
  void* tmp = bucket_order ? const_cast<void*>(r) : pool_alloc(esz * (size_t)sp.N);
  const int gsh = run_group_shift(sp.nb1);
  if (!bucket_order) {
    ScopedStage stg("unpermute.slot_to_bucket_order", st, 1, sp.N);
    if (sp.dtype == 0)
      slot_to_bucket_order_kernel<float><<<div_up(sp.N, 1024), 256, 0, st>>>((const float*)r, sp.B, sp.N, (float*)tmp);
    else
      slot_to_bucket_order_kernel<double><<<div_up(sp.N, 1024), 256, 0, st>>>((const double*)r, sp.B, sp.N, (double*)tmp);
  }
  {
    ScopedStage stg("unpermute.from_bucket_order", st, 1, sp.N);
    if (sp.dtype == 0)
      from_bucket_order_kernel<float><<<sp.ntiles1, kMaxBuckets, 0, st>>>((const float*)tmp, sp.A, sp.N, sp.tileStart,
                                                                          sp.ntiles1, sp.nb1, gsh, (float*)out, ostride);
    else
      from_bucket_order_kernel<double><<<sp.ntiles1, kMaxBuckets, 0, st>>>((const double*)tmp, sp.A, sp.N, sp.tileStart,
                                                                           sp.ntiles1, sp.nb1, gsh, (double*)out, ostride);
  }
  EFGP_CUDA(cudaGetLastError());
  if (!bucket_order) pool_free(tmp);  // stream-ordered reuse
}

// ------------------------------------------------------------------ host driver
static int pick_points_per_thread(int64_t N) {
  // The 2-D spread / interp kernels run one CTA per tile of 256 * P points, two CTAs per SM for the narrow kernels
  // and one for the wide ones: the tile count is made a whole number of waves of 2 * (SM count) CTAs (round 1 rounded
  // P down to a power of two: 763 CTAs = 5.15 waves of 148 paid for 6 at the 8-GPU shard size), with the longest
  // register runs (P <= kMaxP) the wave count allows.  EFGP_MAX_P overrides the cap (A/B measurements).
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        sms <= 0)
      sms = 148;
  }
  static const int kMaxP = [] {
    const char* e = std::getenv("EFGP_MAX_P");
    int v = e ? std::atoi(e) : 0;
    return v >= 1 && v <= 4096 ? v : 512;
  }();
  const int64_t wave = 2 * (int64_t)sms;
  const int64_t per_wave = wave * kTileThreads * kMaxP;
  const int64_t k = std::max<int64_t>(1, (N + per_wave - 1) / per_wave);
  const int64_t tiles = k * wave;
  int64_t P = (N + tiles * kTileThreads - 1) / (tiles * kTileThreads);
  if (P < 1) P = 1;
  if (P > kMaxP) P = kMaxP;
  return (int)P;
}

template <typename T, int D>
static void build_td(SortedPoints& sp, const T* x, const T* ypay, int64_t N, const FoldParams<T>& p,
                     const SortGeom& g, cudaStream_t st, bool carry) {
  if (D != 2) ypay = nullptr;   // only the 2-D record has a spare word
  carry = carry && sizeof(T) == 8;
  sp.has_orig = carry;
  uint32_t* RI = carry ? (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)N) : nullptr;
  const int ntiles1 = div_up(N, kSortTile);
  const int64_t nhist = (int64_t)g.nb1 * ntiles1;
  const int64_t nbin = (int64_t)g.nb1 * g.nbins2;
  // scratch (returned to the pool at the end; reuse is stream-ordered)
  uint32_t* tileHist = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)nhist);
  uint32_t* tileStart = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)nhist);   // kept: sp.tileStart
  sp.tileStart = tileStart;
  sp.ntiles1 = ntiles1;
  sp.nb1 = g.nb1;
  uint32_t* meta = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)(2 * (g.nb1 + 1) + 3 * nbin));
  uint32_t* bucketStart = meta;
  uint32_t* tile2Base = meta + (g.nb1 + 1);
  uint32_t* binCount = tile2Base + (g.nb1 + 1);
  uint32_t* binStart = binCount + nbin;
  uint32_t* binCursor = binStart + nbin;
  T* R[3] = {nullptr, nullptr, nullptr};
  for (int a = 0; a < D; ++a) R[a] = (T*)pool_alloc(sizeof(T) * (size_t)N);
  T* RY = ypay ? (T*)pool_alloc(sizeof(T) * (size_t)N) : nullptr;
  sp.has_payload = ypay != nullptr;
  uint16_t* K2 = (uint16_t*)pool_alloc(sizeof(uint16_t) * (size_t)N);
  size_t scan_bytes = 0, sb2 = 0;
  EFGP_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, tileHist, tileStart, (int)nhist, st));
  EFGP_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, sb2, binCount, binStart, (int)nbin, st));
  scan_bytes = std::max(scan_bytes, sb2);
  void* scan_tmp = pool_alloc(scan_bytes);

  sp.A = (uint16_t*)pool_alloc(sizeof(uint16_t) * (size_t)N);
  sp.B = (uint32_t*)pool_alloc(sizeof(uint32_t) * (size_t)N);
  sp.RW = rec_words(D);
  sp.rec = pool_alloc(sizeof(T) * (size_t)sp.RW * (size_t)sp.npos);
  T* rec = (T*)sp.rec;

  {
    ScopedStage stg("sort.bucket_hist_kernel", st, 1, N);
    bucket_hist_kernel<T, D><<<ntiles1, kL1Threads, 0, st>>>(x, N, p, g, ntiles1, tileHist);
  }
  {
    ScopedStage stg("sort.scans", st, 0, 0);
    size_t tb = scan_bytes;
    EFGP_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp, tb, tileHist, tileStart, (int)nhist, st));
    EFGP_CUDA(cudaMemsetAsync(binCount, 0, sizeof(uint32_t) * (size_t)nbin, st));
    EFGP_CUDA(cudaMemsetAsync(binCursor, 0, sizeof(uint32_t) * (size_t)nbin, st));
  }
  {
    ScopedStage stg(ypay ? "sort.bucket_scatter_kernel_pay" : "sort.bucket_scatter_kernel", st, 1, N);
    const int ns = D + (ypay ? 1 : 0);
    const size_t smem = (size_t)ns * kSortTile * sizeof(T) + 2 * kSortTile * sizeof(uint16_t) +
                        3 * kMaxBuckets * sizeof(uint32_t) + (carry ? kSortTile * sizeof(uint32_t) : 0);
    const int smem_max = (int)(4 * kSortTile * sizeof(double) + 2 * kSortTile * sizeof(uint16_t) +
                               3 * kMaxBuckets * sizeof(uint32_t) + kSortTile * sizeof(uint32_t));
    static bool attr_set = false;
    if (!attr_set) {
      EFGP_CUDA(cudaFuncSetAttribute(bucket_scatter_kernel<T, D, D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     smem_max));
      EFGP_CUDA(cudaFuncSetAttribute(bucket_scatter_kernel<T, D, D + 1>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
      attr_set = true;
    }
    // (A persistent variant that prefetched the next tile's inputs during the write-out was measured slower:
    // 1.81 vs 1.31 ms at N = 1e8 -- three independent CTAs per SM overlap the phases better than one CTA's
    // software pipeline.)
    if (ypay) {
      bucket_scatter_kernel<T, D, D + 1><<<ntiles1, kL1Threads, smem, st>>>(
          x, ypay, N, p, g, ntiles1, tileStart, R[0], R[1], RY, nullptr, K2, sp.A, RI);   // D == 2: word 2 = payload
    } else {
      bucket_scatter_kernel<T, D, D><<<ntiles1, kL1Threads, smem, st>>>(x, nullptr, N, p, g, ntiles1, tileStart,
                                                                          R[0], R[1], R[2], nullptr, K2, sp.A, RI);
    }
  }
  const int ntiles2_max = ntiles1 + g.nb1;
  {
    ScopedStage stg("sort.bin_hist_kernel", st, 2, N);
    level2_tiles_kernel<<<1, kMaxBuckets, 0, st>>>(tileStart, ntiles1, g.nb1, N, bucketStart, tile2Base);
    bin_hist_kernel<<<div_up(ntiles2_max, kHistTiles), kSortThreads, g.nbins2 * sizeof(uint32_t), st>>>(
        K2, bucketStart, tile2Base, g, binCount);
  }
  {
    ScopedStage stg("sort.scans", st, 0, 0);
    size_t tb = scan_bytes;
    EFGP_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp, tb, binCount, binStart, (int)nbin, st));
  }
  {
    ScopedStage stg(ypay ? "sort.bin_scatter_kernel_pay" : "sort.bin_scatter_kernel", st, 2, N);
    static bool attr_set2 = false;
    if (!attr_set2) {
      EFGP_CUDA(cudaFuncSetAttribute(bin_scatter_kernel<T, D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)(2 * kMaxBins2 * sizeof(uint32_t))));
      attr_set2 = true;
    }
    bin_scatter_kernel<T, D><<<ntiles2_max, kSortThreads, 2 * g.nbins2 * sizeof(uint32_t), st>>>(
        R[0], R[1], R[2], RY, K2, bucketStart, tile2Base, g, binStart, binCursor, rec, sp.B, RI);
    if (sp.npos > N)
      pad_slots_kernel<T, D><<<div_up(sp.npos - N, 256), 256, 0, st>>>(N, sp.npos, g.P, rec);
  }
  EFGP_CUDA(cudaGetLastError());
  pool_free(scan_tmp);
  if (RI) pool_free(RI);
  pool_free(K2);
  if (RY) pool_free(RY);
  for (int a = 0; a < D; ++a) pool_free(R[a]);
  pool_free(meta);
  pool_free(tileHist);
}

template <typename T>
static void build_t(SortedPoints& sp, const T* x, const T* ypay, int64_t N, int dim, double h, const double* xcen,
                    const int64_t* nf, int odd, cudaStream_t st, bool carry) {
  FoldParams<T> p;
  p.h = (T)h;
  p.shift = odd ? T(0.5) : T(0);
  double cells = 1;
  for (int a = 0; a < kMaxDim; ++a) {
    p.xcen[a] = a < dim ? (T)xcen[a] : T(0);
    p.nf[a] = a < dim ? (T)nf[a] : T(1);
    p.inf[a] = a < dim ? (int)nf[a] : 1;
    if (a < dim) cells *= (double)nf[a];
  }
  for (int a = 0; a < kMaxDim; ++a) p.ntile[a] = a < dim ? (p.inf[a] + kTile3 - 1) / kTile3 : 1;
  if (dim == 3) cells = (double)p.ntile[0] * p.ntile[1] * p.ntile[2] * (kTile3 * kTile3 * kTile3);  // key space
  EFGP_CHECK(cells < 4.0e9, "fine grid has too many cells for 32-bit cell indices");
  EFGP_CHECK(N < (int64_t)2147483000, "N must be < 2^31");
  const uint32_t ncell = (uint32_t)cells;

  sp.P = dim == 3 ? 1 : pick_points_per_thread(N);  // 3-D uses the per-point kernels: keep rank order
  const int64_t tile_len = (int64_t)kTileThreads * sp.P;
  sp.ntiles = (N + tile_len - 1) / tile_len;
  sp.npos = sp.ntiles * tile_len;

  SortGeom g;
  // Level-1 bucket count.  Fewer buckets = longer per-tile runs (2048 / nb1 points: fuller sectors in the bucket
  // scatter and in the un-permute), more buckets = smaller level-2 windows and histograms.  Measured at N = 1e8
  // on 288^2 cells: 256 / 128 / 64 buckets -> 9.39 / 9.03 / 9.08 ms per solve.  Grids with more than
  // 128 * kMaxBins2 cells take up to 256 buckets to keep the level-2 sort exact.  EFGP_SORT_BUCKETS overrides.
  static const int nb1_env = [] {
    const char* e = std::getenv("EFGP_SORT_BUCKETS");
    int v = e ? std::atoi(e) : 0;
    return v >= 1 && v <= kMaxBuckets ? v : 0;
  }();
  int nb1_want = nb1_env ? nb1_env
                         : (int)std::min<uint32_t>(kMaxBuckets, std::max<uint32_t>(kDefaultBuckets, (ncell + kMaxBins2 - 1) / kMaxBins2));
  int nb1 = (int)std::min<uint32_t>((uint32_t)nb1_want, ncell);
  g.cpb = (ncell + nb1 - 1) / nb1;
  g.nb1 = (int)((ncell + g.cpb - 1) / g.cpb);
  g.g = (g.cpb + kMaxBins2 - 1) / kMaxBins2;          // > 1 only for grids above 2M cells (3-D): coarse order
  if (dim == 3) {
    // keep tiles contiguous in the final order: level-2 bins must not straddle tiles, so buckets hold whole
    // tiles and a bin is a power-of-two fraction of a tile (or, for huge grids, a group of whole tiles)
    const uint32_t tcells = kTile3 * kTile3 * kTile3;
    g.cpb = (g.cpb + tcells - 1) / tcells * tcells;
    g.nb1 = (int)((ncell + g.cpb - 1) / g.cpb);
    uint32_t gmin = (g.cpb + kMaxBins2 - 1) / kMaxBins2;
    if (gmin <= tcells) {
      g.g = 1;
      while (g.g < gmin) g.g <<= 1;
    } else {
      g.g = (gmin + tcells - 1) / tcells * tcells;
    }
  }
  g.nbins2 = (int)((g.cpb + g.g - 1) / g.g);
  g.cpb_inv = g.cpb > 1 ? (uint32_t)((1ull << 32) / g.cpb) : 0u;
  g.g_inv = g.g > 1 ? (uint32_t)((1ull << 32) / g.g) : 0u;
  g.P = sp.P;
  g.P_inv = sp.P > 1 ? (uint32_t)((1ull << 32) / (uint32_t)sp.P) : 0u;
  g.tile_inv = (uint32_t)((1ull << 32) / ((uint32_t)kTileThreads * (uint32_t)sp.P));
  if (dim == 1) build_td<T, 1>(sp, x, ypay, N, p, g, st, carry);
  if (dim == 2) build_td<T, 2>(sp, x, ypay, N, p, g, st, carry);
  if (dim == 3) build_td<T, 3>(sp, x, ypay, N, p, g, st, carry);
}

void build_sorted_points(SortedPoints& sp, const void* x, const void* payload, int64_t N, int dim, int dtype,
                         double h, const double* xcen, const int64_t* nf, int odd, cudaStream_t st, bool carry_index) {
  sp.dim = dim;
  sp.dtype = dtype;
  sp.N = N;
  sp.odd = odd;
  for (int a = 0; a < kMaxDim; ++a) sp.nf[a] = a < dim ? nf[a] : 1;
  if (N == 0) return;
  if (dtype == 0)
    build_t<float>(sp, (const float*)x, (const float*)payload, N, dim, h, xcen, nf, odd, st, carry_index);
  else
    build_t<double>(sp, (const double*)x, (const double*)payload, N, dim, h, xcen, nf, odd, st, carry_index);
}

}  // namespace efgp
