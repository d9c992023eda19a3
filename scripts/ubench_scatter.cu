// Micro-benchmark (design probe, not product code): ONE-level counting sort of N points into ~83k cells with
// global atomic cursors and full-sector (32-byte) record stores, against the two-level sort of points.cu.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gpurun_out/ubench_scatter scripts/ubench_scatter.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("cuda error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ void st256(double* p, double a, double b, double c, double d) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}

__global__ void gen(double* x, double* y, int64_t N, int clustered) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  uint64_t s = (uint64_t)i * 0x9E3779B97F4A7C15ull + 12345;
  s ^= s >> 29; s *= 0xBF58476D1CE4E5B9ull; s ^= s >> 32;
  uint64_t s2 = s * 0x94D049BB133111EBull; s2 ^= s2 >> 31;
  double u = (double)(s >> 11) * (1.0 / 9007199254740992.0), v = (double)(s2 >> 11) * (1.0 / 9007199254740992.0);
  if (clustered) { u = 0.5 + 0.02 * (u - 0.5); v = 0.5 + 0.02 * (v - 0.5); }
  x[2 * i] = u; x[2 * i + 1] = v; y[i] = u - v;
}

template <bool AGG>
__global__ void __launch_bounds__(256) hist(const double* __restrict__ x, int64_t N, int nf, uint32_t* cnt) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  double2 p = reinterpret_cast<const double2*>(x)[i];
  uint32_t cell = (uint32_t)(p.x * nf) * nf + (uint32_t)(p.y * nf);
  if (AGG) {
    unsigned m = __match_any_sync(__activemask(), cell);
    if ((threadIdx.x & 31) == __ffs(m) - 1) atomicAdd(&cnt[cell], (uint32_t)__popc(m));
  } else {
    atomicAdd(&cnt[cell], 1u);
  }
}

__global__ void scan1(const uint32_t* cnt, uint32_t* start, int n) {   // single-thread-block serial scan (not timed)
  if (threadIdx.x == 0) { uint32_t a = 0; for (int i = 0; i < n; ++i) { start[i] = a; a += cnt[i]; } }
}

template <bool AGG, int PER>
__global__ void __launch_bounds__(256) scatter(const double* __restrict__ x, const double* __restrict__ y, int64_t N, int nf,
                                              uint32_t* cursor, double* __restrict__ rec, uint32_t* __restrict__ A) {
  int64_t i0 = ((int64_t)blockIdx.x * blockDim.x) * PER + threadIdx.x;
  double2 p[PER]; double yv[PER]; uint32_t cell[PER], pos[PER];
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    int64_t i = i0 + (int64_t)k * 256;
    if (i < N) { p[k] = reinterpret_cast<const double2*>(x)[i]; yv[k] = y[i]; }
  }
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    int64_t i = i0 + (int64_t)k * 256;
    if (i < N) {
      cell[k] = (uint32_t)(p[k].x * nf) * nf + (uint32_t)(p[k].y * nf);
      if (AGG) {
        unsigned m = __match_any_sync(__activemask(), cell[k]);
        int leader = __ffs(m) - 1, lane = threadIdx.x & 31;
        uint32_t b = 0;
        if (lane == leader) b = atomicAdd(&cursor[cell[k]], (uint32_t)__popc(m));
        b = __shfl_sync(m, b, leader);
        pos[k] = b + __popc(m & ((1u << lane) - 1));
      } else {
        pos[k] = atomicAdd(&cursor[cell[k]], 1u);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    int64_t i = i0 + (int64_t)k * 256;
    if (i < N) {
      st256(rec + (int64_t)pos[k] * 4, p[k].x * nf, p[k].y * nf, __longlong_as_double((long long)i), yv[k]);
      A[i] = pos[k];
    }
  }
}

__global__ void __launch_bounds__(256) gather(const double* __restrict__ tmp, const uint32_t* __restrict__ A, int64_t N,
                                             double* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N) out[i] = tmp[A[i]];
}

// alternative un-permute: scatter in sorted order, out[orig(pos)] = tmp[pos] with orig read from the record
__global__ void __launch_bounds__(256) scatter_out(const double* __restrict__ tmp, const double* __restrict__ rec, int64_t N,
                                                  double* __restrict__ out) {
  int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (a < N) out[__double_as_longlong(rec[a * 4 + 2])] = tmp[a];
}

int main(int argc, char** argv) {
  int64_t N = argc > 1 ? (int64_t)atof(argv[1]) : 100000000;
  int nf = argc > 2 ? atoi(argv[2]) : 288;
  int clustered = argc > 3 ? atoi(argv[3]) : 0;
  int ncell = nf * nf;
  double *x, *y, *rec, *tmp, *out; uint32_t *cnt, *start, *A;
  CK(cudaMalloc(&x, N * 16)); CK(cudaMalloc(&y, N * 8)); CK(cudaMalloc(&rec, N * 32)); CK(cudaMalloc(&tmp, N * 8));
  CK(cudaMalloc(&out, N * 8)); CK(cudaMalloc(&cnt, ncell * 4)); CK(cudaMalloc(&start, ncell * 4)); CK(cudaMalloc(&A, N * 4));
  int nb = (int)((N + 255) / 256);
  gen<<<nb, 256>>>(x, y, N, clustered);
  CK(cudaMemset(tmp, 0, N * 8));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto timeit = [&](const char* name, auto fn, double bytes_per_pt) {
    float best = 1e30f;
    for (int r = 0; r < 4; ++r) {
      CK(cudaMemset(cnt, 0, ncell * 4));
      if (name[0] == 's' && name[1] == 'c' && name[2] == 'a') CK(cudaMemcpy(cnt, start, ncell * 4, cudaMemcpyDeviceToDevice));
      CK(cudaEventRecord(e0)); fn(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (r > 0 && ms < best) best = ms;
    }
    CK(cudaGetLastError());
    printf("%-28s %8.3f ms   %7.0f GB/s (algorithmic %g B/pt)\n", name, best, bytes_per_pt * N / best * 1e-6, bytes_per_pt);
  };
  printf("N=%lld nf=%d cells=%d clustered=%d\n", (long long)N, nf, ncell, clustered);
  timeit("hist", [&] { hist<false><<<nb, 256>>>(x, N, nf, cnt); }, 16);
  timeit("hist_agg", [&] { hist<true><<<nb, 256>>>(x, N, nf, cnt); }, 16);
  hist<false><<<nb, 256>>>(x, N, nf, cnt);   // leave real counts (timeit zeroes first; redo)
  CK(cudaMemset(cnt, 0, ncell * 4)); hist<false><<<nb, 256>>>(x, N, nf, cnt);
  scan1<<<1, 32>>>(cnt, start, ncell); CK(cudaDeviceSynchronize());
  // scatter: timeit copies `start` into cnt (used as the cursor array holding absolute positions)
  timeit("scatter_per1", [&] { scatter<false, 1><<<nb, 256>>>(x, y, N, nf, cnt, rec, A); }, 60);
  timeit("scatter_per4", [&] { scatter<false, 4><<<(nb + 3) / 4, 256>>>(x, y, N, nf, cnt, rec, A); }, 60);
  timeit("scatter_per4_agg", [&] { scatter<true, 4><<<(nb + 3) / 4, 256>>>(x, y, N, nf, cnt, rec, A); }, 60);
  timeit("scatter_per8", [&] { scatter<false, 8><<<(nb + 7) / 8, 256>>>(x, y, N, nf, cnt, rec, A); }, 60);
  timeit("gather out[i]=tmp[A[i]]", [&] { gather<<<nb, 256>>>(tmp, A, N, out); }, 20);
  timeit("gscatter out[orig[a]]=tmp[a]", [&] { scatter_out<<<nb, 256>>>(tmp, rec, N, out); }, 24);
  // sanity: A is a permutation (sum of A == N(N-1)/2) -- checked on a sample
  std::vector<uint32_t> h(1000); CK(cudaMemcpy(h.data(), A, 4000, cudaMemcpyDeviceToHost));
  printf("A[0..3] = %u %u %u %u\n", h[0], h[1], h[2], h[3]);
  return 0;
}
