// Counter-based Rademacher probes: Philox-4x32-10 (Salmon et al., SC'11) keyed by the caller's seed and counted by
// (slot of the point in the sorted set, probe block).  One call yields 128 bits = the signs of 128 consecutive probes
// at one point, so a kernel that handles NB <= 128 probes of one block evaluates the generator ONCE per point.
// The probes of the gradient's Hutchinson estimator (efgpnd.py:179-181: Z in {+-1}^{T x N}) are regenerated inside the
// spread kernel (type 1) and inside the interpolation epilogue (sum_n Z * alpha) and never stored.
#pragma once
#include <cstdint>

namespace efgp {

__host__ __device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
  const uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
  const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
  const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
  c[1] = (uint32_t)p1;
  c[3] = (uint32_t)p0;
  c[0] = n0;
  c[2] = n2;
}

// 128 random bits for (seed, slot, block)
__host__ __device__ __forceinline__ void philox_bits(uint64_t seed, uint64_t slot, uint32_t block, uint32_t (&out)[4]) {
  uint32_t c[4] = {(uint32_t)slot, (uint32_t)(slot >> 32), block, 0x45464750u /* "EFGP" */};
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) out[q] = c[q];
}

// sign of probe t at `slot`: +1 / -1
template <typename T>
__host__ __device__ __forceinline__ T philox_sign_from_bits(const uint32_t (&bits)[4], int t) {
  const int r = t & 127;
  return ((bits[r >> 5] >> (r & 31)) & 1u) ? T(1) : T(-1);
}

struct ProbeSpec {
  uint64_t seed;
  int t0;         // first probe of the launch
  int nb;         // probes of the launch (all inside one 128-probe block)
  int use_orig;   // counter = ORIGINAL index of the point (64-bit records carry it: results do not depend on the order
                  // inside a cell, on the number of ranks or on the run) instead of its slot in the sorted set
};

// the Philox counter of a point: upper half of the record word that also holds the level-1 position, or the slot
template <typename T>
__host__ __device__ __forceinline__ uint64_t probe_counter(T word, int64_t slot, int use_orig) {
  if constexpr (sizeof(T) == 8) {
    if (use_orig) {
      unsigned long long bits;
#ifdef __CUDA_ARCH__
      bits = (unsigned long long)__double_as_longlong((double)word);
#else
      double w = (double)word;
      __builtin_memcpy(&bits, &w, 8);
#endif
      return (uint64_t)(bits >> 32);
    }
  }
  return (uint64_t)slot;
}

}  // namespace efgp
