// Per-thread cp.async (LDGSTS) ring for streaming the tile-transposed point arrays.
//
// Each thread prefetches ITS OWN next kStages points into private shared-memory slots
// (slot = [stage][array][thread]); because the global layout is tile-transposed, one stage of
// one warp is 32 consecutive words (fully coalesced).  The prefetch depth costs no registers,
// which matters for the 2-D spread/interp kernels whose W*W footprint already fills the
// register file, and needs no barriers: a thread only ever reads slots it filled itself
// (cp.async.wait_group is per thread).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace efgp {

constexpr int kRingStages = 6;

template <int BYTES>
__device__ __forceinline__ void cp_async_word(void* smem_dst, const void* gsrc) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], %2;\n" ::"r"(s), "l"(gsrc), "n"(BYTES) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

}  // namespace efgp
