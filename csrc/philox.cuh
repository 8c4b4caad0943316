// Philox4x32-10 (Salmon et al. 2011).  Host twin: utils/philox.py.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace dapb {

struct Philox4 { uint32_t x, y, z, w; };

__host__ __device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2,
                                                          uint32_t c3, uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
#ifdef __CUDA_ARCH__
    uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
#else
    uint64_t p0 = (uint64_t)M0 * c0, p1 = (uint64_t)M1 * c2;
    uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
    uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
#endif
    uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += W0; k1 += W1;
  }
  return Philox4{c0, c1, c2, c3};
}

// Bernoulli row mask of the SGD samplers: keep iff (word0 >> 8) < threshold (24-bit).
// counter = (row_lo, row_hi, iteration, 0x5347), key = seed.
__host__ __device__ __forceinline__ bool philox_keep_row(uint64_t row, uint32_t iteration,
                                                         uint64_t seed, uint32_t threshold24) {
  Philox4 r = philox4x32_10((uint32_t)row, (uint32_t)(row >> 32), iteration, 0x5347u,
                            (uint32_t)seed, (uint32_t)(seed >> 32));
  return (r.x >> 8) < threshold24;
}

}  // namespace dapb
