// F7 -- Monte-Carlo pi: on-device Philox sampler fused with warp/block/grid reduction.
// ref: randomized_algorithm/monte_carlo.py:17-20 (is_accept: x,y ~ U(-1,1), x^2+y^2 <= 1) and :28
// (map + reduce(add)).  The reference draws with CPython's global Mersenne Twister, unseeded and
// per worker process; here sample i uses Philox4x32-10 counter (i/2) -- each call yields two
// points -- so any sub-range can be recomputed on the host (utils/philox.py).
//
// Exactness: a 32-bit word u is read as the signed integer s = (int32)u, i.e. the coordinate
// s / 2^31 in [-1, 1).  Acceptance x^2 + y^2 <= 1 is evaluated as sx^2 + sy^2 <= 2^62 in 64-bit
// integer arithmetic, so the device count is bit-reproducible on the host.
#include "common.cuh"
#include "philox.cuh"

using namespace dapb;

namespace {

__device__ __forceinline__ unsigned accept_point(uint32_t ux, uint32_t uy) {
  const long long sx = (int)ux, sy = (int)uy;
  const unsigned long long r2 = (unsigned long long)(sx * sx) + (unsigned long long)(sy * sy);
  return r2 <= (1ull << 62);
}

__global__ void __launch_bounds__(256) mc_pi_kernel(uint64_t first, uint64_t n, uint64_t seed,
                                                     unsigned long long* count_io) {
  // counters [c0, c1): counter k covers samples 2k and 2k+1
  const uint64_t c_first = first >> 1, c_last = (first + n + 1) >> 1;   // exclusive
  const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  unsigned long long local = 0;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t k = c_first + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; k < c_last; k += stride) {
    Philox4 r = philox4x32_10((uint32_t)k, (uint32_t)(k >> 32), 0u, 0x4D43u, k0, k1);
    const uint64_t s0 = 2 * k, s1 = 2 * k + 1;
    if (s0 >= first && s0 < first + n) local += accept_point(r.x, r.y);
    if (s1 >= first && s1 < first + n) local += accept_point(r.z, r.w);
  }
  local = warp_sum(local);
  __shared__ unsigned long long s_part[8];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) s_part[wid] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += s_part[w];
    if (t) atomicAdd(count_io, t);
  }
}

}  // namespace

extern "C" int dapb_mc_pi(uint64_t first, uint64_t n, uint64_t seed, unsigned long long* count_io, void* stream) {
  DAPB_REQUIRE(count_io != nullptr, "mc_pi: null counter");
  if (n == 0) return 0;
  const uint64_t counters = (n + 3) / 2;
  uint64_t blocks = (counters + 255) / 256;
  const uint64_t cap = (uint64_t)dapb::sm_count() * 8;      // 8 resident CTAs of 256 threads per SM
  if (blocks > cap) blocks = cap;
  mc_pi_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(first, n, seed, count_io);
  DAPB_LAUNCH_CHECK();
  return 0;
}
