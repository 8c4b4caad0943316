// Shared device/host helpers for the dapb200 native library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cstdint>
#include <cstdio>
#include <cstring>

#include "../include/dapb200.h"

namespace dapb {

// ------------------------------------------------------------------ error plumbing
void set_last_error(const char* fmt, ...);

#define DAPB_CUDA_CHECK(expr)                                                            \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      ::dapb::set_last_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,               \
                             cudaGetErrorString(_e));                                    \
      return (int)_e;                                                                    \
    }                                                                                    \
  } while (0)

#define DAPB_REQUIRE(cond, ...)                                                          \
  do {                                                                                   \
    if (!(cond)) {                                                                       \
      ::dapb::set_last_error(__VA_ARGS__);                                               \
      return DAPB_ERR_INVALID;                                                           \
    }                                                                                    \
  } while (0)

#define DAPB_LAUNCH_CHECK() DAPB_CUDA_CHECK(cudaGetLastError())

int sm_count();

// ------------------------------------------------------------------ small device utils
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// splitmix64 finaliser; Python twin: parallel/partition.py::hash_owner_u64
__host__ __device__ __forceinline__ uint64_t hash64(uint64_t x) {
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__host__ __device__ __forceinline__ uint32_t hash_owner(uint64_t key, uint32_t world) {
  return (uint32_t)((hash64(key) >> 33) & 0x7FFFFFFFull) % world;
}

// streaming 16-byte load that does not pollute L1 (data is read exactly once)
__device__ __forceinline__ uint4 ld_stream_16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
// same, with an L2 cache-policy operand (createpolicy encodings, as used for TMA cache hints): evict_last for data
// that will be read once more soon, evict_first for its last use
__device__ __forceinline__ uint4 ld_stream_16_hint(const void* p, uint64_t policy) {
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p), "l"(policy));
  return r;
}
__device__ __forceinline__ uint4 ld_stream_16_keep(const void* p) { return ld_stream_16_hint(p, 0x14F0000000000000ull); }
__device__ __forceinline__ uint4 ld_stream_16_last(const void* p) { return ld_stream_16_hint(p, 0x12F0000000000000ull); }
__device__ __forceinline__ float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

// sigmoid variants of the reference: 1/(exp(-z)+1+eps)  (ssgd.py:24 eps=0, ma.py:26 eps=1e-6)
__device__ __forceinline__ float sigmoid_ref(float z, float eps) {
  return 1.0f / (__expf(-z) + 1.0f + eps);   // exp overflow -> inf -> 0, never NaN
}
__device__ __forceinline__ double sigmoid_ref(double z, double eps) {
  return 1.0 / (exp(-z) + 1.0 + eps);
}

}  // namespace dapb
