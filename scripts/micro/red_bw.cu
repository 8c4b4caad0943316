// Microbenchmarks that decide the k-means accumulation design:
//  (1) red.global.add.v4.f32 throughput into a [1024 x 128] fp32 table (random rows), vs scalar red
//  (2) does 3-input min.f32 exist on sm_100a (compile-time), throughput of packed-min epilogue
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("ERR %s line %d\n", cudaGetErrorString(e), __LINE__); return 1;}}while(0)

__device__ __forceinline__ void red_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__global__ void k_red_v4(float* table, int k, int64_t rows_per_warp, uint32_t seed, int copies) {
  int lane = threadIdx.x & 31;
  int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  uint32_t s = seed ^ (uint32_t)(warp * 2654435761u);
  float* base = table + (size_t)(blockIdx.x % copies) * k * 128;
  for (int64_t r = 0; r < rows_per_warp; ++r) {
    s = s * 1664525u + 1013904223u;
    int c = (s >> 10) % k;
    red_v4(base + (size_t)c * 128 + lane * 4, 1.f, 2.f, 3.f, 4.f);
  }
}
__global__ void k_red_s(float* table, int k, int64_t rows_per_warp, uint32_t seed, int copies) {
  int lane = threadIdx.x & 31;
  int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  uint32_t s = seed ^ (uint32_t)(warp * 2654435761u);
  float* base = table + (size_t)(blockIdx.x % copies) * k * 128;
  for (int64_t r = 0; r < rows_per_warp; ++r) {
    s = s * 1664525u + 1013904223u;
    int c = (s >> 10) % k;
#pragma unroll
    for (int j = 0; j < 4; ++j) atomicAdd(base + (size_t)c * 128 + j * 32 + lane, 1.f);
  }
}
// packed-min epilogue ALU throughput: each thread reduces `n` values (from registers regenerated cheaply)
__global__ void k_min2(const float* in, float* out, int n) {
  float m = 3.0e38f;
  float v = in[threadIdx.x & 31];
  for (int i = 0; i < n; i += 4) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float x = __uint_as_float((__float_as_uint(v) & 0xfffffc00u) | (uint32_t)(i + j));
      m = fminf(m, x);
      v += 1.0f;
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = m;
}
__global__ void k_min3(const float* in, float* out, int n) {
  float m = 3.0e38f;
  float v = in[threadIdx.x & 31];
  for (int i = 0; i < n; i += 4) {
    float x0 = __uint_as_float((__float_as_uint(v) & 0xfffffc00u) | (uint32_t)(i)); v += 1.0f;
    float x1 = __uint_as_float((__float_as_uint(v) & 0xfffffc00u) | (uint32_t)(i + 1)); v += 1.0f;
    float x2 = __uint_as_float((__float_as_uint(v) & 0xfffffc00u) | (uint32_t)(i + 2)); v += 1.0f;
    float x3 = __uint_as_float((__float_as_uint(v) & 0xfffffc00u) | (uint32_t)(i + 3)); v += 1.0f;
    float t;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(t) : "f"(x0), "f"(x1), "f"(m));
    asm("min.f32 %0, %1, %2, %3;" : "=f"(m) : "f"(x2), "f"(x3), "f"(t));
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = m;
}
int main() {
  int k = 1024;
  for (int copies : {1, 8}) {
    float* table; CK(cudaMalloc(&table, (size_t)copies * k * 128 * 4)); CK(cudaMemset(table, 0, (size_t)copies * k * 128 * 4));
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    int blocks = 148 * 4, threads = 256; int64_t rpw = 4096;
    double rows = (double)blocks * threads / 32 * rpw;
    for (int which = 0; which < 2; ++which) {
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        if (which == 0) k_red_v4<<<blocks, threads>>>(table, k, rpw, 1234 + rep, copies);
        else k_red_s<<<blocks, threads>>>(table, k, rpw, 1234 + rep, copies);
        cudaEventRecord(b); CK(cudaEventSynchronize(b));
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (rep) printf("%s copies=%d: %.3f ms, %.1f M rows/s, %.1f GB/s payload\n", which == 0 ? "red.v4.f32" : "red.f32 x4", copies, ms, rows / ms / 1e3, rows * 512 / ms / 1e6);
      }
    }
    cudaFree(table);
  }
  float *in, *out; CK(cudaMalloc(&in, 128)); CK(cudaMemset(in, 0, 128)); CK(cudaMalloc(&out, 148 * 8 * 128 * 4));
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  int n = 1 << 20;
  for (int which = 0; which < 2; ++which)
    for (int wps : {1, 2}) {          // warps per SMSP
      int threads = 128 * wps;
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(a);
        if (which == 0) k_min2<<<148, threads>>>(in, out, n); else k_min3<<<148, threads>>>(in, out, n);
        cudaEventRecord(b); CK(cudaEventSynchronize(b));
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (rep) printf("%s warps/SMSP=%d: %.3f ms -> %.2f G elem/s/SM, %.2f elem/clk/SM @1.9GHz\n", which == 0 ? "pack+min2" : "pack+min3", wps, ms,
                        (double)n * threads / ms / 1e6, (double)n * threads / ms / 1e6 / 1.9);
      }
    }
  return 0;
}
