// Layout shared by the k-means kernels that PRODUCE the tensor-core operands (kmeans_tc.cu: prepare, kmeans.cu: finalize)
// and the assignment kernel that consumes them.
#pragma once
#include <cuda_bf16.h>
#include <cstdint>

namespace dapb {

// cnorm buffer: [0, k_pad) biased norms (fp32), [k_pad, k_pad + 8) aux words (max norm, bias C1), then the NORM SLABS:
// for every centroid j a K = 16 bf16 row (hi, mid, lo, 0 x 13) with hi + mid + lo == cnorm[j] exactly, stored in the
// canonical K-major no-swizzle UMMA form -- 8-row groups of 256 B, the two 16-byte K halves of a group 128 B apart --
// so that a chunk of 128 centroids is one contiguous 4 KB bulk copy into shared memory.
__host__ __device__ inline size_t kmeans_cnorm_floats(int k_pad) { return (size_t)k_pad + 8 + 8 * (size_t)k_pad; }

__device__ __forceinline__ void kmeans_norm_slab_store(float* cnorm, int k_pad, int j, float v) {
  const __nv_bfloat16 h = __float2bfloat16(v);
  const float r1 = v - __bfloat162float(h);
  const __nv_bfloat16 m = __float2bfloat16(r1);
  const __nv_bfloat16 l = __float2bfloat16(r1 - __bfloat162float(m));
  char* slab = reinterpret_cast<char*>(cnorm + k_pad + 8) + (size_t)(j >> 3) * 256 + (size_t)(j & 7) * 16;
  uint4 lo4 = make_uint4(0u, 0u, 0u, 0u);
  lo4.x = (uint32_t)__bfloat16_as_ushort(h) | ((uint32_t)__bfloat16_as_ushort(m) << 16);
  lo4.y = (uint32_t)__bfloat16_as_ushort(l);
  *reinterpret_cast<uint4*>(slab) = lo4;                                       // K = 0 .. 7
  *reinterpret_cast<uint4*>(slab + 128) = make_uint4(0u, 0u, 0u, 0u);          // K = 8 .. 15
}

}  // namespace dapb
