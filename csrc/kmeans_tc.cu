// F3 (tensor-core path) -- k-means nearest-centroid assignment on the 5th-gen tensor cores, fused with
// the per-centroid sum / count accumulation.
//
// ref math: machine_learning/k-means.py:20-28 (argmin_c ||x - c||, ties -> lowest index) and :57-63
// (per-cluster vector sum and count).  argmin_c ||x - c||^2 == argmin_c (||c||^2 - 2 x.c); the cross term
// for a 128-row tile against 64 centroids is one bf16 `tcgen05.mma` chain (M=128, N=64, K=d) whose fp32
// accumulator lives in TMEM, the centroid operand holds -2c and ||c||^2 is added in the epilogue.
//
// Persistent, warp-specialised, one CTA per SM (384 threads):
//   warp 0   TMA producer A : 128-row tiles of X (HBM -> smem, 128B swizzle), ring of 5 slots
//   warp 3   TMA producer B : 64-centroid chunks of the prepared centroid operand (L2 -> smem), ring of 3
//   warp 1   MMA issuer     : for each super tile (4 row tiles) x each centroid chunk: 8 x tcgen05.mma
//   warp 2   TMEM allocator (512 columns = 8 accumulator buffers of 64 columns)
//   warps 4-7  arg-min epilogue : tcgen05.ld -> (+||c||^2) -> value with the centroid index packed into
//              the 10 low mantissa bits -> running min per row; the TMEM buffer is released as soon as it
//              has been copied to registers
//   warps 8-11 scatter : red.global.add.v4.f32 of every row of the tile (read back from the swizzled
//              smem tile) into sums[assignment], counts += 1, then the A slot is released
// A super tile keeps 4 row tiles resident while all centroid chunks stream through, so the centroid
// operand is re-read from L2 once per 512 rows instead of once per 128.
//
// Near-ties: packing the index costs 10 mantissa bits (relative 2^-13 of the value), which is below the
// error already introduced by bf16 inputs; exact tie semantics need the CUDA-core path (kmeans.cu).
#include "tc_common.cuh"

using namespace dapb;
using namespace dapb::tc;

namespace {

constexpr int kTM = 128;          // rows per tile (UMMA M)
constexpr int kNB = 64;           // centroids per chunk (UMMA N)
constexpr int kSuper = 4;         // row tiles per super tile
constexpr int kASlots = 5;
constexpr int kBStages = 3;
constexpr int kTBufs = 8;         // TMEM accumulator buffers (8 x 64 columns = 512)
constexpr int kThreads = 384;

template <int KD>
struct SmemLayout {
  static constexpr int kHalves = KD / 64;
  static constexpr uint32_t kATile = kTM * KD * 2;                 // bytes
  static constexpr uint32_t kBChunk = kNB * KD * 2;
  static constexpr uint32_t kAOff = 0;
  static constexpr uint32_t kBOff = kAOff + kASlots * kATile;
  static constexpr uint32_t kCnOff = kBOff + kBStages * kBChunk;   // float[1024]
  static constexpr uint32_t kAsOff = kCnOff + 1024 * 4;            // int[kASlots][128]
  static constexpr uint32_t kBarOff = kAsOff + kASlots * kTM * 4;
  static constexpr int kNumBars = 2 * kASlots + 2 * kBStages + 2 * kTBufs + kASlots;
  static constexpr uint32_t kTmemOff = kBarOff + kNumBars * 8;
  static constexpr uint32_t kTotal = kTmemOff + 16;
};

struct KmTcArgs {
  int64_t n;              // rows
  int k;                  // centroids (multiple of 64, <= 1024)
  const float* cnorm;     // [k]
  float* sums;            // [k, KD]
  int* counts;            // [k]
  int* assign;            // [n] or null
};

template <int KD>
__global__ void __launch_bounds__(kThreads, 1)
kmeans_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_c, KmTcArgs a) {
  using L = SmemLayout<KD>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // dynamic smem is only guaranteed 16-byte aligned: round up to the 1024 B the 128B swizzle needs
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
  uint64_t* a_full = bars;
  uint64_t* a_empty = a_full + kASlots;
  uint64_t* b_full = a_empty + kASlots;
  uint64_t* b_empty = b_full + kBStages;
  uint64_t* t_full = b_empty + kBStages;
  uint64_t* t_empty = t_full + kTBufs;
  uint64_t* as_full = t_empty + kTBufs;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(smem + L::kTmemOff);
  float* cn_s = reinterpret_cast<float*>(smem + L::kCnOff);
  int* assign_s = reinterpret_cast<int*>(smem + L::kAsOff);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t n_tiles = (a.n + kTM - 1) / kTM;
  const int64_t n_super = (n_tiles + kSuper - 1) / kSuper;
  const int n_chunks = a.k / kNB;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kASlots; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 5); mbar_init(&as_full[i], 4); }
    for (int i = 0; i < kBStages; ++i) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 1); }
    for (int i = 0; i < kTBufs; ++i) { mbar_init(&t_full[i], 1); mbar_init(&t_empty[i], 4); }
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(tmem_holder, 512); tmem_relinquish(); }
  for (int i = threadIdx.x; i < a.k; i += kThreads) cn_s[i] = a.cnorm[i];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0) {
    // ===================== TMA producer: X tiles =====================
    if (lane == 0) {
      prefetch_tensormap(&map_x);
      uint32_t g = 0;                                    // tiles loaded by this CTA so far
      for (int64_t s = blockIdx.x; s < n_super; s += gridDim.x) {
        const int64_t tile0 = s * kSuper;
        const int nt = (int)min((int64_t)kSuper, n_tiles - tile0);
        for (int t = 0; t < nt; ++t, ++g) {
          const uint32_t slot = g % kASlots, use = g / kASlots;
          mbar_wait(&a_empty[slot], (use & 1) ^ 1);
          mbar_arrive_expect_tx(&a_full[slot], L::kATile);
          uint8_t* dst = smem + L::kAOff + slot * L::kATile;
          const int row0 = (int)((tile0 + t) * kTM);
#pragma unroll
          for (int h = 0; h < L::kHalves; ++h)
            tma_load_2d_hint(dst + h * (kTM * 128), &map_x, &a_full[slot], h * 64, row0, kEvictFirst);
        }
      }
    }
  } else if (warp == 3) {
    // ===================== TMA producer: centroid chunks =====================
    if (lane == 0) {
      prefetch_tensormap(&map_c);
      uint32_t j = 0;
      for (int64_t s = blockIdx.x; s < n_super; s += gridDim.x) {
        for (int c = 0; c < n_chunks; ++c, ++j) {
          const uint32_t st = j % kBStages, use = j / kBStages;
          mbar_wait(&b_empty[st], (use & 1) ^ 1);
          mbar_arrive_expect_tx(&b_full[st], L::kBChunk);
          uint8_t* dst = smem + L::kBOff + st * L::kBChunk;
#pragma unroll
          for (int h = 0; h < L::kHalves; ++h)
            tma_load_2d_hint(dst + h * (kNB * 128), &map_c, &b_full[st], h * 64, c * kNB, kEvictLast);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(kTM, kNB, 0, 0);
      uint32_t g0 = 0, j = 0, q = 0;                     // tile / chunk / accumulator counters
      for (int64_t s = blockIdx.x; s < n_super; s += gridDim.x) {
        const int64_t tile0 = s * kSuper;
        const int nt = (int)min((int64_t)kSuper, n_tiles - tile0);
        for (int c = 0; c < n_chunks; ++c, ++j) {
          const uint32_t st = j % kBStages;
          mbar_wait(&b_full[st], (j / kBStages) & 1);
          const uint32_t b_addr = smem_u32(smem + L::kBOff + st * L::kBChunk);
          for (int t = 0; t < nt; ++t, ++q) {
            const uint32_t g = g0 + t, slot = g % kASlots;
            if (c == 0) mbar_wait(&a_full[slot], (g / kASlots) & 1);
            const uint32_t buf = q % kTBufs;
            mbar_wait(&t_empty[buf], ((q / kTBufs) & 1) ^ 1);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(smem + L::kAOff + slot * L::kATile);
            const uint32_t d_tmem = tmem_base + buf * kNB;
#pragma unroll
            for (int h = 0; h < L::kHalves; ++h)
#pragma unroll
              for (int ks = 0; ks < 4; ++ks) {
                const uint64_t da = make_smem_desc_sw128(a_addr + h * (kTM * 128) + ks * 32, 16, 1024);
                const uint64_t db = make_smem_desc_sw128(b_addr + h * (kNB * 128) + ks * 32, 16, 1024);
                umma_bf16(d_tmem, da, db, idesc, (h | ks) != 0);
              }
            umma_commit(&t_full[buf]);
            if (c == n_chunks - 1) umma_commit(&a_empty[slot]);     // tensor core is done reading this X tile
          }
          umma_commit(&b_empty[st]);
        }
        g0 += nt;
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ===================== arg-min epilogue =====================
    const int qd = warp & 3;                              // TMEM lane quarter this warp may access
    const int row = qd * 32 + lane;
    uint32_t g0 = 0, q = 0;
    for (int64_t s = blockIdx.x; s < n_super; s += gridDim.x) {
      const int64_t tile0 = s * kSuper;
      const int nt = (int)min((int64_t)kSuper, n_tiles - tile0);
      float best[kSuper];
#pragma unroll
      for (int t = 0; t < kSuper; ++t) best[t] = 3.0e38f;
      for (int c = 0; c < n_chunks; ++c) {
#pragma unroll
        for (int t = 0; t < kSuper; ++t) {
          if (t < nt) {
            const uint32_t buf = q % kTBufs;
            mbar_wait(&t_full[buf], (q / kTBufs) & 1);
            tc_fence_after();
            uint32_t v[kNB];
            const uint32_t taddr = tmem_base + ((uint32_t)(qd * 32) << 16) + buf * kNB;
            tmem_ld_32x32b_x32(taddr, v);
            tmem_ld_32x32b_x32(taddr + 32, v + 32);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&t_empty[buf]);      // accumulator copied out: MMA may overwrite it
            float m = best[t];
            const float4* cn4 = reinterpret_cast<const float4*>(cn_s + c * kNB);
#pragma unroll
            for (int jj = 0; jj < kNB / 4; ++jj) {
              const float4 cn = cn4[jj];
              const uint32_t idx = (uint32_t)(c * kNB + jj * 4);
              float x0 = __uint_as_float(v[jj * 4 + 0]) + cn.x;
              float x1 = __uint_as_float(v[jj * 4 + 1]) + cn.y;
              float x2 = __uint_as_float(v[jj * 4 + 2]) + cn.z;
              float x3 = __uint_as_float(v[jj * 4 + 3]) + cn.w;
              x0 = __uint_as_float((__float_as_uint(x0) & 0xfffffc00u) | (idx + 0));
              x1 = __uint_as_float((__float_as_uint(x1) & 0xfffffc00u) | (idx + 1));
              x2 = __uint_as_float((__float_as_uint(x2) & 0xfffffc00u) | (idx + 2));
              x3 = __uint_as_float((__float_as_uint(x3) & 0xfffffc00u) | (idx + 3));
              m = fminf(fminf(m, x0), fminf(x1, fminf(x2, x3)));
            }
            best[t] = m;
            ++q;
          }
        }
      }
#pragma unroll
      for (int t = 0; t < kSuper; ++t) {
        if (t < nt) {
          const uint32_t g = g0 + t, slot = g % kASlots;
          const int cidx = (int)(__float_as_uint(best[t]) & 1023u);
          assign_s[slot * kTM + row] = cidx;
          const int64_t grow = (tile0 + t) * kTM + row;
          if (a.assign && grow < a.n) a.assign[grow] = cidx;
          __syncwarp();
          if (lane == 0) mbar_arrive(&as_full[slot]);
        }
      }
      g0 += nt;
    }
  } else if (warp >= 8) {
    // ===================== scatter: sums[assign[row]] += x[row], counts += 1 =====================
    const int qd = warp & 3;
    uint32_t g = 0;
    for (int64_t s = blockIdx.x; s < n_super; s += gridDim.x) {
      const int64_t tile0 = s * kSuper;
      const int nt = (int)min((int64_t)kSuper, n_tiles - tile0);
      for (int t = 0; t < nt; ++t, ++g) {
        const uint32_t slot = g % kASlots, par = (g / kASlots) & 1;
        mbar_wait(&a_full[slot], par);                      // acquire the TMA-written tile
        mbar_wait(&as_full[slot], par);
        const uint8_t* tile = smem + L::kAOff + slot * L::kATile;
        const int64_t row_base = (tile0 + t) * kTM;
        for (int i = 0; i < 32; ++i) {
          const int r = qd * 32 + i;
          if (row_base + r >= a.n) break;
          const int cidx = assign_s[slot * kTM + r];
#pragma unroll
          for (int p = 0; p < KD / 128 + (KD < 128 ? 1 : 0); ++p) {
            const int col = p * 128 + lane * 4;             // 4 consecutive features per lane
            if (col < KD) {
              const int h = col >> 6, cb = (col & 63) * 2;  // 64-column half, byte offset inside the 128 B row
              const int chunk = cb >> 4, within = cb & 15;
              const uint2 raw = *reinterpret_cast<const uint2*>(tile + h * (kTM * 128) + r * 128 +
                                                                 ((chunk ^ (r & 7)) << 4) + within);
              float* dst = a.sums + (size_t)cidx * KD + col;
              asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "f"(bf16_lo(raw.x)),
                           "f"(bf16_hi(raw.x)), "f"(bf16_lo(raw.y)), "f"(bf16_hi(raw.y)) : "memory");
            }
          }
          if (lane == 0) atomicAdd(a.counts + cidx, 1);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&a_empty[slot]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

// centres fp32 [k, d] -> tensor-core operand bf16(-2 c) [k, d] and ||bf16(c)||^2 (the norm of the ROUNDED centre,
// so that the assignment is the exact nearest neighbour w.r.t. one consistent set of centres)
__global__ void kmeans_tc_prep_kernel(const float* __restrict__ C, int k, int d, __nv_bfloat16* __restrict__ Bp,
                                      float* __restrict__ cnorm) {
  const int lane = threadIdx.x & 31;
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (c >= k) return;
  float s = 0.f;
  for (int j = lane; j < d; j += 32) {
    const __nv_bfloat16 cb = __float2bfloat16(C[(size_t)c * d + j]);
    const float cf = __bfloat162float(cb);
    s += cf * cf;
    Bp[(size_t)c * d + j] = __float2bfloat16(-2.0f * cf);       // exact: power-of-two scaling
  }
  s = warp_sum(s);
  if (lane == 0) cnorm[c] = s;
}

template <int KD>
int launch_km(const CUtensorMap& mx, const CUtensorMap& mc, const KmTcArgs& a, int grid, cudaStream_t st) {
  constexpr uint32_t smem = SmemLayout<KD>::kTotal + 1024;
  static_assert(smem <= 227 * 1024, "k-means tensor-core kernel exceeds shared memory");
  DAPB_CUDA_CHECK(cudaFuncSetAttribute(kmeans_tc_kernel<KD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kmeans_tc_kernel<KD><<<grid, kThreads, smem, st>>>(mx, mc, a);
  DAPB_LAUNCH_CHECK();
  return 0;
}

}  // namespace

extern "C" int dapb_kmeans_tc_prepare(const float* centers, int32_t k, int32_t d, void* b_operand, float* cnorm,
                                      void* stream) {
  DAPB_REQUIRE(centers && b_operand && cnorm && k >= 1 && d >= 1, "kmeans_tc_prepare: bad arguments");
  const int blocks = (k * 32 + 255) / 256;
  kmeans_tc_prep_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(centers, k, d, (__nv_bfloat16*)b_operand, cnorm);
  DAPB_LAUNCH_CHECK();
  return 0;
}

// X: [n, d] bf16 row-major (16-byte aligned rows); b_operand / cnorm from dapb_kmeans_tc_prepare.
// sums [k, d] fp32 and counts [k] are accumulated atomically (zero them first); assign may be NULL.
extern "C" int dapb_kmeans_assign_tc(const void* X, int64_t n, int32_t d, const void* b_operand, const float* cnorm,
                                     int32_t k, int32_t* assign, float* sums, int32_t* counts, int32_t grid,
                                     void* stream) {
  DAPB_REQUIRE(X && b_operand && cnorm && sums && counts, "kmeans_assign_tc: null argument");
  DAPB_REQUIRE(d == 64 || d == 128, "kmeans_assign_tc: d must be 64 or 128 (got %d)", d);
  DAPB_REQUIRE(k >= 64 && k <= 1024 && k % 64 == 0, "kmeans_assign_tc: k must be a multiple of 64 in [64, 1024] (got %d)", k);
  DAPB_REQUIRE(n >= 1 && n < (1ll << 31) - 256, "kmeans_assign_tc: n out of range for 32-bit TMA coordinates");
  CUtensorMap mx, mc;
  int rc = make_tensor_map_bf16_2d(&mx, X, (uint64_t)n, (uint64_t)d, (uint64_t)d, 64, kTM);
  if (rc) return rc;
  rc = make_tensor_map_bf16_2d(&mc, b_operand, (uint64_t)k, (uint64_t)d, (uint64_t)d, 64, kNB);
  if (rc) return rc;
  KmTcArgs a{n, k, cnorm, sums, counts, assign};
  if (grid <= 0) grid = dapb::sm_count();
  const int64_t n_super = ((n + kTM - 1) / kTM + kSuper - 1) / kSuper;
  if (grid > n_super) grid = (int)n_super;
  cudaStream_t st = (cudaStream_t)stream;
  if (d == 64) return launch_km<64>(mx, mc, a, grid, st);
  return launch_km<128>(mx, mc, a, grid, st);
}
