// F3 -- k-means: nearest-centroid assignment fused with per-centroid sum/count accumulation,
// and a finalize kernel that fuses the cross-GPU reduce-scatter, the mean update and the centroid
// all-gather (pushes over NVLink peer memory, csrc/comm.cuh).
//
// ref: machine_learning/k-means.py:20-28 (closest_center: strict '<' => lowest index wins ties),
//      :57-63 (assign + reduceByKey of (sum, count)), :66-71 (mean; clusters that received no point
//      keep their old centre).
//
// This file holds the any-shape CUDA-core kernel (fp32 / bf16 / fp64 points) used for the
// reference's toy data, tests and odd shapes.  The tcgen05 tensor-core assignment kernel for the
// benchmark shape lives in kmeans_tc.cu and feeds the same finalize kernel.
#include "comm.cuh"
#include "kmeans_common.cuh"

using namespace dapb;

namespace {

template <typename TX> __device__ __forceinline__ double to_acc(TX v);
template <> __device__ __forceinline__ double to_acc<double>(double v) { return v; }
template <> __device__ __forceinline__ double to_acc<float>(float v) { return (double)v; }
template <> __device__ __forceinline__ double to_acc<__nv_bfloat16>(__nv_bfloat16 v) { return (double)__bfloat162float(v); }

// TA = accumulation / centre type (double for fp64 points, float otherwise)
template <typename TX, typename TA>
__global__ void __launch_bounds__(256) kmeans_assign_simple_kernel(const TX* __restrict__ X, int64_t n, int d,
                                                                    const TA* __restrict__ C, int k,
                                                                    int* __restrict__ assign,
                                                                    TA* __restrict__ sums, int* __restrict__ counts) {
  extern __shared__ __align__(16) unsigned char sm_raw[];
  TA* Cs = reinterpret_cast<TA*>(sm_raw);
  const bool in_smem = (size_t)k * d * sizeof(TA) <= 96 * 1024;
  if (in_smem) {
    for (int i = threadIdx.x; i < k * d; i += blockDim.x) Cs[i] = C[i];
    __syncthreads();
  }
  const TA* Cc = in_smem ? Cs : C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const TX* x = X + i * d;
    int best = 0;
    TA best_d = (TA)INFINITY;
    for (int c = 0; c < k; ++c) {
      TA dist = 0;
      for (int j = 0; j < d; ++j) {
        TA diff = (TA)to_acc<TX>(x[j]) - Cc[c * d + j];
        dist += diff * diff;
      }
      if (dist < best_d) { best_d = dist; best = c; }      // strict '<' (ref k-means.py:25)
    }
    if (assign) assign[i] = best;
    if (sums) {
      for (int j = 0; j < d; ++j) atomicAdd(&sums[(int64_t)best * d + j], (TA)to_acc<TX>(x[j]));
      atomicAdd(&counts[best], 1);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Finalize (one CTA per rank):
//   1. reduce-scatter: rows of the partial [k, d] sums / [k] counts owned by rank o are pushed into
//      slot[my_rank] of rank o's heap;
//   2. the owner adds the `world` slots in rank order, divides (count > 0) or keeps the old centre;
//   3. all-gather: the owner pushes its finished rows into every rank's gathered-centre buffer;
//   4. every rank copies the gathered buffer into `centers`.
// Payload layout at payload_off (parity p = epoch & 1):
//   RS  : [2][world][rows_max][d + 1]  TA      AG : [2][k][d] TA
// Flags: RS uses [flag_base, flag_base+world), AG uses [flag_base+world, flag_base+2*world).
// ------------------------------------------------------------------------------------------------
// counts travel through the TA-typed heap slots as exact integers: for float the int's BIT PATTERN is stored and the
// owner adds integers (a float sum would stop being exact above 2^24 points per cluster); double holds them exactly.
template <typename TA> __device__ __forceinline__ TA enc_count(int c);
template <> __device__ __forceinline__ float enc_count<float>(int c) { return __int_as_float(c); }
template <> __device__ __forceinline__ double enc_count<double>(int c) { return (double)c; }
__device__ __forceinline__ int dec_count(float v) { return __float_as_int(v); }
__device__ __forceinline__ int dec_count(double v) { return (int)v; }

// Optional tail of the finalize kernel: emit the tensor-core assignment kernel's operands for the NEXT iteration
// (csrc/kmeans_tc.cu: bf16(-2c) rows, biased norms ||bf16(c)||^2 + C1), so that a Lloyd iteration is two launches
// (assign, finalize) and no separate prepare pass exists.
struct KmTcEmit {
  __nv_bfloat16* b_operand;   // [k_pad, d] or null
  float* cnorm;               // [9 * k_pad + 8], see kmeans_common.cuh
  int k_pad;
  float xnorm_max;
};

template <typename TA>
__device__ void kmeans_emit_tc_operand(const TA* centers, int k, int d, const KmTcEmit& e) {
  __shared__ float s_wmax[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  float wmax = 0.f;
  for (int c = warp; c < e.k_pad; c += nw) {
    if (c >= k) {                                     // padding centroid: zero operand row, norm set below
      for (int j = lane; j < d; j += 32) e.b_operand[(size_t)c * d + j] = __float2bfloat16(0.f);
      continue;
    }
    float s = 0.f;
    for (int j = lane; j < d; j += 32) {
      const __nv_bfloat16 cb = __float2bfloat16((float)centers[(size_t)c * d + j]);
      const float cf = __bfloat162float(cb);
      s += cf * cf;
      e.b_operand[(size_t)c * d + j] = __float2bfloat16(-2.0f * cf);       // exact: power-of-two scaling
    }
    s = warp_sum(s);
    if (lane == 0) e.cnorm[c] = s;
    wmax = fmaxf(wmax, s);
  }
  if (lane == 0) s_wmax[warp] = wmax;
  __syncthreads();
  float cmax = 0.f;
  for (int w = 0; w < nw; ++w) cmax = fmaxf(cmax, s_wmax[w]);
  // same bias rule as kmeans_tc_bias_kernel (csrc/kmeans_tc.cu)
  const float M = 1.01f * (cmax + 2.0f * sqrtf(e.xnorm_max * cmax)) + 1e-30f;
  int ex;
  frexpf(M, &ex);
  const float C1 = ldexpf(1.5f, ex + 1);
  for (int j = tid; j < e.k_pad; j += blockDim.x) {
    const float v = j < k ? e.cnorm[j] + C1 : C1 + ldexpf(0.9999f, ex);
    e.cnorm[j] = v;
    kmeans_norm_slab_store(e.cnorm, e.k_pad, j, v);
  }
  if (tid == 0) { e.cnorm[e.k_pad] = cmax; e.cnorm[e.k_pad + 1] = C1; }
}

template <typename TA>
__global__ void __launch_bounds__(1024) kmeans_finalize_kernel(DapbComm c, TA* __restrict__ sums,
                                                                int* __restrict__ counts, TA* __restrict__ centers,
                                                                int k, int d, uint64_t payload_off,
                                                                uint32_t flag_base, uint64_t epoch,
                                                                int* __restrict__ counts_out, int clear_partials,
                                                                KmTcEmit emit) {
  const int tid = threadIdx.x, nt = blockDim.x, W = c.world, me = c.rank;
  const int rows_max = (k + W - 1) / W;
  const int stride = d + 1;
  const int r_lo = min(k, me * rows_max), r_hi = min(k, r_lo + rows_max);
  if (W == 1) {
    for (int i = tid; i < k * d; i += nt) {
      const int row = i / d, cnt = counts[row];
      if (cnt > 0) centers[i] = sums[i] / (TA)cnt;          // ref k-means.py:66-71
      if (clear_partials) sums[i] = 0;
    }
    __syncthreads();
    for (int i = tid; i < k; i += nt) {
      if (counts_out) counts_out[i] = counts[i];
      if (clear_partials) counts[i] = 0;
    }
    if (emit.b_operand) { __syncthreads(); kmeans_emit_tc_operand<TA>(centers, k, d, emit); }
    return;
  }
  const uint64_t par = epoch & 1;
  const uint64_t rs_elems = (uint64_t)W * rows_max * stride;
  const uint64_t ag_base_bytes = 2 * rs_elems * sizeof(TA);
  // ---- 1. push my partial rows to their owners -------------------------------------------------
  for (int o = 0; o < W; ++o) {
    const int o_lo = min(k, o * rows_max), o_hi = min(k, o_lo + rows_max);
    TA* slot = reinterpret_cast<TA*>(comm_payload(c, o, payload_off)) + par * rs_elems + (uint64_t)me * rows_max * stride;
    const int n_el = (o_hi - o_lo) * stride;
    for (int i = tid; i < n_el; i += nt) {
      const int row = i / stride, col = i - row * stride;
      slot[i] = col < d ? sums[(int64_t)(o_lo + row) * d + col] : enc_count<TA>(counts[o_lo + row]);
    }
  }
  block_signal_all(c, me, flag_base, epoch);
  block_wait_all(c, me, flag_base, epoch);
  // ---- 2+3. reduce my rows in rank order, update, push to every rank (multimem.st when the heap is multicast-mapped) ----
  const TA* mine = reinterpret_cast<const TA*>(comm_payload(c, me, payload_off)) + par * rs_elems;
  const uint64_t ag_off = payload_off + ag_base_bytes + par * (uint64_t)k * stride * sizeof(TA);
  const int n_rows = r_hi - r_lo;
  for (int i = tid; i < n_rows * d; i += nt) {
    const int row = i / d, col = i - row * d;
    TA s = 0;
    int cnt = 0;
    for (int r = 0; r < W; ++r) {
      const TA* sl = mine + (uint64_t)r * rows_max * stride + (uint64_t)row * stride;
      s += __ldcg(sl + col);
      cnt += dec_count(__ldcg(sl + d));
    }
    const TA v = cnt > 0 ? s / (TA)cnt : centers[(int64_t)(r_lo + row) * d + col];
    push_all<TA>(c, ag_off, (uint64_t)(r_lo + row) * stride + col, v);
    if (col == 0) push_all<TA>(c, ag_off, (uint64_t)(r_lo + row) * stride + d, enc_count<TA>(cnt));
  }
  block_signal_all(c, me, flag_base + W, epoch);
  block_wait_all(c, me, flag_base + W, epoch);
  // ---- 4. adopt the gathered centres --------------------------------------------------------------
  const TA* ag = reinterpret_cast<const TA*>(comm_payload(c, me, ag_off));
  for (int i = tid; i < k * d; i += nt) {
    const int row = i / d, col = i - row * d;
    centers[i] = __ldcg(ag + (int64_t)row * stride + col);
    if (clear_partials) sums[i] = 0;
  }
  for (int i = tid; i < k; i += nt) {
    if (counts_out) counts_out[i] = dec_count(__ldcg(ag + (int64_t)i * stride + d));
    if (clear_partials) counts[i] = 0;
  }
  if (emit.b_operand) { __syncthreads(); kmeans_emit_tc_operand<TA>(centers, k, d, emit); }
}

template <typename TX, typename TA>
int launch_assign(const void* X, int64_t n, int d, const void* C, int k, int* assign, void* sums, int* counts,
                  cudaStream_t s) {
  size_t smem = (size_t)k * d * sizeof(TA) <= 96 * 1024 ? (size_t)k * d * sizeof(TA) : 0;
  DAPB_CUDA_CHECK(cudaFuncSetAttribute(kmeans_assign_simple_kernel<TX, TA>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
  int64_t blocks = (n + 255) / 256;
  const int64_t cap = (int64_t)dapb::sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  kmeans_assign_simple_kernel<TX, TA><<<(unsigned)blocks, 256, smem, s>>>(
      (const TX*)X, n, d, (const TA*)C, k, assign, (TA*)sums, counts);
  DAPB_LAUNCH_CHECK();
  return 0;
}

}  // namespace

extern "C" int dapb_kmeans_assign(const void* X, int64_t n, int32_t d, int32_t dtype, const void* centers,
                                  int32_t k, int32_t* assign, void* sums, int32_t* counts, void* stream) {
  DAPB_REQUIRE(n >= 0 && d >= 1 && k >= 1, "kmeans_assign: bad shape");
  DAPB_REQUIRE((sums == nullptr) == (counts == nullptr), "kmeans_assign: sums and counts go together");
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case 0: return launch_assign<float, float>(X, n, d, centers, k, assign, sums, counts, s);
    case 1: return launch_assign<__nv_bfloat16, float>(X, n, d, centers, k, assign, sums, counts, s);
    case 2: return launch_assign<double, double>(X, n, d, centers, k, assign, sums, counts, s);
    default: DAPB_REQUIRE(false, "kmeans_assign: dtype must be 0 (fp32), 1 (bf16) or 2 (fp64)");
  }
  return 0;
}

extern "C" uint64_t dapb_kmeans_payload_bytes(int32_t world, int32_t k, int32_t d, int32_t elem_bytes) {
  const uint64_t rows_max = (k + world - 1) / world;
  return (2ull * world * rows_max * (d + 1) + 2ull * k * (d + 1)) * elem_bytes;
}

static int kmeans_finalize_impl(const DapbComm* comm, void* sums, int32_t* counts, void* centers, int32_t k,
                                int32_t d, int32_t f64, uint64_t payload_off, uint32_t flag_base,
                                uint64_t epoch, int32_t* counts_out, int32_t clear_partials, KmTcEmit emit, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "kmeans_finalize: bad comm");
  DAPB_REQUIRE(epoch >= 1, "kmeans_finalize: epoch starts at 1");
  if (comm->world > 1) {
    uint64_t need = DAPB_FLAG_BYTES + payload_off + dapb_kmeans_payload_bytes(comm->world, k, d, f64 ? 8 : 4);
    DAPB_REQUIRE(need <= comm->heap_bytes, "kmeans_finalize: heap too small (%llu > %llu)",
                 (unsigned long long)need, (unsigned long long)comm->heap_bytes);
  }
  cudaStream_t s = (cudaStream_t)stream;
  if (f64)
    kmeans_finalize_kernel<double><<<1, 1024, 0, s>>>(*comm, (double*)sums, counts, (double*)centers, k, d,
                                                      payload_off, flag_base, epoch, counts_out, clear_partials, emit);
  else
    kmeans_finalize_kernel<float><<<1, 1024, 0, s>>>(*comm, (float*)sums, counts, (float*)centers, k, d,
                                                     payload_off, flag_base, epoch, counts_out, clear_partials, emit);
  DAPB_LAUNCH_CHECK();
  return 0;
}

extern "C" int dapb_kmeans_finalize(const DapbComm* comm, void* sums, int32_t* counts, void* centers, int32_t k,
                                    int32_t d, int32_t f64, uint64_t payload_off, uint32_t flag_base,
                                    uint64_t epoch, int32_t* counts_out, int32_t clear_partials, void* stream) {
  KmTcEmit e;
  memset(&e, 0, sizeof(e));
  return kmeans_finalize_impl(comm, sums, counts, centers, k, d, f64, payload_off, flag_base, epoch, counts_out,
                              clear_partials, e, stream);
}

// Same, and the finished centres are also written as the tensor-core assignment kernel's operands (b_operand / cnorm as
// in dapb_kmeans_tc_prepare), so the next dapb_kmeans_assign_tc needs no prepare launch.  fp32 centres only.
extern "C" int dapb_kmeans_finalize_tc(const DapbComm* comm, void* sums, int32_t* counts, void* centers, int32_t k,
                                       int32_t d, uint64_t payload_off, uint32_t flag_base, uint64_t epoch,
                                       int32_t* counts_out, int32_t clear_partials, void* b_operand, float* cnorm,
                                       float xnorm_max, void* stream) {
  DAPB_REQUIRE(b_operand && cnorm, "kmeans_finalize_tc: operand buffers required");
  DAPB_REQUIRE(xnorm_max >= 0.f && xnorm_max < 1e30f, "kmeans_finalize_tc: xnorm_max must be a finite bound on ||x||^2");
  KmTcEmit e;
  e.b_operand = (__nv_bfloat16*)b_operand; e.cnorm = cnorm; e.k_pad = (k + 127) / 128 * 128; e.xnorm_max = xnorm_max;
  return kmeans_finalize_impl(comm, sums, counts, centers, k, d, 0, payload_off, flag_base, epoch, counts_out,
                              clear_partials, e, stream);
}
