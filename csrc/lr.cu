// F1 / F2 -- logistic-regression optimiser kernels (SSGD, full-batch GD, MA, BMUF, EASGD).
//
// ref math: optimization/ssgd.py:23-33 (sigmoid, per-row gradient), :97-105 (sample, sum, update);
//           machine_learning/logistic_regression.py:82-85 (sum, not mean);
//           optimization/ma.py:39-43,96-106; optimization/bmuf.py:111-114;
//           optimization/easgd.py:41-45,106.
//
// Two kernels:
//  * lr_parity_kernel  -- the reference's default fp64 problem, the WHOLE training loop in one
//    persistent cooperative launch (one CTA per logical Spark partition, in-kernel exchange).
//  * lr_step_kernel    -- one optimiser step at benchmark scale (bf16/fp32 rows streamed once
//    from HBM, forward + backward from the same registers, deterministic grid reduction, then the
//    last CTA performs the cross-GPU all-reduce over peer memory and the optimiser update).
#include <cstdlib>
#include <cooperative_groups.h>

#include <mutex>

#include "comm.cuh"
#include "philox.cuh"
#include "tc_common.cuh"

using namespace dapb;

// =================================================================================================
// Parity kernel
// =================================================================================================
namespace {

__device__ __forceinline__ double reg_grad(double w, int reg_type, double a) {
  // ref ssgd.py:36-47
  if (reg_type == DAPB_REG_L2) return w;
  double s = (w > 0.0) - (w < 0.0);
  if (reg_type == DAPB_REG_L1) return s;
  if (reg_type == DAPB_REG_ELASTIC) return a * s + (1.0 - a) * w;
  return 0.0;
}

__global__ void __launch_bounds__(256) lr_parity_kernel(DapbLrParityArgs a, DapbComm c) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int D = a.D;
  double* w = reinterpret_cast<double*>(smem_raw);   // global / centre weights (replicated)
  double* wl = w + D;                                 // this partition's replica
  double* contrib = wl + D;                           // what this partition contributes this round
  double* dlt = contrib + D;                          // BMUF block momentum
  double* rlist = dlt + D;                            // [max_part_rows] residuals of kept rows
  int* idx = reinterpret_cast<int*>(rlist + a.max_part_rows);   // [max_part_rows] kept row ids
  __shared__ int s_cnt, s_correct;

  const int p_local = blockIdx.x;
  const int gp = a.part_base + p_local;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = blockDim.x >> 5;
  const int64_t r0 = a.part_off[p_local], r1 = a.part_off[p_local + 1];
  const int n_p = (int)(r1 - r0);
  const bool local_algo = a.algo >= DAPB_LR_MA;
  const int stride = D + 1;                           // slot = D values + count

  for (int d = tid; d < D; d += blockDim.x) {
    w[d] = a.w[d];
    wl[d] = a.ws ? a.ws[(int64_t)p_local * D + d] : a.w[d];
    dlt[d] = a.delta_w ? a.delta_w[d] : 0.0;
  }
  __syncthreads();

  for (int t = 0; t < a.T; ++t) {
    // ---- 1. compact the kept rows of this partition, in row order (one warp) -----------------
    if (wid == 0) {
      int cnt = 0;
      const uint8_t* m = a.masks ? a.masks + (int64_t)t * a.n_local + r0 : nullptr;
      for (int base = 0; base < n_p; base += 32) {
        int i = base + lane;
        bool keep = i < n_p && (m == nullptr || m[i] != 0);
        unsigned b = __ballot_sync(0xffffffffu, keep);
        if (keep) idx[cnt + __popc(b & ((1u << lane) - 1))] = i;
        cnt += __popc(b);
      }
      if (lane == 0) s_cnt = cnt;
    }
    __syncthreads();
    const int cnt = s_cnt;

    // ---- 2. local work ------------------------------------------------------------------------
    if (local_algo && a.algo != DAPB_LR_EASGD)
      for (int d = tid; d < D; d += blockDim.x) wl[d] = w[d];          // ref ma.py:96 reset
    __syncthreads();
    const int n_inner = local_algo ? (a.algo == DAPB_LR_EASGD ? 1 : a.n_local_iters) : 1;
    for (int li = 0; li < n_inner; ++li) {
      const double* wf = local_algo ? wl : w;
      // residuals r_i = sigma(x_i . w) - y_i   (ref ssgd.py:33: -(y - f))
      for (int k = wid; k < cnt; k += nw) {
        const double* x = a.X + (r0 + idx[k]) * (int64_t)D;
        double s = 0.0;
        for (int d = lane; d < D; d += 32) s += x[d] * wf[d];
        s = warp_sum(s);
        if (lane == 0) rlist[k] = -(a.y[r0 + idx[k]] - sigmoid_ref(s, a.eps));
      }
      __syncthreads();
      // ordered left fold over kept rows (Spark seqOp order), one thread per feature
      for (int d = tid; d < D; d += blockDim.x) {
        double g = 0.0;
        for (int k = 0; k < cnt; ++k)
          g = __dadd_rn(g, __dmul_rn(rlist[k], a.X[(r0 + idx[k]) * (int64_t)D + d]));
        if (!local_algo) {
          contrib[d] = g;
        } else if (cnt > 0) {                  // empty sample: the reference raises (ma.py:41); we skip
          double gm = g / (double)cnt;         // ref ma.py:42 np.mean
          double nv = wl[d] - a.eta * gm;      // ref ma.py:43
          if (a.algo == DAPB_LR_EASGD) nv -= a.alpha * (wl[d] - w[d]);   // ref easgd.py:45
          contrib[d] = nv;                     // staged, committed to wl after the barrier
        } else {
          contrib[d] = wl[d];
        }
      }
      __syncthreads();
      if (local_algo) {
        for (int d = tid; d < D; d += blockDim.x) wl[d] = contrib[d];
        __syncthreads();
      }
    }

    // ---- 3. exchange: push (contrib, cnt) to slot[gp] on every rank, partition-ordered sum ----
    const uint64_t epoch = a.epoch_base + (uint64_t)t + 1;
    const uint64_t par = (uint64_t)(epoch & 1) * a.P_global * stride;
    for (int r = 0; r < c.world; ++r) {
      double* slot = reinterpret_cast<double*>(comm_payload(c, r, a.payload_off)) + par + (uint64_t)gp * stride;
      for (int d = tid; d < D; d += blockDim.x) slot[d] = contrib[d];
      if (tid == 0) slot[D] = (double)cnt;
    }
    __syncthreads();
    if (tid == 0) __threadfence_system();
    __syncthreads();
    if (tid < c.world) st_release_sys(comm_flags(c, tid) + a.flag_base + gp, epoch);
    for (int q = tid; q < a.P_global; q += blockDim.x)
      comm_wait_flag(comm_flags(c, c.rank) + a.flag_base + q, epoch);
    __syncthreads();
    const double* slots = reinterpret_cast<const double*>(comm_payload(c, c.rank, a.payload_off)) + par;
    double cnt_tot = 0.0;
    for (int q = 0; q < a.P_global; ++q) cnt_tot += __ldcg(slots + (uint64_t)q * stride + D);

    // ---- 4. replicated update (identical on every CTA of every rank) --------------------------
    for (int d = tid; d < D; d += blockDim.x) {
      double S = __ldcg(slots + d);
      for (int q = 1; q < a.P_global; ++q) S = __dadd_rn(S, __ldcg(slots + (uint64_t)q * stride + d));
      double wd = w[d];
      switch (a.algo) {
        case DAPB_LR_SSGD:
          if (cnt_tot > 0.0) wd -= a.eta * (S / cnt_tot + a.lam * reg_grad(wd, a.reg_type, a.reg_alpha));
          break;                                                          // ref ssgd.py:105
        case DAPB_LR_GD: wd -= a.eta * S; break;                          // ref logistic_regression.py:85
        case DAPB_LR_MA: wd = S / (double)a.P_global; break;              // ref ma.py:106
        case DAPB_LR_BMUF: {                                              // ref bmuf.py:111-114
          double w_avg = S / (double)a.P_global;
          double nd = a.mu * dlt[d] + a.zeta * (w_avg - wd);
          dlt[d] = nd;
          wd = wd + nd;
        } break;
        default:                                                          // ref easgd.py:106
          wd = (1.0 - a.beta) * wd + a.beta * S / (double)a.P_global;
      }
      w[d] = wd;
    }
    __syncthreads();

    // ---- 5. metrics by the CTA that owns global partition 0 (ref ssgd.py:107-112) -------------
    if (gp == 0) {
      if (tid == 0) s_correct = 0;
      __syncthreads();
      int ok = 0;
      for (int i = wid; i < a.n_test; i += nw) {
        const double* x = a.X_test + (int64_t)i * D;
        double s = 0.0;
        for (int d = lane; d < D; d += 32) s += x[d] * w[d];
        s = warp_sum(s);
        int pred = sigmoid_ref(s, a.eps) < 0.5 ? 0 : 1;
        ok += (pred == a.y_test[i]);
      }
      if (lane == 0 && ok) atomicAdd(&s_correct, ok);
      __syncthreads();
      if (tid == 0) {
        a.accs[t] = (double)s_correct / (double)a.n_test;
        a.batch_sizes[t] = (int)cnt_tot;
      }
      if (a.w_hist)
        for (int d = tid; d < D; d += blockDim.x) a.w_hist[(int64_t)t * D + d] = w[d];
      __syncthreads();
    }
  }

  for (int d = tid; d < D; d += blockDim.x) {
    if (p_local == 0) {
      a.w[d] = w[d];
      if (a.delta_w) a.delta_w[d] = dlt[d];
    }
    if (a.ws) a.ws[(int64_t)p_local * D + d] = wl[d];
  }
}

}  // namespace

extern "C" uint64_t dapb_lr_parity_payload_bytes(int32_t P_global, int32_t D) {
  return (uint64_t)2 * P_global * (D + 1) * sizeof(double);
}

extern "C" int dapb_lr_parity_run(const DapbLrParityArgs* args, const DapbComm* comm, void* stream) {
  DAPB_REQUIRE(args && comm, "lr_parity_run: null argument");
  DapbLrParityArgs a = *args;
  DapbComm c = *comm;
  DAPB_REQUIRE(c.world >= 1 && c.world <= DAPB_MAX_RANKS, "lr_parity_run: bad world %d", c.world);
  DAPB_REQUIRE(a.P_local >= 1 && a.P_local <= dapb::sm_count(),
               "lr_parity_run: P_local=%d must be in [1, #SM] (co-resident CTAs)", a.P_local);
  DAPB_REQUIRE(a.P_global >= a.P_local && a.D >= 1 && a.T >= 0, "lr_parity_run: bad sizes");
  DAPB_REQUIRE(a.algo >= DAPB_LR_SSGD && a.algo <= DAPB_LR_EASGD, "lr_parity_run: bad algo %d", a.algo);
  uint64_t need = DAPB_FLAG_BYTES + a.payload_off + dapb_lr_parity_payload_bytes(a.P_global, a.D);
  DAPB_REQUIRE(need <= c.heap_bytes, "lr_parity_run: heap too small (%llu > %llu)",
               (unsigned long long)need, (unsigned long long)c.heap_bytes);
  DAPB_REQUIRE((a.flag_base + a.P_global) * 8ull <= DAPB_FLAG_BYTES, "lr_parity_run: flag range");
  size_t smem = (size_t)4 * a.D * sizeof(double) + (size_t)a.max_part_rows * (sizeof(double) + sizeof(int)) + 16;
  DAPB_REQUIRE(smem <= 200 * 1024, "lr_parity_run: partition too large for the parity kernel (%zu B smem); "
               "use dapb_lr_step for large data", smem);
  DAPB_CUDA_CHECK(cudaFuncSetAttribute(lr_parity_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  void* kargs[] = {&a, &c};
  DAPB_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)lr_parity_kernel, dim3(a.P_local), dim3(256), kargs, smem,
                                              (cudaStream_t)stream));
  return 0;
}

// =================================================================================================
// Benchmark-scale step kernel
// =================================================================================================
namespace {

constexpr int kStepThreads = 256;
constexpr int kStepWarps = kStepThreads / 32;

template <typename T> struct RowVec;   // 16-byte vector of T -> floats
template <> struct RowVec<float> {
  static constexpr int kElems = 4;
  __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
    f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y);
    f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
  }
};
template <> struct RowVec<__nv_bfloat16> {
  static constexpr int kElems = 8;
  __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
    f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
    f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
  }
};

__device__ __forceinline__ float reg_grad_f(float w, int reg_type, float a) {
  if (reg_type == DAPB_REG_L2) return w;
  float s = (float)((w > 0.f) - (w < 0.f));
  if (reg_type == DAPB_REG_L1) return s;
  if (reg_type == DAPB_REG_ELASTIC) return a * s + (1.f - a) * w;
  return 0.f;
}

// ---------------------------------------------------------------------------------------------------------------
// Grid finish, slice-parallel: the LAST kFinSlices CTAs to arrive each take one column slice of the model and run, for
// that slice only: the deterministic sum of the G partial rows -> the in-kernel all-reduce over NVLink peers -> the
// optimiser update.  (One CTA doing all of it was a 40-50 us serial tail per step: 1.2 MB of partials through one SM,
// then one all-reduce; eight slices shrink both eight-fold and run their eight small all-reduces concurrently.)
//   * election: every CTA takes a ticket after publishing its partial row; tickets >= G - S are finalizers, slice =
//     ticket - (G - S).  A finalizer spins until the ticket reaches G (all rows published; the stragglers it waits for
//     are running or schedulable because at most S CTAs ever spin).
//   * every finalizer sums the G counts itself (the update of each slice needs the global batch size).
//   * finalizers bump the ticket once more when done; the one that brings it to G + S resets it for the next launch.
//   * programmatic dependent launch: finalizers let the NEXT step kernel start (griddepcontrol.launch_dependents);
//     that kernel samples and L2-prefetches its first rows, then blocks in griddepcontrol.wait until this grid has
//     completed and flushed, i.e. until the new w is visible.
// Payload per slice: 2 (parity) x world x (cps + 4) floats at payload_off + slice * that; flags flag_base + slice*world + r.
// sm: shared scratch of >= 5 * blockDim.x + cps + 8 floats.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kFinSlices = 8;
__host__ __device__ __forceinline__ int lr_cols_per_slice(int D) { return (((D + kFinSlices - 1) / kFinSlices) + 3) & ~3; }

__device__ __forceinline__ unsigned ld_acquire_gpu_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ void lr_finish_slice(const DapbLrStepArgs& a, const DapbComm& c, float* sm, int slice) {
  const int D = a.D, tid = threadIdx.x, nt = blockDim.x;
  const int G = gridDim.x;
  const int cps = lr_cols_per_slice(D);
  const int c0 = min(D, slice * cps), c1 = min(D, c0 + cps), nc = c1 - c0;
  float* g_s = sm;                         // [cps + 4]: slice of the gradient, then the count in two 16-bit halves
  float* red = sm + cps + 8;               // [nt * 4] cross-row-group scratch
  __shared__ int s_cnt_local;
  unsigned long long t0 = 0;
  if (a.stats && slice == 0 && tid == 0) t0 = global_timer_ns();

  // ---- deterministic sum of the G partial rows over this slice's columns (fixed thread layout, fixed order) ----
  const bool vec = (D & 3) == 0 && ((reinterpret_cast<uintptr_t>(a.partials) & 15) == 0);
  const int unit = vec ? 4 : 1;
  const int nvec = nc / unit;              // nc is a multiple of 4 whenever vec (cps and D are)
  for (int v0 = 0; v0 < nvec; v0 += 32) {
    const int cw = min(32, nvec - v0);     // vectors handled in this pass
    const int RG = nt / cw;                // row groups
    const int rg = tid / cw, cv = tid - rg * cw;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (rg < RG) {
      const int col = c0 + (v0 + cv) * unit;
      if (vec) {
#pragma unroll 8
        for (int b = rg; b < G; b += RG) {
          const float4 v = __ldcg(reinterpret_cast<const float4*>(a.partials + (int64_t)b * D + col));
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
      } else {
#pragma unroll 8
        for (int b = rg; b < G; b += RG) acc.x += __ldcg(a.partials + (int64_t)b * D + col);
      }
      reinterpret_cast<float4*>(red)[rg * cw + cv] = acc;
    }
    __syncthreads();
    if (tid < cw) {
      float4 t = reinterpret_cast<float4*>(red)[tid];
      for (int r = 1; r < RG; ++r) {
        const float4 v = reinterpret_cast<float4*>(red)[r * cw + tid];
        t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
      }
      const int o = (v0 + tid) * unit;
      g_s[o] = t.x;
      if (vec) { g_s[o + 1] = t.y; g_s[o + 2] = t.z; g_s[o + 3] = t.w; }
    }
    __syncthreads();
  }
  if (tid < 32) {                          // batch size of this rank (every slice needs it)
    int cnt = 0;
    for (int b = tid; b < G; b += 32) cnt += __ldcg(a.counts + b);
    cnt = warp_sum(cnt);
    if (tid == 0) s_cnt_local = cnt;
  }
  __syncthreads();
  const int cnt_local = s_cnt_local;
  if (a.stats && slice == 0 && tid == 0) { a.stats[0] = t0; a.stats[1] = global_timer_ns(); }

  const uint64_t off = a.payload_off + (uint64_t)slice * 2 * c.world * (cps + 4) * sizeof(float);
  const uint32_t fb = a.flag_base + (uint32_t)slice * c.world;

  if (a.fin == DAPB_LR_FIN_GRAD_ONLY) {
    for (int d = tid; d < nc; d += nt) a.g_out[c0 + d] = g_s[d];
    if (slice == 0 && tid == 0) {
      if (a.cnt_out) *a.cnt_out = cnt_local;
      if (a.cnt_accum) *a.cnt_accum += (unsigned long long)cnt_local;
    }
    return;
  }

  if (a.fin == DAPB_LR_FIN_SSGD) {
    // all-reduce [g slice | cnt in two 16-bit halves]: the halves are summed exactly in fp32
    if (tid == 0) {
      g_s[nc] = (float)(cnt_local & 0xffff);
      g_s[nc + 1] = (float)(cnt_local >> 16);
    }
    __syncthreads();
    block_allreduce_oneshot<float>(c, c.rank, off, fb, a.epoch, g_s, g_s, nc + 2, cps + 4);
    if (a.stats && slice == 0 && tid == 0) a.stats[2] = global_timer_ns();
    const int cnt_i = (int)g_s[nc] + ((int)g_s[nc + 1] << 16);
    const float cnt_tot = (float)cnt_i;
    for (int d = tid; d < nc; d += nt) {
      float wd = a.w[c0 + d], g = g_s[d];
      if (a.g_out) a.g_out[c0 + d] = g;
      if (a.use_sum) wd -= a.eta * g;                                         // ref logistic_regression.py:85
      else if (cnt_tot > 0.f) wd -= a.eta * (g / cnt_tot + a.lam * reg_grad_f(wd, a.reg_type, a.reg_alpha));
      a.w[c0 + d] = wd;                                                       // ref ssgd.py:105
    }
    if (slice == 0 && tid == 0) {
      if (a.cnt_out) *a.cnt_out = cnt_i;
      if (a.cnt_accum) *a.cnt_accum += (unsigned long long)cnt_i;
      if (a.stats) a.stats[3] = global_timer_ns();
    }
    return;
  }

  // LOCAL / LOCAL_SYNC : replica step  (ref ma.py:42-43, easgd.py:45)
  for (int d = tid; d < nc; d += nt) {
    float wl = a.w_local[c0 + d];
    if (cnt_local > 0) {
      float nv = wl - a.eta * (g_s[d] / (float)cnt_local);
      if (a.algo == DAPB_LR_EASGD) nv -= a.alpha * (wl - a.w[c0 + d]);
      wl = nv;
    }
    g_s[d] = wl;
    if (a.fin == DAPB_LR_FIN_LOCAL) a.w_local[c0 + d] = wl;
  }
  if (slice == 0 && tid == 0) {
    if (a.cnt_out) *a.cnt_out = cnt_local;
    if (a.cnt_accum) *a.cnt_accum += (unsigned long long)cnt_local;
  }
  if (a.fin == DAPB_LR_FIN_LOCAL) return;
  __syncthreads();
  // model all-reduce fused with the MA average / BMUF filter / EASGD centre update
  // (ref ma.py:104-106, bmuf.py:109-114, easgd.py:104-106)
  float* sum_s = red;                      // nt * 4 >= cps floats for every supported block size
  block_allreduce_oneshot<float>(c, c.rank, off, fb, a.epoch, g_s, sum_s, nc, cps + 4);
  if (a.stats && slice == 0 && tid == 0) a.stats[2] = global_timer_ns();
  const float invP = 1.f / (float)a.n_replicas_total;
  for (int d = tid; d < nc; d += nt) {
    float S = sum_s[d], wd = a.w[c0 + d];
    if (a.algo == DAPB_LR_MA) wd = S * invP;
    else if (a.algo == DAPB_LR_BMUF) {
      float nd = a.mu * a.delta_w[c0 + d] + a.zeta * (S * invP - wd);
      a.delta_w[c0 + d] = nd;
      wd += nd;
    } else wd = (1.f - a.beta) * wd + a.beta * S * invP;
    a.w[c0 + d] = wd;
    a.w_local[c0 + d] = a.reset_replica ? wd : g_s[d];
  }
  if (a.stats && slice == 0 && tid == 0) a.stats[3] = global_timer_ns();
}

// CTA-collective tail of every step kernel: publish order is partial row -> __threadfence -> ticket.
__device__ void lr_grid_finish(const DapbLrStepArgs& a, const DapbComm& c, float* sm) {
  __shared__ unsigned s_ticket;
  const unsigned G = gridDim.x;
  const unsigned S = G < (unsigned)kFinSlices ? G : (unsigned)kFinSlices;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_ticket = atomicAdd(a.ticket, 1u);
  __syncthreads();
  const unsigned ticket = s_ticket;
  if (ticket + S < G) return;                          // not one of the last S arrivals
  pdl_launch_dependents();
  if (threadIdx.x == 0) {
    unsigned spins = 0;
    while (ld_acquire_gpu_u32(a.ticket) < G) if (++spins > 64) __nanosleep(40);
  }
  __syncthreads();
  // slices beyond S (G < kFinSlices) are folded onto the last finalizer
  const int first = (int)(ticket - (G - S));
  if (S == (unsigned)kFinSlices) lr_finish_slice(a, c, sm, first);
  else for (int sl = first; sl < kFinSlices; sl += (int)S) { lr_finish_slice(a, c, sm, sl); __syncthreads(); }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned t = atomicAdd(a.ticket, 1u);
    if (t == G + S - 1) *a.ticket = 0;                 // ready for the next launch
  }
}

// Fast path: D == 32 * EPL (each lane owns EPL features, EPL/kElems 16-byte vectors per row).
// PF > 0: the sampling decision of the chunk this warp will process PF iterations later is taken now and every kept row
// is requested from HBM with one cp.async.bulk.prefetch.L2 (no registers, no completion to wait for), so the demand
// loads find their rows in L2.  Two sampled bf16 rows per warp are only 64 KB in flight per SM -- not enough to cover
// the HBM latency at 6.5 TB/s; the prefetch lifts bf16 SSGD from 86 % to 94 % of the measured copy bandwidth.  fp32
// rows (twice the bytes in flight) do not profit and keep PF = 0.
template <typename T, int EPL, int PF>
__global__ void __launch_bounds__(kStepThreads, 2) lr_step_kernel(DapbLrStepArgs a, DapbComm c) {
  constexpr int kE = RowVec<T>::kElems;
  constexpr int kV = EPL / kE;                     // vectors per lane per row
  constexpr int D = 32 * EPL;
  extern __shared__ __align__(16) float smem[];   // [D (weights) | max(kStepWarps*D, 2*D+32)]
  __shared__ int s_cnt;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;

  // forward weights in shared memory, permuted to this kernel's lane order so that every lane
  // reads its own float4s conflict-free: w4[(j*kQ + q)*32 + lane] = w[(j*32 + lane)*kE + 4q .. +4)
  constexpr int kQ = kE / 4;
  float4* w4 = reinterpret_cast<float4*>(smem);
  float* red = smem + D;                               // [kStepWarps][D] reduction scratch
  float acc[EPL];
#pragma unroll
  for (int i = 0; i < EPL; ++i) acc[i] = 0.f;
  int kept = 0;
  if (tid == 0) s_cnt = 0;

  const int64_t n_chunks = (a.n_rows + 31) >> 5;
  const int64_t warp_g = (int64_t)blockIdx.x * kStepWarps + wid;
  const int64_t warp_n = (int64_t)gridDim.x * kStepWarps;
  const char* Xb = reinterpret_cast<const char*>(a.X);
  const bool all_rows = a.mask == nullptr && a.threshold24 >= (1u << 24);

  auto keep_of = [&](int64_t chn) -> bool {
    const int64_t row = (chn << 5) + lane;
    bool keep = row < a.n_rows;
    if (keep && !all_rows)
      keep = a.mask ? (a.mask[row] != 0)
                    : philox_keep_row((uint64_t)(a.row_base + row), a.iteration, a.seed, a.threshold24);
    return keep;
  };
  bool kq[PF > 0 ? PF : 1];                            // decisions for the next PF chunks of this warp
#pragma unroll
  for (int i = 0; i < PF; ++i) {
    const int64_t chn = warp_g + i * warp_n;
    kq[i] = chn < n_chunks ? keep_of(chn) : false;
    if (i > 0 && kq[i]) {
      const char* pn = Xb + ((chn << 5) + lane) * (int64_t)(D * sizeof(T));
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(pn), "r"((uint32_t)(D * sizeof(T))) : "memory");
    }
  }
  // Everything above (sampling decisions, L2 prefetch of the first rows) is independent of the model; from here on
  // the previous step's w is needed: with programmatic dependent launch this kernel may have started while the previous
  // step's finalizers were still exchanging gradients -- wait until that grid has completed and its writes are visible.
  pdl_wait();
  for (int i = tid; i < D / 4; i += kStepThreads) {
    const int ln = i & 31, jq = i >> 5, j = jq / kQ, q = jq % kQ;
    w4[i] = *reinterpret_cast<const float4*>(a.w_fwd + (j * 32 + ln) * kE + 4 * q);
  }
  __syncthreads();
  for (int64_t ch = warp_g; ch < n_chunks; ch += warp_n) {
    bool keep;
    if (PF > 0) {
      keep = kq[0];
#pragma unroll
      for (int i = 0; i + 1 < PF; ++i) kq[i] = kq[i + 1];
      const int64_t chn = ch + PF * warp_n;
      bool kn = false;
      if (chn < n_chunks) {
        kn = keep_of(chn);
        if (kn) {
          const char* pn = Xb + ((chn << 5) + lane) * (int64_t)(D * sizeof(T));
          asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(pn), "r"((uint32_t)(D * sizeof(T))) : "memory");
        }
      }
      kq[PF - 1] = kn;
    } else {
      keep = keep_of(ch);
    }
    unsigned bits = __ballot_sync(0xffffffffu, keep);
    kept += __popc(bits);
    while (bits) {
      // two sampled rows in flight per warp (2 * kV independent 16-byte loads per lane)
      const int b0 = __ffs(bits) - 1; bits &= bits - 1;
      const int b1 = bits ? __ffs(bits) - 1 : -1; if (b1 >= 0) bits &= bits - 1;
      const int64_t ra = (ch << 5) + b0, rb = (ch << 5) + (b1 >= 0 ? b1 : b0);
      uint4 va[kV], vb[kV];
      const char* pa = Xb + ra * (int64_t)(D * sizeof(T));
      const char* pb = Xb + rb * (int64_t)(D * sizeof(T));
#pragma unroll
      for (int j = 0; j < kV; ++j) va[j] = ld_stream_16(pa + (size_t)(j * 32 + lane) * 16);
#pragma unroll
      for (int j = 0; j < kV; ++j) vb[j] = ld_stream_16(pb + (size_t)(j * 32 + lane) * 16);
      const float ya = a.y[ra], yb = a.y[rb];
      float xa[EPL], dota = 0.f;
#pragma unroll
      for (int j = 0; j < kV; ++j) RowVec<T>::unpack(va[j], xa + j * kE);
      float dotb = 0.f;
#pragma unroll
      for (int j = 0; j < kV; ++j) {
        float xb[kE], wv[kE];
        RowVec<T>::unpack(vb[j], xb);
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
          const float4 t4 = w4[(j * kQ + q) * 32 + lane];
          wv[4 * q] = t4.x; wv[4 * q + 1] = t4.y; wv[4 * q + 2] = t4.z; wv[4 * q + 3] = t4.w;
        }
#pragma unroll
        for (int e = 0; e < kE; ++e) {
          dota = fmaf(xa[j * kE + e], wv[e], dota);
          dotb = fmaf(xb[e], wv[e], dotb);
        }
      }
      dota = warp_sum(dota);
      dotb = warp_sum(dotb);
      const float resa = sigmoid_ref(dota, a.eps) - ya;                    // ref ssgd.py:33
      const float resb = b1 >= 0 ? sigmoid_ref(dotb, a.eps) - yb : 0.f;
#pragma unroll
      for (int i = 0; i < EPL; ++i) acc[i] = fmaf(resa, xa[i], acc[i]);
#pragma unroll
      for (int j = 0; j < kV; ++j) {
        float xb[kE];
        RowVec<T>::unpack(vb[j], xb);
#pragma unroll
        for (int e = 0; e < kE; ++e) acc[j * kE + e] = fmaf(resb, xb[e], acc[j * kE + e]);
      }
    }
  }

  // ---- CTA reduction: warps -> smem -> one partial row per CTA --------------------------------
  __syncthreads();
#pragma unroll
  for (int j = 0; j < kV; ++j)
#pragma unroll
    for (int e = 0; e < kE; ++e) red[wid * D + (j * 32 + lane) * kE + e] = acc[j * kE + e];
  if (lane == 0 && kept) atomicAdd(&s_cnt, kept);
  __syncthreads();
  for (int d = tid; d < D; d += kStepThreads) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < kStepWarps; ++w) s += red[w * D + d];
    a.partials[(int64_t)blockIdx.x * D + d] = s;
  }
  if (tid == 0) a.counts[blockIdx.x] = s_cnt;
  lr_grid_finish(a, c, red);
}

// Generic path (any D <= 8192): per-warp accumulators in shared memory.
template <typename T>
__global__ void __launch_bounds__(kStepThreads) lr_step_generic_kernel(DapbLrStepArgs a, DapbComm c) {
  extern __shared__ __align__(16) float smem[];       // [kStepWarps][D] (+ finalize scratch)
  __shared__ int s_cnt;
  const int D = a.D, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  pdl_wait();
  float* acc = smem + (int64_t)wid * D;
  for (int d = lane; d < D; d += 32) acc[d] = 0.f;
  if (tid == 0) s_cnt = 0;
  __syncthreads();
  int kept = 0;
  const T* X = reinterpret_cast<const T*>(a.X);
  const int64_t warp_g = (int64_t)blockIdx.x * kStepWarps + wid, warp_n = (int64_t)gridDim.x * kStepWarps;
  const bool all_rows = a.mask == nullptr && a.threshold24 >= (1u << 24);
  for (int64_t row = warp_g; row < a.n_rows; row += warp_n) {
    bool keep = all_rows ? true
                : a.mask ? (a.mask[row] != 0)
                         : philox_keep_row((uint64_t)(a.row_base + row), a.iteration, a.seed, a.threshold24);
    if (!keep) continue;
    ++kept;
    const T* x = X + row * (int64_t)D;
    float s = 0.f;
    for (int d = lane; d < D; d += 32) s = fmaf((float)x[d], a.w_fwd[d], s);
    s = warp_sum(s);
    const float r = sigmoid_ref(s, a.eps) - a.y[row];
    for (int d = lane; d < D; d += 32) acc[d] = fmaf(r, (float)x[d], acc[d]);
  }
  if (lane == 0 && kept) atomicAdd(&s_cnt, kept);
  __syncthreads();
  for (int d = tid; d < D; d += kStepThreads) {
    float s = 0.f;
    for (int w = 0; w < kStepWarps; ++w) s += smem[(int64_t)w * D + d];
    a.partials[(int64_t)blockIdx.x * D + d] = s;
  }
  if (tid == 0) a.counts[blockIdx.x] = s_cnt;
  lr_grid_finish(a, c, smem);
}

// Launch with programmatic stream serialization: a step kernel queued behind another step kernel may begin (sample,
// prefetch) as soon as the earlier grid's finalizers signal griddepcontrol.launch_dependents.  DAPB_LR_PDL=0 disables.
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kernel)(KArgs...), int grid, int block, size_t smem, cudaStream_t s, Args... args) {
  static const bool use_pdl = !(getenv("DAPB_LR_PDL") && atoi(getenv("DAPB_LR_PDL")) == 0);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = use_pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, args...);
}

// shared-memory floats the grid finish needs behind the kernel's own scratch pointer (see lr_finish_slice)
inline size_t lr_finish_floats(int D, int threads) { return (size_t)lr_cols_per_slice(D) + 8 + 4 * (size_t)threads + 16; }

template <typename T, int EPL>
int launch_fast(const DapbLrStepArgs& a, const DapbComm& c, int grid, cudaStream_t s) {
  constexpr int D = 32 * EPL;
  size_t red = kStepWarps * D > 2 * D + 32 ? kStepWarps * D : 2 * D + 32;
  if (red < lr_finish_floats(D, kStepThreads)) red = lr_finish_floats(D, kStepThreads);
  size_t smem = sizeof(float) * (size_t)(D + red);
  // prefetch distance: 16-bit rows 1 chunk ahead by default, 32-bit rows none (measured); DAPB_LR_PREFETCH=0|1|2 overrides
  static const int pf_env = getenv("DAPB_LR_PREFETCH") ? atoi(getenv("DAPB_LR_PREFETCH")) : -1;
  const int pf = pf_env >= 0 ? pf_env : (sizeof(T) == 2 ? 1 : 0);
#define DAPB_LAUNCH_PF(N)                                                                                              \
  {                                                                                                                     \
    DAPB_CUDA_CHECK(cudaFuncSetAttribute(lr_step_kernel<T, EPL, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    DAPB_CUDA_CHECK(launch_pdl(lr_step_kernel<T, EPL, N>, grid, kStepThreads, smem, s, a, c));                          \
  }
  if (pf >= 2) DAPB_LAUNCH_PF(2) else if (pf == 1) DAPB_LAUNCH_PF(1) else DAPB_LAUNCH_PF(0)
#undef DAPB_LAUNCH_PF
  DAPB_LAUNCH_CHECK();
  return 0;
}
template <typename T>
int launch_generic(const DapbLrStepArgs& a, const DapbComm& c, int grid, cudaStream_t s) {
  size_t fl = kStepWarps * a.D > 2 * a.D + 32 ? kStepWarps * a.D : 2 * a.D + 32;
  if (fl < lr_finish_floats(a.D, kStepThreads)) fl = lr_finish_floats(a.D, kStepThreads);
  size_t smem = sizeof(float) * fl;
  DAPB_REQUIRE(smem <= 220 * 1024, "lr_step: D=%d too large for the generic kernel", a.D);
  DAPB_CUDA_CHECK(cudaFuncSetAttribute(lr_step_generic_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  DAPB_CUDA_CHECK(launch_pdl(lr_step_generic_kernel<T>, grid, kStepThreads, smem, s, a, c));
  DAPB_LAUNCH_CHECK();
  return 0;
}

}  // namespace

extern "C" uint64_t dapb_lr_step_payload_bytes(int32_t world, int32_t D) {
  return (uint64_t)kFinSlices * 2 * world * (lr_cols_per_slice(D) + 4) * sizeof(float);
}

extern "C" int dapb_lr_step(const DapbLrStepArgs* args, const DapbComm* comm, void* stream) {
  DAPB_REQUIRE(args && comm, "lr_step: null argument");
  DapbLrStepArgs a = *args;
  DapbComm c = *comm;
  DAPB_REQUIRE(a.dtype == 0 || a.dtype == 1, "lr_step: dtype must be 0 (fp32) or 1 (bf16)");
  DAPB_REQUIRE(a.D >= 1 && a.n_rows >= 0, "lr_step: bad shape");
  DAPB_REQUIRE(a.fin >= DAPB_LR_FIN_GRAD_ONLY && a.fin <= DAPB_LR_FIN_LOCAL_SYNC, "lr_step: bad fin");
  DAPB_REQUIRE(a.partials && a.counts && a.ticket && a.w_fwd, "lr_step: missing scratch / weights");
  if (a.fin == DAPB_LR_FIN_SSGD || a.fin == DAPB_LR_FIN_LOCAL_SYNC) {
    DAPB_REQUIRE(c.world >= 1 && c.world <= DAPB_MAX_RANKS, "lr_step: bad world");
    DAPB_REQUIRE(a.epoch >= 1, "lr_step: epoch starts at 1");
    if (c.world > 1) {
      uint64_t need = DAPB_FLAG_BYTES + a.payload_off + dapb_lr_step_payload_bytes(c.world, a.D);
      DAPB_REQUIRE(need <= c.heap_bytes, "lr_step: heap too small");
      DAPB_REQUIRE((a.flag_base + (uint64_t)DAPB_LR_STEP_FLAGS) * 8 <= DAPB_FLAG_BYTES, "lr_step: flag range");
    }
  }
  int grid = a.grid > 0 ? a.grid : 2 * dapb::sm_count();
  a.grid = grid;
  cudaStream_t s = (cudaStream_t)stream;
  // the fast kernels read X, w_fwd (float4) and the partial rows (float4) with 16-byte vector loads
  const bool aligned = ((reinterpret_cast<uintptr_t>(a.X) | reinterpret_cast<uintptr_t>(a.w_fwd) |
                         reinterpret_cast<uintptr_t>(a.partials)) & 15) == 0;
  if (aligned && a.dtype == 1) {
    switch (a.D) {
      case 256: return launch_fast<__nv_bfloat16, 8>(a, c, grid, s);
      case 512: return launch_fast<__nv_bfloat16, 16>(a, c, grid, s);
      case 1024: return launch_fast<__nv_bfloat16, 32>(a, c, grid, s);
      default: break;
    }
  } else if (aligned && a.dtype == 0) {
    switch (a.D) {
      case 128: return launch_fast<float, 4>(a, c, grid, s);
      case 256: return launch_fast<float, 8>(a, c, grid, s);
      case 512: return launch_fast<float, 16>(a, c, grid, s);
      case 1024: return launch_fast<float, 32>(a, c, grid, s);
      default: break;
    }
  }
  return a.dtype == 1 ? launch_generic<__nv_bfloat16>(a, c, grid, s) : launch_generic<float>(a, c, grid, s);
}

// =================================================================================================
// Tensor-core step kernel (bf16, D = 1024): forward z = Xw and backward g = X^T r as tcgen05.mma chains
// =================================================================================================
// The GEMV-shaped problem is mapped onto the tensor cores by padding N to 16 and spending the padding on
// precision: the weight operand holds (bf16(w), bf16(w - bf16(w))) and the residual operand (bf16(r), bf16(r - bf16(r))),
// so both products carry ~16 mantissa bits although every MMA input is bf16.
//
// A 128-row x 1024-feature tile is 256 KB and does not fit in shared memory, so it is streamed twice in slabs
// of 128 features (32 KB): pass F (forward, K-major A) comes from HBM, pass B (backward, the same slabs read as
// MN-major A) hits L2.  Sampled rows are not contiguous, so the slabs are gathered by loader warps
// (16-byte LDG -> 128B-swizzled STS, then fence.proxy.async) instead of TMA boxes.  Roles per CTA (768 threads):
//   warp 0      sampler: Philox Bernoulli decisions -> lists of 128 selected row indices (ring of 3)
//   warp 1      MMA issuer, software-pipelined as F(0) F(1) B(0) F(2) B(1) ... so the sigmoid epilogue of tile t
//               overlaps the forward MMAs of tile t+1
//   warp 2      TMEM allocator (256 columns: 2 x 16 for z, 8 x 16 for the gradient blocks)
//   warps 4-7   epilogue: tcgen05.ld z -> r = sigmoid(z) - y -> residual operand in swizzled smem; at the end
//               tcgen05.ld of the gradient accumulators -> this CTA's partial row
//   warps 8-23  two loader groups of 8 warps working on alternating slabs (64 KB in flight per SM)
// The partial rows go through the same deterministic grid reduction + in-kernel peer all-reduce + optimiser
// update (lr_grid_finish) as the CUDA-core kernel.
namespace {

constexpr int kTcThreads = 768;
constexpr int kTcD = 1024;
constexpr int kTcTM = 128;                 // rows per tile
constexpr int kTcFB = 128;                 // features per slab
constexpr int kTcSlabs = kTcD / kTcFB;     // 8 slabs per pass
constexpr int kTcRing = 5;                 // slab ring slots (32 KB each)
constexpr int kTcLists = 3;
constexpr uint32_t kTcSlabBytes = kTcTM * kTcFB * 2;                       // 32 KB
constexpr uint32_t kTcWOff = kTcRing * kTcSlabBytes;                       // weight operand: 16 x [16 rows x 128 B]
constexpr uint32_t kTcWBytes = (kTcD / 64) * 16 * 128;                     // 32 KB
constexpr uint32_t kTcROff = kTcWOff + kTcWBytes;                          // residual operand: 2 x (2 x [16 x 128 B])
constexpr uint32_t kTcRBytes = 2 * 16 * 128;                               // 4 KB per buffer
constexpr uint32_t kTcListOff = kTcROff + 2 * kTcRBytes;                   // 3 x (n_valid, pad[3], rows[128])
constexpr uint32_t kTcListBytes = (4 + kTcTM) * 4;
constexpr uint32_t kTcBarOff = kTcListOff + kTcLists * kTcListBytes;
constexpr int kTcNumBars = 2 * kTcRing + 2 * kTcLists + 2 + 2 + 2 + 2 + 1;
constexpr uint32_t kTcTmemOff = kTcBarOff + kTcNumBars * 8;
constexpr uint32_t kTcTotal = kTcTmemOff + 16;

using namespace dapb::tc;

__global__ void __launch_bounds__(kTcThreads, 1) lr_tc_kernel(DapbLrStepArgs a, DapbComm c, uint32_t smem_bytes) {
  extern __shared__ __align__(1024) uint8_t tc_smem_raw[];
  const uint32_t pad = (1024u - (smem_u32(tc_smem_raw) & 1023u)) & 1023u;
  uint8_t* smem = tc_smem_raw + pad;
  if (pad + kTcTotal > smem_bytes) __trap();
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kTcBarOff);
  uint64_t* slot_full = bars;                       // [ring]  8 arrivals (one loader group)
  uint64_t* slot_empty = slot_full + kTcRing;       // [ring]  1 (MMA commit)
  uint64_t* list_full = slot_empty + kTcRing;       // [lists] 1 (sampler)
  uint64_t* list_empty = list_full + kTcLists;      // [lists] 20 (16 loader warps + 4 epilogue warps)
  uint64_t* z_full = list_empty + kTcLists;         // [2] 1 (MMA commit)
  uint64_t* z_empty = z_full + 2;                   // [2] 4 (epilogue warps)
  uint64_t* r_full = z_empty + 2;                   // [2] 4 (epilogue warps)
  uint64_t* r_empty = r_full + 2;                   // [2] 1 (MMA commit)
  uint64_t* g_done = r_empty + 2;                   // 1
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(smem + kTcTmemOff);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, tid = threadIdx.x;
  pdl_wait();                                        // the previous step's w must be visible (see lr_grid_finish)

  if (tid == 0) {
    for (int i = 0; i < kTcRing; ++i) { mbar_init(&slot_full[i], 8); mbar_init(&slot_empty[i], 1); }
    for (int i = 0; i < kTcLists; ++i) { mbar_init(&list_full[i], 1); mbar_init(&list_empty[i], 20); }
    for (int i = 0; i < 2; ++i) { mbar_init(&z_full[i], 1); mbar_init(&z_empty[i], 4); mbar_init(&r_full[i], 4); mbar_init(&r_empty[i], 1); }
    mbar_init(g_done, 1);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(tmem_holder, 256); tmem_relinquish(); }
  // weight operand (K-major, 128B swizzle): row 0 = bf16(w), row 1 = bf16(w - bf16(w)), rows 2..15 = 0
  for (int i = tid; i < (int)(kTcWBytes + 2 * kTcRBytes) / 16; i += kTcThreads)
    reinterpret_cast<uint4*>(smem + kTcWOff)[i] = make_uint4(0u, 0u, 0u, 0u);
  __syncthreads();
  for (int k = tid; k < kTcD; k += kTcThreads) {
    const float wv = a.w_fwd[k];
    const __nv_bfloat16 hi = __float2bfloat16(wv);
    const __nv_bfloat16 lo = __float2bfloat16(wv - __bfloat162float(hi));
    const int slab = k >> 6, kk = k & 63, chunk = kk >> 3, within = (kk & 7) * 2;
    uint8_t* base = smem + kTcWOff + slab * 2048;
    *reinterpret_cast<__nv_bfloat16*>(base + 0 * 128 + ((chunk ^ 0) << 4) + within) = hi;
    *reinterpret_cast<__nv_bfloat16*>(base + 1 * 128 + ((chunk ^ 1) << 4) + within) = lo;
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  const uint32_t tmem_z = tmem_base, tmem_g = tmem_base + 32;
  auto list_ptr = [&](int t) { return reinterpret_cast<int*>(smem + kTcListOff + (t % kTcLists) * kTcListBytes); };

  if (warp == 0) {
    // ===================== sampler =====================
    const int64_t n_chunks = (a.n_rows + 31) >> 5;
    const bool all_rows = a.mask == nullptr && a.threshold24 >= (1u << 24);
    int t = 0, fill = 0, kept_total = 0;
    mbar_wait(&list_empty[0], 1);                        // fresh barrier: passes immediately
    int* lst = list_ptr(0);
    for (int64_t ch = blockIdx.x; ch < n_chunks; ch += gridDim.x) {
      const int64_t row = (ch << 5) + lane;
      bool keep = row < a.n_rows;
      if (keep && !all_rows)
        keep = a.mask ? (a.mask[row] != 0)
                      : philox_keep_row((uint64_t)(a.row_base + row), a.iteration, a.seed, a.threshold24);
      unsigned bits = __ballot_sync(0xffffffffu, keep);
      kept_total += __popc(bits);
      while (bits) {
        const int avail = __popc(bits), take = min(avail, kTcTM - fill);
        const int rank = __popc(bits & ((1u << lane) - 1));
        if (((bits >> lane) & 1u) && rank < take) lst[4 + fill + rank] = (int)row;
        fill += take;
        if (take == avail) bits = 0;
        else for (int i = 0; i < take; ++i) bits &= bits - 1;        // drop the `take` lowest set bits (tile boundary)
        if (fill == kTcTM) {
          if (lane == 0) lst[0] = kTcTM;
          __syncwarp();
          if (lane == 0) mbar_arrive(&list_full[t % kTcLists]);
          ++t; fill = 0;
          mbar_wait(&list_empty[t % kTcLists], ((t / kTcLists) & 1) ^ 1);
          lst = list_ptr(t);
        }
      }
    }
    if (fill > 0) {                                       // ragged last tile: pad with "no row"
      for (int i = fill + lane; i < kTcTM; i += 32) lst[4 + i] = -1;
      if (lane == 0) lst[0] = fill;
      __syncwarp();
      if (lane == 0) mbar_arrive(&list_full[t % kTcLists]);
      ++t;
      mbar_wait(&list_empty[t % kTcLists], ((t / kTcLists) & 1) ^ 1);
      lst = list_ptr(t);
    }
    if (lane == 0) lst[0] = -1;                           // terminator
    __syncwarp();
    if (lane == 0) { mbar_arrive(&list_full[t % kTcLists]); a.counts[blockIdx.x] = kept_total; }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc_f = make_idesc_bf16(kTcTM, 16, 0, 0);           // A K-major, B K-major
      constexpr uint32_t idesc_b = make_idesc_bf16(kTcFB, 16, 1, 0);           // A MN-major (X^T), B K-major
      const uint64_t tmpl_k = make_smem_desc_sw128(0, 16, 1024);
      const uint64_t tmpl_mn = make_smem_desc_sw128(0, kTcTM * 128, 1024);     // LBO = next 64-feature half
      const uint32_t ring = smem_u32(smem), wop = smem_u32(smem + kTcWOff), rop = smem_u32(smem + kTcROff);
      uint32_t n = 0;                                                          // slab sequence counter
      auto fwd = [&](int t) {
        const int zb = t & 1;
        mbar_wait(&z_empty[zb], ((t >> 1) & 1) ^ 1);
        tc_fence_after();
        for (int fb = 0; fb < kTcSlabs; ++fb, ++n) {
          const uint32_t slot = n % kTcRing;
          mbar_wait(&slot_full[slot], (n / kTcRing) & 1);
          tc_fence_after();
          const uint32_t sa = ring + slot * kTcSlabBytes;
#pragma unroll
          for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
              const uint64_t da = tmpl_k | (uint64_t)(((sa + h * (kTcTM * 128) + ks * 32) & 0x3FFFF) >> 4);
              const uint64_t db = tmpl_k | (uint64_t)(((wop + (fb * 2 + h) * 2048 + ks * 32) & 0x3FFFF) >> 4);
              umma_bf16(tmem_z + zb * 16, da, db, idesc_f, (fb | h | ks) != 0);
            }
          umma_commit(&slot_empty[slot]);
        }
        umma_commit(&z_full[zb]);
      };
      auto bwd = [&](int t) {
        const int rb = t & 1;
        mbar_wait(&r_full[rb], (t >> 1) & 1);
        tc_fence_after();
        for (int fb = 0; fb < kTcSlabs; ++fb, ++n) {
          const uint32_t slot = n % kTcRing;
          mbar_wait(&slot_full[slot], (n / kTcRing) & 1);
          tc_fence_after();
          const uint32_t sa = ring + slot * kTcSlabBytes;
#pragma unroll
          for (int kr = 0; kr < 8; ++kr) {                                    // 16 rows per MMA
            const uint64_t da = tmpl_mn | (uint64_t)(((sa + kr * 2048) & 0x3FFFF) >> 4);
            const uint64_t db = tmpl_k | (uint64_t)(((rop + rb * kTcRBytes + (kr >> 2) * 2048 + (kr & 3) * 32) & 0x3FFFF) >> 4);
            umma_bf16(tmem_g + fb * 16, da, db, idesc_b, (t | kr) != 0);
          }
          umma_commit(&slot_empty[slot]);
        }
        umma_commit(&r_empty[rb]);
      };
      mbar_wait(&list_full[0], 0);
      if (list_ptr(0)[0] >= 0) {
        fwd(0);
        for (int t = 1;; ++t) {
          mbar_wait(&list_full[t % kTcLists], (t / kTcLists) & 1);
          const int nv = list_ptr(t)[0];
          if (nv >= 0) fwd(t);
          bwd(t - 1);
          if (nv < 0) break;
        }
      }
      umma_commit(g_done);
    }
  } else if (warp >= 4 && warp < 8) {
    // ===================== epilogue =====================
    const int qd = warp & 3, row_in_tile = qd * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(qd * 32) << 16;
    int n_tiles = 0;
    for (int t = 0;; ++t) {
      mbar_wait(&list_full[t % kTcLists], (t / kTcLists) & 1);
      const int* lst = list_ptr(t);
      if (lst[0] < 0) break;
      n_tiles = t + 1;
      const int row = lst[4 + row_in_tile];
      const float yv = row >= 0 ? a.y[row] : 0.f;
      const int zb = t & 1;
      mbar_wait(&z_full[zb], (t >> 1) & 1);
      tc_fence_after();
      uint32_t v0, v1;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(v0), "=r"(v1) : "r"(tmem_z + lane_addr + zb * 16) : "memory");
      tmem_ld_wait();
      tc_fence_before();
      const float z = __uint_as_float(v0) + __uint_as_float(v1);
      const float r = row >= 0 ? sigmoid_ref(z, a.eps) - yv : 0.f;                 // ref ssgd.py:33
      const __nv_bfloat16 hi = __float2bfloat16(r);
      const __nv_bfloat16 lo = __float2bfloat16(r - __bfloat162float(hi));
      mbar_wait(&r_empty[zb], ((t >> 1) & 1) ^ 1);
      {
        const int slab = row_in_tile >> 6, kk = row_in_tile & 63, chunk = kk >> 3, within = (kk & 7) * 2;
        uint8_t* base = smem + kTcROff + zb * kTcRBytes + slab * 2048;
        *reinterpret_cast<__nv_bfloat16*>(base + 0 * 128 + ((chunk ^ 0) << 4) + within) = hi;
        *reinterpret_cast<__nv_bfloat16*>(base + 1 * 128 + ((chunk ^ 1) << 4) + within) = lo;
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) { mbar_arrive(&r_full[zb]); mbar_arrive(&z_empty[zb]); mbar_arrive(&list_empty[t % kTcLists]); }
    }
    // gradient accumulators -> this CTA's partial row (column 0 = hi part, column 1 = lo part)
    mbar_wait(g_done, 0);
    tc_fence_after();
    float* prow = a.partials + (int64_t)blockIdx.x * kTcD;
    for (int fb = 0; fb < kTcSlabs; ++fb) {
      float g = 0.f;
      if (n_tiles > 0) {
        uint32_t v0, v1;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(v0), "=r"(v1) : "r"(tmem_g + lane_addr + fb * 16) : "memory");
        tmem_ld_wait();
        g = __uint_as_float(v0) + __uint_as_float(v1);
      }
      prow[fb * kTcFB + row_in_tile] = g;
    }
    tc_fence_before();
  } else if (warp >= 8) {
    // ===================== loaders: gather rows -> 128B-swizzled slabs =====================
    const int grp = (warp - 8) >> 3;                       // two groups of 8 warps, alternating slabs
    const int j = ((warp - 8) & 7) * 32 + lane;            // 0..255 inside the group
    const int cc = j & 15, r_off = j >> 4;                 // 16-byte chunk of the 256-byte row segment, first row
    const int h = cc >> 3, c8 = cc & 7;
    const char* Xb = reinterpret_cast<const char*>(a.X);
    uint32_t n = 0;
    auto phase = [&](int t, bool first_pass) {
      const int* lst = list_ptr(t);
      int rows[8];
#pragma unroll
      for (int it = 0; it < 8; ++it) rows[it] = lst[4 + r_off + 16 * it];
      for (int fb = 0; fb < kTcSlabs; ++fb, ++n) {
        if ((int)(n & 1) != grp) continue;
        const uint32_t slot = n % kTcRing;
        mbar_wait(&slot_empty[slot], ((n / kTcRing) & 1) ^ 1);
        uint4 v[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const char* src = Xb + (int64_t)rows[it] * (kTcD * 2) + fb * (kTcFB * 2) + cc * 16;
          // forward pass: keep the line in L2 (evict_last), the backward pass reads it once more and releases it
          v[it] = rows[it] < 0 ? make_uint4(0u, 0u, 0u, 0u) : (first_pass ? ld_stream_16_keep(src) : ld_stream_16_last(src));
        }
        uint8_t* dst = smem + slot * kTcSlabBytes + h * (kTcTM * 128);
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int i = r_off + 16 * it;
          *reinterpret_cast<uint4*>(dst + i * 128 + ((c8 ^ (i & 7)) << 4)) = v[it];
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(&slot_full[slot]);
      }
    };
    mbar_wait(&list_full[0], 0);
    if (list_ptr(0)[0] >= 0) {
      phase(0, true);
      for (int t = 1;; ++t) {
        mbar_wait(&list_full[t % kTcLists], (t / kTcLists) & 1);
        const int nv = list_ptr(t)[0];
        if (nv >= 0) phase(t, true);
        phase(t - 1, false);                               // backward pass of the previous tile (L2 hits)
        __syncwarp();
        if (lane == 0) mbar_arrive(&list_empty[(t - 1) % kTcLists]);
        if (nv < 0) break;
      }
    }
  }

  __threadfence();
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 256);
  lr_grid_finish(a, c, reinterpret_cast<float*>(smem));
}

}  // namespace

// Tensor-core variant of dapb_lr_step: same arguments and semantics, bf16 rows with D == 1024 only.
extern "C" int dapb_lr_step_tc(const DapbLrStepArgs* args, const DapbComm* comm, void* stream) {
  DAPB_REQUIRE(args && comm, "lr_step_tc: null argument");
  DapbLrStepArgs a = *args;
  DapbComm c = *comm;
  DAPB_REQUIRE(a.dtype == 1 && a.D == kTcD, "lr_step_tc: needs bf16 rows with D == 1024 (got dtype %d, D %d)", a.dtype, a.D);
  DAPB_REQUIRE((reinterpret_cast<uintptr_t>(a.X) & 15) == 0, "lr_step_tc: X must be 16-byte aligned");
  DAPB_REQUIRE(a.n_rows < (1ll << 31), "lr_step_tc: more than 2^31 rows per shard");
  DAPB_REQUIRE(a.fin >= DAPB_LR_FIN_GRAD_ONLY && a.fin <= DAPB_LR_FIN_LOCAL_SYNC, "lr_step_tc: bad fin");
  DAPB_REQUIRE(a.partials && a.counts && a.ticket && a.w_fwd, "lr_step_tc: missing scratch / weights");
  if ((a.fin == DAPB_LR_FIN_SSGD || a.fin == DAPB_LR_FIN_LOCAL_SYNC) && c.world > 1) {
    DAPB_REQUIRE(a.epoch >= 1, "lr_step_tc: epoch starts at 1");
    uint64_t need = DAPB_FLAG_BYTES + a.payload_off + dapb_lr_step_payload_bytes(c.world, a.D);
    DAPB_REQUIRE(need <= c.heap_bytes, "lr_step_tc: heap too small");
  }
  int grid = a.grid > 0 ? a.grid : dapb::sm_count();
  a.grid = grid;
  const uint32_t smem = kTcTotal + 1024;
  DAPB_CUDA_CHECK(cudaFuncSetAttribute(lr_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  DAPB_CUDA_CHECK(launch_pdl(lr_tc_kernel, grid, kTcThreads, (size_t)smem, (cudaStream_t)stream, a, c, smem));
  DAPB_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Accuracy (ref ssgd.py:107-110): pred = sigmoid(x.w) < 0.5 ? 0 : 1; acc = mean(pred == y)
// ---------------------------------------------------------------------------------------------
namespace {
template <typename T>
__global__ void lr_eval_kernel(const T* X, const float* y, int64_t n, int D, const float* w, float eps,
                               float* acc_out, uint32_t* scratch) {
  const int lane = threadIdx.x & 31;
  const int64_t warp_g = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t warp_n = ((int64_t)gridDim.x * blockDim.x) >> 5;
  unsigned ok = 0;
  for (int64_t row = warp_g; row < n; row += warp_n) {
    const T* x = X + row * (int64_t)D;
    float s = 0.f;
    for (int d = lane; d < D; d += 32) s = fmaf((float)x[d], w[d], s);
    s = warp_sum(s);
    float pred = sigmoid_ref(s, eps) < 0.5f ? 0.f : 1.f;
    ok += (pred == y[row]);
  }
  if (lane == 0 && ok) atomicAdd(&scratch[0], ok);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned t = atomicAdd(&scratch[1], 1u);
    if (t == gridDim.x - 1) {
      __threadfence();
      *acc_out = (float)((double)atomicAdd(&scratch[0], 0u) / (double)n);
      scratch[0] = 0; scratch[1] = 0;
    }
  }
}
}  // namespace

// Bernoulli keep flags of one iteration as a byte mask (the same Philox decision the step kernels take in-line).
// Not used by the fused path -- it lets the PyTorch+NCCL BASELINE sample the identical mini-batch at memory speed.
namespace {
__global__ void philox_mask_kernel(int64_t row_base, int64_t n, uint64_t seed, uint32_t iteration, uint32_t threshold24,
                                   uint8_t* out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = threshold24 >= (1u << 24) ? 1 : (uint8_t)philox_keep_row((uint64_t)(row_base + i), iteration, seed, threshold24);
}
}  // namespace
extern "C" int dapb_philox_mask(int64_t row_base, int64_t n_rows, uint64_t seed, uint32_t iteration, uint32_t threshold24,
                                uint8_t* out, void* stream) {
  DAPB_REQUIRE(n_rows >= 0 && out != nullptr, "philox_mask: bad arguments");
  if (n_rows == 0) return 0;
  int64_t blocks = (n_rows + 255) / 256;
  if (blocks > dapb::sm_count() * 16) blocks = dapb::sm_count() * 16;
  philox_mask_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(row_base, n_rows, seed, iteration, threshold24, out);
  DAPB_LAUNCH_CHECK();
  return 0;
}

extern "C" int dapb_lr_eval(const void* X, const float* y, int64_t n_rows, int32_t D, int32_t dtype,
                            const float* w, float eps, float* acc_out, uint32_t* scratch2, void* stream) {
  DAPB_REQUIRE(n_rows > 0 && D > 0 && scratch2 && acc_out, "lr_eval: bad arguments");
  int grid = (int)((n_rows * 32 + 255) / 256);
  if (grid > 4 * dapb::sm_count()) grid = 4 * dapb::sm_count();
  if (dtype == 1)
    lr_eval_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)X, y, n_rows, D, w, eps, acc_out, scratch2);
  else
    lr_eval_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)X, y, n_rows, D, w, eps, acc_out, scratch2);
  DAPB_LAUNCH_CHECK();
  return 0;
}

// =================================================================================================
// Allocation-free gradient entry point
// =================================================================================================
// Some users of this library are not in a position to allocate device memory: a host process whose
// own allocator has already reserved the whole GPU for the resident data shard (the 64 GiB shards of
// the benchmark configuration leave no room to negotiate), a serving / control loop that must not hit
// the device-wide synchronisation that cudaMalloc and cudaFree imply, or a plug-in callback that is
// simply forbidden from calling allocating CUDA APIs.  For them the mini-batch gradient is offered
// once more with NO caller-provided scratch: the per-CTA partial sums, per-CTA counts and the
// last-CTA ticket of the step kernel live in the file-scope device array below.
//
// Cost and limits: the array occupies kNoAllocScratchBytes (about 1.16 MiB) on EVERY GPU that loads
// this library, from load until the process exits.  It is sized for D <= DAPB_NOALLOC_MAX_D features
// and a grid of at most DAPB_NOALLOC_MAX_GRID CTAs; larger requests are refused.  All calls share it
// and it keeps whatever the previous call left behind, so the ticket word is cleared on the stream
// before each launch (partials and counts are fully rewritten by every launch before they are read).
// Two calls whose kernels can overlap on the GPU (different streams, or different host threads on the
// same device) race on the array -- serialise them.
namespace {

constexpr size_t kNoAllocTicketBytes = 256;
constexpr size_t kNoAllocCountsBytes = ((size_t)DAPB_NOALLOC_MAX_GRID * sizeof(int32_t) + 255) / 256 * 256;
constexpr size_t kNoAllocPartialBytes = (size_t)DAPB_NOALLOC_MAX_GRID * DAPB_NOALLOC_MAX_D * sizeof(float);
constexpr size_t kNoAllocScratchBytes = kNoAllocTicketBytes + kNoAllocCountsBytes + kNoAllocPartialBytes;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

// In host code `g_noalloc_scratch` is only a stand-in symbol, not a device address; the real
// address is looked up once per device (with that device current) and cached here.
constexpr int kMaxDevices = 64;
static unsigned char* g_noalloc_dev_ptr[kMaxDevices] = {nullptr};
static int g_noalloc_grid[kMaxDevices] = {0};

static std::mutex g_noalloc_mutex;

int noalloc_setup(int dev, unsigned char** out, int* grid_out) {
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_dev_ptr[dev] == nullptr) {
    unsigned char* p = nullptr;
    DAPB_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    int sms = 0;
    DAPB_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    int grid = 2 * sms;
    if (grid > DAPB_NOALLOC_MAX_GRID) grid = DAPB_NOALLOC_MAX_GRID;
    g_noalloc_grid[dev] = grid;
    g_noalloc_dev_ptr[dev] = p;
  }
  *out = g_noalloc_dev_ptr[dev];
  *grid_out = g_noalloc_grid[dev];
  return 0;
}

}  // namespace

extern "C" int dapb_lr_grad_noalloc(const void* X, const float* y, int64_t n_rows, int64_t row_base, int32_t D,
                                    int32_t dtype, const uint8_t* mask, uint64_t seed, uint32_t iteration,
                                    uint32_t threshold24, float eps, const float* w, float* g_out,
                                    int32_t* cnt_out, void* stream) {
  DAPB_REQUIRE(X && y && w && g_out && cnt_out, "lr_grad_noalloc: null argument");
  DAPB_REQUIRE(dtype == 0 || dtype == 1, "lr_grad_noalloc: dtype must be 0 (fp32) or 1 (bf16)");
  DAPB_REQUIRE(n_rows >= 0, "lr_grad_noalloc: negative row count");
  DAPB_REQUIRE(D >= 1 && D <= DAPB_NOALLOC_MAX_D,
               "lr_grad_noalloc: D=%d is outside [1, %d], the size the built-in scratch array was compiled for; "
               "use dapb_lr_step with caller-provided scratch instead", D, DAPB_NOALLOC_MAX_D);
  int dev = 0;
  DAPB_CUDA_CHECK(cudaGetDevice(&dev));
  DAPB_REQUIRE(dev >= 0 && dev < kMaxDevices, "lr_grad_noalloc: device ordinal %d not supported", dev);
  unsigned char* base = nullptr;
  int grid = 0;
  int rc = noalloc_setup(dev, &base, &grid);      // first call on a device: one cudaGetSymbolAddress
  if (rc != 0) return rc;

  DapbLrStepArgs a;
  memset(&a, 0, sizeof(a));
  a.X = X; a.y = y; a.n_rows = n_rows; a.row_base = row_base; a.D = D; a.dtype = dtype;
  a.mask = mask; a.seed = seed; a.iteration = iteration; a.threshold24 = threshold24; a.eps = eps;
  a.w_fwd = w;
  a.ticket = reinterpret_cast<uint32_t*>(base);
  a.counts = reinterpret_cast<int32_t*>(base + kNoAllocTicketBytes);
  a.partials = reinterpret_cast<float*>(base + kNoAllocTicketBytes + kNoAllocCountsBytes);
  a.grid = grid;
  a.fin = DAPB_LR_FIN_GRAD_ONLY;
  a.g_out = g_out; a.cnt_out = cnt_out;
  DapbComm c;
  memset(&c, 0, sizeof(c));
  c.world = 1;
  // the array keeps the previous call's contents: the ticket must start at zero
  DAPB_CUDA_CHECK(cudaMemsetAsync(a.ticket, 0, sizeof(uint32_t), (cudaStream_t)stream));
  return dapb_lr_step(&a, &c, stream);
}

extern "C" uint64_t dapb_lr_grad_noalloc_scratch_bytes(void) { return (uint64_t)kNoAllocScratchBytes; }
