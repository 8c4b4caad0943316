// F4 -- PageRank: CSR(A^T) SpMV fused with the damping / teleport update and with the all-gather of
// the contribution vector (every finished entry is pushed straight into every peer's copy over
// NVLink; no separate collective, no host round trip).
//
// ref: graph_computation/pagerank.py:21-25 (computeContribs: rank / out-degree to every neighbour),
//      :52-57 (join, flatMap, reduceByKey(add), mapValues(q/n + (1-q)*rank)).
// Reference quirks kept: n = number of distinct SOURCE vertices; a vertex that
// receives no contribution in an iteration disappears (rank 0, contributes nothing afterwards);
// sinks receive rank but their mass is dropped.  Vector form:
//      s[v] = sum_{u->v} c[u];  r[v] = s[v] > 0 ? q/n + (1-q) s[v] : 0;  c'[v] = r[v] / outdeg[v].
//
// Work split: rows are grouped on the host into blocks of ~2048 non-zeros (models/pagerank.py), so the skewed
// R-MAT degree distribution is balanced by construction.
//
// Sharding: rank g owns destination vertices [v0, v1) and the CSR rows of A^T for them; the
// contribution vector c (all V entries) is replicated in the symmetric heap, double-buffered on the
// iteration parity: iteration t reads buffer t&1 and pushes into buffer (t+1)&1 of every rank.
#include <cstdlib>
#include "comm.cuh"

using namespace dapb;

namespace {

constexpr int kPrThreads = 256;
constexpr int kPrCap = 4096;         // gathered contributions staged per row block (16 KB of shared memory)

struct PrArgs {
  const int64_t* row_ptr;   // [nv_local + 1], offsets into col_idx
  const int32_t* col_idx;   // [nnz_local] source vertex ids (global)
  const float* inv_outdeg;  // [nv_local] 1/outdeg of the OWNED vertices (0 for sinks)
  const int32_t* blocks;    // [n_blocks + 1] row boundaries: block b covers rows [blocks[b], blocks[b+1])
  int32_t n_blocks;
  int64_t nv_local, v0, V;
  float base, damp;         // q/n and (1-q)
  float* ranks_local;       // [nv_local] out
  uint64_t c_off;           // payload offset of the double-buffered contribution vector [2][V] float
  uint32_t flag_base;       // `world` flags
  uint64_t epoch;           // iteration number, starts at 1
  uint32_t* ticket;
  int32_t wait_prev;        // wait for epoch-1 from every rank before reading
  int32_t skip_empty;       // rows without in-edges already hold 0 in BOTH buffers (from the 3rd iteration on)
  int32_t no_push;          // DAPB_PR_NOPUSH=1: keep contributions local (TIMING EXPERIMENTS ONLY, results are wrong)
};

__device__ __forceinline__ void pr_finish_row_io(const PrArgs& a, const DapbComm& c, float* c_next_local,
                                                  int64_t row, float s, float inv_outdeg) {
  const float r = s > 0.f ? a.base + a.damp * s : 0.f;          // ref pagerank.py:57 + presence rule
  a.ranks_local[row] = r;
  const float cn = r * inv_outdeg;                               // ref pagerank.py:21-25
  const int64_t v = a.v0 + row;
  if (c.world == 1 || a.no_push) c_next_local[v] = cn;
  else if (inv_outdeg != 0.f) push_all<float>(c, a.c_off, (a.epoch & 1) * (uint64_t)a.V + (uint64_t)v, cn);
}
__device__ __forceinline__ void pr_finish_row(const PrArgs& a, const DapbComm& c, float* c_next_local,
                                               int64_t row, float s) {
  pr_finish_row_io(a, c, c_next_local, row, s, a.inv_outdeg[row]);
}

// CSR-stream SpMV: a CTA takes a block of consecutive rows holding <= kPrCap non-zeros, gathers ALL their
// contributions with perfectly parallel, coalesced index reads (many independent loads in flight per thread),
// stages them in shared memory, and then one thread per row sums its segment and runs the fused epilogue
// (consecutive threads -> consecutive vertices -> coalesced local and peer stores).  Blocks that contain a
// row longer than the staging buffer fall back to a CTA-wide strided sum per row.
__global__ void __launch_bounds__(kPrThreads) pagerank_step_kernel(PrArgs a, DapbComm c) {
  __shared__ float vals[kPrCap];
  __shared__ float s_red[kPrThreads / 32];
  __shared__ bool s_last;
  const int tid = threadIdx.x;
  // iteration `epoch` reads parity (epoch-1)&1 and writes parity epoch&1
  const float* c_cur = reinterpret_cast<const float*>(comm_payload(c, c.rank, a.c_off)) + (((a.epoch - 1) & 1) * a.V);
  float* c_next_local = reinterpret_cast<float*>(comm_payload(c, c.rank, a.c_off)) + ((a.epoch & 1) * a.V);

  if (c.world > 1 && a.wait_prev) {                  // all peers' pushes of the previous iteration landed?
    if (tid < c.world) comm_wait_flag(comm_flags(c, c.rank) + a.flag_base + tid, a.epoch - 1);
    __syncthreads();
  }

  for (int blk = blockIdx.x; blk < a.n_blocks; blk += gridDim.x) {
    const int r0 = a.blocks[blk], r1 = a.blocks[blk + 1];
    if (r0 >= r1) continue;
    const int64_t e0 = a.row_ptr[r0], e1 = a.row_ptr[r1];
    const int nnz = (int)min(e1 - e0, (int64_t)(kPrCap + 1));
    if (nnz <= kPrCap) {
      const int32_t* ci = a.col_idx + e0;
      int i = tid;
      for (; i + 3 * kPrThreads < nnz; i += 4 * kPrThreads) {          // 4 independent gathers in flight
        const int32_t u0 = ci[i], u1 = ci[i + kPrThreads], u2 = ci[i + 2 * kPrThreads], u3 = ci[i + 3 * kPrThreads];
        const float f0 = __ldcg(c_cur + u0), f1 = __ldcg(c_cur + u1), f2 = __ldcg(c_cur + u2), f3 = __ldcg(c_cur + u3);
        vals[i] = f0; vals[i + kPrThreads] = f1; vals[i + 2 * kPrThreads] = f2; vals[i + 3 * kPrThreads] = f3;
      }
      for (; i < nnz; i += kPrThreads) vals[i] = __ldcg(c_cur + ci[i]);
      __syncthreads();
      for (int r = r0 + tid; r < r1; r += kPrThreads) {
        const int b0 = (int)(a.row_ptr[r] - e0), b1 = (int)(a.row_ptr[r + 1] - e0);
        if (b1 == b0 && a.skip_empty) continue;          // rank and contribution stay 0: nothing to store or push
        float s = 0.f;
        for (int j = b0; j < b1; ++j) s += vals[j];
        pr_finish_row(a, c, c_next_local, r, s);
      }
      __syncthreads();
    } else {
      for (int r = r0; r < r1; ++r) {                                   // block with a very long row
        const int64_t b0 = a.row_ptr[r], b1 = a.row_ptr[r + 1];
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        int64_t e = b0 + tid;
        for (; e + 3 * kPrThreads < b1; e += 4 * kPrThreads) {
          s0 += __ldcg(c_cur + a.col_idx[e]);
          s1 += __ldcg(c_cur + a.col_idx[e + kPrThreads]);
          s2 += __ldcg(c_cur + a.col_idx[e + 2 * kPrThreads]);
          s3 += __ldcg(c_cur + a.col_idx[e + 3 * kPrThreads]);
        }
        for (; e < b1; e += kPrThreads) s0 += __ldcg(c_cur + a.col_idx[e]);
        float s = warp_sum((s0 + s1) + (s2 + s3));
        if ((tid & 31) == 0) s_red[tid >> 5] = s;
        __syncthreads();
        if (tid == 0) {
          float t = 0.f;
#pragma unroll
          for (int w = 0; w < kPrThreads / 32; ++w) t += s_red[w];
          pr_finish_row(a, c, c_next_local, r, t);
        }
        __syncthreads();
      }
    }
  }

  // ---- completion: last CTA tells every rank that this rank's slice of iteration `epoch` is in ----
  __threadfence_system();
  __syncthreads();
  if (tid == 0) {
    unsigned t = atomicAdd(a.ticket, 1u);
    s_last = (t == gridDim.x - 1);
    if (s_last) *a.ticket = 0;
  }
  __syncthreads();
  if (!s_last || c.world == 1) return;
  if (tid == 0) __threadfence_system();
  __syncthreads();
  if (tid < c.world) st_release_sys(comm_flags(c, tid) + a.flag_base + c.rank, a.epoch);
}

// =================================================================================================
// Persistent warp-granular kernel (dapb_pagerank_run): ALL iterations of a run in ONE cooperative launch.
//
// Every WARP owns a row block of <= 512 (non-zeros + rows): 16 independent gathers per lane are in flight before the
// first is consumed, the staging buffer is warp-private (__syncwarp only) and warps never wait for each other.
// Rows too long for that are taken out of the main CSR on the host ("hub" rows, R-MAT has a few thousand): their
// non-zeros are cut into 2048-element chunks, one warp per chunk writes a partial sum and bumps the row's counter; the
// warp that completes a row adds the partials IN CHUNK ORDER (deterministic) and runs the same fused epilogue.
// Work items (hub chunks first, then row blocks) are handed out through an atomic counter, four at a time.
//
// Cross-GPU protocol per iteration `e` (no host involvement, no extra launch):
//   1. every CTA spins on this rank's LOCAL flag words until all W ranks (this one included) published e - 1
//      -- the own flag doubles as the grid-wide barrier between iterations;
//   2. SpMV from parity (e-1)&1; each finished vertex that HAS out-edges is pushed into parity e&1 of every rank's
//      contribution vector: one multimem.st through the NVLS multicast mapping (the switch replicates it, 4 B of
//      egress per vertex) or, without multicast, W unicast stores (sinks contribute 0 forever and are never pushed);
//   3. fence.sys, CTA ticket; the last CTA publishes e to every rank.
// Value type float (benchmarks) or double (parity with the reference's printed ranks, pagerank.py:60-68).
// =================================================================================================
constexpr int kWarpCap = 512;
constexpr int kWarpsPerCta = kPrThreads / 32;
// work items per atomic are chosen per launch (see pagerank_run_impl; DAPB_PR_GRAB overrides): an item keeps a warp busy
// for tens of microseconds once 48 warps share an SM's gather bandwidth, so the end-of-iteration tail is one GRAB long
constexpr int kHubDeep = 16;         // measured: 8 -> 16 gathers in flight per lane +5..8 %, 32 no better

template <typename VT>
struct PrRunArgs {
  const int64_t* row_ptr;   // [nv_local + 1] CSR WITHOUT the hub rows' non-zeros
  const int32_t* col_idx;
  const VT* inv_outdeg;     // [nv_local] float: 1/outdeg, double: outdeg of the owned vertices (0 for sinks)
  const int32_t* blocks;    // [n_blocks + 1] row blocks of <= kWarpCap (non-zeros + rows)
  int32_t n_blocks;
  int64_t nv_local, v0, V;
  VT base, damp;            // q/n and (1-q)
  VT* ranks_local;          // [nv_local] out
  uint64_t c_off;           // payload offset of the contribution vector [2][V] VT
  uint32_t flag_base;
  uint64_t epoch0;          // epoch of the first iteration of this launch
  int32_t n_iters;
  int32_t it0;              // iterations done since init (rows without in-edges are skipped once it >= 2)
  int32_t wait_first;       // wait for epoch0 - 1 before the first iteration (0 right after init + host barrier)
  int32_t skip_empty_ok;
  uint32_t* sync;           // [4] ticket, work counter parity 0 / 1, pad -- zero at launch
  int32_t grab;             // work items taken per atomic
  unsigned long long* stats;  // optional [n_iters][4] globaltimer stamps: enter, wait passed, last ticket, published
  // hub rows
  const int32_t* hub_rows; const int32_t* hub_col; const int64_t* chunk_start; const int64_t* chunk_end;
  const int32_t* hub_chunk_ptr; const int32_t* chunk_row; VT* chunk_partial; uint32_t* row_done;
  int32_t n_hub, n_chunks;
};

template <typename VT>
__device__ __forceinline__ bool pr_is_hub(const PrRunArgs<VT>& a, int32_t row) {
  int lo = 0, hi = a.n_hub;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    const int32_t v = a.hub_rows[mid];
    if (v == row) return true;
    if (v < row) lo = mid + 1; else hi = mid;
  }
  return false;
}

// Value-type policy.  float: `deg` holds 1/outdeg and a*b+c may contract to an FMA.  double (parity mode): `deg` holds
// outdeg itself and the arithmetic is the reference's, operation by operation -- q/n + (1-q)*s with a separately rounded
// product (pagerank.py:57) and rank / len(neighbors) as a true division (pagerank.py:24) -- so the printed digits match.
__device__ __forceinline__ float pr_rank(float base, float damp, float s) { return base + damp * s; }
__device__ __forceinline__ double pr_rank(double base, double damp, double s) { return __dadd_rn(base, __dmul_rn(damp, s)); }
__device__ __forceinline__ float pr_contrib(float r, float deg) { return r * deg; }
__device__ __forceinline__ double pr_contrib(double r, double deg) { return __ddiv_rn(r, deg); }

// fused epilogue of one finished row (ref pagerank.py:57 + presence rule, :21-25)
template <typename VT>
__device__ __forceinline__ void pr_emit(const PrRunArgs<VT>& a, const DapbComm& c, uint64_t next_off, int64_t row, VT s,
                                        VT deg) {
  const VT r = s > VT(0) ? pr_rank(a.base, a.damp, s) : VT(0);
  a.ranks_local[row] = r;
  if (deg != VT(0)) push_all<VT>(c, a.c_off, next_off + (uint64_t)(a.v0 + row), pr_contrib(r, deg));
}

// lane-strided sum of c_cur[col[e]] for e in [b0, b1), up to kDeep gathers in flight per lane (predicated in the last
// batch: a plain remainder loop would keep ONE gather in flight, and most hub rows are shorter than 256);
// result valid on every lane
template <int kDeep, typename VT>
__device__ __forceinline__ VT pr_warp_row_sum(const VT* __restrict__ c_cur, const int32_t* __restrict__ col,
                                              int64_t b0, int64_t b1, int lane) {
  VT s0 = 0, s1 = 0, s2 = 0, s3 = 0;
  int64_t e = b0 + lane;
  for (; e + (kDeep - 1) * 32 < b1; e += kDeep * 32) {
    int32_t u[kDeep];
#pragma unroll
    for (int j = 0; j < kDeep; ++j) u[j] = col[e + j * 32];
    VT f[kDeep];
#pragma unroll
    for (int j = 0; j < kDeep; ++j) f[j] = __ldcg(c_cur + u[j]);
    _Pragma("unroll") for (int j = 0; j < kDeep; j += 4) { s0 += f[j]; s1 += f[j + 1]; s2 += f[j + 2]; s3 += f[j + 3]; }
  }
  if (e < b1) {                                            // warp-uniform tail of fewer than kDeep full strides
    int32_t u[kDeep];
#pragma unroll
    for (int j = 0; j < kDeep; ++j) u[j] = e + j * 32 < b1 ? col[e + j * 32] : -1;
    VT f[kDeep];
#pragma unroll
    for (int j = 0; j < kDeep; ++j) f[j] = u[j] >= 0 ? __ldcg(c_cur + u[j]) : VT(0);
    _Pragma("unroll") for (int j = 0; j < kDeep; j += 4) { s0 += f[j]; s1 += f[j + 1]; s2 += f[j + 2]; s3 += f[j + 3]; }
  }
  VT s = (s0 + s1) + (s2 + s3);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  return s;
}

template <typename VT>
__global__ void __launch_bounds__(kPrThreads, sizeof(VT) == 4 ? 6 : 4) pagerank_run_kernel(PrRunArgs<VT> a, DapbComm c) {
  __shared__ VT vals_all[kWarpsPerCta][kWarpCap];
  __shared__ bool s_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  VT* vals = vals_all[warp];
  const int n_items = a.n_chunks + a.n_blocks;
  const uint64_t* my_flags = comm_flags(c, c.rank) + a.flag_base;

  for (int it = 0; it < a.n_iters; ++it) {
    const uint64_t epoch = a.epoch0 + it;
    const bool skip_empty = a.skip_empty_ok && (a.it0 + it >= 2);
    const VT* c_cur = reinterpret_cast<const VT*>(comm_payload(c, c.rank, a.c_off)) + (((epoch - 1) & 1) * a.V);
    const uint64_t next_off = (epoch & 1) * (uint64_t)a.V;
    if (a.stats && blockIdx.x == 0 && tid == 0) a.stats[it * 4 + 0] = global_timer_ns();
    if (it > 0 || a.wait_first) {                      // every rank (and every CTA of this one) finished epoch - 1?
      if (tid < c.world) comm_wait_flag(my_flags + tid, epoch - 1);
      __syncthreads();
    }
    if (a.stats && blockIdx.x == 0 && tid == 0) a.stats[it * 4 + 1] = global_timer_ns();
    uint32_t* work = a.sync + 1 + (epoch & 1);

    for (;;) {
      int first = 0;
      if (lane == 0) first = (int)atomicAdd(work, (uint32_t)a.grab);
      first = __shfl_sync(0xffffffffu, first, 0);
      if (first >= n_items) break;
      const int last = min(first + a.grab, n_items);
      for (int item = first; item < last; ++item) {
        if (item < a.n_chunks) {
          // ---- one chunk of a hub row ----
          const VT s = pr_warp_row_sum<kHubDeep>(c_cur, a.hub_col, a.chunk_start[item], a.chunk_end[item], lane);
          if (lane == 0) {
            const int hi = a.chunk_row[item];
            const int ch0 = a.hub_chunk_ptr[hi], ch1 = a.hub_chunk_ptr[hi + 1];
            if (ch1 - ch0 == 1) {
              // the common case (half a million rows of 65..2048 in-edges at scale 24): the row IS the chunk -- no
              // partial, no fence, no counter.  (ncu showed the two __threadfence of the general path, executed once per
              // chunk, as the top stall of the kernel: MEMBAR.SC.GPU + CCTL.IVALL, 16 % stall_membar; with peers they
              // also wait for the warp's outstanding NVLink stores.)
              const int32_t row = a.hub_rows[hi];
              pr_emit(a, c, next_off, row, s, a.inv_outdeg[row]);
              continue;
            }
            a.chunk_partial[item] = s;
            __threadfence();
            const uint32_t done = atomicAdd(a.row_done + hi, 1u);
            if (done == (uint32_t)(ch1 - ch0 - 1)) {                 // this warp completed the row
              __threadfence();
              VT t = 0;
              for (int ch = ch0; ch < ch1; ++ch) t += __ldcg(a.chunk_partial + ch);
              a.row_done[hi] = 0;
              const int32_t row = a.hub_rows[hi];
              pr_emit(a, c, next_off, row, t, a.inv_outdeg[row]);
            }
          }
          continue;
        }
        // ---- one row block ----
        const int blk = item - a.n_chunks;
        const int r0 = a.blocks[blk], r1 = a.blocks[blk + 1];
        if (r0 >= r1) continue;
        const int64_t e0 = a.row_ptr[r0], e1 = a.row_ptr[r1];
        if (e1 - e0 <= kWarpCap) {
          const int nnz = (int)(e1 - e0);
          const int32_t* ci = a.col_idx + e0;
          // row extents and 1/outdeg of the first 32 rows are requested BEFORE the gathers so that their latency overlaps
          int r = r0 + lane;
          int64_t p0 = 0, p1 = 0;
          VT io = 0;
          if (r < r1) { p0 = a.row_ptr[r]; p1 = a.row_ptr[r + 1]; io = a.inv_outdeg[r]; }
          int32_t u[kWarpCap / 32];
#pragma unroll
          for (int j = 0; j < kWarpCap / 32; ++j) u[j] = lane + j * 32 < nnz ? ci[lane + j * 32] : -1;
          VT f[kWarpCap / 32];
#pragma unroll
          for (int j = 0; j < kWarpCap / 32; ++j) f[j] = u[j] >= 0 ? __ldcg(c_cur + u[j]) : VT(0);   // 16 gathers in flight
#pragma unroll
          for (int j = 0; j < kWarpCap / 32; ++j) vals[lane + j * 32] = f[j];
          __syncwarp();
          while (r < r1) {
            const int rn = r + 32;                                         // next row of this lane: prefetch its extents
            int64_t q0 = 0, q1 = 0;
            VT ion = 0;
            if (rn < r1) { q0 = a.row_ptr[rn]; q1 = a.row_ptr[rn + 1]; ion = a.inv_outdeg[rn]; }
            const int b0 = (int)(p0 - e0), b1 = (int)(p1 - e0);
            if (!(b1 == b0 && (skip_empty || pr_is_hub(a, r)))) {          // else: stays 0 / written by the hub path
              VT s = 0;
              for (int j = b0; j < b1; ++j) s += vals[j];
              pr_emit(a, c, next_off, r, s, io);
            }
            r = rn; p0 = q0; p1 = q1; io = ion;
          }
          __syncwarp();
        } else {                                                            // block with a row of 512 .. hub_threshold non-zeros
          for (int r = r0; r < r1; ++r) {
            const int64_t b0 = a.row_ptr[r], b1 = a.row_ptr[r + 1];
            if (b1 == b0 && (skip_empty || pr_is_hub(a, r))) continue;
            const VT s = pr_warp_row_sum<8>(c_cur, a.col_idx, b0, b1, lane);
            if (lane == 0) pr_emit(a, c, next_off, r, s, a.inv_outdeg[r]);
          }
        }
      }
    }

    // ---- completion: the last CTA tells every rank (this one included) that this rank's slice of `epoch` is in ----
    __threadfence_system();
    __syncthreads();
    if (tid == 0) {
      const unsigned t = atomicAdd(a.sync, 1u);
      s_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (s_last) {
      if (tid == 0) {
        if (a.stats) a.stats[it * 4 + 2] = global_timer_ns();
        a.sync[0] = 0;
        a.sync[1 + ((epoch + 1) & 1)] = 0;             // next iteration's work counter
        __threadfence_system();
      }
      __syncthreads();
      if (tid < c.world) st_release_sys(comm_flags(c, tid) + a.flag_base + c.rank, epoch);
      if (a.stats && tid == 0) a.stats[it * 4 + 3] = global_timer_ns();
    }
  }
}

// contribution vector of iteration 0: c[v] = (1/n) / outdeg[v] for source vertices, written by the owner of v into
// parity 0 of every rank's copy (ref pagerank.py:47); parity 1 is cleared, so entries that are never pushed again
// (sinks, and from the third iteration on rows without in-edges) read 0 in both buffers whatever the region held.
template <typename VT>
__global__ void pagerank_init_kernel(DapbComm c, const VT* inv_outdeg, int64_t nv_local, int64_t v0, int64_t V,
                                     VT r0, uint64_t c_off) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nv_local; i += (int64_t)gridDim.x * blockDim.x) {
    const VT io = inv_outdeg[i];
    push_all<VT>(c, c_off, (uint64_t)(v0 + i), io > VT(0) ? pr_contrib(r0, io) : VT(0));
    push_all<VT>(c, c_off, (uint64_t)(V + v0 + i), VT(0));
  }
}

}  // namespace

extern "C" int dapb_pagerank_init(const DapbComm* comm, const void* inv_outdeg, int64_t nv_local, int64_t v0,
                                  int64_t V, double r0, uint64_t c_off, int32_t f64, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "pagerank_init: bad comm");
  uint64_t need = DAPB_FLAG_BYTES + c_off + 2ull * V * (f64 ? 8 : 4);
  DAPB_REQUIRE(need <= comm->heap_bytes, "pagerank_init: heap too small (%llu > %llu)", (unsigned long long)need,
               (unsigned long long)comm->heap_bytes);
  int blocks = (int)((nv_local + 255) / 256);
  if (blocks > dapb::sm_count() * 8) blocks = dapb::sm_count() * 8;
  if (blocks < 1) blocks = 1;
  if (f64)
    pagerank_init_kernel<double><<<blocks, 256, 0, (cudaStream_t)stream>>>(*comm, (const double*)inv_outdeg, nv_local, v0, V, r0, c_off);
  else
    pagerank_init_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>(*comm, (const float*)inv_outdeg, nv_local, v0, V, (float)r0, c_off);
  DAPB_LAUNCH_CHECK();
  return 0;
}

extern "C" int dapb_pagerank_step(const DapbComm* comm, const int64_t* row_ptr, const int32_t* col_idx,
                                  const float* inv_outdeg, const int32_t* blocks, int32_t n_blocks,
                                  int64_t nv_local, int64_t v0, int64_t V, float base,
                                  float damp, float* ranks_local, uint64_t c_off, uint32_t flag_base,
                                  uint64_t epoch, uint32_t* ticket, int32_t wait_prev, int32_t skip_empty,
                                  int32_t grid, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "pagerank_step: bad comm");
  DAPB_REQUIRE(epoch >= 1 && ticket != nullptr, "pagerank_step: epoch starts at 1, ticket required");
  DAPB_REQUIRE(blocks != nullptr && n_blocks >= 0, "pagerank_step: row blocks required (see build_sharded_graph)");
  PrArgs a;
  a.row_ptr = row_ptr; a.col_idx = col_idx; a.inv_outdeg = inv_outdeg; a.blocks = blocks; a.n_blocks = n_blocks;
  a.nv_local = nv_local; a.v0 = v0; a.V = V; a.base = base; a.damp = damp; a.ranks_local = ranks_local;
  a.c_off = c_off; a.flag_base = flag_base; a.epoch = epoch; a.ticket = ticket; a.wait_prev = wait_prev; a.skip_empty = skip_empty;
  { static const int np = getenv("DAPB_PR_NOPUSH") ? atoi(getenv("DAPB_PR_NOPUSH")) : 0; a.no_push = np; }
  if (grid <= 0) grid = dapb::sm_count() * 8;
  if (grid > n_blocks) grid = n_blocks > 0 ? n_blocks : 1;
  pagerank_step_kernel<<<grid, kPrThreads, 0, (cudaStream_t)stream>>>(a, *comm);
  DAPB_LAUNCH_CHECK();
  return 0;
}

namespace {
template <typename VT>
int pagerank_run_impl(const DapbPrRunArgs& r, const DapbComm& c, cudaStream_t st) {
  PrRunArgs<VT> a;
  a.row_ptr = r.row_ptr; a.col_idx = r.col_idx; a.inv_outdeg = (const VT*)r.inv_outdeg; a.blocks = r.blocks;
  a.n_blocks = r.n_blocks; a.nv_local = r.nv_local; a.v0 = r.v0; a.V = r.V; a.base = (VT)r.base; a.damp = (VT)r.damp;
  a.ranks_local = (VT*)r.ranks_local; a.c_off = r.c_off; a.flag_base = r.flag_base; a.epoch0 = r.epoch0;
  a.n_iters = r.n_iters; a.it0 = r.it0; a.wait_first = r.wait_first; a.skip_empty_ok = r.skip_empty_ok;
  a.sync = r.sync; a.stats = r.stats;
  a.hub_rows = r.hub.hub_rows; a.hub_col = r.hub.hub_col; a.chunk_start = r.hub.chunk_start; a.chunk_end = r.hub.chunk_end;
  a.hub_chunk_ptr = r.hub.hub_chunk_ptr; a.chunk_row = r.hub.chunk_row; a.chunk_partial = (VT*)r.hub.chunk_partial;
  a.row_done = r.hub.row_done; a.n_hub = r.hub.n_hub; a.n_chunks = r.hub.n_chunks;
  int per_sm = 0;
  DAPB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pagerank_run_kernel<VT>, kPrThreads, 0));
  DAPB_REQUIRE(per_sm >= 1, "pagerank_run: kernel does not fit on an SM");
  int grid = per_sm * dapb::sm_count();                 // every CTA must be resident: they wait for each other's flag
  if (r.grid > 0 && r.grid < grid) grid = r.grid;
  const int64_t items = (int64_t)a.n_chunks + a.n_blocks;
  {
    // items per atomic: one when a warp sees few items per iteration (the tail of the iteration is one grab long), more
    // when there are hundreds per warp (otherwise ~3e5 same-address atomics per millisecond become the bottleneck:
    // scale 26 on 8 GPUs measured 1.12 ms with grab 1 against 0.98 ms with grab 4)
    static const int g_env = getenv("DAPB_PR_GRAB") ? atoi(getenv("DAPB_PR_GRAB")) : 0;
    const int64_t per_warp = items / ((int64_t)grid * kWarpsPerCta) + 1;
    const int64_t g_auto = per_warp / 16 < 1 ? 1 : (per_warp / 16 > 8 ? 8 : per_warp / 16);
    a.grab = g_env > 0 ? g_env : (int)g_auto;
  }
  const int64_t need = (items + kWarpsPerCta * a.grab - 1) / (kWarpsPerCta * a.grab);
  if (grid > need) grid = need > 0 ? (int)need : 1;
  DAPB_CUDA_CHECK(cudaMemsetAsync(a.sync, 0, 4 * sizeof(uint32_t), st));
  DapbComm cc = c;
  void* kargs[] = {&a, &cc};
  DAPB_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)pagerank_run_kernel<VT>, dim3(grid), dim3(kPrThreads), kargs, 0, st));
  return 0;
}
}  // namespace

// Persistent run: n_iters iterations in one cooperative launch (see pagerank_run_kernel).
extern "C" int dapb_pagerank_run(const DapbPrRunArgs* args, const DapbComm* comm, void* stream) {
  DAPB_REQUIRE(args && comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "pagerank_run: bad comm");
  const DapbPrRunArgs& r = *args;
  DAPB_REQUIRE(r.epoch0 >= 1 && r.n_iters >= 1 && r.sync != nullptr, "pagerank_run: epoch starts at 1, sync words required");
  DAPB_REQUIRE(r.blocks != nullptr && r.n_blocks >= 0, "pagerank_run: row blocks required (see build_sharded_graph)");
  DAPB_REQUIRE(r.hub.n_hub >= 0 && r.hub.n_chunks >= 0 && (r.hub.n_hub == 0 || (r.hub.hub_rows && r.hub.hub_col &&
               r.hub.chunk_start && r.hub.chunk_end && r.hub.hub_chunk_ptr && r.hub.chunk_partial && r.hub.chunk_row &&
               r.hub.row_done)), "pagerank_run: incomplete hub description");
  DAPB_REQUIRE(DAPB_FLAG_BYTES + r.c_off + 2ull * r.V * (r.f64 ? 8 : 4) <= comm->heap_bytes, "pagerank_run: heap too small");
  return r.f64 ? pagerank_run_impl<double>(r, *comm, (cudaStream_t)stream)
               : pagerank_run_impl<float>(r, *comm, (cudaStream_t)stream);
}
