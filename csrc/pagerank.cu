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
  if (c.world == 1 || a.no_push) {
    c_next_local[v] = cn;
  } else {
#pragma unroll 1
    for (int g = 0; g < c.world; ++g)
      reinterpret_cast<float*>(comm_payload(c, g, a.c_off))[((a.epoch & 1) * a.V) + v] = cn;
  }
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
// Warp-granular variant (dapb_pagerank_step_v2).  The CTA-wide kernel above spends ~40 % of its warp time in
// the two __syncthreads per row block (ncu: stall_barrier) with only 4 gathers in flight per thread.  Here every
// WARP owns a row block of <= 512 non-zeros: 16 independent gathers per lane are in flight before the first is
// consumed, the staging buffer is warp-private (__syncwarp only), and warps never wait for each other.
// Rows too long for that are taken out of the main CSR on the host ("hub" rows, R-MAT has a few thousand):
// their non-zeros are cut into 2048-element chunks, one warp per chunk writes a partial sum
// (pagerank_hub_partial_kernel), and one thread per hub row adds the partials IN CHUNK ORDER (deterministic) and
// runs the same fused epilogue (pagerank_hub_finish_kernel).  Launch order hub partial -> hub finish -> main, so
// the completion flag published by the main kernel's last CTA covers the hub pushes as well.
// =================================================================================================
constexpr int kWarpCap = 512;
constexpr int kWarpsPerCta = kPrThreads / 32;

__device__ __forceinline__ bool pr_is_hub(const DapbPrHub& h, int32_t row) {
  int lo = 0, hi = h.n_hub;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    const int32_t v = h.hub_rows[mid];
    if (v == row) return true;
    if (v < row) lo = mid + 1; else hi = mid;
  }
  return false;
}

__device__ __forceinline__ void pr_wait_prev(const PrArgs& a, const DapbComm& c) {
  if (c.world > 1 && a.wait_prev) {                  // all peers' pushes of the previous iteration landed?
    if (threadIdx.x < (unsigned)c.world) comm_wait_flag(comm_flags(c, c.rank) + a.flag_base + threadIdx.x, a.epoch - 1);
    __syncthreads();
  }
}

// lane-strided sum of c_cur[col[e]] for e in [b0, b1), up to kDeep gathers in flight per lane (predicated in the last
// batch: a plain remainder loop would keep ONE gather in flight, and most hub rows are shorter than 256);
// result valid on every lane
template <int kDeep>   // gathers in flight per lane
__device__ __forceinline__ float pr_warp_row_sum(const float* __restrict__ c_cur, const int32_t* __restrict__ col,
                                                 int64_t b0, int64_t b1, int lane) {
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  int64_t e = b0 + lane;
  for (; e + (kDeep - 1) * 32 < b1; e += kDeep * 32) {
    int32_t u[kDeep];
#pragma unroll
    for (int j = 0; j < kDeep; ++j) u[j] = col[e + j * 32];
    float f[kDeep];
#pragma unroll
    for (int j = 0; j < kDeep; ++j) f[j] = __ldcg(c_cur + u[j]);
    _Pragma("unroll") for (int j = 0; j < kDeep; j += 4) { s0 += f[j]; s1 += f[j + 1]; s2 += f[j + 2]; s3 += f[j + 3]; }
  }
  if (e < b1) {                                            // warp-uniform tail of fewer than 8 full strides
    int32_t u[kDeep];
#pragma unroll
    for (int j = 0; j < kDeep; ++j) u[j] = e + j * 32 < b1 ? col[e + j * 32] : -1;
    float f[kDeep];
#pragma unroll
    for (int j = 0; j < kDeep; ++j) f[j] = u[j] >= 0 ? __ldcg(c_cur + u[j]) : 0.f;
    _Pragma("unroll") for (int j = 0; j < kDeep; j += 4) { s0 += f[j]; s1 += f[j + 1]; s2 += f[j + 2]; s3 += f[j + 3]; }
  }
  return warp_sum((s0 + s1) + (s2 + s3));
}

__global__ void __launch_bounds__(kPrThreads) pagerank_warp_kernel(PrArgs a, DapbPrHub h, DapbComm c) {
  __shared__ float vals_all[kWarpsPerCta][kWarpCap];
  __shared__ bool s_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float* vals = vals_all[warp];
  const float* c_cur = reinterpret_cast<const float*>(comm_payload(c, c.rank, a.c_off)) + (((a.epoch - 1) & 1) * a.V);
  float* c_next_local = reinterpret_cast<float*>(comm_payload(c, c.rank, a.c_off)) + ((a.epoch & 1) * a.V);
  pr_wait_prev(a, c);

  const int n_warps = gridDim.x * kWarpsPerCta;
  for (int blk = blockIdx.x * kWarpsPerCta + warp; blk < a.n_blocks; blk += n_warps) {
    const int r0 = a.blocks[blk], r1 = a.blocks[blk + 1];
    if (r0 >= r1) continue;
    const int64_t e0 = a.row_ptr[r0], e1 = a.row_ptr[r1];
    if (e1 - e0 <= kWarpCap) {
      const int nnz = (int)(e1 - e0);
      const int32_t* ci = a.col_idx + e0;
      // row extents and 1/outdeg of the first 32 rows are requested BEFORE the gathers so that their latency overlaps
      int r = r0 + lane;
      int64_t p0 = 0, p1 = 0;
      float io = 0.f;
      if (r < r1) { p0 = a.row_ptr[r]; p1 = a.row_ptr[r + 1]; io = a.inv_outdeg[r]; }
      int32_t u[kWarpCap / 32];
#pragma unroll
      for (int j = 0; j < kWarpCap / 32; ++j) u[j] = lane + j * 32 < nnz ? ci[lane + j * 32] : -1;
      float f[kWarpCap / 32];
#pragma unroll
      for (int j = 0; j < kWarpCap / 32; ++j) f[j] = u[j] >= 0 ? __ldcg(c_cur + u[j]) : 0.f;   // 16 gathers in flight
#pragma unroll
      for (int j = 0; j < kWarpCap / 32; ++j) vals[lane + j * 32] = f[j];
      __syncwarp();
      while (r < r1) {
        const int rn = r + 32;                                         // next row of this lane: prefetch its extents
        int64_t q0 = 0, q1 = 0;
        float ion = 0.f;
        if (rn < r1) { q0 = a.row_ptr[rn]; q1 = a.row_ptr[rn + 1]; ion = a.inv_outdeg[rn]; }
        const int b0 = (int)(p0 - e0), b1 = (int)(p1 - e0);
        if (!(b1 == b0 && (a.skip_empty || pr_is_hub(h, r)))) {        // else: stays 0 / written by the hub kernels
          float s = 0.f;
          for (int j = b0; j < b1; ++j) s += vals[j];
          pr_finish_row_io(a, c, c_next_local, r, s, io);
        }
        r = rn; p0 = q0; p1 = q1; io = ion;
      }
      __syncwarp();
    } else {                                                            // block with a row of 512 .. hub_threshold non-zeros
      for (int r = r0; r < r1; ++r) {
        const int64_t b0 = a.row_ptr[r], b1 = a.row_ptr[r + 1];
        if (b1 == b0 && (a.skip_empty || pr_is_hub(h, r))) continue;
        const float s = pr_warp_row_sum<8>(c_cur, a.col_idx, b0, b1, lane);
        if (lane == 0) pr_finish_row(a, c, c_next_local, r, s);
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

constexpr int kHubDeep = 16;     // measured: 8 -> 16 gathers in flight per lane +5..8 %, 32 no better
__global__ void __launch_bounds__(kPrThreads) pagerank_hub_partial_kernel(PrArgs a, DapbPrHub h, DapbComm c) {
  const int lane = threadIdx.x & 31;
  const float* c_cur = reinterpret_cast<const float*>(comm_payload(c, c.rank, a.c_off)) + (((a.epoch - 1) & 1) * a.V);
  pr_wait_prev(a, c);
  const int n_warps = gridDim.x * kWarpsPerCta;
  for (int ch = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5); ch < h.n_chunks; ch += n_warps) {
    const float s = pr_warp_row_sum<kHubDeep>(c_cur, h.hub_col, h.chunk_start[ch], h.chunk_end[ch], lane);
    if (lane == 0) h.chunk_partial[ch] = s;
  }
}

__global__ void pagerank_hub_finish_kernel(PrArgs a, DapbPrHub h, DapbComm c) {
  float* c_next_local = reinterpret_cast<float*>(comm_payload(c, c.rank, a.c_off)) + ((a.epoch & 1) * a.V);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= h.n_hub) return;
  float s = 0.f;
  for (int ch = h.hub_chunk_ptr[i]; ch < h.hub_chunk_ptr[i + 1]; ++ch) s += h.chunk_partial[ch];
  pr_finish_row(a, c, c_next_local, h.hub_rows[i], s);
}

// contribution vector of iteration 0: c[v] = (1/n) / outdeg[v] for source vertices, written by the
// owner of v into every rank's buffer 0 (ref pagerank.py:47).
__global__ void pagerank_init_kernel(DapbComm c, const float* inv_outdeg, int64_t nv_local, int64_t v0, int64_t V,
                                     float r0, uint64_t c_off) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nv_local; i += (int64_t)gridDim.x * blockDim.x) {
    const float io = inv_outdeg[i];
    const float cn = io > 0.f ? r0 * io : 0.f;
    for (int g = 0; g < c.world; ++g) reinterpret_cast<float*>(comm_payload(c, g, c_off))[v0 + i] = cn;
  }
}

}  // namespace

extern "C" int dapb_pagerank_init(const DapbComm* comm, const float* inv_outdeg, int64_t nv_local, int64_t v0,
                                  int64_t V, float r0, uint64_t c_off, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "pagerank_init: bad comm");
  uint64_t need = DAPB_FLAG_BYTES + c_off + 2ull * V * sizeof(float);
  DAPB_REQUIRE(need <= comm->heap_bytes, "pagerank_init: heap too small (%llu > %llu)", (unsigned long long)need,
               (unsigned long long)comm->heap_bytes);
  int blocks = (int)((nv_local + 255) / 256);
  if (blocks > dapb::sm_count() * 8) blocks = dapb::sm_count() * 8;
  if (blocks < 1) blocks = 1;
  pagerank_init_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(*comm, inv_outdeg, nv_local, v0, V, r0, c_off);
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

// Warp-granular step (see pagerank_warp_kernel): `blocks` must hold row blocks of <= 512 non-zeros + rows built on the
// CSR WITHOUT the hub rows' non-zeros (hub rows have length 0 there); `hub` describes the hub rows (may be empty).
extern "C" int dapb_pagerank_step_v2(const DapbComm* comm, const int64_t* row_ptr, const int32_t* col_idx,
                                     const float* inv_outdeg, const int32_t* blocks, int32_t n_blocks,
                                     int64_t nv_local, int64_t v0, int64_t V, const DapbPrHub* hub, float base,
                                     float damp, float* ranks_local, uint64_t c_off, uint32_t flag_base,
                                     uint64_t epoch, uint32_t* ticket, int32_t wait_prev, int32_t skip_empty,
                                     int32_t grid, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "pagerank_step_v2: bad comm");
  DAPB_REQUIRE(epoch >= 1 && ticket != nullptr && hub != nullptr, "pagerank_step_v2: epoch starts at 1, ticket and hub required");
  DAPB_REQUIRE(blocks != nullptr && n_blocks >= 0, "pagerank_step_v2: row blocks required (see build_sharded_graph)");
  DAPB_REQUIRE(hub->n_hub >= 0 && hub->n_chunks >= 0 && (hub->n_hub == 0 || (hub->hub_rows && hub->hub_col &&
               hub->chunk_start && hub->chunk_end && hub->hub_chunk_ptr && hub->chunk_partial)),
               "pagerank_step_v2: incomplete hub description");
  PrArgs a;
  a.row_ptr = row_ptr; a.col_idx = col_idx; a.inv_outdeg = inv_outdeg; a.blocks = blocks; a.n_blocks = n_blocks;
  a.nv_local = nv_local; a.v0 = v0; a.V = V; a.base = base; a.damp = damp; a.ranks_local = ranks_local;
  a.c_off = c_off; a.flag_base = flag_base; a.epoch = epoch; a.ticket = ticket; a.wait_prev = wait_prev; a.skip_empty = skip_empty;
  { static const int np = getenv("DAPB_PR_NOPUSH") ? atoi(getenv("DAPB_PR_NOPUSH")) : 0; a.no_push = np; }
  const DapbPrHub h = *hub;
  cudaStream_t st = (cudaStream_t)stream;
  const int cap = dapb::sm_count() * 8;
  if (h.n_hub > 0) {
    int g1 = (h.n_chunks + kWarpsPerCta - 1) / kWarpsPerCta;
    if (g1 > cap) g1 = cap;
    pagerank_hub_partial_kernel<<<g1, kPrThreads, 0, st>>>(a, h, *comm);
    DAPB_LAUNCH_CHECK();
    pagerank_hub_finish_kernel<<<(h.n_hub + 127) / 128, 128, 0, st>>>(a, h, *comm);
    DAPB_LAUNCH_CHECK();
  }
  if (grid <= 0) grid = cap;
  const int need = (n_blocks + kWarpsPerCta - 1) / kWarpsPerCta;
  if (grid > need) grid = need > 0 ? need : 1;
  pagerank_warp_kernel<<<grid, kPrThreads, 0, st>>>(a, h, *comm);
  DAPB_LAUNCH_CHECK();
  return 0;
}
