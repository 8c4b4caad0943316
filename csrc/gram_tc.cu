// F5 (tensor-core path) -- Gram matrix G = A^T A of a row-sharded tall-skinny bf16 matrix on the 5th-gen
// tensor cores, with the row-split reduction done by vector atomics in L2 and the cross-GPU reduction done
// from inside the same kernel over NVLink peer memory.
//
// ref math: matrix_computation/matrix_decomposition.py:27 (XtX = mat^T mat) -- the reference builds this k x k
// Gram once per row on the driver; BASELINE.json config 5 scales it to a 2^27 x 512 matrix.
//
// G[i, j] = sum_r A[r, i] A[r, j].  The contraction runs over ROWS, so both MMA operands are the same
// shared-memory tile [KR rows x columns] read in MN-major form (the M / N index is the contiguous one):
//   TMA loads the tile as boxes of [KR rows x 64 columns] with the 128-byte swizzle; in UMMA terms each box
//   is an MN-major SW128 atom column: 8-row groups 1024 B apart (SBO), 64-column groups one box apart (LBO).
//
// SYRK schedule.  G is symmetric, so only the 128 x 128 blocks (i, j) with i <= j are computed -- 10 of 16 for
// P = 512 -- and off-diagonal blocks are written twice by the epilogue (as is and transposed).  The blocks are dealt
// to ROLES of equal work (P = 512: six roles of one M128 x N256 MMA chain each, see make_roles); a CTA is (role,
// row split), loads only the column strips its blocks touch (2 or 3 of 4), and keeps its <= 2 accumulators
// (128 lanes x 256 columns) in TMEM for the whole kernel, so there is ONE epilogue: TMEM -> red.global.add into G.
// All roles of a row split cost the same per stage, walk the same rows at the same pace, and so HBM sees every row
// once (L2 serves the other roles).
//
// Cross-GPU reduction (world > 1), same launch, whole grid: the row-split atomics go into a G-sized buffer in the
// symmetric heap; after a grid barrier + rank barrier every rank reduces ITS 1/W slice of G over all ranks
// (multimem.ld_reduce through the NVLS mapping, or W peer loads added in rank order) and broadcasts the sums
// (multimem.st, or W peer stores); rank barrier; every rank copies the finished G out.  All CTAs take part, so the
// 1 MB exchange is spread over the whole GPU instead of one CTA per strip.
#include <cooperative_groups.h>
#include <cstdlib>
#include "comm.cuh"
#include "tc_common.cuh"

using namespace dapb;
using namespace dapb::tc;

namespace {

namespace cg = cooperative_groups;

constexpr int kKR = 64;            // rows per pipeline stage
constexpr int kStages = 4;
constexpr int kGThreads = 256;     // warp 0 TMA, warp 1 MMA, warp 2 TMEM alloc, warps 4-7 epilogue
constexpr int kMaxRoles = 6;

// One role = up to two MMA groups on up to three loaded column strips (see the schedule above).
struct GramRole {
  int n_strips;            // column strips (128 columns each) this role loads, in ascending order
  int strips[3];
  int n_groups;            // MMA groups per K step
  int a_pos[2], b_pos[2];  // operand positions within the loaded strips (B spans b_pos, b_pos + 1 when n == 256)
  int n[2];                // N of the group's MMA: 128 or 256
  int tmem_col[2];         // accumulator column of the group
  int n_out;               // 128 x 128 output blocks
  int out_i[2], out_j[2];  // block coordinates (strip units)
  int out_col[2];          // TMEM column of output block k
};

struct GramTcArgs {
  int64_t n_rows;
  int P;                  // columns (multiple of 128, <= 512)
  int splits;             // row splits (grid = n_roles * splits)
  int n_roles;
  float* G;               // [P, P] fp32 result
  float* G_acc;           // where the row-split atomics accumulate: G itself (world == 1) or the heap buffer
  uint64_t payload_off;   // symmetric-heap offset of that buffer
  uint32_t flag_base;
  uint64_t epoch;
  int do_allreduce;
  uint32_t smem_bytes;
  int* progress;          // [splits][kMaxRoles] stage counters of the pacing protocol (zeroed by the host wrapper)
  int pace_every, pace_window;
  int prefetch;           // L2 prefetch distance in stages (0 = off)
  int dbg;                // DAPB_GRAM_DBG=1: per-CTA cycle counters into progress[512 + 4 * blockIdx.x ..] (>> 10)
  GramRole roles[kMaxRoles];
};

// Optional pacing / L2 prefetch (both OFF by default -- measured, they do not pay):
//  * pacing: every pace_every stages a producer publishes its stage counter and waits until the slowest role of its row
//    split is within pace_window stages, so that the roles of a split keep hitting each other's rows in L2;
//  * prefetch: role 0 of every split issues cp.async.bulk.prefetch.tensor.L2 `prefetch` stages ahead, full width.
// With the six equal-cost roles of make_roles the roles stay in lockstep by themselves (cycle counters: 9.961 / 9.965 /
// 9.961 / 9.965 / 9.961 M cycles per role) and the kernel is bound by the tensor core's MN-major operand reads (an
// M128 x N256 x K16 MMA measures 201 cycles against a 128-cycle floor), not by DRAM: pacing on cost +40 %, the prefetch
// instructions slowed role 0's producer (+48 % on that role).  DAPB_GRAM_PACE_EVERY / _WINDOW / DAPB_GRAM_PREFETCH switch
// them on for experiments; profiles/gram_tc_roles_r2.jsonl and profiles/ncu_gram_tc_syrk_r2.md hold the measurements.
constexpr int kPaceEvery = 0;      // defaults (0 = off); DAPB_GRAM_PACE_EVERY / DAPB_GRAM_PACE_WINDOW override
constexpr int kPaceWindow = 32;
__device__ __forceinline__ void pace_store(int* p, int v) {
  asm volatile("st.relaxed.gpu.global.s32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int pace_load(const int* p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(kGThreads, 1)
gram_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ GramTcArgs a, DapbComm c) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t pad = (1024u - (smem_u32(smem_raw) & 1023u)) & 1023u;
  uint8_t* smem = smem_raw + pad;
  const int P = a.P;
  const int role_id = blockIdx.x % a.n_roles, split = blockIdx.x / a.n_roles;
  const GramRole& role = a.roles[role_id];
  const uint32_t box_bytes = kKR * 128;                   // one [KR x 64] box
  const uint32_t strip_bytes = 2 * box_bytes;             // one 128-column strip of a stage
  const uint32_t stage_bytes = 3 * strip_bytes;           // room for the widest role
  const uint32_t load_bytes = (uint32_t)role.n_strips * strip_bytes;
  const uint32_t bar_off = kStages * stage_bytes;
  if (pad + bar_off + 128 > a.smem_bytes) __trap();
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + bar_off);
  uint64_t* empty = full + kStages;
  uint64_t* acc_done = empty + kStages;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(acc_done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t n_tiles = (a.n_rows + kKR - 1) / kKR;
  const int64_t t_begin = (n_tiles * split) / a.splits, t_end = (n_tiles * (split + 1)) / a.splits;
  constexpr uint32_t tmem_cols = 256;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(acc_done, 1);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(tmem_holder, tmem_cols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0) {
    if (lane == 0) {
      prefetch_tensormap(&map_a);
      if (role_id == 0 && a.prefetch > 0)                  // prologue of the L2 prefetch: the first kPrefetch stages
        for (int64_t t = t_begin; t < t_end && t < t_begin + a.prefetch; ++t)
          for (int b = 0; b < P / 64; ++b) tma_prefetch_2d(&map_a, b * 64, (int)(t * kKR));
      int* prog = a.progress + split * kMaxRoles;
      uint32_t it = 0;
      for (int64_t t = t_begin; t < t_end; ++t, ++it) {
        const uint32_t st = it % kStages;
        if (a.pace_every > 0 && a.n_roles > 1 && (it % (uint32_t)a.pace_every) == 0) {
          pace_store(prog + role_id, (int)it);
          for (;;) {
            int slowest = 0x7fffffff;
            for (int r = 0; r < a.n_roles; ++r) slowest = min(slowest, pace_load(prog + r));
            if (slowest + a.pace_window >= (int)it) break;
            __nanosleep(200);
          }
        }
        // optional L2 prefetch by role 0 (off by default, see above)
        if (role_id == 0 && a.prefetch > 0 && t + a.prefetch < t_end) {
          for (int b = 0; b < P / 64; ++b) tma_prefetch_2d(&map_a, b * 64, (int)((t + a.prefetch) * kKR));
        }
        mbar_wait(&empty[st], ((it / kStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&full[st], load_bytes);
        uint8_t* dst = smem + st * stage_bytes;
        for (int p = 0; p < role.n_strips; ++p)
#pragma unroll
          for (int b = 0; b < 2; ++b)
            tma_load_2d(dst + (p * 2 + b) * box_bytes, &map_a, &full[st], role.strips[p] * 128 + b * 64, (int)(t * kKR));
      }
      pace_store(prog + role_id, 0x7fffffff);              // done: never hold anybody back
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc0 = make_idesc_bf16(128, role.n[0], 1, 1);   // both operands MN-major
      const uint32_t idesc1 = make_idesc_bf16(128, role.n_groups > 1 ? role.n[1] : 128, 1, 1);
      const uint64_t tmpl = make_smem_desc_sw128(0, box_bytes, 1024);  // LBO = next 64-column box, SBO = next 8 rows
      uint32_t it = 0;
      long long w_full = 0;
      const long long t_mma0 = clock64();
      for (int64_t t = t_begin; t < t_end; ++t, ++it) {
        const uint32_t st = it % kStages;
        if (a.dbg) { const long long c0 = clock64(); mbar_wait(&full[st], (it / kStages) & 1); w_full += clock64() - c0; }
        else mbar_wait(&full[st], (it / kStages) & 1);
        tc_fence_after();
        const uint32_t base = smem_u32(smem + st * stage_bytes);
#pragma unroll
        for (int ks = 0; ks < kKR / 16; ++ks) {
          for (int g = 0; g < role.n_groups; ++g) {
            const uint32_t a_addr = base + (uint32_t)role.a_pos[g] * strip_bytes + ks * 2048;
            const uint32_t b_addr = base + (uint32_t)role.b_pos[g] * strip_bytes + ks * 2048;
            const uint64_t da = tmpl | (uint64_t)((a_addr & 0x3FFFF) >> 4);
            const uint64_t db = tmpl | (uint64_t)((b_addr & 0x3FFFF) >> 4);
            umma_bf16(tmem_base + role.tmem_col[g], da, db, g == 0 ? idesc0 : idesc1, (it | ks) != 0);
          }
        }
        umma_commit(&empty[st]);
      }
      umma_commit(acc_done);
      if (a.dbg) {
        int* d = a.progress + 512 + 4 * blockIdx.x;
        d[0] = role_id; d[1] = (int)((clock64() - t_mma0) >> 10); d[2] = (int)(w_full >> 10);
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM blocks -> atomics into G (and the mirrored block) =====================
    const int qd = warp & 3;
    const bool has_work = t_end > t_begin;
    if (has_work) {
      mbar_wait(acc_done, 0);
      tc_fence_after();
      for (int k = 0; k < role.n_out; ++k) {
        const int bi = role.out_i[k], bj = role.out_j[k];
        const int gi = bi * 128 + qd * 32 + lane;             // row of G owned by this thread
        float* grow = a.G_acc + (size_t)gi * P + bj * 128;
        float* gcol = a.G_acc + (size_t)(bj * 128) * P + gi;  // mirrored element (j, gi): lanes hit consecutive addresses
        for (int j0 = 0; j0 < 128; j0 += 32) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(qd * 32) << 16) + role.out_col[k] + j0, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(grow + j0 + j), "f"(__uint_as_float(v[j])),
                         "f"(__uint_as_float(v[j + 1])), "f"(__uint_as_float(v[j + 2])), "f"(__uint_as_float(v[j + 3]))
                         : "memory");
          if (bi != bj) {
#pragma unroll
            for (int j = 0; j < 32; ++j) atomicAdd(gcol + (size_t)(j0 + j) * P, __uint_as_float(v[j]));
          }
        }
      }
    }
  }

  __threadfence();
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, tmem_cols);
  if (!a.do_allreduce || c.world == 1) return;

  // ---- grid-wide two-shot all-reduce of the heap copy of G (cooperative launch: every CTA is resident) ----
  cg::grid_group grid = cg::this_grid();
  const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, gth = (int64_t)gridDim.x * blockDim.x;
  __threadfence_system();
  grid.sync();                                               // every row split of this rank is in
  if (blockIdx.x == 0) block_barrier(c, c.rank, a.flag_base, 2 * a.epoch - 1);   // ... and of every other rank
  grid.sync();
  grid_twoshot_allreduce_f32(c, a.payload_off, P * P, gtid, gth);
  __threadfence_system();
  grid.sync();
  if (blockIdx.x == 0) block_barrier(c, c.rank, a.flag_base, 2 * a.epoch);       // all slices broadcast everywhere
  grid.sync();
  const float4* src = reinterpret_cast<const float4*>(a.G_acc);
  float4* dst = reinterpret_cast<float4*>(a.G);
  for (int64_t i = gtid; i < (int64_t)P * P / 4; i += gth) dst[i] = __ldcg(src + i);
}

// Deal the upper-triangular blocks of an S x S block grid to roles of (nearly) equal work.
int make_roles(int S, GramRole* r) {
  memset(r, 0, sizeof(GramRole) * kMaxRoles);
  auto pair256 = [](GramRole& q, int i, int j0) {            // blocks (i, j0), (i, j0 + 1) as ONE N = 256 MMA
    q.n_groups = 1; q.n[0] = 256; q.tmem_col[0] = 0;
    q.n_out = 2; q.out_i[0] = q.out_i[1] = i; q.out_j[0] = j0; q.out_j[1] = j0 + 1; q.out_col[0] = 0; q.out_col[1] = 128;
    if (j0 == i) { q.n_strips = 2; q.strips[0] = i; q.strips[1] = i + 1; q.a_pos[0] = 0; q.b_pos[0] = 0; }
    else { q.n_strips = 3; q.strips[0] = i; q.strips[1] = j0; q.strips[2] = j0 + 1; q.a_pos[0] = 0; q.b_pos[0] = 1; }
  };
  if (S == 1) {
    r[0].n_strips = 1; r[0].strips[0] = 0; r[0].n_groups = 1; r[0].n[0] = 128; r[0].n_out = 1;
    return 1;
  }
  if (S == 2) {                                              // {00,01} and {11}: 2 : 1, small case
    pair256(r[0], 0, 0);
    r[1].n_strips = 1; r[1].strips[0] = 1; r[1].n_groups = 1; r[1].n[0] = 128; r[1].n_out = 1; r[1].out_i[0] = r[1].out_j[0] = 1;
    return 2;
  }
  // P = 512.  Measured (per-role cycle counters, profiles/gram_tc_roles_r2.jsonl): MN-major operand reads make an
  // M128 x N256 MMA cost ~200 cycles and an M128 x N128 one ~170, so a role built from two N = 128 MMAs is 1.6x slower
  // than one N = 256 MMA covering the same two blocks.  The 10 blocks cannot be tiled exactly by (strip, adjacent strip
  // pair) products, so six N = 256 roles are used and two of their twelve blocks (12 and 23, computed twice) are simply
  // not written by the duplicate: 12 block products instead of the 16 of the full matrix, all roles equally fast.
  auto strip_pair = [](GramRole& q, int a_strip, int b0, bool write0, bool write1) {    // A = a_strip, B = strips b0, b0 + 1
    q.n_groups = 1; q.n[0] = 256; q.tmem_col[0] = 0;
    if (a_strip == b0) { q.n_strips = 2; q.strips[0] = b0; q.strips[1] = b0 + 1; q.a_pos[0] = 0; q.b_pos[0] = 0; }
    else if (a_strip == b0 + 1) { q.n_strips = 2; q.strips[0] = b0; q.strips[1] = b0 + 1; q.a_pos[0] = 1; q.b_pos[0] = 0; }
    else if (a_strip < b0) { q.n_strips = 3; q.strips[0] = a_strip; q.strips[1] = b0; q.strips[2] = b0 + 1; q.a_pos[0] = 0; q.b_pos[0] = 1; }
    else { q.n_strips = 3; q.strips[0] = b0; q.strips[1] = b0 + 1; q.strips[2] = a_strip; q.a_pos[0] = 2; q.b_pos[0] = 0; }
    q.n_out = 0;
    for (int t = 0; t < 2; ++t) {
      if (!(t == 0 ? write0 : write1)) continue;
      q.out_i[q.n_out] = a_strip; q.out_j[q.n_out] = b0 + t; q.out_col[q.n_out] = 128 * t; ++q.n_out;
    }
  };
  strip_pair(r[0], 0, 0, true, true);      // 00 01
  strip_pair(r[1], 0, 2, true, true);      // 02 03
  strip_pair(r[2], 1, 1, true, true);      // 11 12
  strip_pair(r[3], 1, 2, false, true);     // (12) 13
  strip_pair(r[4], 2, 2, true, true);      // 22 23
  strip_pair(r[5], 3, 2, false, true);     // (32) 33
  return 6;
}

}  // namespace

extern "C" uint64_t dapb_gram_tc_payload_bytes(int32_t world, int32_t P) {
  (void)world;
  return (uint64_t)P * P * sizeof(float);
}

// A: [n_rows, P] bf16 row-major (P in {128, 256, 512}); G: [P, P] fp32 (overwritten); `tickets`: scratch of
// DAPB_GRAM_TC_SCRATCH_WORDS 32-bit words (cleared here on the stream).
// With do_allreduce the result is summed over all ranks of `comm` (needs dapb_gram_tc_payload_bytes of heap and
// `world` flags from flag_base; TWO barrier epochs per call: 2*epoch-1 and 2*epoch).
extern "C" int dapb_gram_tc(const DapbComm* comm, const void* A, int64_t n_rows, int32_t P, float* G, uint32_t* tickets,
                            int32_t splits, uint64_t payload_off, uint32_t flag_base, uint64_t epoch,
                            int32_t do_allreduce, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "gram_tc: bad comm");
  DAPB_REQUIRE(A && G && tickets, "gram_tc: null argument");
  DAPB_REQUIRE(P == 128 || P == 256 || P == 512, "gram_tc: P must be 128, 256 or 512 (got %d)", P);
  DAPB_REQUIRE(n_rows >= 1 && n_rows < (1ll << 31) - 256, "gram_tc: n_rows out of range for 32-bit TMA coordinates");
  DAPB_REQUIRE((reinterpret_cast<uintptr_t>(G) & 15) == 0, "gram_tc: G must be 16-byte aligned");
  const bool xchg = do_allreduce && comm->world > 1;
  if (xchg) {
    DAPB_REQUIRE(epoch >= 1, "gram_tc: epoch starts at 1");
    DAPB_REQUIRE((payload_off & 15) == 0 && (P * P) % (4 * comm->world) == 0, "gram_tc: payload alignment");
    uint64_t need = DAPB_FLAG_BYTES + payload_off + dapb_gram_tc_payload_bytes(comm->world, P);
    DAPB_REQUIRE(need <= comm->heap_bytes, "gram_tc: heap too small (%llu > %llu)", (unsigned long long)need,
                 (unsigned long long)comm->heap_bytes);
  }
  CUtensorMap map;
  int rc = make_tensor_map_bf16_2d(&map, A, (uint64_t)n_rows, (uint64_t)P, (uint64_t)P, 64, kKR);
  if (rc) return rc;
  GramTcArgs a;
  memset(&a, 0, sizeof(a));
  a.n_roles = make_roles(P / 128, a.roles);
  if (splits <= 0) splits = dapb::sm_count() / a.n_roles;
  const int64_t n_tiles = (n_rows + kKR - 1) / kKR;
  if (splits > n_tiles) splits = (int)n_tiles;
  if (splits < 1) splits = 1;
  if (a.n_roles * splits > dapb::sm_count()) splits = dapb::sm_count() / a.n_roles;   // one resident CTA per SM
  cudaStream_t st = (cudaStream_t)stream;
  a.n_rows = n_rows; a.P = P; a.splits = splits; a.G = G;
  a.G_acc = xchg ? reinterpret_cast<float*>(reinterpret_cast<char*>(comm->heap[comm->rank]) + DAPB_FLAG_BYTES + payload_off) : G;
  a.payload_off = payload_off; a.flag_base = flag_base; a.epoch = epoch; a.do_allreduce = do_allreduce;
  DAPB_CUDA_CHECK(cudaMemsetAsync(a.G_acc, 0, (size_t)P * P * sizeof(float), st));
  a.progress = reinterpret_cast<int*>(tickets);
  {
    static const int pe = getenv("DAPB_GRAM_PACE_EVERY") ? atoi(getenv("DAPB_GRAM_PACE_EVERY")) : kPaceEvery;
    static const int pw = getenv("DAPB_GRAM_PACE_WINDOW") ? atoi(getenv("DAPB_GRAM_PACE_WINDOW")) : kPaceWindow;
    a.pace_every = pe; a.pace_window = pw;
    static const int pf = getenv("DAPB_GRAM_PREFETCH") ? atoi(getenv("DAPB_GRAM_PREFETCH")) : 0;
    a.prefetch = pf;
    static const int dg = getenv("DAPB_GRAM_DBG") ? atoi(getenv("DAPB_GRAM_DBG")) : 0;
    a.dbg = dg;
  }
  DAPB_CUDA_CHECK(cudaMemsetAsync(tickets, 0, (size_t)DAPB_GRAM_TC_SCRATCH_WORDS * sizeof(uint32_t), st));
  const uint32_t smem = kStages * 3 * 2 * kKR * 128 + 128 + 1024;
  a.smem_bytes = smem;
  DAPB_CUDA_CHECK(cudaFuncSetAttribute(gram_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  DapbComm cc = *comm;
  void* kargs[] = {&map, &a, &cc};
  // cooperative in every case: the pacing protocol (and the grid-wide all-reduce) need all CTAs resident
  DAPB_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)gram_tc_kernel, dim3(a.n_roles * splits), dim3(kGThreads), kargs, smem, st));
  return 0;
}
