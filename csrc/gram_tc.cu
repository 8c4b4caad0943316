// F5 (tensor-core path) -- Gram matrix G = A^T A of a row-sharded tall-skinny bf16 matrix on the 5th-gen
// tensor cores, with the row-split reduction done by vector atomics in L2 and the cross-GPU reduction done
// from inside the same kernel over NVLink peer memory.
//
// ref math: matrix_computation/matrix_decomposition.py:27 (XtX = mat^T mat) -- the reference builds this k x k
// Gram once per row on the driver; BASELINE.json config 5 scales it to a 2^27 x 512 matrix.
//
// G[i, j] = sum_r A[r, i] A[r, j].  The contraction runs over ROWS, so both MMA operands are the same
// shared-memory tile [KR rows x P columns] read in MN-major form (the M / N index is the contiguous one):
//   TMA loads the tile as P/64 boxes of [KR rows x 64 columns] with the 128-byte swizzle; in UMMA terms each box
//   is an MN-major SW128 atom column: 8-row groups 1024 B apart (SBO), 64-column groups one box apart (LBO).
//   A operand = columns [128*strip, +128) (2 boxes), B operand = 256 columns (4 boxes), K = 16 rows per MMA.
// CTA (strip, split) owns the 128 x P strip of G for its share of the rows; the fp32 accumulator (128 lanes x P
// columns) stays in TMEM for the whole kernel, so there is ONE epilogue: TMEM -> red.global.add.v4.f32 into G.
// The last CTA of each strip (ticket) then all-reduces that strip across the GPUs (one-shot push, rank-ordered
// sum) -- no separate collective launch.
#include "comm.cuh"
#include "tc_common.cuh"

using namespace dapb;
using namespace dapb::tc;

namespace {

constexpr int kKR = 64;            // rows per pipeline stage
constexpr int kStages = 3;
constexpr int kGThreads = 256;     // warp 0 TMA, warp 1 MMA, warp 2 TMEM alloc, warps 4-7 epilogue

struct GramTcArgs {
  int64_t n_rows;
  int P;                  // columns (multiple of 128, <= 512)
  int splits;             // row splits (grid = strips * splits)
  float* G;               // [P, P] fp32, zeroed by the host wrapper
  uint32_t* tickets;      // [P / 128] zeroed words
  uint64_t payload_off;   // symmetric-heap region for the per-strip all-reduce
  uint32_t flag_base;
  uint64_t epoch;
  int do_allreduce;
  uint32_t smem_bytes;
};

__global__ void __launch_bounds__(kGThreads, 1)
gram_tc_kernel(const __grid_constant__ CUtensorMap map_a, GramTcArgs a, DapbComm c) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t pad = (1024u - (smem_u32(smem_raw) & 1023u)) & 1023u;
  uint8_t* smem = smem_raw + pad;
  const int P = a.P;
  const uint32_t box_bytes = kKR * 128;                   // one [KR x 64] box
  const uint32_t stage_bytes = (uint32_t)(P / 64) * box_bytes;
  const uint32_t bar_off = kStages * stage_bytes;
  if (pad + bar_off + 128 > a.smem_bytes) __trap();
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + bar_off);
  uint64_t* empty = full + kStages;
  uint64_t* acc_done = empty + kStages;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(acc_done + 1);
  __shared__ bool s_last;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int strips = P / 128;
  const int strip = blockIdx.x % strips, split = blockIdx.x / strips;
  const int64_t n_tiles = (a.n_rows + kKR - 1) / kKR;
  const int64_t t_begin = (n_tiles * split) / a.splits, t_end = (n_tiles * (split + 1)) / a.splits;
  const uint32_t tmem_cols = P <= 128 ? 128 : (P <= 256 ? 256 : 512);

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
      uint32_t it = 0;
      for (int64_t t = t_begin; t < t_end; ++t, ++it) {
        const uint32_t st = it % kStages;
        mbar_wait(&empty[st], ((it / kStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&full[st], stage_bytes);
        uint8_t* dst = smem + st * stage_bytes;
        for (int b = 0; b < P / 64; ++b)
          tma_load_2d(dst + b * box_bytes, &map_a, &full[st], b * 64, (int)(t * kKR));
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const int n_per = P < 256 ? P : 256;                // N of one MMA
      const int n_blk = P / n_per;
      const uint32_t idesc = make_idesc_bf16(128, n_per, 1, 1);        // both operands MN-major
      const uint64_t tmpl = make_smem_desc_sw128(0, box_bytes, 1024);  // LBO = next 64-column box, SBO = next 8 rows
      uint32_t it = 0;
      for (int64_t t = t_begin; t < t_end; ++t, ++it) {
        const uint32_t st = it % kStages;
        mbar_wait(&full[st], (it / kStages) & 1);
        tc_fence_after();
        const uint32_t base = smem_u32(smem + st * stage_bytes);
        const uint32_t a_addr = base + (uint32_t)(strip * 2) * box_bytes;
#pragma unroll
        for (int ks = 0; ks < kKR / 16; ++ks) {
          const uint64_t da = tmpl | (uint64_t)(((a_addr + ks * 2048) & 0x3FFFF) >> 4);
          for (int nb = 0; nb < n_blk; ++nb) {
            const uint32_t b_addr = base + (uint32_t)(nb * (n_per / 64)) * box_bytes + ks * 2048;
            const uint64_t db = tmpl | (uint64_t)((b_addr & 0x3FFFF) >> 4);
            umma_bf16(tmem_base + nb * n_per, da, db, idesc, (it | ks) != 0);
          }
        }
        umma_commit(&empty[st]);
      }
      umma_commit(acc_done);
    }
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM strip -> atomics into G =====================
    const int qd = warp & 3;
    const bool has_work = t_end > t_begin;
    if (has_work) {
      mbar_wait(acc_done, 0);
      tc_fence_after();
      const int gi = strip * 128 + qd * 32 + lane;          // row of G owned by this thread
      float* grow = a.G + (size_t)gi * P;
      for (int j0 = 0; j0 < P; j0 += 32) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(qd * 32) << 16) + j0, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(grow + j0 + j), "f"(__uint_as_float(v[j])),
                       "f"(__uint_as_float(v[j + 1])), "f"(__uint_as_float(v[j + 2])), "f"(__uint_as_float(v[j + 3]))
                       : "memory");
      }
    }
  }

  // ---- per-strip completion: the last CTA of a strip reduces it across the GPUs ---------------------
  __threadfence();
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, tmem_cols);
  if (threadIdx.x == 0) {
    const unsigned t = atomicAdd(a.tickets + strip, 1u);
    s_last = (t == (unsigned)a.splits - 1);
    if (s_last) a.tickets[strip] = 0;
  }
  __syncthreads();
  if (!s_last || !a.do_allreduce || c.world == 1) return;
  __threadfence();
  float* gs = a.G + (size_t)strip * 128 * P;
  const int n = 128 * P;
  const uint64_t off = a.payload_off + (uint64_t)strip * 2 * c.world * n * sizeof(float);
  block_allreduce_oneshot<float>(c, c.rank, off, a.flag_base + strip * DAPB_MAX_RANKS, a.epoch, gs, gs, n);
}

}  // namespace

extern "C" uint64_t dapb_gram_tc_payload_bytes(int32_t world, int32_t P) {
  return (uint64_t)(P / 128) * 2 * world * 128 * P * sizeof(float);
}

// A: [n_rows, P] bf16 row-major (P in {128, 256, 384, 512}); G: [P, P] fp32 (overwritten); tickets: P/128 zeroed words.
// With do_allreduce the result is summed over all ranks of `comm` (needs dapb_gram_tc_payload_bytes of heap and
// (P/128) * DAPB_MAX_RANKS flags from flag_base).
extern "C" int dapb_gram_tc(const DapbComm* comm, const void* A, int64_t n_rows, int32_t P, float* G, uint32_t* tickets,
                            int32_t splits, uint64_t payload_off, uint32_t flag_base, uint64_t epoch,
                            int32_t do_allreduce, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "gram_tc: bad comm");
  DAPB_REQUIRE(A && G && tickets, "gram_tc: null argument");
  DAPB_REQUIRE(P >= 128 && P <= 512 && P % 128 == 0 && (P <= 256 || P == 512), "gram_tc: P must be 128, 256 or 512 (got %d)", P);
  DAPB_REQUIRE(n_rows >= 1 && n_rows < (1ll << 31) - 256, "gram_tc: n_rows out of range for 32-bit TMA coordinates");
  if (do_allreduce && comm->world > 1) {
    DAPB_REQUIRE(epoch >= 1, "gram_tc: epoch starts at 1");
    uint64_t need = DAPB_FLAG_BYTES + payload_off + dapb_gram_tc_payload_bytes(comm->world, P);
    DAPB_REQUIRE(need <= comm->heap_bytes, "gram_tc: heap too small (%llu > %llu)", (unsigned long long)need,
                 (unsigned long long)comm->heap_bytes);
  }
  CUtensorMap map;
  int rc = make_tensor_map_bf16_2d(&map, A, (uint64_t)n_rows, (uint64_t)P, (uint64_t)P, 64, kKR);
  if (rc) return rc;
  const int strips = P / 128;
  if (splits <= 0) splits = dapb::sm_count() / strips;
  const int64_t n_tiles = (n_rows + kKR - 1) / kKR;
  if (splits > n_tiles) splits = (int)n_tiles;
  if (splits < 1) splits = 1;
  cudaStream_t st = (cudaStream_t)stream;
  DAPB_CUDA_CHECK(cudaMemsetAsync(G, 0, (size_t)P * P * sizeof(float), st));
  GramTcArgs a{n_rows, P, splits, G, tickets, payload_off, flag_base, epoch, do_allreduce, 0};
  const uint32_t smem = kStages * (uint32_t)(P / 64) * kKR * 128 + 128 + 1024;
  a.smem_bytes = smem;
  DAPB_CUDA_CHECK(cudaFuncSetAttribute(gram_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  gram_tc_kernel<<<strips * splits, kGThreads, smem, st>>>(map, a, *comm);
  DAPB_LAUNCH_CHECK();
  return 0;
}
