// F3 (tensor-core path) -- k-means nearest-centroid assignment on the 5th-gen tensor cores, fused with
// the per-centroid sum / count accumulation.
//
// ref math: machine_learning/k-means.py:20-28 (argmin_c ||x - c||, ties -> lowest index) and :57-63
// (per-cluster vector sum and count).  argmin_c ||x - c||^2 == argmin_c (||c||^2 - 2 x.c); the cross term
// for a 128-row tile against 128 centroids is one bf16 `tcgen05.mma` chain (M=128, N=128, K=d) whose fp32
// accumulator lives in TMEM.  The centroid operand holds -2c, and the biased norm ||c_j||^2 + C1 enters the accumulator
// through the tensor core as well: a ninth, K = 16 MMA multiplies a constant slab of ones (three 1s per row) with a
// slab holding the norm split into three bf16 terms (hi + mid + lo = the fp32 value), and is issued FIRST with
// accumulate = 0.  So the tensor core itself delivers the biased distance C1 + ||c||^2 - 2 x.c and the epilogue neither
// loads norms nor writes TMEM (an earlier version pre-loaded the accumulators with tcgen05.st from broadcast LDS.128:
// those loads alone took 47 % of the shared-memory pipe next to the tensor core's own operand reads).
// C1 = 1.5 * 2^E is chosen (dapb_kmeans_tc_prepare, Cauchy-Schwarz bound from max ||x||^2 and max ||c||^2) so that
// every biased distance lies in [2^E, 2^(E+1)): one common exponent, i.e. the bit patterns are fixed-point numbers
// and key = bits * 32 + column-in-group orders like (distance, column) in unsigned arithmetic.  The epilogue is
// then one IMAD (FMA pipe) per distance and half a 3-input integer min (ALU pipe).
//
// Persistent, warp-specialised, one CTA per SM (512 threads):
//   warp 0   three producer lanes: lane 0 TMA of 128-row tiles of X (HBM -> smem, 128B swizzle, 4 slots = one super
//            tile), lane 1 TMA of the prepared centroid operand in 16 KB half-chunks (L2 -> smem, ring of 5 = 2.5 chunks:
//            with 2 chunks the ~1.5 us L2 latency was exposed), lane 2 bulk copies of the 4 KB norm slabs (ring of 3)
//   warps 1,3  MMA issuers : for each super tile (4 row tiles) x each centroid chunk: the norm MMA (accumulate = 0) and
//            8 x tcgen05.mma per row tile; warp 1 issues row tiles 0 and 2, warp 3 row tiles 1 and 3
//   warp 2   TMEM allocator (512 columns = 4 accumulator buffers of 128 columns)
//   warps 4-7 and 12-15  arg-min epilogue (row tiles 0-1 / 2-3 of the super tile) : tcgen05.ld (32 columns at a time,
//            next load in flight while the current one is reduced); the buffer goes back to the MMA issuers as soon as
//            its fourth part is in registers
//   warps 8-11 scatter : red.global.add.v4.f32 of every row (re-read from L2) into sums[assignment], counts += 1
// A super tile keeps 4 row tiles resident while all centroid chunks stream through, so the centroid operand
// is re-read from L2 once per 512 rows.  N = 128 per MMA keeps the shared-memory operand traffic
// ((M + N) * 32 B per 64-cycle MMA = 128 B/clk) within the SM's shared-memory bandwidth; N = 64 was measured
// at 59 cycles per MMA instead of 32.  The last super tile is always processed in full (tiles beyond the data
// re-read the last tile, outputs are masked).
//
// What bounds it: every distance must leave TMEM as 4 bytes and TMEM reads run at ~64 B/clk per SM (the Blackwell
// micro-architecture notes this build was given list "LDTM throughput" at that rate): a 128 x 1024 tile is 512 KB = 8192
// cycles of tcgen05.ld against 4608 cycles of MMA (K is only 128), and the kernel measures ~7700 cycles per tile.  The other
// candidate of round 1 -- the shared-memory data pipe, 99 % busy in ncu with 47 % of it broadcast LDS.128 norm loads --
// was eliminated in round 2: with the norm fed through the tensor core the epilogue touches neither shared memory nor TMEM
// writes, and the time per iteration stayed at 34 ms (profiles/kmeans_tc_norm_mma_r2.md).  Three earlier epilogue
// instruction mixes (88, 88 and 62 instructions per 32 columns) had measured 7700-7900 cycles per tile as well.
//
// Near-ties: distances are compared at a resolution of 2^(E-23) <= 4 M * 2^-23 (M = bound on |distance|), far below
// the error already introduced by bf16 inputs; exact tie semantics need the CUDA-core path (kmeans.cu).
#include "tc_common.cuh"
#include "kmeans_common.cuh"

using namespace dapb;
using namespace dapb::tc;

namespace {

constexpr int kTM = 128;          // rows per tile (UMMA M)
constexpr int kNB = 128;          // centroids per chunk (UMMA N)
constexpr int kSuper = 4;         // row tiles per super tile
constexpr int kASlots = 4;
constexpr int kBSlots = 5;        // ring of centroid HALF-chunks (128 centroids x 64 features = 16 KB each)
constexpr int kExtStages = 3;     // ring of norm slabs (128 centroids x K = 16, no swizzle, 4 KB each)
constexpr uint32_t kExtBytes = kNB * 32;
constexpr int kTBufs = 4;         // TMEM accumulator buffers (4 x 128 columns = 512)
constexpr int kThreads = 512;        // 16 warps: 4 service warps, 2 x 4 arg-min warps, 4 scatter warps
constexpr int kParts = kNB / 32;  // 32-column epilogue steps per accumulator

template <int KD>
struct SmemLayout {
  static constexpr int kHalves = KD / 64;
  static constexpr uint32_t kATile = kTM * KD * 2;                 // bytes
  static constexpr uint32_t kBHalf = kNB * 128;                    // one half-chunk: 128 centroids x 64 features
  static constexpr uint32_t kAOff = 0;
  static constexpr uint32_t kBOff = kAOff + kASlots * kATile;
  static constexpr uint32_t kExtOff = kBOff + kBSlots * kBHalf;    // [kExtStages][4 KB] norm slabs (K-major, no swizzle)
  static constexpr uint32_t kOnesOff = kExtOff + kExtStages * kExtBytes;   // [128 rows x K = 16] ones slab, same layout
  static constexpr uint32_t kAsOff = kOnesOff + kTM * 32;          // uint16[kASlots][128]
  static constexpr uint32_t kBarOff = kAsOff + kASlots * kTM * 2;
  static constexpr int kNumBars = 4 * kASlots + 2 * kBSlots + 2 * kTBufs + 2 * kExtStages;
  static constexpr uint32_t kTmemOff = kBarOff + kNumBars * 8;
  static constexpr uint32_t kTotal = kTmemOff + 16;
};

struct KmTcArgs {
  const __nv_bfloat16* X; // [n, KD] bf16 points: tensor-core operand (and source of the sums unless X32 is set)
  const float* X32;       // optional [n, KD] fp32 originals: the per-cluster sums are taken from these
  int64_t n;              // rows
  int k;                  // centroids incl. padding (multiple of 128, <= 16384)
  const float* cnorm;     // [k + 8 + 8 k]: biased norms, 8 aux words, then the K = 16 norm slabs (see kmeans_norm_slab)
  float* sums;            // [k, KD]
  int* counts;            // [k]
  int* assign;            // [n] or null
  long long* dbg;         // optional [64] cycle counters written by CTA 0 (profiling aid), or null
  uint32_t smem_bytes;    // dynamic shared memory actually provided
  uint32_t key_mul;       // 32, passed at run time so that bits * 32 + idx stays ONE IMAD (FMA pipe) instead of LEA/LOP3
};

// cycle accounting for the profiling aid: adds the cycles spent in stmt to acc when dbg is set
#define DAPB_T(acc, stmt) do { long long _t0 = prof ? clock64() : 0; stmt; if (prof) acc += clock64() - _t0; } while (0)

__device__ __forceinline__ void bulk_copy_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

template <int KD>
__global__ void __launch_bounds__(kThreads, 1)
kmeans_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_c, KmTcArgs a) {
  using L = SmemLayout<KD>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t pad = (1024u - (smem_u32(smem_raw) & 1023u)) & 1023u;   // 128B swizzle needs 1024 B alignment
  uint8_t* smem = smem_raw + pad;                                          // (keeps the shared address space)
  if (pad + L::kTotal > a.smem_bytes) {
    if (threadIdx.x == 0) printf("dapb kmeans_tc: shared memory layout does not fit (pad %u)\n", pad);
    __trap();
  }
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
  uint64_t* a_full = bars;
  uint64_t* a_empty = a_full + kASlots;
  uint64_t* as_full = a_empty + kASlots;
  uint64_t* as_empty = as_full + kASlots;
  uint64_t* b_full = as_empty + kASlots;
  uint64_t* b_empty = b_full + kBSlots;
  uint64_t* t_full = b_empty + kBSlots;
  uint64_t* t_empty = t_full + kTBufs;
  uint64_t* cn_full = t_empty + kTBufs;
  uint64_t* cn_empty = cn_full + kExtStages;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(smem + L::kTmemOff);
  uint16_t* assign_s = reinterpret_cast<uint16_t*>(smem + L::kAsOff);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool prof = a.dbg != nullptr && blockIdx.x == 0;
  long long pc0 = 0, pc1 = 0, pc2 = 0, pc3 = 0, pc4 = 0;
  const long long t_start = prof ? clock64() : 0;
  const int64_t n_tiles = (a.n + kTM - 1) / kTM;
  const int64_t n_super = (n_tiles + kSuper - 1) / kSuper;
  const int n_chunks = a.k / kNB;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kASlots; ++i) {
      mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); mbar_init(&as_full[i], 4); mbar_init(&as_empty[i], 4);
    }
    for (int i = 0; i < kBSlots; ++i) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 2); }   // two MMA issuers
    for (int i = 0; i < kTBufs; ++i) { mbar_init(&t_full[i], 1); mbar_init(&t_empty[i], 4); }
    for (int i = 0; i < kExtStages; ++i) { mbar_init(&cn_full[i], 1); mbar_init(&cn_empty[i], 2); }
    fence_barrier_init();
  }
  // ones slab: row r = (1, 1, 1, 0 x 13) bf16 in the canonical K-major no-swizzle form (8-row x 16-byte core matrices;
  // the two K halves of a row group 128 B apart, row groups 256 B apart)
  for (int i = threadIdx.x; i < kTM * 2; i += kThreads) {
    const int r = i >> 1, kc = i & 1;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (kc == 0) { v.x = 0x3F803F80u; v.y = 0x00003F80u; }          // bf16 1.0 = 0x3F80
    *reinterpret_cast<uint4*>(smem + L::kOnesOff + (r >> 3) * 256 + kc * 128 + (r & 7) * 16) = v;
  }
  fence_proxy_async();
  if (warp == 2) { tmem_alloc(tmem_holder, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0 && lane == 0) {
    // ===================== TMA producer: X tiles (warp 0, lane 0) =====================
    {
      prefetch_tensormap(&map_x);
      uint32_t g = 0;                                    // tiles loaded by this CTA so far
      for (int64_t s = blockIdx.x; s < n_super; s += gridDim.x) {
        const int64_t tile0 = s * kSuper;
        constexpr int nt = kSuper;            // the last super tile is processed in full, rows beyond n are masked
        for (int t = 0; t < nt; ++t, ++g) {
          const uint32_t slot = g % kASlots, use = g / kASlots;
          DAPB_T(pc0, mbar_wait(&a_empty[slot], (use & 1) ^ 1));
          mbar_arrive_expect_tx(&a_full[slot], L::kATile);
          uint8_t* dst = smem + L::kAOff + slot * L::kATile;
          const int row0 = (int)(min(tile0 + t, n_tiles - 1) * kTM);   // tiles beyond the data re-read the last one
#pragma unroll
          for (int h = 0; h < L::kHalves; ++h)          // default L2 policy: the scatter warps re-read these rows
            tma_load_2d(dst + h * (kTM * 128), &map_x, &a_full[slot], h * 64, row0);
        }
      }
    }
  } else if (warp == 0 && lane == 1) {
    // ===================== TMA producer: centroid chunks (warp 0, lane 1) =====================
    {
      prefetch_tensormap(&map_c);
      uint32_t j = 0;                                    // half-chunks loaded so far
      for (int64_t s = blockIdx.x; s < n_super; s += gridDim.x) {
        for (int c = 0; c < n_chunks; ++c) {
#pragma unroll
          for (int h = 0; h < L::kHalves; ++h, ++j) {
            const uint32_t st = j % kBSlots, use = j / kBSlots;
            DAPB_T(pc0, mbar_wait(&b_empty[st], (use & 1) ^ 1));
            mbar_arrive_expect_tx(&b_full[st], L::kBHalf);
            tma_load_2d_hint(smem + L::kBOff + st * L::kBHalf, &map_c, &b_full[st], h * 64, c * kNB, kEvictLast);
          }
        }
      }
    }
  } else if (warp == 0 && lane == 2) {
    // ===================== producer: K = 16 norm slabs for the MMA issuers (warp 0, lane 2) =====================
    {
      const char* ext = reinterpret_cast<const char*>(a.cnorm + a.k + 8);
      uint32_t j = 0;
      for (int64_t s = blockIdx.x; s < n_super; s += gridDim.x) {
        for (int c = 0; c < n_chunks; ++c, ++j) {
          const uint32_t st = j % kExtStages, use = j / kExtStages;
          mbar_wait(&cn_empty[st], (use & 1) ^ 1);
          mbar_arrive_expect_tx(&cn_full[st], kExtBytes);
          bulk_copy_g2s(smem + L::kExtOff + st * kExtBytes, ext + (size_t)c * kExtBytes, kExtBytes, &cn_full[st]);
        }
      }
    }
  } else if (warp == 1 || warp == 3) {
    // ===================== MMA issuers: warp 1 takes row tiles 0 and 2 of every super tile, warp 3 tiles 1 and 3 =====================
    // (one issuing thread spends ~600 cycles per accumulator on waits, descriptors and commits -- more than the
    //  512 cycles the tensor core needs for it -- so two threads alternate)
    {
      constexpr uint32_t idesc = make_idesc_bf16(kTM, kNB, 0, 0);
      const uint64_t desc_tmpl = make_smem_desc_sw128(0, 16, 1024);
      const uint32_t a_base = smem_u32(smem + L::kAOff), b_base = smem_u32(smem + L::kBOff);
      // K-major, no swizzle: LBO (K-adjacent core matrices) = 128 B, SBO (row groups) = 256 B
      const uint64_t nosw_tmpl = ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(256 >> 4) << 32) | ((uint64_t)1 << 46);
      const uint64_t d_ones = nosw_tmpl | (uint64_t)((smem_u32(smem + L::kOnesOff) & 0x3FFFF) >> 4);
      const uint32_t ext_base = smem_u32(smem + L::kExtOff);
      const int mma_id = warp == 1 ? 0 : 1;
      uint32_t g0 = 0, j = 0, q0 = 0;                    // tiles / chunks / accumulators before this super tile
      // only lane 0 runs the loop: 32 lanes polling the mbarriers measurably slows the whole SM down (18.1 vs 8.9 ms)
      const int64_t n_super_mma = lane == 0 ? n_super : 0;
      for (int64_t s = blockIdx.x; s < n_super_mma; s += gridDim.x) {
        const int64_t tile0 = s * kSuper;
        constexpr int nt = kSuper;            // the last super tile is processed in full, rows beyond n are masked
        for (int c = 0; c < n_chunks; ++c, ++j) {
          const uint32_t est = j % kExtStages;
          uint32_t bst[L::kHalves];
          DAPB_T(pc0, {
            mbar_wait(&cn_full[est], (j / kExtStages) & 1);
#pragma unroll
            for (int h = 0; h < L::kHalves; ++h) {
              const uint32_t jh = j * L::kHalves + h;
              bst[h] = jh % kBSlots;
              mbar_wait(&b_full[bst[h]], (jh / kBSlots) & 1);
            }
          });
          const uint64_t d_ext = nosw_tmpl | (uint64_t)(((ext_base + est * kExtBytes) & 0x3FFFF) >> 4);
          for (int t = mma_id; t < nt; t += 2) {
            const uint32_t g = g0 + t, slot = g % kASlots;
            const uint32_t q = q0 + (uint32_t)(c * nt + t);
            if (c == 0) DAPB_T(pc1, mbar_wait(&a_full[slot], (g / kASlots) & 1));
            const uint32_t buf = q % kTBufs;
            DAPB_T(pc2, mbar_wait(&t_empty[buf], ((q / kTBufs) & 1) ^ 1));   // drained by the arg-min warps (first use passes)
            tc_fence_after();
            // descriptors differ only in the 14-bit start-address field: build the bases once per step and
            // add compile-time offsets (units of 16 bytes) for the K steps
            const uint64_t da0 = desc_tmpl | (uint64_t)(((a_base + slot * L::kATile) & 0x3FFFF) >> 4);
            const uint32_t d_tmem = tmem_base + buf * kNB;
            DAPB_T(pc3, {
              umma_bf16(d_tmem, d_ones, d_ext, idesc, 0u);           // accumulator := ||c||^2 + C1 in every row
#pragma unroll
              for (int h = 0; h < L::kHalves; ++h) {
                const uint64_t db0 = desc_tmpl | (uint64_t)(((b_base + bst[h] * L::kBHalf) & 0x3FFFF) >> 4);
#pragma unroll
                for (int ks = 0; ks < 4; ++ks)
                  umma_bf16(d_tmem, da0 + (uint64_t)((h * (kTM * 128) + ks * 32) >> 4), db0 + (uint64_t)((ks * 32) >> 4), idesc, 1u);
              }
            });
            DAPB_T(pc4, {
              umma_commit(&t_full[buf]);
              if (c == n_chunks - 1) umma_commit(&a_empty[slot]);   // tensor core is done reading this X tile
            });
          }
#pragma unroll
          for (int h = 0; h < L::kHalves; ++h) umma_commit(&b_empty[bst[h]]);   // this issuer is done with the centroid chunk
          umma_commit(&cn_empty[est]);
        }
        g0 += nt;
        q0 += (uint32_t)(n_chunks * nt);
      }
    }
  } else if ((warp >= 4 && warp < 8) || warp >= 12) {
    // ===================== arg-min epilogue =====================
    // Two groups of four warps: group 0 (warps 4-7) reduces row tiles 0 and 1 of every super tile, group 1
    // (warps 12-15) tiles 2 and 3, so every scheduler has two arg-min warps to interleave.  Within a warp the
    // work is software-pipelined at 32-column granularity: the next TMEM load is in flight while the current 32
    // values per row are reduced.
    const int qd = warp & 3;                              // TMEM lane quarter this warp may access
    const int grp = warp >= 12 ? 1 : 0;
    const int t_lo = grp * 2;                             // first row tile of this group
    const int row = qd * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(qd * 32) << 16);
    const uint32_t kmul = a.key_mul;
    uint32_t g0 = 0, q0 = 0;                              // tiles / accumulators before this super tile
    uint32_t vA[32], vB[32];

    auto issue_load = [&](uint32_t qq, int part, uint32_t* v) {
      const uint32_t buf = qq % kTBufs;
      if (part == 0) {
        DAPB_T(pc0, mbar_wait(&t_full[buf], (qq / kTBufs) & 1));
        tc_fence_after();
      }
      tmem_ld_32x32b_x32(lane_addr + buf * kNB + part * 32, v);
    };
    // What comes back from TMEM already is C1 + ||c||^2 - 2 x.c (the norm MMA opened the accumulation), so the epilogue
    // is pure register work: no norm add, no shared-memory operand, no TMEM write.
    auto release_buf = [&](uint32_t buf) {                // drained into registers: hand it back to the MMA issuers
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&t_empty[buf]);
    };
    // The prepared norms carry a bias C1 = 1.5 * 2^E that puts every biased distance C1 + ||c||^2 - 2 x.c into the
    // binade [2^E, 2^(E+1)) (dapb_kmeans_tc_prepare chooses E from the Cauchy-Schwarz bound), so the bit patterns
    // are fixed-point numbers with a common exponent and key = bits * 32 + column is ordered like (distance, column)
    // in unsigned arithmetic (the 5 bits shifted out are constant).  One IMAD per element on the FMA pipe and half
    // a 3-input integer min on the ALU pipe; which 32-column group won is tracked per row.
#define DAPB_PACK(dst, x, J) asm("mad.lo.u32 %0, %1, %2, " #J ";" : "=r"(dst) : "r"(x), "r"(kmul))
    auto reduce = [&](const uint32_t* v, int group, uint32_t m, int& m_group) -> uint32_t {
      uint32_t hm = 0xffffffffu;
#define DAPB_RED4(JJ, J0, J1, J2, J3) {                                          \
        uint32_t p0, p1, p2, p3;                                                 \
        DAPB_PACK(p0, v[JJ * 4 + 0], J0);                                        \
        DAPB_PACK(p1, v[JJ * 4 + 1], J1);                                        \
        DAPB_PACK(p2, v[JJ * 4 + 2], J2);                                        \
        DAPB_PACK(p3, v[JJ * 4 + 3], J3);                                        \
        hm = min(min(hm, p0), min(p1, min(p2, p3))); }
      DAPB_RED4(0, 0, 1, 2, 3) DAPB_RED4(1, 4, 5, 6, 7) DAPB_RED4(2, 8, 9, 10, 11) DAPB_RED4(3, 12, 13, 14, 15)
      DAPB_RED4(4, 16, 17, 18, 19) DAPB_RED4(5, 20, 21, 22, 23) DAPB_RED4(6, 24, 25, 26, 27) DAPB_RED4(7, 28, 29, 30, 31)
#undef DAPB_RED4
      // across groups only the DISTANCE part of the key is compared (key >> 5): on an exact tie the earlier group, i.e.
      // the lower global centroid index, keeps the row -- the reference's strict '<' scan (k-means.py:25)
      if ((hm >> 5) < (m >> 5)) { m = hm; m_group = group; }
      return m;
    };

    // number of super tiles of this CTA (all are processed as FULL super tiles: rows beyond n are masked on output)
    for (int64_t s = blockIdx.x; s < n_super; s += gridDim.x) {
      const int64_t tile0 = s * kSuper;
      uint32_t best0 = 0xffffffffu, best1 = 0xffffffffu;
      int bgrp0 = 0, bgrp1 = 0;
      // accumulator index of (chunk c, row tile t) in this super tile: q0 + c * kSuper + t, TMEM buffer = t
      issue_load(q0 + t_lo, 0, vA);
      for (int c = 0; c < n_chunks; ++c) {
        auto drain = [&](const int tt, uint32_t& bst, int& bg) {
          const uint32_t qq = q0 + c * kSuper + t_lo + tt, buf = qq % kTBufs;
          const uint32_t qn = tt == 0 ? qq + 1 : qq + kSuper - 1;
          const bool has_next = tt == 0 || c + 1 < n_chunks;
          DAPB_T(pc1, tmem_ld_wait()); issue_load(qq, 1, vB);
          bst = reduce(vA, c * kParts + 0, bst, bg);
          DAPB_T(pc1, tmem_ld_wait()); issue_load(qq, 2, vA);
          bst = reduce(vB, c * kParts + 1, bst, bg);
          DAPB_T(pc1, tmem_ld_wait()); issue_load(qq, 3, vB);
          bst = reduce(vA, c * kParts + 2, bst, bg);
          DAPB_T(pc1, tmem_ld_wait()); release_buf(buf);       // all 128 columns of this accumulator are in registers
          if (has_next) issue_load(qn, 0, vA);
          bst = reduce(vB, c * kParts + 3, bst, bg);
        };
        drain(0, best0, bgrp0);
        drain(1, best1, bgrp1);
      }
#pragma unroll
      for (int tt = 0; tt < 2; ++tt) {
        const int t = t_lo + tt;
        const uint32_t g = g0 + t, slot = g % kASlots;
        const int cidx = (tt == 0 ? bgrp0 : bgrp1) * 32 + (int)((tt == 0 ? best0 : best1) & 31u);
        mbar_wait(&as_empty[slot], ((g / kASlots) & 1) ^ 1);     // scatter warps are done with the previous use
        assign_s[slot * kTM + row] = (uint16_t)cidx;
        const int64_t grow = (tile0 + t) * kTM + row;
        if (a.assign && grow < a.n) a.assign[grow] = cidx;
        __syncwarp();
        if (lane == 0) mbar_arrive(&as_full[slot]);
      }
      g0 += kSuper;
      q0 += (uint32_t)(n_chunks * kSuper);
    }
#undef DAPB_PACK
  } else if (warp >= 8 && warp < 12) {
    // ===================== scatter: sums[assign[row]] += x[row], counts += 1 =====================
    const int qd = warp & 3;
    uint32_t g = 0;
    for (int64_t s = blockIdx.x; s < n_super; s += gridDim.x) {
      const int64_t tile0 = s * kSuper;
      constexpr int nt = kSuper;
      for (int t = 0; t < nt; ++t, ++g) {
        const uint32_t slot = g % kASlots, par = (g / kASlots) & 1;
        DAPB_T(pc0, mbar_wait(&as_full[slot], par));
        const int64_t row_base = (tile0 + t) * kTM + qd * 32;
        const int my_c = (int)assign_s[slot * kTM + qd * 32 + lane];
        __syncwarp();
        if (lane == 0) mbar_arrive(&as_empty[slot]);        // assignments are in registers
        if (row_base + lane < a.n) atomicAdd(a.counts + my_c, 1);
        const int n_rows = (int)max((int64_t)0, min((int64_t)32, a.n - row_base));
        constexpr int kP = (KD + 127) / 128;                // 128-feature passes per row (1 for d <= 128)
        const bool lane_on = lane * 4 < KD || kP > 1;
        // batches of 8 rows: all loads of a batch are issued before the first atomic, so one L2 round trip is
        // paid per batch instead of per row
        if (a.X32 != nullptr) {
          // fp32 points: the bf16 shadow copy decided the assignment, the sums are accumulated from the ORIGINAL rows
          // (ref k-means.py:62-63 sums the points themselves), 16 bytes per lane per row
          for (int i0 = 0; i0 < n_rows; i0 += 8) {
            float4 raw4[8][kP];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              const float* xr = a.X32 + (row_base + min(i0 + u, n_rows - 1)) * (int64_t)KD;
#pragma unroll
              for (int p = 0; p < kP; ++p) {
                const int col = p * 128 + lane * 4;
                raw4[u][p] = (col < KD) ? __ldg(reinterpret_cast<const float4*>(xr + col)) : make_float4(0.f, 0.f, 0.f, 0.f);
              }
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              const int cidx = __shfl_sync(0xffffffffu, my_c, min(i0 + u, 31));
              if (i0 + u < n_rows && lane_on) {
#pragma unroll
                for (int p = 0; p < kP; ++p) {
                  const int col = p * 128 + lane * 4;
                  if (col < KD) {
                    float* dst = a.sums + (size_t)cidx * KD + col;
                    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "f"(raw4[u][p].x), "f"(raw4[u][p].y),
                                 "f"(raw4[u][p].z), "f"(raw4[u][p].w) : "memory");
                  }
                }
              }
            }
          }
          continue;
        }
        for (int i0 = 0; i0 < n_rows; i0 += 8) {
          uint2 raw[8][kP];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const __nv_bfloat16* xr = a.X + (row_base + min(i0 + u, n_rows - 1)) * (int64_t)KD;
#pragma unroll
            for (int p = 0; p < kP; ++p) {
              const int col = p * 128 + lane * 4;           // 4 consecutive features per lane
              raw[u][p] = (col < KD) ? __ldg(reinterpret_cast<const uint2*>(xr + col)) : make_uint2(0u, 0u);
            }
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int cidx = __shfl_sync(0xffffffffu, my_c, min(i0 + u, 31));
            if (i0 + u < n_rows && lane_on) {
#pragma unroll
              for (int p = 0; p < kP; ++p) {
                const int col = p * 128 + lane * 4;
                if (col < KD) {
                  float* dst = a.sums + (size_t)cidx * KD + col;
                  asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "f"(bf16_lo(raw[u][p].x)),
                               "f"(bf16_hi(raw[u][p].x)), "f"(bf16_lo(raw[u][p].y)), "f"(bf16_hi(raw[u][p].y)) : "memory");
                }
              }
            }
          }
        }
      }
    }
  }

  if (prof && lane == 0 && (warp <= 4 || warp == 8)) {   // one representative warp per role
    long long* d = a.dbg + (warp == 8 ? 5 : warp) * 8;
    d[0] = pc0; d[1] = pc1; d[2] = pc2; d[3] = pc3; d[5] = pc4; d[4] = clock64() - t_start;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

// centres fp32 [k, d] -> tensor-core operand bf16(-2 c) [k, d] and ||bf16(c)||^2 (the norm of the ROUNDED centre,
// so that the assignment is the exact nearest neighbour w.r.t. one consistent set of centres); cnorm[k] receives
// max_j ||c_j||^2 (as float bits; norms are >= 0 so the integer order is the float order)
__global__ void kmeans_tc_prep_kernel(const float* __restrict__ C, int k, int k_pad, int d, __nv_bfloat16* __restrict__ Bp,
                                      float* __restrict__ cnorm) {
  const int lane = threadIdx.x & 31;
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (c >= k_pad) return;
  if (c >= k) {                                   // padding centroid: zero operand row, norm set by the bias kernel
    for (int j = lane; j < d; j += 32) Bp[(size_t)c * d + j] = __float2bfloat16(0.f);
    return;
  }
  float s = 0.f;
  for (int j = lane; j < d; j += 32) {
    const __nv_bfloat16 cb = __float2bfloat16(C[(size_t)c * d + j]);
    const float cf = __bfloat162float(cb);
    s += cf * cf;
    Bp[(size_t)c * d + j] = __float2bfloat16(-2.0f * cf);       // exact: power-of-two scaling
  }
  s = warp_sum(s);
  if (lane == 0) {
    cnorm[c] = s;
    atomicMax(reinterpret_cast<int*>(cnorm + k_pad), __float_as_int(s));
  }
}

// cnorm[j] += C1 with C1 = 1.5 * 2^E, 2^(E-1) > M >= |cn_j - 2 x.c_j| for every row x and centre j:
// M = cn_max + 2 sqrt(xn_max * cn_max) (Cauchy-Schwarz) with 1 % slack for the fp32 accumulation.
// Then C1 + (cn_j - 2 x.c_j) lies in [2^E, 2^(E+1)): one common exponent, resolution 2^(E-23) <= M * 2^-20.
// Padding centroids (k .. k_pad) get C1 + 0.9999 * 2^(E-1): inside the binade and above every real distance
// (those stay below M / 1.01), so they never win.
__global__ void kmeans_tc_bias_kernel(float* __restrict__ cnorm, int k, int k_pad, float xnorm_max) {
  const float cmax = cnorm[k_pad];
  const float M = 1.01f * (cmax + 2.0f * sqrtf(xnorm_max * cmax)) + 1e-30f;
  int e;
  frexpf(M, &e);                                  // M = f * 2^e, f in [0.5, 1)  =>  M < 2^e
  const float C1 = ldexpf(1.5f, e + 1);           // E = e + 1: 2^(E-1) = 2^e > M
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < k_pad) {
    const float v = j < k ? cnorm[j] + C1 : C1 + ldexpf(0.9999f, e);
    cnorm[j] = v;
    kmeans_norm_slab_store(cnorm, k_pad, j, v);       // what the assignment kernel's norm MMA reads
  }
  if (j == 0) cnorm[k_pad + 1] = C1;
}

template <int KD>
int launch_km(const CUtensorMap& mx, const CUtensorMap& mc, KmTcArgs a, int grid, cudaStream_t st) {
  constexpr uint32_t smem = (SmemLayout<KD>::kTotal + 1024 <= 227 * 1024) ? SmemLayout<KD>::kTotal + 1024 : 227 * 1024;
  static_assert(SmemLayout<KD>::kTotal <= 227 * 1024, "k-means tensor-core kernel exceeds shared memory");
  a.smem_bytes = smem;
  DAPB_CUDA_CHECK(cudaFuncSetAttribute(kmeans_tc_kernel<KD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kmeans_tc_kernel<KD><<<grid, kThreads, smem, st>>>(mx, mc, a);
  DAPB_LAUNCH_CHECK();
  return 0;
}

}  // namespace

// k_pad = k rounded up to a multiple of 128 (the MMA's N); b_operand: [k_pad, d] bf16; cnorm: 9 * k_pad + 8 floats
// (k_pad biased norms, the running max and the bias, then the K = 16 norm slabs, see kmeans_common.cuh); xnorm_max >= max_i ||x_i||^2 over every row the
// assignment kernel will see (a looser bound only costs resolution).
extern "C" int dapb_kmeans_tc_prepare(const float* centers, int32_t k, int32_t d, void* b_operand, float* cnorm,
                                      float xnorm_max, void* stream) {
  DAPB_REQUIRE(centers && b_operand && cnorm && k >= 1 && d >= 1, "kmeans_tc_prepare: bad arguments");
  DAPB_REQUIRE(xnorm_max >= 0.f && xnorm_max < 1e30f, "kmeans_tc_prepare: xnorm_max must be a finite bound on ||x||^2");
  const int k_pad = (k + kNB - 1) / kNB * kNB;
  DAPB_CUDA_CHECK(cudaMemsetAsync(cnorm + k_pad, 0, 8 * sizeof(float), (cudaStream_t)stream));
  const int blocks = (k_pad * 32 + 255) / 256;
  kmeans_tc_prep_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(centers, k, k_pad, d, (__nv_bfloat16*)b_operand, cnorm);
  DAPB_LAUNCH_CHECK();
  kmeans_tc_bias_kernel<<<(k_pad + 255) / 256, 256, 0, (cudaStream_t)stream>>>(cnorm, k, k_pad, xnorm_max);
  DAPB_LAUNCH_CHECK();
  return 0;
}

// X: [n, d] bf16 row-major (16-byte aligned rows); b_operand / cnorm from dapb_kmeans_tc_prepare.
// sums [k, d] fp32 and counts [k] are accumulated atomically (zero them first); assign may be NULL.
extern "C" int dapb_kmeans_assign_tc(const void* X, int64_t n, int32_t d, const void* b_operand, const float* cnorm,
                                     int32_t k, int32_t* assign, float* sums, int32_t* counts, int32_t grid,
                                     void* stream) {
  return dapb_kmeans_assign_tc_dbg(X, n, d, b_operand, cnorm, k, assign, sums, counts, grid, nullptr, stream);
}

// Same, with an optional device buffer of 64 int64 cycle counters filled by CTA 0 (per-role wait times):
// [role*8 + i], roles 0 X-producer, 1 MMA issuer, 3 centroid-producer, 4 arg-min warp, 5 scatter warp;
// i = 0..2 wait counters, 3 MMA issue, 5 commits, 4 = total cycles.  Used by scripts/kmeans_tc_probe.py.
static int kmeans_assign_tc_impl(const void* X, const float* X32, int64_t n, int32_t d, const void* b_operand,
                                 const float* cnorm, int32_t k, int32_t* assign, float* sums,
                                 int32_t* counts, int32_t grid, long long* dbg, void* stream);
extern "C" int dapb_kmeans_assign_tc_dbg(const void* X, int64_t n, int32_t d, const void* b_operand,
                                         const float* cnorm, int32_t k, int32_t* assign, float* sums,
                                         int32_t* counts, int32_t grid, long long* dbg, void* stream) {
  return kmeans_assign_tc_impl(X, nullptr, n, d, b_operand, cnorm, k, assign, sums, counts, grid, dbg, stream);
}
// fp32 points on the tensor-core path: X_bf16 is a bf16 copy of X32 (same shape) that feeds the distance MMAs, X32 the
// original rows from which the per-cluster sums are accumulated.  Assignments are nearest centroids of the ROUNDED
// points (they differ from the fp32 arg-min only on near-ties below bf16 resolution).
extern "C" int dapb_kmeans_assign_tc_f32(const void* X_bf16, const float* X32, int64_t n, int32_t d, const void* b_operand,
                                         const float* cnorm, int32_t k, int32_t* assign, float* sums,
                                         int32_t* counts, int32_t grid, void* stream) {
  DAPB_REQUIRE(X32 != nullptr && (reinterpret_cast<uintptr_t>(X32) & 15) == 0, "kmeans_assign_tc_f32: X32 must be 16-byte aligned");
  return kmeans_assign_tc_impl(X_bf16, X32, n, d, b_operand, cnorm, k, assign, sums, counts, grid, nullptr, stream);
}
static int kmeans_assign_tc_impl(const void* X, const float* X32, int64_t n, int32_t d, const void* b_operand,
                                 const float* cnorm, int32_t k, int32_t* assign, float* sums,
                                 int32_t* counts, int32_t grid, long long* dbg, void* stream) {
  DAPB_REQUIRE(X && b_operand && cnorm && sums && counts, "kmeans_assign_tc: null argument");
  DAPB_REQUIRE(d == 64 || d == 128, "kmeans_assign_tc: d must be 64 or 128 (got %d)", d);
  DAPB_REQUIRE(k >= 1 && k <= 16384, "kmeans_assign_tc: k must be in [1, 16384] (got %d)", k);
  k = (k + kNB - 1) / kNB * kNB;                  // padding centroids of dapb_kmeans_tc_prepare never win
  DAPB_REQUIRE(n >= 1 && n < (1ll << 31) - 256, "kmeans_assign_tc: n out of range for 32-bit TMA coordinates");
  CUtensorMap mx, mc;
  int rc = make_tensor_map_bf16_2d(&mx, X, (uint64_t)n, (uint64_t)d, (uint64_t)d, 64, kTM);
  if (rc) return rc;
  rc = make_tensor_map_bf16_2d(&mc, b_operand, (uint64_t)k, (uint64_t)d, (uint64_t)d, 64, kNB);
  if (rc) return rc;
  KmTcArgs a{(const __nv_bfloat16*)X, X32, n, k, cnorm, sums, counts, assign, dbg, 0, 32u};
  if (grid <= 0) grid = dapb::sm_count();
  const int64_t n_super = ((n + kTM - 1) / kTM + kSuper - 1) / kSuper;
  if (grid > n_super) grid = (int)n_super;
  cudaStream_t st = (cudaStream_t)stream;
  if (d == 64) return launch_km<64>(mx, mc, a, grid, st);
  return launch_km<128>(mx, mc, a, grid, st);
}
