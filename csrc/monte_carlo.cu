// F7 -- Monte-Carlo pi: on-device Philox sampler fused with warp/block/grid reduction.
// ref: randomized_algorithm/monte_carlo.py:17-20 (is_accept: x,y ~ U(-1,1), x^2+y^2 <= 1) and :28
// (map + reduce(add)).  The reference draws with CPython's global Mersenne Twister, unseeded and
// per worker process; here sample i uses Philox4x32-10 counter (i/2) -- each call yields two
// points -- so any sub-range can be recomputed on the host (utils/philox.py).
//
// Exactness: a 32-bit word u is read as the signed integer s = (int32)u, i.e. the coordinate
// s / 2^31 in [-1, 1).  Acceptance x^2 + y^2 <= 1 is evaluated as sx^2 + sy^2 <= 2^62 in 64-bit
// integer arithmetic, so the device count is bit-reproducible on the host.
#include "comm.cuh"
#include "philox.cuh"

using namespace dapb;

namespace {

__device__ __forceinline__ unsigned accept_point(uint32_t ux, uint32_t uy) {
  const long long sx = (int)ux, sy = (int)uy;
  const unsigned long long r2 = (unsigned long long)(sx * sx) + (unsigned long long)(sy * sy);
  return r2 <= (1ull << 62);
}

// Cross-GPU part (F7 "fused as last-block peer store", ref monte_carlo.py:28 reduce(add)): when `x.total_out` is set the
// CTA that finishes last pushes this rank's count into slot [parity][rank] of every rank's heap (one multimem.st with
// NVLS, else W peer stores), publishes the epoch flag, waits for the W flags and adds the W slots -- the job-wide total
// is on every GPU when the kernel ends, no NCCL call.
struct McXchg {
  unsigned long long* total_out;   // null: single-GPU behaviour
  uint32_t* ticket;                // 1 word, zero before first use
  uint64_t payload_off;            // 2 * world u64 slots
  uint32_t flag_base;              // `world` flags
  uint64_t epoch;
};

__global__ void __launch_bounds__(256) mc_pi_kernel(uint64_t first, uint64_t n, uint64_t seed,
                                                     unsigned long long* count_io, McXchg x, DapbComm c) {
  // counters [c0, c1): counter k covers samples 2k and 2k+1
  const uint64_t c_first = first >> 1, c_last = (first + n + 1) >> 1;   // exclusive
  const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  unsigned long long local = 0;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t k = c_first + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; k < c_last; k += stride) {
    Philox4 r = philox4x32_10((uint32_t)k, (uint32_t)(k >> 32), 0u, 0x4D43u, k0, k1);
    const uint64_t s0 = 2 * k, s1 = 2 * k + 1;
    if (s0 >= first && s0 < first + n) local += accept_point(r.x, r.y);
    if (s1 >= first && s1 < first + n) local += accept_point(r.z, r.w);
  }
  local = warp_sum(local);
  __shared__ unsigned long long s_part[8];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) s_part[wid] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += s_part[w];
    if (t) atomicAdd(count_io, t);
  }
  if (x.total_out == nullptr) return;
  __shared__ bool s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned t = atomicAdd(x.ticket, 1u);
    s_last = (t == gridDim.x - 1);
    if (s_last) *x.ticket = 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const uint64_t par = (x.epoch & 1) * (uint64_t)c.world;
  if (threadIdx.x == 0) {
    const unsigned long long mine = atomicAdd(count_io, 0ull);
    push_all<uint64_t>(c, x.payload_off, par + (uint64_t)c.rank, (uint64_t)mine);
  }
  block_barrier(c, c.rank, x.flag_base, x.epoch);
  if (threadIdx.x == 0) {
    const uint64_t* slots = reinterpret_cast<const uint64_t*>(comm_payload(c, c.rank, x.payload_off)) + par;
    unsigned long long t = 0;
    for (int r = 0; r < c.world; ++r) t += __ldcg(slots + r);
    *x.total_out = t;
  }
}

}  // namespace

extern "C" int dapb_mc_pi(uint64_t first, uint64_t n, uint64_t seed, unsigned long long* count_io, void* stream) {
  DAPB_REQUIRE(count_io != nullptr, "mc_pi: null counter");
  if (n == 0) return 0;
  const uint64_t counters = (n + 3) / 2;
  uint64_t blocks = (counters + 255) / 256;
  const uint64_t cap = (uint64_t)dapb::sm_count() * 8;      // 8 resident CTAs of 256 threads per SM
  if (blocks > cap) blocks = cap;
  McXchg x;
  memset(&x, 0, sizeof(x));
  DapbComm c;
  memset(&c, 0, sizeof(c));
  c.world = 1;
  mc_pi_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(first, n, seed, count_io, x, c);
  DAPB_LAUNCH_CHECK();
  return 0;
}

// Same sampler, plus the job-wide total on every rank (see McXchg).  count_io: this rank's running count (zero it to
// get this call's local count), total_out: device u64 that receives sum over ranks of count_io after this call.
extern "C" int dapb_mc_pi_allreduce(const DapbComm* comm, uint64_t first, uint64_t n, uint64_t seed,
                                    unsigned long long* count_io, unsigned long long* total_out, uint32_t* ticket,
                                    uint64_t payload_off, uint32_t flag_base, uint64_t epoch, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "mc_pi_allreduce: bad comm");
  DAPB_REQUIRE(count_io && total_out && ticket && epoch >= 1, "mc_pi_allreduce: null argument / epoch starts at 1");
  DAPB_REQUIRE(DAPB_FLAG_BYTES + payload_off + 2ull * comm->world * 8 <= comm->heap_bytes, "mc_pi_allreduce: heap too small");
  const uint64_t counters = (n + 3) / 2;
  uint64_t blocks = (counters + 255) / 256;
  const uint64_t cap = (uint64_t)dapb::sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  McXchg x;
  x.total_out = total_out; x.ticket = ticket; x.payload_off = payload_off; x.flag_base = flag_base; x.epoch = epoch;
  mc_pi_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(first, n, seed, count_io, x, *comm);
  DAPB_LAUNCH_CHECK();
  return 0;
}
