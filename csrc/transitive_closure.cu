// F6 -- transitive closure: join the frontier with the edge relation and de-duplicate against a
// device hash set of packed 64-bit (x, z) pairs; on several GPUs candidate pairs are pushed into the
// hash-owner's queue in peer memory (all-to-all by hash) and inserted there.
//
// ref: graph_computation/transitive_closure.py:27 (edges reversed for the join), :35 (paths (y,z)
// joined with edges (x,y) -> new path (x,z)), :37 (union + distinct), :38-40 (stop when the count is
// unchanged).  The reference re-joins ALL paths every round; joining only the pairs discovered in
// the previous round (semi-naive evaluation) yields the same set after the same number of rounds
// because every older pair's extensions were already produced.
#include "comm.cuh"

using namespace dapb;

namespace {

constexpr uint64_t kEmpty = 0xFFFFFFFFFFFFFFFFull;

// returns true iff `key` was not present (and is now inserted)
__device__ __forceinline__ bool hs_insert(unsigned long long* table, uint64_t mask, uint64_t key) {
  uint64_t slot = hash64(key) & mask;
  for (uint64_t probe = 0; probe <= mask; ++probe) {
    unsigned long long old = atomicCAS(&table[slot], (unsigned long long)kEmpty, (unsigned long long)key);
    if (old == kEmpty) return true;
    if (old == key) return false;
    slot = (slot + 1) & mask;
  }
  return false;   // table full: caller guarantees load factor <= 0.5, unreachable
}

struct TcArgs {
  const int64_t* in_ptr;       // [V + 1] CSR over the join key y: in-neighbours x of y (edges (x, y))
  const int32_t* in_src;       // [E]
  const unsigned long long* frontier;   // [n_frontier] packed (y << 32 | z)
  int64_t n_frontier;
  unsigned long long* table;   // local hash-set shard (capacity = mask + 1)
  uint64_t mask;
  unsigned long long* next_frontier;    // [next_cap]
  unsigned long long* next_count;       // 1 counter
  // multi-GPU: per-source queues in the owner's heap
  uint64_t q_off;              // payload offset of queues: [world (src rank)][1 + q_cap] u64 (slot 0 = tail)
  uint64_t q_cap;
};

__global__ void __launch_bounds__(256) tc_expand_kernel(TcArgs a, DapbComm c) {
  const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = i0; i < a.n_frontier; i += stride) {
    const uint64_t yz = a.frontier[i];
    const uint32_t y = (uint32_t)(yz >> 32), z = (uint32_t)yz;
    const int64_t e0 = a.in_ptr[y], e1 = a.in_ptr[y + 1];
    for (int64_t e = e0; e < e1; ++e) {
      const uint64_t key = ((uint64_t)(uint32_t)a.in_src[e] << 32) | z;          // new path (x, z)
      if (c.world == 1) {
        if (hs_insert(a.table, a.mask, key)) {
          unsigned long long pos = atomicAdd(a.next_count, 1ull);
          a.next_frontier[pos] = key;
        }
      } else {
        const int owner = (int)hash_owner(key, (uint32_t)c.world);
        // queue written only by this rank: tail lives with the data in the owner's heap
        unsigned long long* q = reinterpret_cast<unsigned long long*>(comm_payload(c, owner, a.q_off)) +
                                (uint64_t)c.rank * (1 + a.q_cap);
        unsigned long long pos = atomicAdd(a.next_count + 1 + owner, 1ull);      // local per-owner tail
        if (pos < a.q_cap) q[1 + pos] = key;
      }
    }
  }
}

// After tc_expand on every rank: publish the per-owner tails (one CTA), then cross-rank barrier.
__global__ void tc_publish_kernel(TcArgs a, DapbComm c, uint32_t flag_base, uint64_t epoch) {
  if (threadIdx.x < c.world) {
    unsigned long long* q = reinterpret_cast<unsigned long long*>(comm_payload(c, threadIdx.x, a.q_off)) +
                            (uint64_t)c.rank * (1 + a.q_cap);
    q[0] = a.next_count[1 + threadIdx.x];
    a.next_count[1 + threadIdx.x] = 0;
  }
  block_barrier(c, c.rank, flag_base, epoch);
}

// Owner side: insert everything the peers queued for me; new pairs form my next frontier.
__global__ void __launch_bounds__(256) tc_drain_kernel(TcArgs a, DapbComm c) {
  for (int src = 0; src < c.world; ++src) {
    const unsigned long long* q = reinterpret_cast<const unsigned long long*>(comm_payload(c, c.rank, a.q_off)) +
                                  (uint64_t)src * (1 + a.q_cap);
    const uint64_t n = __ldcg(q);
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
      const uint64_t key = __ldcg(q + 1 + i);
      if (hs_insert(a.table, a.mask, key)) {
        unsigned long long pos = atomicAdd(a.next_count, 1ull);
        a.next_frontier[pos] = key;
      }
    }
  }
}

__global__ void tc_seed_kernel(const unsigned long long* keys, int64_t n, unsigned long long* table, uint64_t mask,
                               unsigned long long* out, unsigned long long* out_count, int world, int rank) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t key = keys[i];
    if (world > 1 && (int)hash_owner(key, (uint32_t)world) != rank) continue;
    if (hs_insert(table, mask, key)) {
      unsigned long long pos = atomicAdd(out_count, 1ull);
      out[pos] = key;
    }
  }
}

int grid_for(int64_t n) {
  int64_t b = (n + 255) / 256;
  const int64_t cap = (int64_t)dapb::sm_count() * 8;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace

// Insert the initial edge set (each rank keeps the keys it owns) -> first frontier.
extern "C" int dapb_tc_seed(const unsigned long long* keys, int64_t n, unsigned long long* table, uint64_t capacity,
                            unsigned long long* out, unsigned long long* out_count, int32_t world, int32_t rank,
                            void* stream) {
  DAPB_REQUIRE(capacity >= 2 && (capacity & (capacity - 1)) == 0, "tc_seed: capacity must be a power of two");
  tc_seed_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(keys, n, table, capacity - 1, out, out_count, world, rank);
  DAPB_LAUNCH_CHECK();
  return 0;
}

// One join round on this rank's frontier.  world == 1: candidates are inserted directly.
// world > 1: candidates are pushed to their hash owners; call dapb_tc_exchange then dapb_tc_drain.
// counters: [0] = size of next frontier, [1 + r] = scratch tail for owner r (zero on entry).
extern "C" int dapb_tc_expand(const DapbComm* comm, const int64_t* in_ptr, const int32_t* in_src,
                              const unsigned long long* frontier, int64_t n_frontier, unsigned long long* table,
                              uint64_t capacity, unsigned long long* next_frontier, unsigned long long* counters,
                              uint64_t q_off, uint64_t q_cap, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "tc_expand: bad comm");
  DAPB_REQUIRE(capacity >= 2 && (capacity & (capacity - 1)) == 0, "tc_expand: capacity must be a power of two");
  if (comm->world > 1) {
    uint64_t need = DAPB_FLAG_BYTES + q_off + (uint64_t)comm->world * (1 + q_cap) * 8;
    DAPB_REQUIRE(need <= comm->heap_bytes, "tc_expand: heap too small for queues (%llu > %llu)",
                 (unsigned long long)need, (unsigned long long)comm->heap_bytes);
  }
  TcArgs a{in_ptr, in_src, frontier, n_frontier, table, capacity - 1, next_frontier, counters, q_off, q_cap};
  tc_expand_kernel<<<grid_for(n_frontier), 256, 0, (cudaStream_t)stream>>>(a, *comm);
  DAPB_LAUNCH_CHECK();
  return 0;
}

extern "C" int dapb_tc_exchange(const DapbComm* comm, unsigned long long* counters, uint64_t q_off, uint64_t q_cap,
                                uint32_t flag_base, uint64_t epoch, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "tc_exchange: bad comm");
  TcArgs a{};
  a.next_count = counters; a.q_off = q_off; a.q_cap = q_cap;
  tc_publish_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a, *comm, flag_base, epoch);
  DAPB_LAUNCH_CHECK();
  return 0;
}

extern "C" int dapb_tc_drain(const DapbComm* comm, unsigned long long* table, uint64_t capacity,
                             unsigned long long* next_frontier, unsigned long long* counters, uint64_t q_off,
                             uint64_t q_cap, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "tc_drain: bad comm");
  TcArgs a{};
  a.table = table; a.mask = capacity - 1; a.next_frontier = next_frontier; a.next_count = counters;
  a.q_off = q_off; a.q_cap = q_cap;
  tc_drain_kernel<<<dapb::sm_count() * 4, 256, 0, (cudaStream_t)stream>>>(a, *comm);
  DAPB_LAUNCH_CHECK();
  return 0;
}
