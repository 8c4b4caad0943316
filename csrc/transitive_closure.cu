// F6 -- transitive closure: join the frontier with the edge relation and de-duplicate against a
// device hash set of packed 64-bit (x, z) pairs; on several GPUs candidate pairs are pushed into the
// hash-owner's queue in peer memory (all-to-all by hash) and inserted there.
//
// ref: graph_computation/transitive_closure.py:27 (edges reversed for the join), :35 (paths (y,z)
// joined with edges (x,y) -> new path (x,z)), :37 (union + distinct), :38-40 (stop when the count is
// unchanged).  The reference re-joins ALL paths every round; joining only the pairs discovered in
// the previous round (semi-naive evaluation) yields the same set after the same number of rounds
// because every older pair's extensions were already produced.
#include <cooperative_groups.h>
#include "comm.cuh"

using namespace dapb;
namespace cg = cooperative_groups;

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

// ---------------------------------------------------------------------------------------------
// Persistent closure kernel: ALL rounds in one cooperative launch.  Per round
//   expand   every CTA joins its part of the frontier with the in-edge CSR and pushes each candidate
//            (x, z) into the hash owner's queue in peer memory (parity-double-buffered),
//   exchange CTA 0 all-gathers one 16-word message per rank {W queue tails, |frontier|, |known|,
//            table capacity, frontier capacity}: this single exchange is the queue-publish barrier,
//            the termination test (ref :38-40, identical on every rank) and the capacity check,
//   drain    every CTA inserts the candidates queued for this rank; new pairs form the next frontier.
// Three grid syncs and one cross-rank exchange per round, no host involvement.  If any queue,
// table or frontier would overflow, every rank sees it in the same message matrix and leaves
// BEFORE draining (nothing was inserted, so the round can be repeated after the host has grown
// the buffers).
// state words (u64): see TcS below.
// ---------------------------------------------------------------------------------------------
enum TcS { S_CNT0 = 0, S_CNT1 = 1, S_CUR = 2, S_KNOWN = 3, S_TOTAL = 4, S_ROUNDS = 5, S_STATUS = 6, S_EPOCH = 7,
           S_NEED_Q = 8, S_NEED_TABLE = 9, S_NEED_FRONT = 10, S_TAILS = 16 /* .. 16 + DAPB_MAX_RANKS */ };
constexpr int kTcMsg = 16;                 // u64 words per rank message

struct TcRunArgs {
  const int64_t* in_ptr;
  const int32_t* in_src;
  unsigned long long* table;
  uint64_t mask;
  unsigned long long* front[2];
  uint64_t front_cap;
  unsigned long long* state;
  uint64_t q_off;              // queues: [parity 2][src rank W][q_cap] u64 in every rank's heap
  uint64_t q_cap;
  uint64_t msg_off;            // messages: [parity 2][rank W][kTcMsg] u64 in every rank's heap
  uint32_t flag_base;          // W flags
  uint64_t max_rounds;
};

// warp-aggregated slot allocation: lanes of the calling (possibly diverged) group that target the same counter
// take consecutive slots with one atomic
__device__ __forceinline__ unsigned long long agg_alloc(unsigned long long* counters, int which) {
  const unsigned act = __activemask();
  const unsigned peers = __match_any_sync(act, which);
  const int lane = threadIdx.x & 31, leader = __ffs(peers) - 1;
  unsigned long long base = 0;
  if (lane == leader) base = atomicAdd(counters + which, (unsigned long long)__popc(peers));
  base = __shfl_sync(peers, base, leader);
  return base + __popc(peers & ((1u << lane) - 1u));
}

__global__ void __launch_bounds__(256) tc_closure_kernel(TcRunArgs a, DapbComm c) {
  cg::grid_group grid = cg::this_grid();
  volatile unsigned long long* st = a.state;
  __shared__ unsigned long long s_mat[DAPB_MAX_RANKS * kTcMsg];
  const int W = c.world, me = c.rank;
  uint64_t cur = st[S_CUR], epoch = st[S_EPOCH], rounds = st[S_ROUNDS];
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;

  for (;;) {
    const uint64_t par = epoch & 1;            // parity of THIS round's exchange = queue / message buffer
    // ---------------- expand
    {
      const uint64_t n_front = st[S_CNT0 + cur];
      const unsigned long long* fr = a.front[cur];
      for (int64_t i = tid; i < (int64_t)n_front; i += nthreads) {
        const uint64_t yz = fr[i];
        const uint32_t y = (uint32_t)(yz >> 32), z = (uint32_t)yz;
        const int64_t e0 = a.in_ptr[y], e1 = a.in_ptr[y + 1];
        for (int64_t e = e0; e < e1; ++e) {
          const uint64_t key = ((uint64_t)(uint32_t)a.in_src[e] << 32) | z;          // new path (x, z)
          const int owner = W == 1 ? 0 : (int)hash_owner(key, (uint32_t)W);
          const unsigned long long pos = agg_alloc(a.state + S_TAILS, owner);
          if (pos < a.q_cap)
            reinterpret_cast<unsigned long long*>(comm_payload(c, owner, a.q_off))[(par * W + me) * a.q_cap + pos] = key;
        }
      }
    }
    grid.sync();
    // ---------------- exchange (CTA 0)
    if (blockIdx.x == 0) {
      const uint64_t ep = epoch + 1;
      if (threadIdx.x < kTcMsg) {
        unsigned long long v = 0;
        const int t = threadIdx.x;
        if (t < W) v = st[S_TAILS + t];
        else if (t == 8) v = st[S_CNT0 + cur];
        else if (t == 9) v = st[S_KNOWN];
        else if (t == 10) v = a.mask + 1;
        else if (t == 11) v = a.front_cap;
        for (int r = 0; r < W; ++r)
          reinterpret_cast<unsigned long long*>(comm_payload(c, r, a.msg_off))[((ep & 1) * W + me) * kTcMsg + t] = v;
      }
      if (W > 1) {
        block_signal_all(c, me, a.flag_base, ep);
        block_wait_all(c, me, a.flag_base, ep);
      } else {
        __syncthreads();
      }
      if (threadIdx.x < W * kTcMsg)
        s_mat[threadIdx.x] = __ldcg(reinterpret_cast<const unsigned long long*>(comm_payload(c, me, a.msg_off)) +
                                    (ep & 1) * W * kTcMsg + threadIdx.x);
      __syncthreads();
      if (threadIdx.x == 0) {
        uint64_t new_global = 0, total = 0, need_q = 0, need_table = 0, need_front = 0;
        bool grow = false;
        for (int r = 0; r < W; ++r) {
          new_global += s_mat[r * kTcMsg + 8];
          total += s_mat[r * kTcMsg + 9];
          uint64_t incoming = 0;
          for (int s = 0; s < W; ++s) {
            const uint64_t t = s_mat[s * kTcMsg + r];
            incoming += t;
            need_q = t > need_q ? t : need_q;
          }
          const uint64_t want_table = 2 * (s_mat[r * kTcMsg + 9] + incoming);
          if (want_table > s_mat[r * kTcMsg + 10] || incoming > s_mat[r * kTcMsg + 11]) grow = true;
          if (r == me) { need_table = want_table; need_front = incoming; }
        }
        if (need_q > a.q_cap) grow = true;
        st[S_TOTAL] = total;
        st[S_NEED_Q] = need_q; st[S_NEED_TABLE] = need_table; st[S_NEED_FRONT] = need_front;
        uint64_t status = DAPB_TC_RUNNING;
        if ((rounds > 0 && new_global == 0) || rounds >= a.max_rounds) status = DAPB_TC_DONE;
        else if (grow) status = DAPB_TC_GROW;
        st[S_STATUS] = status;
        st[S_EPOCH] = ep;
        st[S_CNT0 + (cur ^ 1)] = 0;             // tail of the frontier this round's drain will build
      }
      if (threadIdx.x < DAPB_MAX_RANKS) st[S_TAILS + threadIdx.x] = 0;
      __threadfence();
    }
    grid.sync();
    ++epoch;
    if (st[S_STATUS] != DAPB_TC_RUNNING) return;   // DONE, or GROW (same frontier is expanded again after growth)
    // ---------------- drain
    {
      const unsigned long long* msg = reinterpret_cast<const unsigned long long*>(comm_payload(c, me, a.msg_off)) +
                                      (epoch & 1) * W * kTcMsg;
      unsigned long long* nxt = a.front[cur ^ 1];
      for (int s = 0; s < W; ++s) {
        const uint64_t n = __ldcg(msg + s * kTcMsg + me);
        const unsigned long long* q = reinterpret_cast<const unsigned long long*>(comm_payload(c, me, a.q_off)) +
                                      (par * W + s) * a.q_cap;
        for (int64_t i = tid; i < (int64_t)n; i += nthreads) {
          const uint64_t key = __ldcg(q + i);
          if (hs_insert(a.table, a.mask, key)) nxt[agg_alloc(a.state + S_CNT0, (int)(cur ^ 1))] = key;
        }
      }
    }
    grid.sync();
    cur ^= 1;
    ++rounds;
    if (tid == 0) {
      st[S_KNOWN] = st[S_KNOWN] + st[S_CNT0 + cur];
      st[S_CUR] = cur;
      st[S_ROUNDS] = rounds;
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

// All rounds of the closure in one cooperative launch (see tc_closure_kernel).  `state` holds DAPB_TC_STATE_WORDS
// u64 words: the caller seeds the first frontier into frontier0 with dapb_tc_seed(out_count = state + 0), sets
// state[3] (= known) to the same count and state[7] to the epoch of the message region, and reads back
// state[4] = global pair count, state[5] = rounds, state[6] = DAPB_TC_DONE / DAPB_TC_GROW (then state[8..10] = needed
// queue / table / frontier capacities of this rank) and state[7] = epoch to store back into the region.
extern "C" int dapb_tc_closure(const DapbComm* comm, const int64_t* in_ptr, const int32_t* in_src,
                               unsigned long long* table, uint64_t capacity, unsigned long long* frontier0,
                               unsigned long long* frontier1, uint64_t front_cap, unsigned long long* state,
                               uint64_t q_off, uint64_t q_cap, uint64_t msg_off, uint32_t flag_base,
                               uint64_t max_rounds, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "tc_closure: bad comm");
  DAPB_REQUIRE(capacity >= 2 && (capacity & (capacity - 1)) == 0, "tc_closure: capacity must be a power of two");
  const uint64_t W = (uint64_t)comm->world;
  DAPB_REQUIRE(DAPB_FLAG_BYTES + q_off + 2 * W * q_cap * 8 <= comm->heap_bytes &&
               DAPB_FLAG_BYTES + msg_off + 2 * W * kTcMsg * 8 <= comm->heap_bytes,
               "tc_closure: heap too small for queues / messages");
  DAPB_REQUIRE((flag_base + W) * 8ull <= DAPB_FLAG_BYTES, "tc_closure: flag range");
  static int per_sm = 0;
  if (!per_sm) {
    DAPB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tc_closure_kernel, 256, 0));
    if (per_sm > 4) per_sm = 4;
    DAPB_REQUIRE(per_sm >= 1, "tc_closure: kernel does not fit on an SM");
  }
  TcRunArgs a{in_ptr, in_src, table, capacity - 1, {frontier0, frontier1}, front_cap, state,
              q_off, q_cap, msg_off, flag_base, max_rounds};
  DapbComm c = *comm;
  void* kargs[] = {&a, &c};
  DAPB_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)tc_closure_kernel, dim3(dapb::sm_count() * per_sm), dim3(256),
                                              kargs, 0, (cudaStream_t)stream));
  return 0;
}
