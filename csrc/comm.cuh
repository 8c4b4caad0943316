// F8 -- device-side communication primitives over NVLink 5 / NVSwitch peer memory.
//
// These replace the reference's Spark data movement: broadcast +
// treeAggregate (optimization/ssgd.py:95-103) -> one-shot all-reduce; reduceByKey + collect
// (machine_learning/k-means.py:62-67) -> reduce-scatter + all-gather; reduceByKey by vertex
// (graph_computation/pagerank.py:57) -> all-gather push; join/distinct shuffles
// (graph_computation/transitive_closure.py:35-37) -> all-to-all push queues.
//
// They are called from INSIDE the compute kernels (one CTA acts for its GPU), never as separate
// launches.  Model: every rank owns a "symmetric heap" (same layout everywhere) that all peers
// have mapped (CUDA IPC); DapbComm carries the mapped base pointers.  All transfers are PUSHES
// (st.global to a peer's heap) followed by a release-store of an epoch number into the
// receiver's flag word; receivers spin with ld.acquire.sys on their LOCAL flag words and read
// payloads with L1-bypassing loads.  Flags only ever grow (epoch numbers), so they never need
// resetting, and payload slots are double-buffered on epoch parity.
#pragma once
#include "common.cuh"

namespace dapb {

__device__ __forceinline__ void st_release_sys(uint64_t* p, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t global_timer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// ---- NVLS multicast (multimem.*) wrappers -----------------------------------------------------------------------
// `p` must be an address inside the multicast mapping DapbComm::mc_heap.  A multimem store is replicated by the
// NVSwitch into every rank's copy of the heap (one NVLink egress instead of W - 1); ld_reduce pulls the W copies and
// adds them inside the switch; red adds the operand into every copy.
__device__ __forceinline__ void multimem_st(float* p, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" :: "l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void multimem_st(double* p, double v) {
  asm volatile("multimem.st.relaxed.sys.global.f64 [%0], %1;" :: "l"(p), "d"(v) : "memory");
}
__device__ __forceinline__ void multimem_st(uint64_t* p, uint64_t v) {
  asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void multimem_st(float4* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float4* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float multimem_ld_reduce_add(const float* p) {
  float v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double multimem_ld_reduce_add(const double* p) {
  double v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_red_add(float* p, float v) {
  asm volatile("multimem.red.relaxed.sys.global.add.f32 [%0], %1;" :: "l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void multimem_red_add(unsigned long long* p, unsigned long long v) {
  asm volatile("multimem.red.relaxed.sys.global.add.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

// Heap layout: [0, DAPB_FLAG_BYTES) = uint64 flag words, rest = payload.
__device__ __forceinline__ uint64_t* comm_flags(const DapbComm& c, int r) {
  return reinterpret_cast<uint64_t*>(c.heap[r]);
}
__device__ __forceinline__ char* comm_payload(const DapbComm& c, int r, uint64_t off) {
  return reinterpret_cast<char*>(c.heap[r]) + DAPB_FLAG_BYTES + off;
}
// multicast alias of the payload (only valid when c.mc_heap != nullptr)
__device__ __forceinline__ char* comm_payload_mc(const DapbComm& c, uint64_t off) {
  return reinterpret_cast<char*>(c.mc_heap) + DAPB_FLAG_BYTES + off;
}
__device__ __forceinline__ uint64_t* comm_flags_mc(const DapbComm& c) {
  return reinterpret_cast<uint64_t*>(c.mc_heap);
}

// Spin until the local flag reaches `epoch`.  A dead peer must not hang the GPU forever
// (SURVEY.md section 5.3): after DAPB_WAIT_TIMEOUT_NS the kernel traps.
__device__ __forceinline__ void comm_wait_flag(const uint64_t* flag, uint64_t epoch) {
  uint64_t t0 = 0;
  unsigned spins = 0;
  while (ld_acquire_sys(flag) < epoch) {
    if ((++spins & 0x3ff) == 0) {
      uint64_t now = global_timer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > DAPB_WAIT_TIMEOUT_NS) {
        printf("dapb: flag wait timed out (flag=%p want=%llu have=%llu)\n", (const void*)flag,
               (unsigned long long)epoch, (unsigned long long)ld_acquire_sys(flag));
        __trap();
      }
      __nanosleep(64);
    }
  }
}

// CTA-collective: publish "my data for `epoch` is in place" to every rank (incl. self).
// All threads' prior peer stores are ordered before the flag via bar.sync + fence.sys.
__device__ __forceinline__ void block_signal_all(const DapbComm& c, int my_rank, uint32_t flag_base,
                                                 uint64_t epoch) {
  __syncthreads();
  if (threadIdx.x == 0) __threadfence_system();
  __syncthreads();
  if (threadIdx.x < (unsigned)c.world)
    st_release_sys(comm_flags(c, threadIdx.x) + flag_base + my_rank, epoch);
}
// Signal one peer only.
__device__ __forceinline__ void block_signal_one(const DapbComm& c, int my_rank, int peer,
                                                 uint32_t flag_base, uint64_t epoch) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    st_release_sys(comm_flags(c, peer) + flag_base + my_rank, epoch);
  }
}
// CTA-collective: wait until every rank has signalled `epoch` on my flags.
__device__ __forceinline__ void block_wait_all(const DapbComm& c, int my_rank, uint32_t flag_base,
                                               uint64_t epoch) {
  if (threadIdx.x < (unsigned)c.world)
    comm_wait_flag(comm_flags(c, my_rank) + flag_base + threadIdx.x, epoch);
  __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// One-shot all-reduce(sum) of n values of T executed by ONE CTA per rank.
//   * every rank pushes its vector into slot[rank] of every peer (world * n * sizeof(T) bytes out),
//   * then sums the `world` slots in RANK ORDER, so all ranks hold bit-identical results
//     (Spark's treeAggregate folds partials in partition order).
// Payload needed at `off`: 2 (parity) * world * n_pad elements, n_pad = n rounded up to 4.
// Flags used: [flag_base, flag_base + world).
// ---------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ uint64_t oneshot_bytes(int world, int n) {
  return (uint64_t)2 * world * ((n + 3) & ~3) * sizeof(T);
}

// `slot_stride` (elements, multiple of 4, >= n): distance between the ranks' slots; 0 = n rounded up to 4.  Callers that
// send messages of different lengths through ONE region pass a fixed stride so that the two parity areas never overlap.
template <typename T>
__device__ void block_allreduce_oneshot(const DapbComm& c, int my_rank, uint64_t off,
                                        uint32_t flag_base, uint64_t epoch, const T* src,
                                        T* dst, int n, int slot_stride = 0) {   // src may alias dst
  const int n_pad = slot_stride > 0 ? slot_stride : ((n + 3) & ~3);
  if (c.world == 1) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
    __syncthreads();
    return;
  }
  const uint64_t par = (epoch & 1) * (uint64_t)c.world * n_pad;
  for (int r = 0; r < c.world; ++r) {
    T* slot = reinterpret_cast<T*>(comm_payload(c, r, off)) + par + (uint64_t)my_rank * n_pad;
    for (int i = threadIdx.x; i < n; i += blockDim.x) slot[i] = src[i];
  }
  block_signal_all(c, my_rank, flag_base, epoch);
  block_wait_all(c, my_rank, flag_base, epoch);
  const T* mine = reinterpret_cast<const T*>(comm_payload(c, my_rank, off)) + par;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    T acc = __ldcg(mine + i);
    for (int r = 1; r < c.world; ++r) acc += __ldcg(mine + (uint64_t)r * n_pad + i);
    dst[i] = acc;
  }
  __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// Barrier across ranks (one CTA per rank).  Flags: [flag_base, flag_base + world).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void block_barrier(const DapbComm& c, int my_rank, uint32_t flag_base,
                                              uint64_t epoch) {
  if (c.world == 1) { __syncthreads(); return; }
  block_signal_all(c, my_rank, flag_base, epoch);
  block_wait_all(c, my_rank, flag_base, epoch);
}

// ---------------------------------------------------------------------------------------------
// All-gather push helper: a thread that produced element `idx` of the gathered vector stores it
// into the same offset of every rank's copy (local copy included): ONE multimem.st when the heap
// has a multicast mapping (the switch replicates it), else W unicast stores issued back to back.
// ---------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void push_all(const DapbComm& c, uint64_t off, uint64_t idx, T v) {
  if (c.mc_heap != nullptr) {
    multimem_st(reinterpret_cast<T*>(comm_payload_mc(c, off)) + idx, v);
    return;
  }
#pragma unroll
  for (int r = 0; r < DAPB_MAX_RANKS; ++r)
    if (r < c.world) reinterpret_cast<T*>(comm_payload(c, r, off))[idx] = v;
}

// ---------------------------------------------------------------------------------------------
// Grid-wide collectives over a payload region that every rank has ALREADY FILLED LOCALLY with its
// contribution (n elements at `off` of its own heap) and published with a barrier:
//   * grid_reduce_scatter_allgather: two-shot all-reduce.  Rank r owns elements [r*seg, (r+1)*seg);
//     it reduces them over all ranks (multimem.ld_reduce, or W peer loads added in RANK ORDER) and
//     broadcasts the sums (multimem.st, or W peer stores), so every rank ends with bit-identical
//     values.  Work is spread over the whole grid (threads stride over float4 groups).
// The caller brackets it with rank barriers (data in place before, results landed after).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void grid_twoshot_allreduce_f32(const DapbComm& c, uint64_t off, int n /* multiple of 4*world */,
                                                           int64_t gtid, int64_t gthreads) {
  const int seg4 = n / 4 / c.world;                   // float4 groups owned by each rank
  const int64_t base4 = (int64_t)c.rank * seg4;
  if (c.mc_heap != nullptr) {
    float4* mc = reinterpret_cast<float4*>(comm_payload_mc(c, off));
    for (int64_t i = gtid; i < seg4; i += gthreads) {
      const float4 v = multimem_ld_reduce_add(mc + base4 + i);
      multimem_st(mc + base4 + i, v);
    }
  } else {
    for (int64_t i = gtid; i < seg4; i += gthreads) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < DAPB_MAX_RANKS; ++r)
        if (r < c.world) {
          const float4 v = __ldcg(reinterpret_cast<const float4*>(comm_payload(c, r, off)) + base4 + i);
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
#pragma unroll
      for (int r = 0; r < DAPB_MAX_RANKS; ++r)
        if (r < c.world) reinterpret_cast<float4*>(comm_payload(c, r, off))[base4 + i] = acc;
    }
  }
}

}  // namespace dapb
