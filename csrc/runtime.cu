// Runtime: error reporting, device info, symmetric heap + CUDA IPC, stand-alone F8 collectives.
// Replaces the reference's implicit Spark substrate: there is no driver
// process, peers map each other's heaps once and kernels exchange data directly.
#include <cooperative_groups.h>
#include <cstdarg>

#include "comm.cuh"

namespace cg = cooperative_groups;

namespace dapb {
static thread_local char g_err[1024] = "";
void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) n = 148;
  }
  return n;
}
}  // namespace dapb

using namespace dapb;

extern "C" const char* dapb_last_error(void) { return dapb::g_err; }
extern "C" int dapb_version(void) { return 100; }

extern "C" int dapb_device_info(int32_t* sms, int32_t* cc_major, int32_t* cc_minor, uint64_t* total_mem) {
  int dev = 0;
  DAPB_CUDA_CHECK(cudaGetDevice(&dev));
  cudaDeviceProp p;
  DAPB_CUDA_CHECK(cudaGetDeviceProperties(&p, dev));
  if (sms) *sms = p.multiProcessorCount;
  if (cc_major) *cc_major = p.major;
  if (cc_minor) *cc_minor = p.minor;
  if (total_mem) *total_mem = (uint64_t)p.totalGlobalMem;
  return 0;
}

extern "C" int dapb_symm_alloc(uint64_t bytes, void** out) {
  DAPB_REQUIRE(out != nullptr && bytes > DAPB_FLAG_BYTES, "symm_alloc: need > %llu bytes",
               (unsigned long long)DAPB_FLAG_BYTES);
  DAPB_CUDA_CHECK(cudaMalloc(out, bytes));
  DAPB_CUDA_CHECK(cudaMemset(*out, 0, bytes));
  DAPB_CUDA_CHECK(cudaDeviceSynchronize());
  return 0;
}
extern "C" int dapb_symm_free(void* ptr) {
  DAPB_CUDA_CHECK(cudaFree(ptr));
  return 0;
}
extern "C" int dapb_ipc_get_handle(void* ptr, void* handle64) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  DAPB_CUDA_CHECK(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle64, &h, 64);
  return 0;
}
extern "C" int dapb_ipc_open_handle(const void* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  DAPB_CUDA_CHECK(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
extern "C" int dapb_ipc_close_handle(void* ptr) {
  DAPB_CUDA_CHECK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

// ------------------------------------------------------------------ stand-alone collectives
template <typename T>
__global__ void allreduce_kernel(DapbComm c, uint64_t off, uint32_t flag_base, uint64_t epoch,
                                 const T* src, T* dst, int n) {
  block_allreduce_oneshot<T>(c, c.rank, off, flag_base, epoch, src, dst, n);
}
__global__ void barrier_kernel(DapbComm c, uint32_t flag_base, uint64_t epoch) {
  block_barrier(c, c.rank, flag_base, epoch);
}

template <typename T>
static int allreduce_impl(const DapbComm* comm, uint64_t off, uint32_t flag_base, uint64_t epoch,
                          const T* src, T* dst, int32_t n, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "allreduce: bad comm");
  DAPB_REQUIRE(epoch >= 1, "allreduce: epoch must start at 1");
  uint64_t need = DAPB_FLAG_BYTES + off + (uint64_t)2 * comm->world * ((n + 3) & ~3) * sizeof(T);
  DAPB_REQUIRE(comm->world == 1 || need <= comm->heap_bytes, "allreduce: heap too small (%llu > %llu)",
               (unsigned long long)need, (unsigned long long)comm->heap_bytes);
  allreduce_kernel<T><<<1, 512, 0, (cudaStream_t)stream>>>(*comm, off, flag_base, epoch, src, dst, n);
  DAPB_LAUNCH_CHECK();
  return 0;
}
extern "C" int dapb_allreduce_f32(const DapbComm* comm, uint64_t off, uint32_t flag_base, uint64_t epoch,
                                  const float* src, float* dst, int32_t n, void* stream) {
  return allreduce_impl<float>(comm, off, flag_base, epoch, src, dst, n, stream);
}
extern "C" int dapb_allreduce_f64(const DapbComm* comm, uint64_t off, uint32_t flag_base, uint64_t epoch,
                                  const double* src, double* dst, int32_t n, void* stream) {
  return allreduce_impl<double>(comm, off, flag_base, epoch, src, dst, n, stream);
}
extern "C" int dapb_barrier(const DapbComm* comm, uint32_t flag_base, uint64_t epoch, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "barrier: bad comm");
  barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*comm, flag_base, epoch);
  DAPB_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------ stand-alone grid-wide collectives (F8)
// Multi-CTA versions for messages beyond one CTA's reach (KBs .. 100s of MBs).  All are cooperative launches (the grid
// synchronises around the cross-rank flag barriers) and use the multicast mapping when the heap has one:
//   all-gather       every element is pushed once (multimem.st) or W times (unicast) by the thread that read it
//   reduce-scatter   owner pulls its segment from all ranks (multimem.ld_reduce, or W peer loads summed in rank order)
//   two-shot all-reduce = reduce-scatter + all-gather in place, bit-identical on all ranks
//   all-to-all       rank r stores block d of its send buffer into slot r of rank d
// Each call consumes TWO barrier epochs on its flag words: 2*epoch - 1 and 2*epoch.
namespace {

__device__ __forceinline__ void grid_rank_barrier(const DapbComm& c, uint32_t flag_base, uint64_t epoch,
                                                  cg::grid_group& grid) {
  grid.sync();
  if (blockIdx.x == 0) block_barrier(c, c.rank, flag_base, epoch);
  grid.sync();
}

template <typename T>
__global__ void allgather_kernel(DapbComm c, uint64_t off, uint32_t flag_base, uint64_t epoch, const T* src,
                                 T* dst, int64_t n) {
  cg::grid_group grid = cg::this_grid();
  const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, gth = (int64_t)gridDim.x * blockDim.x;
  grid_rank_barrier(c, flag_base, 2 * epoch - 1, grid);           // everybody finished reading the previous result
  for (int64_t i = gtid; i < n; i += gth) push_all<T>(c, off, (uint64_t)c.rank * n + i, src[i]);
  __threadfence_system();
  grid_rank_barrier(c, flag_base, 2 * epoch, grid);
  const T* mine = reinterpret_cast<const T*>(comm_payload(c, c.rank, off));
  for (int64_t i = gtid; i < n * c.world; i += gth) dst[i] = __ldcg(mine + i);
}

// mode 0: reduce-scatter (dst[n/world] = my segment), mode 1: all-reduce (dst[n])
__global__ void twoshot_kernel(DapbComm c, uint64_t off, uint32_t flag_base, uint64_t epoch, const float* src,
                               float* dst, int n, int mode) {
  cg::grid_group grid = cg::this_grid();
  const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, gth = (int64_t)gridDim.x * blockDim.x;
  float4* mine = reinterpret_cast<float4*>(comm_payload(c, c.rank, off));
  for (int64_t i = gtid; i < n / 4; i += gth) mine[i] = reinterpret_cast<const float4*>(src)[i];
  __threadfence_system();
  grid_rank_barrier(c, flag_base, 2 * epoch - 1, grid);
  grid_twoshot_allreduce_f32(c, off, n, gtid, gth);
  __threadfence_system();
  grid_rank_barrier(c, flag_base, 2 * epoch, grid);
  if (mode == 1) {
    for (int64_t i = gtid; i < n / 4; i += gth) reinterpret_cast<float4*>(dst)[i] = __ldcg(mine + i);
  } else {
    const int seg4 = n / 4 / c.world;
    for (int64_t i = gtid; i < seg4; i += gth) reinterpret_cast<float4*>(dst)[i] = __ldcg(mine + (int64_t)c.rank * seg4 + i);
  }
}

__global__ void alltoall_kernel(DapbComm c, uint64_t off, uint32_t flag_base, uint64_t epoch,
                                const unsigned long long* src, unsigned long long* dst, int64_t n) {
  cg::grid_group grid = cg::this_grid();
  const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, gth = (int64_t)gridDim.x * blockDim.x;
  grid_rank_barrier(c, flag_base, 2 * epoch - 1, grid);
  for (int d = 0; d < c.world; ++d) {
    unsigned long long* slot = reinterpret_cast<unsigned long long*>(comm_payload(c, d, off)) + (int64_t)c.rank * n;
    for (int64_t i = gtid; i < n; i += gth) slot[i] = src[(int64_t)d * n + i];
  }
  __threadfence_system();
  grid_rank_barrier(c, flag_base, 2 * epoch, grid);
  const unsigned long long* mine = reinterpret_cast<const unsigned long long*>(comm_payload(c, c.rank, off));
  for (int64_t i = gtid; i < n * c.world; i += gth) dst[i] = __ldcg(mine + i);
}

template <typename K>
int coop_launch(K kernel, void** kargs, int threads, cudaStream_t st, int64_t work) {
  int per_sm = 0;
  DAPB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0));
  int64_t grid = (int64_t)per_sm * dapb::sm_count();
  const int64_t need = (work + threads - 1) / threads;
  if (grid > need) grid = need;
  if (grid < 1) grid = 1;
  DAPB_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)kernel, dim3((unsigned)grid), dim3(threads), kargs, 0, st));
  return 0;
}

int check_grid_coll(const DapbComm* comm, uint64_t off, uint64_t bytes, uint64_t epoch, const char* what) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "%s: bad comm", what);
  DAPB_REQUIRE(epoch >= 1, "%s: epoch must start at 1", what);
  DAPB_REQUIRE(DAPB_FLAG_BYTES + off + bytes <= comm->heap_bytes, "%s: heap too small (%llu > %llu)", what,
               (unsigned long long)(DAPB_FLAG_BYTES + off + bytes), (unsigned long long)comm->heap_bytes);
  return 0;
}

}  // namespace

extern "C" int dapb_allgather(const DapbComm* comm, uint64_t off, uint32_t flag_base, uint64_t epoch, const void* src,
                              void* dst, int64_t n_per_rank, int32_t elem_bytes, void* stream) {
  DAPB_REQUIRE(elem_bytes == 4 || elem_bytes == 8, "allgather: 4- or 8-byte elements");
  if (int rc = check_grid_coll(comm, off, (uint64_t)comm->world * n_per_rank * elem_bytes, epoch, "allgather")) return rc;
  DapbComm c = *comm;
  void* kargs[] = {&c, &off, &flag_base, &epoch, &src, &dst, &n_per_rank};
  if (elem_bytes == 4) return coop_launch(allgather_kernel<float>, kargs, 256, (cudaStream_t)stream, n_per_rank * c.world);
  return coop_launch(allgather_kernel<double>, kargs, 256, (cudaStream_t)stream, n_per_rank * c.world);
}
static int twoshot(const DapbComm* comm, uint64_t off, uint32_t flag_base, uint64_t epoch, const float* src, float* dst,
                   int32_t n, int mode, void* stream, const char* what) {
  if (int rc = check_grid_coll(comm, off, (uint64_t)n * 4, epoch, what)) return rc;
  DAPB_REQUIRE(n > 0 && n % (4 * comm->world) == 0, "%s: n must be a multiple of 4 * world", what);
  DAPB_REQUIRE(((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 15) == 0 && (off & 15) == 0, "%s: 16-byte alignment", what);
  DapbComm c = *comm;
  void* kargs[] = {&c, &off, &flag_base, &epoch, &src, &dst, &n, &mode};
  return coop_launch(twoshot_kernel, kargs, 256, (cudaStream_t)stream, n / 4);
}
extern "C" int dapb_allreduce_twoshot_f32(const DapbComm* comm, uint64_t off, uint32_t flag_base, uint64_t epoch,
                                          const float* src, float* dst, int32_t n, void* stream) {
  return twoshot(comm, off, flag_base, epoch, src, dst, n, 1, stream, "allreduce_twoshot");
}
extern "C" int dapb_reduce_scatter_f32(const DapbComm* comm, uint64_t off, uint32_t flag_base, uint64_t epoch,
                                       const float* src, float* dst, int32_t n, void* stream) {
  return twoshot(comm, off, flag_base, epoch, src, dst, n, 0, stream, "reduce_scatter");
}
extern "C" int dapb_alltoall_u64(const DapbComm* comm, uint64_t off, uint32_t flag_base, uint64_t epoch,
                                 const unsigned long long* src, unsigned long long* dst, int64_t n_per_pair, void* stream) {
  if (int rc = check_grid_coll(comm, off, (uint64_t)comm->world * n_per_pair * 8, epoch, "alltoall")) return rc;
  DapbComm c = *comm;
  void* kargs[] = {&c, &off, &flag_base, &epoch, &src, &dst, &n_per_pair};
  return coop_launch(alltoall_kernel, kargs, 256, (cudaStream_t)stream, n_per_pair * c.world);
}

// ------------------------------------------------------------------ cross-GPU flag-protocol stress
// ONE CTA per rank runs `iters` back-to-back one-shot all-reduces on real peers (NVLink ordering, not one GPU's L2),
// with pseudo-random __nanosleep skew between epochs so that fast ranks run a full epoch ahead of slow ones and the
// double-buffered slots / monotonic flags are exercised; every element of every epoch is checked bit for bit.
__global__ void allreduce_stress_kernel(DapbComm c, uint64_t off, uint32_t flag_base, uint64_t epoch0, int n, int iters,
                                        uint32_t seed, int* out_fail) {
  extern __shared__ float sm[];
  float* src = sm;
  float* dst = sm + n;
  int fails = 0;
  uint32_t rng = seed * 2654435761u + (uint32_t)c.rank * 40503u + 1u;
  for (int it = 0; it < iters; ++it) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) src[i] = (float)((c.rank + 1) * (i % 7 + 1) + (it & 1023));
    __syncthreads();
    block_allreduce_oneshot<float>(c, c.rank, off, flag_base, epoch0 + it, src, dst, n);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      float want = 0.f;
      for (int r = 0; r < c.world; ++r) want += (float)((r + 1) * (i % 7 + 1) + (it & 1023));
      if (dst[i] != want) ++fails;
    }
    rng = rng * 1664525u + 1013904223u;
    if ((rng >> 28) == 0) __nanosleep((rng >> 8) & 0x1fff);      // 1 epoch in 16: stall up to 8 us
    __syncthreads();
  }
  if (fails) atomicAdd(out_fail, fails);
}

extern "C" int dapb_allreduce_stress(const DapbComm* comm, uint64_t off, uint32_t flag_base, uint64_t epoch0, int32_t n,
                                     int32_t iters, uint32_t seed, int32_t* out_fail, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "allreduce_stress: bad comm");
  DAPB_REQUIRE(n > 0 && n <= 4096 && iters > 0 && epoch0 >= 1, "allreduce_stress: n in [1,4096], iters > 0, epoch0 >= 1");
  uint64_t need = DAPB_FLAG_BYTES + off + (uint64_t)2 * comm->world * ((n + 3) & ~3) * sizeof(float);
  DAPB_REQUIRE(comm->world == 1 || need <= comm->heap_bytes, "allreduce_stress: heap too small");
  allreduce_stress_kernel<<<1, 256, 2 * n * sizeof(float), (cudaStream_t)stream>>>(*comm, off, flag_base, epoch0, n, iters,
                                                                                   seed, out_fail);
  DAPB_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------ single-GPU protocol test
// `world` CTAs of one cooperative launch play the ranks (B200_PROFILING.md: never emulate ranks
// with separate launches that wait on each other on one GPU).
__global__ void comm_selftest_kernel(char* heaps, uint64_t heap_bytes, int world, int n, int iters,
                                     int* out_fail) {
  extern __shared__ float sm[];
  float* src = sm;
  float* dst = sm + n;
  DapbComm c;
  c.rank = blockIdx.x;
  c.world = world;
  c.heap_bytes = heap_bytes;
  c.mc_heap = nullptr;
  for (int r = 0; r < DAPB_MAX_RANKS; ++r) c.heap[r] = r < world ? heaps + (uint64_t)r * heap_bytes : nullptr;
  int fails = 0;
  for (int it = 1; it <= iters; ++it) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) src[i] = (float)((c.rank + 1) * (i % 7 + 1) + it);
    __syncthreads();
    block_allreduce_oneshot<float>(c, c.rank, 0, 0, (uint64_t)it, src, dst, n);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      float want = 0.f;
      for (int r = 0; r < world; ++r) want += (float)((r + 1) * (i % 7 + 1) + it);
      if (dst[i] != want) ++fails;
    }
    // skew the ranks so fast ones run ahead and exercise the double buffering
    if ((it + c.rank) % 3 == 0) __nanosleep(2000);
    __syncthreads();
  }
  if (fails) atomicAdd(out_fail, fails);
}

extern "C" int dapb_comm_selftest(void* heaps, uint64_t heap_bytes, int32_t world, int32_t n,
                                  int32_t iters, int32_t* out_fail, void* stream) {
  DAPB_REQUIRE(world >= 1 && world <= DAPB_MAX_RANKS, "selftest: world out of range");
  DAPB_REQUIRE(n > 0 && n <= 4096, "selftest: n out of range");
  uint64_t need = DAPB_FLAG_BYTES + (uint64_t)2 * world * ((n + 3) & ~3) * sizeof(float);
  DAPB_REQUIRE(need <= heap_bytes, "selftest: heap too small");
  char* hp = (char*)heaps;
  void* kargs[] = {&hp, &heap_bytes, &world, &n, &iters, &out_fail};
  DAPB_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)comm_selftest_kernel, dim3(world), dim3(256), kargs,
                                              2 * n * sizeof(float), (cudaStream_t)stream));
  return 0;
}
