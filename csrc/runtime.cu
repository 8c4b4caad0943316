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
