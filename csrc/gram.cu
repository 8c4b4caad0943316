// F5 (any-shape path) -- tall-skinny products C = X^T Y over row-sharded operands, reduced across
// GPUs from the epilogue of the same kernel, plus the small dense steps that follow it: SPD solve
// (ALS normal equations), Cholesky factor (Cholesky-QR "R" of A), residual / RMSE.
//
// ref: matrix_computation/matrix_decomposition.py:19-21 (rmse), :24-33 (update: XtX = mat^T mat,
//      ridge lam * mat.shape[0] on the diagonal, Xty = mat^T ratings[i], np.linalg.solve), :51-67.
// The reference recomputes the identical k x k Gram for every row i (:27); here it is computed once
// per half-step and the m (or n) right-hand sides are solved as a batch.
// The tcgen05 tensor-core Gram for the 2^27 x 512 benchmark shape lives in gram_tc.cu.
#include "comm.cuh"

using namespace dapb;

namespace {

constexpr int kTile = 32;       // output tile 32 x 32 per CTA
constexpr int kRows = 32;       // rows staged per step

// C[p, q] (+)= sum_i X[i, p] * Y[i, q]; grid = (ceil(P/32), ceil(Q/32), row_splits)
// TA: accumulation type (double for fp64 inputs, float otherwise).  Partial tiles are added into a
// per-split slab `partial[split][P][Q]`; the last CTA sums the slabs in order (deterministic),
// all-reduces over the peers and writes C.
template <typename T, typename TA>
__global__ void __launch_bounds__(256) gram_tn_kernel(const T* __restrict__ X, const T* __restrict__ Y, int64_t n,
                                                      int P, int Q, int64_t ldx, int64_t ldy,
                                                      TA* __restrict__ partial, TA* __restrict__ C, uint32_t* ticket,
                                                      DapbComm c, uint64_t payload_off, uint32_t flag_base,
                                                      uint64_t epoch, int do_allreduce) {
  __shared__ TA xs[kRows][kTile + 1];
  __shared__ TA ys[kRows][kTile + 1];
  __shared__ bool s_last;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;       // 32 x 8 threads
  const int p0 = blockIdx.x * kTile, q0 = blockIdx.y * kTile;
  const int splits = gridDim.z;
  const int64_t rows_per = ((n + splits - 1) / splits + kRows - 1) / kRows * kRows;
  const int64_t r_begin = (int64_t)blockIdx.z * rows_per, r_end = min(n, r_begin + rows_per);
  TA acc[4] = {0, 0, 0, 0};                                      // thread owns C[p0 + ty*4 + j][q0 + tx]
  for (int64_t r = r_begin; r < r_end; r += kRows) {
    for (int i = ty; i < kRows; i += 8) {
      const int64_t row = r + i;
      xs[i][tx] = (row < r_end && p0 + tx < P) ? (TA)X[row * ldx + p0 + tx] : (TA)0;
      ys[i][tx] = (row < r_end && q0 + tx < Q) ? (TA)Y[row * ldy + q0 + tx] : (TA)0;
    }
    __syncthreads();
#pragma unroll 8
    for (int i = 0; i < kRows; ++i) {
      const TA yv = ys[i][tx];
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[j] += xs[i][ty * 4 + j] * yv;
    }
    __syncthreads();
  }
  TA* slab = partial + (int64_t)blockIdx.z * P * Q;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int p = p0 + ty * 4 + j, q = q0 + tx;
    if (p < P && q < Q) slab[(int64_t)p * Q + q] = acc[j];
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned total = gridDim.x * gridDim.y * gridDim.z;
    unsigned t = atomicAdd(ticket, 1u);
    s_last = (t == total - 1);
    if (s_last) *ticket = 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // ordered sum over the row splits, then (optionally) over the GPUs
  const int total_el = P * Q;
  for (int i = threadIdx.x; i < total_el; i += blockDim.x) {
    TA s = 0;
    for (int z = 0; z < splits; ++z) s += __ldcg(partial + (int64_t)z * total_el + i);
    C[i] = s;
  }
  __syncthreads();
  if (do_allreduce && c.world > 1) {
    __threadfence();
    block_allreduce_oneshot<TA>(c, c.rank, payload_off, flag_base, epoch, C, C, total_el);
  }
}

// out[i, :] = R[i, :] @ V   (R [m, n], V [n, k], k <= 64) -- row-local half of the ALS U-step
template <typename T>
__global__ void __launch_bounds__(256) rows_times_small_kernel(const T* __restrict__ R, const T* __restrict__ V,
                                                               int64_t m, int n, int k, int64_t ldr, int transpose_r,
                                                               T* __restrict__ out) {
  // one warp per output row; lane l accumulates columns l, l+32
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= m) return;
  T acc0 = 0, acc1 = 0;
  for (int j = 0; j < n; ++j) {
    const T r = transpose_r ? R[(int64_t)j * ldr + row] : R[row * ldr + j];
    if (lane < k) acc0 += r * V[(int64_t)j * k + lane];
    if (lane + 32 < k) acc1 += r * V[(int64_t)j * k + lane + 32];
  }
  if (lane < k) out[row * k + lane] = acc0;
  if (lane + 32 < k) out[row * k + lane + 32] = acc1;
}

// Solve (G + ridge*I) x_i = b_i for every row b_i of B [m, k] (k <= 64): LU with partial pivoting in
// shared memory by one CTA (np.linalg.solve == LAPACK gesv, ref matrix_decomposition.py:33), then one
// thread per right-hand side.
template <typename T>
__global__ void __launch_bounds__(256) solve_rows_kernel(const T* __restrict__ G, T ridge, const T* __restrict__ B,
                                                         int64_t m, int k, T* __restrict__ Xout) {
  extern __shared__ __align__(16) unsigned char sm_raw[];
  T* A = reinterpret_cast<T*>(sm_raw);              // [k][k] LU factors
  int* piv = reinterpret_cast<int*>(A + k * k);     // [k]
  __shared__ int s_p;
  const int tid = threadIdx.x;
  for (int i = tid; i < k * k; i += blockDim.x) A[i] = G[i] + ((i / k == i % k) ? ridge : (T)0);
  __syncthreads();
  for (int col = 0; col < k; ++col) {
    if (tid == 0) {
      int p = col;
      T best = fabs(A[col * k + col]);
      for (int r = col + 1; r < k; ++r) {
        T v = fabs(A[r * k + col]);
        if (v > best) { best = v; p = r; }
      }
      piv[col] = p;
      s_p = p;
    }
    __syncthreads();
    const int p = s_p;
    if (p != col)
      for (int j = tid; j < k; j += blockDim.x) { T t = A[col * k + j]; A[col * k + j] = A[p * k + j]; A[p * k + j] = t; }
    __syncthreads();
    const T d = A[col * k + col];
    for (int r = col + 1 + tid; r < k; r += blockDim.x) A[r * k + col] /= d;
    __syncthreads();
    for (int idx = tid; idx < (k - col - 1) * (k - col - 1); idx += blockDim.x) {
      const int r = col + 1 + idx / (k - col - 1), j = col + 1 + idx % (k - col - 1);
      A[r * k + j] -= A[r * k + col] * A[col * k + j];
    }
    __syncthreads();
  }
  // every CTA factored redundantly; rows are distributed over the grid
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + tid; row < m; row += (int64_t)gridDim.x * blockDim.x) {
    T x[64];
    for (int j = 0; j < k; ++j) x[j] = B[row * k + j];
    for (int j = 0; j < k; ++j) { const int p = piv[j]; if (p != j) { T t = x[j]; x[j] = x[p]; x[p] = t; } }
    for (int i = 1; i < k; ++i) { T s = x[i]; for (int j = 0; j < i; ++j) s -= A[i * k + j] * x[j]; x[i] = s; }
    for (int i = k - 1; i >= 0; --i) {
      T s = x[i];
      for (int j = i + 1; j < k; ++j) s -= A[i * k + j] * x[j];
      x[i] = s / A[i * k + i];
    }
    for (int j = 0; j < k; ++j) Xout[row * k + j] = x[j];
  }
}

// sum over local rows of (R[i, j] - U[i] . V[j])^2  (ref matrix_decomposition.py:19-21)
template <typename T>
__global__ void __launch_bounds__(256) sq_residual_kernel(const T* __restrict__ R, const T* __restrict__ U,
                                                          const T* __restrict__ V, int64_t m, int n, int k,
                                                          double* __restrict__ out) {
  double local = 0.0;
  const int64_t total = m * (int64_t)n;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = idx / n;
    const int j = (int)(idx - i * n);
    T dot = 0;
    for (int t = 0; t < k; ++t) dot += U[i * k + t] * V[(int64_t)j * k + t];
    const double diff = (double)R[idx] - (double)dot;
    local += diff * diff;
  }
  local = warp_sum(local);
  if ((threadIdx.x & 31) == 0 && local != 0.0) atomicAdd(out, local);
}

template <typename T, typename TA>
int launch_gram(const DapbComm* comm, const void* X, const void* Y, int64_t n, int P, int Q, int64_t ldx, int64_t ldy,
                void* partial, int splits, void* C, uint32_t* ticket, uint64_t payload_off, uint32_t flag_base,
                uint64_t epoch, int do_allreduce, cudaStream_t s) {
  dim3 grid((P + kTile - 1) / kTile, (Q + kTile - 1) / kTile, splits);
  gram_tn_kernel<T, TA><<<grid, 256, 0, s>>>((const T*)X, (const T*)Y, n, P, Q, ldx, ldy, (TA*)partial, (TA*)C, ticket,
                                             *comm, payload_off, flag_base, epoch, do_allreduce);
  DAPB_LAUNCH_CHECK();
  return 0;
}

}  // namespace

// C[P, Q] = X^T Y summed over all ranks' rows.  dtype: 0 fp32, 1 bf16 (fp32 accumulate), 2 fp64.
// partial: scratch [splits, P, Q] of the accumulation type; ticket: 1 zeroed word.
extern "C" int dapb_gram(const DapbComm* comm, const void* X, const void* Y, int64_t n, int32_t P, int32_t Q,
                         int64_t ldx, int64_t ldy, int32_t dtype, void* partial, int32_t splits, void* C,
                         uint32_t* ticket, uint64_t payload_off, uint32_t flag_base, uint64_t epoch,
                         int32_t do_allreduce, void* stream) {
  DAPB_REQUIRE(comm && comm->world >= 1 && comm->world <= DAPB_MAX_RANKS, "gram: bad comm");
  DAPB_REQUIRE(P >= 1 && Q >= 1 && splits >= 1 && partial && C && ticket, "gram: bad arguments");
  if (do_allreduce && comm->world > 1) {
    DAPB_REQUIRE(epoch >= 1, "gram: epoch starts at 1");
    uint64_t need = DAPB_FLAG_BYTES + payload_off + 2ull * comm->world * ((P * Q + 3) & ~3) * (dtype == 2 ? 8 : 4);
    DAPB_REQUIRE(need <= comm->heap_bytes, "gram: heap too small (%llu > %llu)", (unsigned long long)need,
                 (unsigned long long)comm->heap_bytes);
  }
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case 0: return launch_gram<float, float>(comm, X, Y, n, P, Q, ldx, ldy, partial, splits, C, ticket, payload_off, flag_base, epoch, do_allreduce, s);
    case 1: return launch_gram<__nv_bfloat16, float>(comm, X, Y, n, P, Q, ldx, ldy, partial, splits, C, ticket, payload_off, flag_base, epoch, do_allreduce, s);
    case 2: return launch_gram<double, double>(comm, X, Y, n, P, Q, ldx, ldy, partial, splits, C, ticket, payload_off, flag_base, epoch, do_allreduce, s);
    default: DAPB_REQUIRE(false, "gram: dtype must be 0, 1 or 2");
  }
  return 0;
}

extern "C" int dapb_rows_times_small(const void* R, const void* V, int64_t m, int32_t n, int32_t k, int64_t ldr,
                                     int32_t transpose_r, int32_t f64, void* out, void* stream) {
  DAPB_REQUIRE(k >= 1 && k <= 64, "rows_times_small: k must be in [1, 64]");
  const int64_t blocks = (m * 32 + 255) / 256;
  if (blocks == 0) return 0;
  if (f64) rows_times_small_kernel<double><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const double*)R, (const double*)V, m, n, k, ldr, transpose_r, (double*)out);
  else rows_times_small_kernel<float><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const float*)R, (const float*)V, m, n, k, ldr, transpose_r, (float*)out);
  DAPB_LAUNCH_CHECK();
  return 0;
}

extern "C" int dapb_solve_rows(const void* G, double ridge, const void* B, int64_t m, int32_t k, int32_t f64,
                               void* X, void* stream) {
  DAPB_REQUIRE(k >= 1 && k <= 64, "solve_rows: k must be in [1, 64]");
  int blocks = (int)((m + 255) / 256);
  if (blocks < 1) blocks = 1;
  if (blocks > dapb::sm_count()) blocks = dapb::sm_count();
  const size_t smem = (size_t)k * k * (f64 ? 8 : 4) + k * sizeof(int) + 16;
  if (f64) solve_rows_kernel<double><<<blocks, 256, smem, (cudaStream_t)stream>>>((const double*)G, ridge, (const double*)B, m, k, (double*)X);
  else solve_rows_kernel<float><<<blocks, 256, smem, (cudaStream_t)stream>>>((const float*)G, (float)ridge, (const float*)B, m, k, (float*)X);
  DAPB_LAUNCH_CHECK();
  return 0;
}

extern "C" int dapb_sq_residual(const void* R, const void* U, const void* V, int64_t m, int32_t n, int32_t k,
                                int32_t f64, double* out, void* stream) {
  int64_t blocks = (m * n + 255) / 256;
  if (blocks > dapb::sm_count() * 8) blocks = dapb::sm_count() * 8;
  if (blocks < 1) blocks = 1;
  if (f64) sq_residual_kernel<double><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const double*)R, (const double*)U, (const double*)V, m, n, k, out);
  else sq_residual_kernel<float><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const float*)R, (const float*)U, (const float*)V, m, n, k, out);
  DAPB_LAUNCH_CHECK();
  return 0;
}
