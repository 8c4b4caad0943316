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

// In-place lower Cholesky of a symmetric n x n fp32/fp64 matrix in global memory, blocked right-looking with
// panels of kCb columns: per panel one single-CTA kernel factors the diagonal block in shared memory and solves
// the rows below it, then a grid of 32 x 32 tiles applies the rank-kCb trailing update (only the lower triangle).
// `info` = 1 + index of the first non-positive pivot, or 0.  The strict upper triangle is zeroed at the end.
constexpr int kCb = 32;

template <typename T>
__global__ void __launch_bounds__(1024) chol_panel_kernel(T* __restrict__ A, int n, int j0, int* __restrict__ info) {
  __shared__ T L[kCb][kCb + 1];
  const int tid = threadIdx.x;
  const int w = min(kCb, n - j0);
  for (int i = tid; i < kCb * kCb; i += blockDim.x) {
    const int r = i / kCb, cc = i % kCb;
    L[r][cc] = (r < w && cc < w && cc <= r) ? A[(int64_t)(j0 + r) * n + j0 + cc] : (T)0;
  }
  __syncthreads();
  for (int j = 0; j < w; ++j) {                         // unblocked factorisation of the w x w diagonal block
    if (tid == 0) {
      T d = L[j][j];
      if (!(d > (T)0)) { if (*info == 0) *info = j0 + j + 1; d = (T)1; }
      L[j][j] = sqrt(d);
    }
    __syncthreads();
    if (tid > j && tid < w) L[tid][j] /= L[j][j];
    __syncthreads();
    for (int idx = tid; idx < w * w; idx += blockDim.x) {
      const int r = idx / w, cc = idx % w;
      if (cc > j && cc <= r) L[r][cc] -= L[r][j] * L[cc][j];
    }
    __syncthreads();
  }
  for (int i = tid; i < w * w; i += blockDim.x) {
    const int r = i / w, cc = i % w;
    if (cc <= r) A[(int64_t)(j0 + r) * n + j0 + cc] = L[r][cc];
  }
  // rows below the block: solve  X L11^T = A21  row by row (forward substitution), one thread per row
  for (int r = j0 + w + tid; r < n; r += blockDim.x) {
    T x[kCb];
    for (int cc = 0; cc < w; ++cc) {
      T s = A[(int64_t)r * n + j0 + cc];
      for (int k = 0; k < cc; ++k) s -= x[k] * L[cc][k];
      x[cc] = s / L[cc][cc];
    }
    for (int cc = 0; cc < w; ++cc) A[(int64_t)r * n + j0 + cc] = x[cc];
  }
}

// A22[i, c] -= sum_k L21[i, k] L21[c, k] for the lower triangle of the trailing matrix (tiles of 32 x 32)
template <typename T>
__global__ void __launch_bounds__(256) chol_update_kernel(T* __restrict__ A, int n, int j0, int w) {
  const int bi = blockIdx.y, bj = blockIdx.x;
  if (bj > bi) return;
  __shared__ T Li[32][kCb + 1], Lj[32][kCb + 1];
  const int r0 = j0 + w + bi * 32, c0 = j0 + w + bj * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;         // 32 x 8
  for (int rr = ty; rr < 32; rr += 8) {
    Li[rr][tx] = (r0 + rr < n && tx < w) ? A[(int64_t)(r0 + rr) * n + j0 + tx] : (T)0;
    Lj[rr][tx] = (c0 + rr < n && tx < w) ? A[(int64_t)(c0 + rr) * n + j0 + tx] : (T)0;
  }
  __syncthreads();
  for (int rr = ty; rr < 32; rr += 8) {
    const int r = r0 + rr, cc = c0 + tx;
    if (r < n && cc <= r) {
      T s = 0;
#pragma unroll
      for (int k = 0; k < kCb; ++k) s += Li[rr][k] * Lj[tx][k];
      A[(int64_t)r * n + cc] -= s;
    }
  }
}

template <typename T>
__global__ void chol_zero_upper_kernel(T* __restrict__ A, int n) {
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < (int64_t)n * n; idx += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(idx / n), cc = (int)(idx % n);
    if (cc > i) A[idx] = (T)0;
  }
}

template <typename T>
int cholesky_blocked(T* A, int n, int* info, cudaStream_t st) {
  for (int j0 = 0; j0 < n; j0 += kCb) {
    const int w = n - j0 < kCb ? n - j0 : kCb;
    chol_panel_kernel<T><<<1, 1024, 0, st>>>(A, n, j0, info);
    const int rem = n - j0 - w;
    if (rem > 0) {
      const int nb = (rem + 31) / 32;
      chol_update_kernel<T><<<dim3(nb, nb), 256, 0, st>>>(A, n, j0, w);
    }
  }
  chol_zero_upper_kernel<T><<<64, 256, 0, st>>>(A, n);
  DAPB_LAUNCH_CHECK();
  return 0;
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

extern "C" int dapb_cholesky(void* A, int32_t n, int32_t f64, int32_t* info, void* stream) {
  DAPB_REQUIRE(n >= 1 && n <= 4096 && info, "cholesky: bad arguments");
  if (f64) return cholesky_blocked<double>((double*)A, n, info, (cudaStream_t)stream);
  return cholesky_blocked<float>((float*)A, n, info, (cudaStream_t)stream);
}
