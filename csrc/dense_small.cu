// F5, small dense step on the device -- the part of the decomposition that follows the Gram matrix:
//   * Cholesky  G = L L^T      (R = L^T is the QR factor of the tall-skinny A; also the ALS normal equations)
//   * symmetric eigen-decomposition  G = V diag(lambda) V^T  (singular values of A = sqrt(lambda), right singular vectors V)
// Both are ONE cooperative launch each; nothing here calls cuSOLVER.
//
// ref: matrix_computation/matrix_decomposition.py:24-33 -- the reference hands its k x k system to LAPACK
// (np.linalg.solve) once per row on the driver; BASELINE.json config 5 asks for "a small on-device eigen/QR step" after
// the 512 x 512 Gram matrix.
//
// Cholesky, right-looking, panels of 32 columns.  Per panel: every CTA factors the 32 x 32 diagonal block ITSELF inside
// one warp in shared memory (row per lane, __syncwarp only -- redundant work, but it saves a grid barrier; keeping the
// block in per-lane register arrays instead ended in local memory: 25 us per block), solves its share of the rows below against it, grid barrier, rank-32 update of its share of the trailing
// lower-triangle tiles, grid barrier.  16 panels x 2 barriers for n = 512 instead of 33 kernel launches.
//
// Eigen-solver: block one-sided Jacobi (Hestenes).  W starts as G, V as I; plane rotations make the columns of W = G V
// mutually orthogonal, then lambda_j = v_j . w_j.  Columns are grouped into 2B blocks of b; a round-robin tournament
// pairs the blocks (B pairs per step, 2B - 1 steps per sweep); a CTA pulls its two blocks of W and V into shared memory,
// runs a complete tournament over the 2b columns (one warp per column pair, dot products by warp shuffles), writes them
// back; grid barrier.  Sweeps repeat until the largest |w_p.w_q| / (|w_p||w_q|) seen in a sweep drops below tol.
#include <cooperative_groups.h>
#include <cstdlib>
#include "common.cuh"

namespace cg = cooperative_groups;
using namespace dapb;

namespace {

// ================================================================================================ Cholesky
constexpr int kPb = 32;            // panel width = warp size

// The 32 x 32 diagonal block is factored by ONE warp with the block in registers (row per lane).  Every index must be
// a compile-time constant or the row array lands in local memory (measured: 25 us per block), and nvcc does not unroll
// a 32 x 31 loop nest on request -- template recursion forces it.  Step J: pivot, scale column J, rank-1 update of the
// columns right of it; the column entries travel by shuffles.
// 1 / sqrt(d) and sqrt(d): fp32 takes the SFU estimate plus one Newton step (relative error ~1e-7) instead of the
// IEEE sqrt + divide sequences, which sat on the critical path of every one of the 32 dependent steps.
__device__ __forceinline__ void inv_sqrt(float d, float& sd, float& inv) {
  float r = rsqrtf(d);
  r = r * (1.5f - 0.5f * d * r * r);
  inv = r; sd = d * r;
}
__device__ __forceinline__ void inv_sqrt(double d, double& sd, double& inv) { sd = sqrt(d); inv = 1.0 / sd; }

template <typename T, int J>
struct CholStep {
  static __device__ __forceinline__ void run(T (&a)[kPb], int lane, int w, int& bad) {
    T d = __shfl_sync(0xffffffffu, a[J], J);
    if (J < w && !(d > (T)0)) { if (!bad) bad = J + 1; d = (T)1; }
    T sd, inv;
    inv_sqrt(d, sd, inv);
    if (lane == J) a[J] = sd;
    else if (lane > J) a[J] = a[J] * inv;
#pragma unroll
    for (int c = J + 1; c < kPb; ++c) {
      const T lcj = __shfl_sync(0xffffffffu, a[J], c);       // L[c][J]
      if (lane >= c) a[c] -= a[J] * lcj;
    }
    if constexpr (J + 1 < kPb) CholStep<T, J + 1>::run(a, lane, w, bad);
  }
};
// Forward substitution of one panel row against L11 (shared memory, broadcast reads), same trick.
template <typename T, int CC>
struct SolveStep {
  static __device__ __forceinline__ void run(T (&x)[kPb], const T (*L)[kPb + 1], const T* invd) {
    T s = x[CC];
#pragma unroll
    for (int k = 0; k < CC; ++k) s -= x[k] * L[CC][k];
    x[CC] = s * invd[CC];
    if constexpr (CC + 1 < kPb) SolveStep<T, CC + 1>::run(x, L, invd);
  }
};

// Shared-memory layout of the Cholesky kernel (dynamic: the fp64 variant exceeds the 48 KB static limit).
template <typename T>
struct CholSmem {
  T L11[kPb][kPb + 1];
  T invd[kPb];             // 1 / L11[c][c]
  T Li[kPb][kPb + 1];
  T Lj[kPb][kPb + 1];
};

template <typename T>
__global__ void __launch_bounds__(256) chol_coop_kernel(T* __restrict__ A, int n, int* __restrict__ info, long long* dbg) {
  cg::grid_group grid = cg::this_grid();
  extern __shared__ __align__(16) unsigned char chol_smem_raw[];
  CholSmem<T>& sm = *reinterpret_cast<CholSmem<T>*>(chol_smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gtid = blockIdx.x * blockDim.x + tid, gth = gridDim.x * blockDim.x;
  long long t_fac = 0, t_sol = 0, t_s1 = 0, t_upd = 0, t_s2 = 0;
  const bool prof = dbg != nullptr && blockIdx.x == 0 && tid == 0;
  for (int j0 = 0; j0 < n; j0 += kPb) {
    const int w = min(kPb, n - j0);
    long long c0 = prof ? clock64() : 0;
    // ---- diagonal block: every CTA factors its own copy in shared memory (redundant, but no grid barrier needed
    //      before the panel solve); rows / columns beyond w act as an identity block ----
    if (warp == 0) {
      T a[kPb];
#pragma unroll
      for (int c = 0; c < kPb; ++c)
        a[c] = (lane < w && c < w) ? (c <= lane ? A[(int64_t)(j0 + lane) * n + j0 + c] : (T)0) : (lane == c ? (T)1 : (T)0);
      int bad = 0;
      CholStep<T, 0>::run(a, lane, w, bad);
#pragma unroll
      for (int c = 0; c < kPb; ++c) sm.L11[lane][c] = a[c];
      T dg = a[0];
#pragma unroll
      for (int c = 1; c < kPb; ++c) dg = lane == c ? a[c] : dg;
      sm.invd[lane] = (T)1 / dg;
      if (blockIdx.x == 0) {
        if (bad && lane == 0 && *info == 0) *info = j0 + bad;
#pragma unroll
        for (int c = 0; c < kPb; ++c)
          if (lane < w && c <= lane && c < w) A[(int64_t)(j0 + lane) * n + j0 + c] = a[c];
      }
    }
    __syncthreads();
    if (prof) { const long long c = clock64(); t_fac += c - c0; c0 = c; }
    // ---- rows below the block: solve X L11^T = A21, one thread per row, rows dealt over the whole grid ----
    for (int r = j0 + w + gtid; r < n; r += gth) {
      T x[kPb];
#pragma unroll
      for (int cc = 0; cc < kPb; ++cc) x[cc] = cc < w ? A[(int64_t)r * n + j0 + cc] : (T)0;
      SolveStep<T, 0>::run(x, sm.L11, sm.invd);
#pragma unroll
      for (int cc = 0; cc < kPb; ++cc)
        if (cc < w) A[(int64_t)r * n + j0 + cc] = x[cc];
    }
    if (prof) { const long long c = clock64(); t_sol += c - c0; c0 = c; }
    __threadfence();
    grid.sync();
    if (prof) { const long long c = clock64(); t_s1 += c - c0; c0 = c; }
    // ---- trailing update: A22[i, c] -= sum_k L21[i, k] L21[c, k], lower-triangle tiles of 32 x 32 over the grid ----
    const int rem = n - j0 - w;
    if (rem > 0) {
      const int nb = (rem + 31) / 32;
      const int n_tiles = nb * (nb + 1) / 2;
      const int tx = tid & 31, ty = tid >> 5;              // 32 x 8
      for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        int bi = (int)((sqrtf(8.f * t + 1.f) - 1.f) * 0.5f);
        while ((bi + 1) * (bi + 2) / 2 <= t) ++bi;
        while (bi * (bi + 1) / 2 > t) --bi;
        const int bj = t - bi * (bi + 1) / 2;
        const int r0 = j0 + w + bi * 32, c0t = j0 + w + bj * 32;
        __syncthreads();
        for (int rr = ty; rr < 32; rr += 8) {
          sm.Li[rr][tx] = (r0 + rr < n && tx < w) ? A[(int64_t)(r0 + rr) * n + j0 + tx] : (T)0;
          sm.Lj[rr][tx] = (c0t + rr < n && tx < w) ? A[(int64_t)(c0t + rr) * n + j0 + tx] : (T)0;
        }
        __syncthreads();
        for (int rr = ty; rr < 32; rr += 8) {
          const int r = r0 + rr, cc = c0t + tx;
          if (r < n && cc <= r) {
            T s = 0;
#pragma unroll
            for (int k = 0; k < kPb; ++k) s += sm.Li[rr][k] * sm.Lj[tx][k];
            A[(int64_t)r * n + cc] -= s;
          }
        }
      }
      if (prof) { const long long c = clock64(); t_upd += c - c0; c0 = c; }
      __threadfence();
      grid.sync();
      if (prof) { const long long c = clock64(); t_s2 += c - c0; c0 = c; }
    }
  }
  if (prof) { dbg[0] = t_fac; dbg[1] = t_sol; dbg[2] = t_s1; dbg[3] = t_upd; dbg[4] = t_s2; }
  for (int64_t idx = gtid; idx < (int64_t)n * n; idx += gth) {     // strict upper triangle := 0
    const int i = (int)(idx / n), cc = (int)(idx % n);
    if (cc > i) A[idx] = (T)0;
  }
}

// ================================================================================================ Jacobi eigen-solver
struct EighArgs {
  const void* G;  // [n, n] the input matrix (read-only): final eigenvalues are Rayleigh quotients v^T G v
  void* W;        // [n, n] column-major work matrix, starts as G
  void* V;        // [n, n] column-major, starts as I
  void* evals;    // [n] out, ascending
  void* evecs;    // [n, n] out, row-major, column j = eigenvector j (like torch.linalg.eigh)
  int n, b;       // b columns per block, n % (2 b) == 0
  int max_sweeps;
  float tol;
  unsigned* offmax;   // [2] float bits of the sweep's largest normalised off-diagonal (double-buffered on sweep parity)
  int* sweeps_out;
};

// Round-robin tournament over m players (m even): in step s (0 .. m-2) pair k (0 .. m/2-1) is (x, y).
__device__ __forceinline__ void rr_pair(int m, int s, int k, int& x, int& y) {
  const int r = m - 1;
  x = k == 0 ? m - 1 : (s + k) % r;
  y = (s + r - k) % r;
}

template <typename T>
__global__ void __launch_bounds__(512) eigh_jacobi_kernel(EighArgs a) {
  cg::grid_group grid = cg::this_grid();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int n = a.n, b = a.b, cols = 2 * b;
  T* Ws = reinterpret_cast<T*>(smem_raw);                    // [cols][n]
  T* Vs = Ws + (size_t)cols * n;                              // [cols][n]
  T* W = reinterpret_cast<T*>(a.W);
  T* V = reinterpret_cast<T*>(a.V);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nt = blockDim.x;
  const int nblk = n / b;                                     // players of the outer tournament (even)
  const int npairs = nblk / 2;
  const T tol = (T)a.tol;
  int sweep = 0;
  for (; sweep < a.max_sweeps; ++sweep) {
    float my_off = 0.f;
    for (int step = 0; step < nblk - 1; ++step) {
      for (int pr = blockIdx.x; pr < npairs; pr += gridDim.x) {
        int bp, bq;
        rr_pair(nblk, step, pr, bp, bq);
        // ---- load the two column blocks (columns are contiguous in W / V) ----
        __syncthreads();
        for (int i = tid; i < cols * n; i += nt) {
          const int cl = i / n, row = i - cl * n;
          const int gc = (cl < b ? bp * b + cl : bq * b + (cl - b));
          Ws[i] = W[(size_t)gc * n + row];
          Vs[i] = V[(size_t)gc * n + row];
        }
        __syncthreads();
        // ---- inner tournament over the 2b columns: one warp per pair ----
        for (int is = 0; is < cols - 1; ++is) {
          if (warp < b) {
            int p, q;
            rr_pair(cols, is, warp, p, q);
            if (p > q) { const int t = p; p = q; q = t; }
            T* wp = Ws + (size_t)p * n; T* wq = Ws + (size_t)q * n;
            T al = 0, be = 0, ga = 0;
            for (int r = lane; r < n; r += 32) { const T x = wp[r], y = wq[r]; al += x * x; be += y * y; ga += x * y; }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
              al += __shfl_xor_sync(0xffffffffu, al, o);
              be += __shfl_xor_sync(0xffffffffu, be, o);
              ga += __shfl_xor_sync(0xffffffffu, ga, o);
            }
            const T denom = sqrt(al * be);
            const T off = denom > (T)0 ? fabs(ga) / denom : (T)0;
            my_off = fmaxf(my_off, (float)off);
            if (off > tol) {
              const T zeta = (be - al) / ((T)2 * ga);
              const T t = (zeta >= (T)0 ? (T)1 : (T)-1) / (fabs(zeta) + sqrt((T)1 + zeta * zeta));
              const T c = (T)1 / sqrt((T)1 + t * t), s = c * t;
              T* vp = Vs + (size_t)p * n; T* vq = Vs + (size_t)q * n;
              for (int r = lane; r < n; r += 32) {
                const T x = wp[r], y = wq[r];
                wp[r] = c * x - s * y; wq[r] = s * x + c * y;
                const T u = vp[r], v = vq[r];
                vp[r] = c * u - s * v; vq[r] = s * u + c * v;
              }
            }
          }
          __syncthreads();
        }
        // ---- store back ----
        for (int i = tid; i < cols * n; i += nt) {
          const int cl = i / n, row = i - cl * n;
          const int gc = (cl < b ? bp * b + cl : bq * b + (cl - b));
          W[(size_t)gc * n + row] = Ws[i];
          V[(size_t)gc * n + row] = Vs[i];
        }
      }
      __threadfence();
      grid.sync();
    }
    // ---- convergence: largest normalised off-diagonal of this sweep over the whole grid ----
    unsigned* slot = a.offmax + (sweep & 1);
    if (my_off > 0.f) atomicMax(slot, __float_as_uint(my_off));          // non-negative floats order like their bits
    __threadfence();
    grid.sync();
    const float worst = __uint_as_float(*reinterpret_cast<volatile unsigned*>(slot));
    if (blockIdx.x == 0 && tid == 0) a.offmax[(sweep + 1) & 1] = 0u;       // next sweep's slot (nobody reads it before the next barrier)
    if (worst <= a.tol) { ++sweep; break; }
  }
  grid.sync();
  // ---- eigenvalues lambda_j = v_j . w_j, sorted ascending; every CTA sorts redundantly in shared memory ----
  T* lam = reinterpret_cast<T*>(smem_raw);                    // [n]
  int np2_ = 1;
  while (np2_ < n) np2_ <<= 1;
  int* perm = reinterpret_cast<int*>(lam + np2_);             // [np2]
  // lambda_j = v_j^T G v_j with the ORIGINAL G: W has been through every rotation of every sweep and carries their
  // accumulated rounding, the Rayleigh quotient of a (nearly) orthonormal v_j is accurate to eps * ||G||.  Columns are
  // dealt over the grid, results exchanged through the evals buffer.
  {
    const T* G = reinterpret_cast<const T*>(a.G);
    T* ev = reinterpret_cast<T*>(a.evals);
    const int warps_total = gridDim.x * (nt / 32);
    for (int j = blockIdx.x * (nt / 32) + warp; j < n; j += warps_total) {
      const T* vj = V + (size_t)j * n;
      T acc = 0;
      for (int r = 0; r < n; ++r) {                           // y_r = G[r, :] . v_j  (coalesced row reads)
        T d = 0;
        for (int cI = lane; cI < n; cI += 32) d += G[(size_t)r * n + cI] * vj[cI];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
        acc += vj[r] * d;
      }
      if (lane == 0) ev[j] = acc;
    }
  }
  __threadfence();
  grid.sync();
  for (int j = tid; j < n; j += nt) { lam[j] = reinterpret_cast<const T*>(a.evals)[j]; perm[j] = j; }
  __syncthreads();
  int np2 = 1;
  while (np2 < n) np2 <<= 1;
  for (int j = n + tid; j < np2; j += nt) { lam[j] = (T)INFINITY; perm[j] = j; }      // padding sorts to the end
  __syncthreads();
  for (int k = 2; k <= np2; k <<= 1)                          // bitonic sort of (lam, perm), ascending
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < np2; i += nt) {
        const int l = i ^ j;
        if (l > i) {
          const bool up = (i & k) == 0;
          const T x = lam[i], y = lam[l];
          if ((x > y) == up) { lam[i] = y; lam[l] = x; const int t = perm[i]; perm[i] = perm[l]; perm[l] = t; }
        }
      }
      __syncthreads();
    }
  T* evals = reinterpret_cast<T*>(a.evals);
  T* evecs = reinterpret_cast<T*>(a.evecs);
  grid.sync();                                               // everybody has read the unsorted values
  if (blockIdx.x == 0) for (int j = tid; j < n; j += nt) evals[j] = lam[j];
  for (int64_t idx = (int64_t)blockIdx.x * nt + tid; idx < (int64_t)n * n; idx += (int64_t)gridDim.x * nt) {
    const int row = (int)(idx / n), j = (int)(idx % n);
    evecs[idx] = V[(size_t)perm[j] * n + row];
  }
  if (blockIdx.x == 0 && tid == 0 && a.sweeps_out) *a.sweeps_out = sweep;
}

template <typename T>
int eigh_launch(EighArgs a, cudaStream_t st) {
  // block width: the largest b in {16, 8, 4} whose two column blocks of W and V fit in shared memory
  int b = 0;
  static const int b_env = getenv("DAPB_EIGH_B") ? atoi(getenv("DAPB_EIGH_B")) : 0;
  for (int cand : {16, 8, 4}) {
    if (b_env > 0 && cand > b_env) continue;
    if (a.n % (2 * cand) == 0 && (size_t)4 * cand * a.n * sizeof(T) <= 200 * 1024) { b = cand; break; }
  }
  DAPB_REQUIRE(b > 0, "eigh: n = %d is not supported (needs n %% 8 == 0 and n <= %d)", a.n, (int)(200 * 1024 / (16 * sizeof(T))));
  a.b = b;
  size_t smem = (size_t)4 * b * a.n * sizeof(T);
  const size_t tail = (size_t)a.n * (sizeof(T) + sizeof(int)) * 2;          // sort scratch incl. power-of-two padding (< 2 n)
  if (smem < tail) smem = tail;
  DAPB_CUDA_CHECK(cudaFuncSetAttribute(eigh_jacobi_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int grid = a.n / (2 * b);
  if (grid > dapb::sm_count()) grid = dapb::sm_count();
  void* kargs[] = {&a};
  DAPB_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)eigh_jacobi_kernel<T>, dim3(grid), dim3(512), kargs, smem, st));
  return 0;
}

template <typename T>
__global__ void eigh_init_kernel(const T* __restrict__ G, T* __restrict__ W, T* __restrict__ V, int n, unsigned* offmax) {
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < (int64_t)n * n; idx += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(idx / n), j = (int)(idx % n);
    // symmetrised copy: column-major W == row-major G for a symmetric matrix; averaging removes atomics' asymmetry
    W[idx] = (T)0.5 * (G[idx] + G[(int64_t)j * n + i]);
    V[idx] = i == j ? (T)1 : (T)0;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) { offmax[0] = 0u; offmax[1] = 0u; }
}

}  // namespace

extern "C" int dapb_cholesky_dbg(void* A, int32_t n, int32_t f64, int32_t* info, long long* dbg, void* stream);
extern "C" int dapb_cholesky(void* A, int32_t n, int32_t f64, int32_t* info, void* stream) {
  return dapb_cholesky_dbg(A, n, f64, info, nullptr, stream);
}
// dbg: optional device buffer of 5 int64 cycle counters of CTA 0 (diagonal factor, row solve, barrier, update, barrier)
extern "C" __attribute__((visibility("default"))) int dapb_cholesky_dbg(void* A, int32_t n, int32_t f64, int32_t* info, long long* dbg, void* stream) {
  DAPB_REQUIRE(n >= 1 && n <= 4096 && info && A, "cholesky: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  int per_sm = 0;
  const size_t smem = f64 ? sizeof(CholSmem<double>) : sizeof(CholSmem<float>);
  if (f64) {
    DAPB_CUDA_CHECK(cudaFuncSetAttribute(chol_coop_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    DAPB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chol_coop_kernel<double>, 256, smem));
  } else {
    DAPB_CUDA_CHECK(cudaFuncSetAttribute(chol_coop_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    DAPB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chol_coop_kernel<float>, 256, smem));
  }
  const int nb = (n + 31) / 32;
  int grid = nb * (nb + 1) / 2;                                  // tiles of the first trailing update
  int cap = dapb::sm_count() * (per_sm > 2 ? 2 : per_sm);
  { static const int g_env = getenv("DAPB_CHOL_GRID") ? atoi(getenv("DAPB_CHOL_GRID")) : 0; if (g_env > 0 && g_env < cap) cap = g_env; }
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  void* kargs[] = {&A, &n, &info, &dbg};
  if (f64) DAPB_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)chol_coop_kernel<double>, dim3(grid), dim3(256), kargs, smem, st));
  else DAPB_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)chol_coop_kernel<float>, dim3(grid), dim3(256), kargs, smem, st));
  return 0;
}

// Symmetric eigen-decomposition of G [n, n] (row-major, fp32 or fp64; only its symmetric part is used).
// work: 2 * n * n elements + 16 bytes of scratch; evals [n] ascending; evecs [n, n] row-major with eigenvector j in
// column j.  sweeps_out (device int, optional) receives the number of Jacobi sweeps performed.
extern "C" int dapb_eigh(const void* G, int32_t n, int32_t f64, void* work, void* evals, void* evecs, float tol,
                         int32_t max_sweeps, int32_t* sweeps_out, void* stream) {
  DAPB_REQUIRE(G && work && evals && evecs && n >= 8, "eigh: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t esz = f64 ? 8 : 4;
  EighArgs a;
  a.G = G;
  a.W = work;
  a.V = reinterpret_cast<char*>(work) + (size_t)n * n * esz;
  a.offmax = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(work) + 2 * (size_t)n * n * esz);
  a.evals = evals; a.evecs = evecs; a.n = n; a.b = 0;
  a.max_sweeps = max_sweeps > 0 ? max_sweeps : 20;
  a.tol = tol > 0.f ? tol : (f64 ? 1e-13f : 3e-5f);    // above the rounding noise of an n-term dot product
  a.sweeps_out = sweeps_out;
  if (f64) {
    eigh_init_kernel<double><<<64, 256, 0, st>>>((const double*)G, (double*)a.W, (double*)a.V, n, a.offmax);
    DAPB_LAUNCH_CHECK();
    return eigh_launch<double>(a, st);
  }
  eigh_init_kernel<float><<<64, 256, 0, st>>>((const float*)G, (float*)a.W, (float*)a.V, n, a.offmax);
  DAPB_LAUNCH_CHECK();
  return eigh_launch<float>(a, st);
}
