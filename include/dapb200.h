/* dapb200 -- C ABI of the B200-native (sm_100a) kernels behind distributed_algorithm_pyspark_b200.
 *
 * Every entry point returns 0 on success, a cudaError_t (> 0) or DAPB_ERR_* (< 0) otherwise;
 * dapb_last_error() returns the message of the last failure on the calling thread.
 * Pointers are device pointers unless a comment says "host".  `stream` is a cudaStream_t.
 * Citations "ref:" point at /root/reference (orion-orion/Distributed-Algorithm-PySpark).
 */
#ifndef DAPB200_H_
#define DAPB200_H_

#include <stdint.h>

#if defined(__GNUC__)
#define DAPB_API __attribute__((visibility("default")))
#else
#define DAPB_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define DAPB_MAX_RANKS 8
#define DAPB_FLAG_BYTES 65536ull          /* first 64 KiB of every symmetric heap = uint64 flags */
#define DAPB_WAIT_TIMEOUT_NS 20000000000ull

#define DAPB_ERR_INVALID (-1)
#define DAPB_ERR_UNSUPPORTED (-2)

/* Symmetric-heap view handed to kernels that communicate (csrc/comm.cuh).
 * heap[r] = base of rank r's heap as mapped into THIS process; heap[rank] is local memory.
 * mc_heap = NVLS multicast mapping of the same heap (a store to mc_heap + x lands at heap[r] + x on EVERY rank, a
 * multimem.ld_reduce from it sums the W copies inside the NVSwitch), or NULL when the fabric offers no multicast --
 * every kernel then falls back to W unicast peer stores / loads. */
typedef struct DapbComm {
  int32_t rank;
  int32_t world;
  void* heap[DAPB_MAX_RANKS];
  uint64_t heap_bytes;
  void* mc_heap;
} DapbComm;

/* ------------------------------------------------------------------ runtime ------------ */
DAPB_API const char* dapb_last_error(void);
DAPB_API int dapb_version(void);
DAPB_API int dapb_device_info(int32_t* sm_count, int32_t* cc_major, int32_t* cc_minor, uint64_t* total_mem);
/* Symmetric heap: cudaMalloc'ed (not from a caching allocator, so IPC handles map 1:1), zeroed. */
DAPB_API int dapb_symm_alloc(uint64_t bytes, void** out);
DAPB_API int dapb_symm_free(void* ptr);
DAPB_API int dapb_ipc_get_handle(void* ptr, void* handle64 /* host, 64 bytes */);
DAPB_API int dapb_ipc_open_handle(const void* handle64 /* host */, void** out);
DAPB_API int dapb_ipc_close_handle(void* ptr);
/* F8 self-test: `world` CTAs of ONE cooperative launch act as ranks over `world` heaps carved
 * from `heaps` (each heap_bytes long); runs `iters` all-reduces of n floats and checks them.
 * out_fail (device int32) receives the number of mismatching elements. */
DAPB_API int dapb_comm_selftest(void* heaps, uint64_t heap_bytes, int32_t world, int32_t n, int32_t iters,
                       int32_t* out_fail, void* stream);
/* Cross-GPU stress of the flag protocol: one persistent CTA per rank runs `iters` consecutive one-shot all-reduces of n
 * floats (epochs epoch0 .. epoch0+iters-1) with random in-kernel stalls and checks every result bit for bit;
 * out_fail (device int32, zero it first) receives the number of wrong elements. */
DAPB_API int dapb_allreduce_stress(const DapbComm* comm, uint64_t payload_off, uint32_t flag_base, uint64_t epoch0,
                                   int32_t n, int32_t iters, uint32_t seed, int32_t* out_fail, void* stream);
/* Stand-alone F8 collectives (tests / baselines; production kernels inline them).
 * One CTA per rank; every rank must call with the same epoch, which must grow by EXACTLY 1 per call on a
 * flag_base (payload slots are double-buffered on the epoch's parity, so skipping an epoch would let a fast
 * rank overwrite a slot a slower peer still reads). */
DAPB_API int dapb_allreduce_f32(const DapbComm* comm, uint64_t payload_off, uint32_t flag_base,
                       uint64_t epoch, const float* src, float* dst, int32_t n, void* stream);
DAPB_API int dapb_allreduce_f64(const DapbComm* comm, uint64_t payload_off, uint32_t flag_base,
                       uint64_t epoch, const double* src, double* dst, int32_t n, void* stream);
DAPB_API int dapb_barrier(const DapbComm* comm, uint32_t flag_base, uint64_t epoch, void* stream);
/* Grid-wide stand-alone collectives for large messages (cooperative launches; multimem.* when comm->mc_heap is set,
 * unicast peer loads/stores otherwise).  Each call uses `world` flag words at flag_base and TWO barrier epochs
 * (2*epoch-1, 2*epoch); `epoch` must grow by exactly 1 per call on a flag_base.  Payload at payload_off:
 * all-gather world*n_per_rank elements, two-shot / reduce-scatter n floats (n % (4*world) == 0), all-to-all
 * world*n_per_pair u64.  Replaces: collect / reduceByKey shuffles / join shuffles. */
DAPB_API int dapb_allgather(const DapbComm* comm, uint64_t payload_off, uint32_t flag_base, uint64_t epoch,
                            const void* src, void* dst, int64_t n_per_rank, int32_t elem_bytes, void* stream);
DAPB_API int dapb_allreduce_twoshot_f32(const DapbComm* comm, uint64_t payload_off, uint32_t flag_base, uint64_t epoch,
                                        const float* src, float* dst, int32_t n, void* stream);
DAPB_API int dapb_reduce_scatter_f32(const DapbComm* comm, uint64_t payload_off, uint32_t flag_base, uint64_t epoch,
                                     const float* src, float* dst /* [n / world] */, int32_t n, void* stream);
DAPB_API int dapb_alltoall_u64(const DapbComm* comm, uint64_t payload_off, uint32_t flag_base, uint64_t epoch,
                               const unsigned long long* src, unsigned long long* dst, int64_t n_per_pair, void* stream);

/* ------------------------------------------------------------------ F1/F2 logistic -------
 * ref: optimization/ssgd.py:23-33,93-112; machine_learning/logistic_regression.py:76-92;
 *      optimization/ma.py:93-113; optimization/bmuf.py:98-121; optimization/easgd.py:95-113 */
enum { DAPB_LR_SSGD = 0, DAPB_LR_GD = 1, DAPB_LR_MA = 2, DAPB_LR_BMUF = 3, DAPB_LR_EASGD = 4 };
enum { DAPB_REG_NONE = 0, DAPB_REG_L2 = 1, DAPB_REG_L1 = 2, DAPB_REG_ELASTIC = 3 };

/* Whole training loop of the reference's default (small, fp64) configuration in ONE persistent
 * cooperative kernel: one CTA per logical partition, per-iteration exchange through the
 * symmetric heap (in-kernel all-reduce, partition-ordered sum), test accuracy on device. */
typedef struct DapbLrParityArgs {
  const double* X;          /* [n_local, D] rows of this rank's partitions, bias column included */
  const double* y;          /* [n_local] */
  const int64_t* part_off;  /* [P_local + 1] local row offsets */
  const uint8_t* masks;     /* [T, n_local] keep flags (ref ssgd.py:97 sampler), NULL = all rows */
  const double* X_test;     /* [n_test, D] */
  const int32_t* y_test;    /* [n_test] */
  int64_t n_local;
  int32_t n_test, D, P_local, P_global, part_base, T;
  int32_t algo, n_local_iters, reg_type, max_part_rows;
  double eta, lam, reg_alpha, mu, zeta, alpha, beta, eps;
  double* w;                /* [D] in/out (replicated) */
  double* ws;               /* [P_local, D] in/out replicas (EASGD persists them, ref easgd.py:88) */
  double* delta_w;          /* [D] in/out (BMUF, ref bmuf.py:95) */
  double* accs;             /* [T] out, written by the rank owning partition 0 */
  int32_t* batch_sizes;     /* [T] out */
  double* w_hist;           /* [T, D] out or NULL */
  uint64_t payload_off;     /* needs 2*P_global*(D+1) doubles */
  uint32_t flag_base;       /* needs P_global flags */
  uint64_t epoch_base;      /* flags must be < epoch_base+1 on entry; kernel uses epoch_base+t+1 */
} DapbLrParityArgs;
DAPB_API int dapb_lr_parity_run(const DapbLrParityArgs* args, const DapbComm* comm, void* stream);
DAPB_API uint64_t dapb_lr_parity_payload_bytes(int32_t P_global, int32_t D);

/* One optimiser step at benchmark scale, ONE launch: Bernoulli row sample (Philox or explicit
 * mask) -> forward z = Xw -> r = sigmoid(z) - y -> backward g = X^T r -> deterministic grid
 * reduction -> [last 8 CTAs, one column slice each] in-kernel all-reduce over NVLink peers -> optimiser update. */
enum { DAPB_LR_FIN_GRAD_ONLY = 0,   /* g_out[D], cnt_out only                                  */
       DAPB_LR_FIN_SSGD = 1,        /* all-reduce(g,cnt); w -= eta*(g/cnt + lam*reg(w))  (or sum: GD) */
       DAPB_LR_FIN_LOCAL = 2,       /* w_p -= eta*g/cnt + alpha*(w_p - centre)                 */
       DAPB_LR_FIN_LOCAL_SYNC = 3   /* LOCAL, then all-reduce(w_p) + MA / BMUF / EASGD rule    */ };
typedef struct DapbLrStepArgs {
  const void* X;            /* [n_rows, D] row-major, dtype 0 = fp32, 1 = bf16 */
  const float* y;           /* [n_rows] */
  int64_t n_rows;
  int64_t row_base;         /* global index of row 0 (Philox counter) */
  int32_t D, dtype;
  const uint8_t* mask;      /* optional explicit keep flags [n_rows]; NULL -> Philox */
  uint64_t seed;
  uint32_t iteration;
  uint32_t threshold24;     /* keep iff (philox>>8) < threshold24; >= 1<<24 keeps every row */
  float eps;                /* sigmoid variant (ref ma.py:26) */
  const float* w_fwd;       /* [D] weights used in the forward pass (replica or global) */
  /* scratch (caller-allocated, reusable): */
  float* partials;          /* [grid, D] */
  int32_t* counts;          /* [grid] */
  uint32_t* ticket;         /* 1 word, zero before first use */
  int32_t grid;             /* CTAs to launch (0 = 2 per SM) */
  /* finalize */
  int32_t fin, algo, reg_type, use_sum;   /* use_sum=1: GD (ref logistic_regression.py:85) */
  float eta, lam, reg_alpha, mu, zeta, alpha, beta;
  int32_t n_replicas_total; /* P (= n_slices) for the model average */
  float* g_out;             /* [D]  (GRAD_ONLY; also receives the reduced gradient otherwise) */
  int32_t* cnt_out;         /* [1]  global sampled count of this step */
  float* w;                 /* [D] in/out global / centre weights */
  float* w_local;           /* [D] in/out replica (LOCAL / LOCAL_SYNC) */
  float* delta_w;           /* [D] in/out (BMUF) */
  int32_t reset_replica;    /* LOCAL_SYNC: copy new w into w_local (MA/BMUF, ref ma.py:96) */
  uint64_t payload_off;     /* needs dapb_lr_step_payload_bytes(world, D) */
  uint32_t flag_base;       /* needs DAPB_LR_STEP_FLAGS flags (8 column slices x up to 8 ranks) */
  uint64_t epoch;           /* +1 per call on this flag_base (SSGD / LOCAL_SYNC calls) */
  unsigned long long* cnt_accum;  /* optional: += the step's batch size (global for SSGD, local otherwise) */
  unsigned long long* stats;      /* optional [4] %globaltimer ns of slice 0: finish entered, partials reduced,
                                     all-reduce done, update done */
} DapbLrStepArgs;
#define DAPB_LR_STEP_FLAGS 64
DAPB_API uint64_t dapb_lr_step_payload_bytes(int32_t world, int32_t D);
/* Alignment: X, w_fwd and partials 16-byte aligned select the vectorised kernels (D in {128,256,512,1024}); any other
 * alignment (4-byte minimum) or D falls back to the generic kernel -- never a fault.
 * Launched with programmatic stream serialization: back-to-back step kernels on one stream overlap the next step's
 * sampling / L2 prefetch with this step's cross-GPU exchange (the kernel waits before it reads w). */
DAPB_API int dapb_lr_step(const DapbLrStepArgs* args, const DapbComm* comm, void* stream);
/* Tensor-core variant (tcgen05.mma forward K-major + backward MN-major, TMEM accumulators, hi/lo bf16 split of w
 * and of the residual): bf16 rows, D == 1024, grid defaults to one CTA per SM (partials must hold `grid` rows). */
DAPB_API int dapb_lr_step_tc(const DapbLrStepArgs* args, const DapbComm* comm, void* stream);
/* out[i] = 1 iff row row_base+i is in the mini-batch of `iteration` (same Philox rule as the step kernels).  Serves the
 * PyTorch+NCCL baseline, which must train on the identical batches (ref ssgd.py:97 `sample(False, f, 42 + t)`). */
DAPB_API int dapb_philox_mask(int64_t row_base, int64_t n_rows, uint64_t seed, uint32_t iteration, uint32_t threshold24,
                              uint8_t* out, void* stream);
/* Test-set accuracy with the reference's 0.5 threshold (ref ssgd.py:107-110). acc_out: [1] float. */
DAPB_API int dapb_lr_eval(const void* X, const float* y, int64_t n_rows, int32_t D, int32_t dtype,
                 const float* w, float eps, float* acc_out, uint32_t* scratch2, void* stream);

/* Mini-batch gradient WITHOUT any caller-provided scratch and without a single allocation inside.
 * (No counterpart in the reference -- it exists for hosts that cannot allocate device memory: a process
 * whose own allocator already owns the whole GPU, a latency-critical loop that must avoid the device-wide
 * synchronisation of cudaMalloc/cudaFree, or a plug-in hook that may not call allocating CUDA APIs.)
 * Same math and same kernels as dapb_lr_step(fin = DAPB_LR_FIN_GRAD_ONLY): g_out[D] receives
 * sum over kept rows of (sigmoid(x.w) - y) x, cnt_out[1] the number of kept rows.  The grid-reduction
 * scratch is a built-in device array of dapb_lr_grad_noalloc_scratch_bytes() (~1.16 MiB) that every GPU
 * loading this library holds until exit.  Limits: 1 <= D <= DAPB_NOALLOC_MAX_D, otherwise DAPB_ERR_INVALID
 * with a message.  The first call on each device looks the array's address up (cudaGetSymbolAddress);
 * make that first call outside of any stream capture.  All calls share the array: calls that may overlap
 * on the GPU (other streams / threads) race -- serialise them.
 * When the memory is actually taken: with CUDA's lazy module loading (the default) the array is backed when this
 * library's logistic module is first used on a device -- the first call of this function or of any dapb_lr_* entry
 * point -- not when the .so is loaded.  A host whose allocator is about to claim the rest of the GPU must therefore
 * make that first call BEFORE it does so (or run with CUDA_MODULE_LOADING=EAGER); afterwards no call allocates.
 * The scratch address is looked up for the CURRENT device: `stream`, X, y, w and the outputs must belong to it.
 * X, w 16-byte aligned select the vectorised kernels; other alignments run the generic kernel (never a fault). */
#define DAPB_NOALLOC_MAX_D 1024
#define DAPB_NOALLOC_MAX_GRID 296
DAPB_API int dapb_lr_grad_noalloc(const void* X, const float* y, int64_t n_rows, int64_t row_base, int32_t D,
                                  int32_t dtype, const uint8_t* mask, uint64_t seed, uint32_t iteration,
                                  uint32_t threshold24, float eps, const float* w, float* g_out,
                                  int32_t* cnt_out, void* stream);
DAPB_API uint64_t dapb_lr_grad_noalloc_scratch_bytes(void);

/* ------------------------------------------------------------------ F7 Monte-Carlo pi ----
 * ref: randomized_algorithm/monte_carlo.py:17-31.  Sample i in [first, first+n) comes from
 * Philox counter i/2; a point is accepted iff x^2+y^2 <= 1 evaluated exactly in integers.
 * count_io (device uint64) is atomically incremented. */
DAPB_API int dapb_mc_pi(uint64_t first, uint64_t n, uint64_t seed, unsigned long long* count_io, void* stream);
/* Same, fused with the cross-GPU total (ref :28 reduce(add)): the last CTA pushes this rank's count to every peer
 * (multimem.st / peer stores), waits for the others and writes the sum over ranks to total_out (device u64).
 * Needs 2*world u64 of payload, `world` flags, a zeroed ticket word; epoch +1 per call. */
DAPB_API int dapb_mc_pi_allreduce(const DapbComm* comm, uint64_t first, uint64_t n, uint64_t seed,
                                  unsigned long long* count_io, unsigned long long* total_out, uint32_t* ticket,
                                  uint64_t payload_off, uint32_t flag_base, uint64_t epoch, void* stream);

/* ------------------------------------------------------------------ F3 k-means ------------
 * ref: machine_learning/k-means.py:20-28,57-71.  dtype: 0 fp32, 1 bf16, 2 fp64 points; centres and
 * sums are float (dtype 0/1) or double (dtype 2).  Ties -> lowest index; sums/counts accumulate
 * atomically (zero them first); assign may be NULL. */
DAPB_API int dapb_kmeans_assign(const void* X, int64_t n, int32_t d, int32_t dtype, const void* centers,
                                int32_t k, int32_t* assign, void* sums, int32_t* counts, void* stream);
/* Fused reduce-scatter(sum,count) -> mean (empty cluster keeps its centre, ref :70-71) -> all-gather. */
DAPB_API int dapb_kmeans_finalize(const DapbComm* comm, void* sums, int32_t* counts, void* centers, int32_t k,
                                  int32_t d, int32_t f64, uint64_t payload_off, uint32_t flag_base,
                                  uint64_t epoch, int32_t* counts_out, int32_t clear_partials, void* stream);
DAPB_API uint64_t dapb_kmeans_payload_bytes(int32_t world, int32_t k, int32_t d, int32_t elem_bytes);
/* Same for fp32 centres, and the finished centres are also emitted as the operands of dapb_kmeans_assign_tc (b_operand,
 * cnorm: see dapb_kmeans_tc_prepare) -- a Lloyd iteration on the tensor-core path is then two launches: assign, finalize. */
DAPB_API int dapb_kmeans_finalize_tc(const DapbComm* comm, void* sums, int32_t* counts, void* centers, int32_t k,
                                     int32_t d, uint64_t payload_off, uint32_t flag_base, uint64_t epoch,
                                     int32_t* counts_out, int32_t clear_partials, void* b_operand, float* cnorm,
                                     float xnorm_max, void* stream);
/* Tensor-core assignment (tcgen05.mma + TMEM + TMA), bf16 points, d in {64,128}, 1 <= k <= 16384; b_operand and cnorm
 * are sized for k_pad = k rounded up to a multiple of 128 ([k_pad, d] bf16 and 9 * k_pad + 8 floats: the biased norms,
 * 8 aux words and the K = 16 bf16 norm slabs that feed the norm to the tensor core, csrc/kmeans_common.cuh).
 * dapb_kmeans_tc_prepare turns fp32 centres into the bf16 operand (-2c) [k,d] and the BIASED norms
 * ||bf16(c)||^2 + C1 in cnorm[0..k) (C1 makes all distances share one exponent so that the
 * arg-min runs on integer keys, see csrc/kmeans_tc.cu).  xnorm_max: an upper bound on ||x||^2 over all rows. */
DAPB_API int dapb_kmeans_tc_prepare(const float* centers, int32_t k, int32_t d, void* b_operand, float* cnorm,
                                    float xnorm_max, void* stream);
DAPB_API int dapb_kmeans_assign_tc(const void* X, int64_t n, int32_t d, const void* b_operand, const float* cnorm,
                                   int32_t k, int32_t* assign, float* sums, int32_t* counts, int32_t grid,
                                   void* stream);
/* fp32 points on the same tensor-core kernel: X_bf16 = bf16 copy of X32 (feeds the MMAs), sums come from the fp32 rows. */
DAPB_API int dapb_kmeans_assign_tc_f32(const void* X_bf16, const float* X32, int64_t n, int32_t d, const void* b_operand,
                                       const float* cnorm, int32_t k, int32_t* assign, float* sums, int32_t* counts,
                                       int32_t grid, void* stream);
DAPB_API int dapb_kmeans_assign_tc_dbg(const void* X, int64_t n, int32_t d, const void* b_operand,
                                       const float* cnorm, int32_t k, int32_t* assign, float* sums,
                                       int32_t* counts, int32_t grid, long long* dbg, void* stream);

/* ------------------------------------------------------------------ F4 PageRank -----------
 * ref: graph_computation/pagerank.py:21-25,44-57.  Contribution vector [2][V] lives in the
 * symmetric heap at c_off; iteration `epoch` reads parity (epoch-1)&1 and pushes into parity epoch&1
 * on every rank.  Host must barrier between dapb_pagerank_init and the first step (wait_prev = 0).
 * skip_empty = 1 (allowed from the 3rd step after init) skips rows without in-edges: both buffers already hold 0.
 * dapb_pagerank_step is the simple CTA-per-row-block kernel (one launch per iteration, float only). */
DAPB_API int dapb_pagerank_init(const DapbComm* comm, const void* inv_outdeg, int64_t nv_local, int64_t v0,
                                int64_t V, double r0, uint64_t c_off, int32_t f64, void* stream);
DAPB_API int dapb_pagerank_step(const DapbComm* comm, const int64_t* row_ptr, const int32_t* col_idx,
                                const float* inv_outdeg, const int32_t* blocks, int32_t n_blocks,
                                int64_t nv_local, int64_t v0, int64_t V, float base,
                                float damp, float* ranks_local, uint64_t c_off, uint32_t flag_base,
                                uint64_t epoch, uint32_t* ticket, int32_t wait_prev, int32_t skip_empty,
                                int32_t grid, void* stream);

/* Persistent warp-granular PageRank: n_iters iterations in ONE cooperative launch; between iterations the ranks
 * synchronise through flag words in the heap, contributions travel by multimem.st (or unicast peer stores).
 * One warp per row block of <= 512 (non-zeros + rows); rows longer than the host's hub threshold are removed from
 * the main CSR (length 0 there) and described by DapbPrHub: their non-zeros live in hub_col, cut into chunks
 * [chunk_start, chunk_end) that belong to hub rows hub_chunk_ptr[i] .. hub_chunk_ptr[i+1].
 * Values (inv_outdeg, ranks_local, chunk_partial, the [2][V] contribution vector at c_off) are float, or double when
 * f64 = 1 (parity with the 17-digit ranks the reference prints, ref pagerank.py:60-68).  With f64 = 1 `inv_outdeg`
 * holds the out-degree ITSELF (the kernel divides, like the reference) and no multiply-add is contracted. */
typedef struct DapbPrHub {
  const int32_t* hub_rows;      /* [n_hub] local row ids, ascending */
  const int32_t* hub_col;       /* non-zeros (source ids) of all hub rows, row after row */
  const int64_t* chunk_start;   /* [n_chunks] offsets into hub_col */
  const int64_t* chunk_end;     /* [n_chunks] */
  const int32_t* hub_chunk_ptr; /* [n_hub + 1] */
  void* chunk_partial;          /* [n_chunks] scratch (value type) */
  int32_t n_hub, n_chunks;
  const int32_t* chunk_row;     /* [n_chunks] hub row index of each chunk */
  uint32_t* row_done;           /* [n_hub] scratch, zero before first use */
} DapbPrHub;
typedef struct DapbPrRunArgs {
  const int64_t* row_ptr; const int32_t* col_idx; const void* inv_outdeg; const int32_t* blocks;
  int32_t n_blocks, f64;
  int64_t nv_local, v0, V;
  double base, damp;            /* q/n and 1-q */
  void* ranks_local;            /* [nv_local] out */
  uint64_t c_off;               /* payload offset of the contribution vector */
  uint32_t flag_base;           /* `world` flags */
  uint64_t epoch0;              /* epoch of the first iteration (>= 1; grows by 1 per iteration, also across launches) */
  int32_t n_iters, it0;         /* it0: iterations done since dapb_pagerank_init */
  int32_t wait_first;           /* 0 right after init + host barrier, else 1 */
  int32_t skip_empty_ok;        /* 1: rows without in-edges are skipped from the third iteration after init */
  int32_t grid;                 /* 0 = as many CTAs as can be co-resident */
  uint32_t* sync;               /* [4] scratch */
  unsigned long long* stats;    /* optional [n_iters][4] %globaltimer stamps (enter, wait passed, last ticket, published) */
  DapbPrHub hub;
} DapbPrRunArgs;
DAPB_API int dapb_pagerank_run(const DapbPrRunArgs* args, const DapbComm* comm, void* stream);

/* ------------------------------------------------------------------ F6 transitive closure --
 * ref: graph_computation/transitive_closure.py:27-40.  Pairs are packed (x << 32 | z); the hash set
 * is open-addressing with 0xFFFF...F as the empty marker (memset 0xFF) and a power-of-two capacity
 * the caller keeps at load factor <= 0.5.  counters: [0] next-frontier size, [1 + r] per-owner tails. */
DAPB_API int dapb_tc_seed(const unsigned long long* keys, int64_t n, unsigned long long* table, uint64_t capacity,
                          unsigned long long* out, unsigned long long* out_count, int32_t world, int32_t rank,
                          void* stream);
DAPB_API int dapb_tc_expand(const DapbComm* comm, const int64_t* in_ptr, const int32_t* in_src,
                            const unsigned long long* frontier, int64_t n_frontier, unsigned long long* table,
                            uint64_t capacity, unsigned long long* next_frontier, unsigned long long* counters,
                            uint64_t q_off, uint64_t q_cap, void* stream);
DAPB_API int dapb_tc_exchange(const DapbComm* comm, unsigned long long* counters, uint64_t q_off, uint64_t q_cap,
                              uint32_t flag_base, uint64_t epoch, void* stream);
DAPB_API int dapb_tc_drain(const DapbComm* comm, unsigned long long* table, uint64_t capacity,
                           unsigned long long* next_frontier, unsigned long long* counters, uint64_t q_off,
                           uint64_t q_cap, void* stream);

/* Persistent variant: every round of the closure in ONE cooperative launch (expand -> in-kernel all-gather of the
 * queue tails / termination test / capacity check -> drain).  See csrc/transitive_closure.cu for the state words. */
#define DAPB_TC_STATE_WORDS 32
#define DAPB_TC_RUNNING 0
#define DAPB_TC_DONE 1
#define DAPB_TC_GROW 2
DAPB_API int dapb_tc_closure(const DapbComm* comm, const int64_t* in_ptr, const int32_t* in_src,
                             unsigned long long* table, uint64_t capacity, unsigned long long* frontier0,
                             unsigned long long* frontier1, uint64_t front_cap, unsigned long long* state,
                             uint64_t q_off, uint64_t q_cap, uint64_t msg_off, uint32_t flag_base,
                             uint64_t max_rounds, void* stream);


/* ------------------------------------------------------------------ F5 Gram / ALS ----------
 * ref: matrix_computation/matrix_decomposition.py:19-33,51-67.  dtype: 0 fp32, 1 bf16, 2 fp64. */
DAPB_API int dapb_gram(const DapbComm* comm, const void* X, const void* Y, int64_t n, int32_t P, int32_t Q,
                       int64_t ldx, int64_t ldy, int32_t dtype, void* partial, int32_t splits, void* C,
                       uint32_t* ticket, uint64_t payload_off, uint32_t flag_base, uint64_t epoch,
                       int32_t do_allreduce, void* stream);
/* Tensor-core Gram (tcgen05.mma, MN-major operands straight from one TMA tile): G = A^T A, A [n_rows, P] bf16,
 * P in {128, 256, 512}.  Only the blocks on and above the diagonal are computed (SYRK schedule) and mirrored by the
 * epilogue; row splits are reduced with atomics, ranks with a grid-wide in-kernel two-shot all-reduce (multimem when the
 * heap is multicast-mapped) that needs dapb_gram_tc_payload_bytes of heap, `world` flags and P*P % (4*world) == 0.
 * `tickets`: DAPB_GRAM_TC_SCRATCH_WORDS words of scratch.  epoch +1 per call. */
#define DAPB_GRAM_TC_SCRATCH_WORDS 1024
DAPB_API int dapb_gram_tc(const DapbComm* comm, const void* A, int64_t n_rows, int32_t P, float* G, uint32_t* tickets,
                          int32_t splits, uint64_t payload_off, uint32_t flag_base, uint64_t epoch,
                          int32_t do_allreduce, void* stream);
DAPB_API uint64_t dapb_gram_tc_payload_bytes(int32_t world, int32_t P);
DAPB_API int dapb_rows_times_small(const void* R, const void* V, int64_t m, int32_t n, int32_t k, int64_t ldr,
                                   int32_t transpose_r, int32_t f64, void* out, void* stream);
DAPB_API int dapb_solve_rows(const void* G, double ridge, const void* B, int64_t m, int32_t k, int32_t f64,
                             void* X, void* stream);
DAPB_API int dapb_sq_residual(const void* R, const void* U, const void* V, int64_t m, int32_t n, int32_t k,
                              int32_t f64, double* out, void* stream);
/* In-place lower Cholesky (strict upper triangle zeroed), ONE cooperative launch (csrc/dense_small.cu); info (device int,
 * zero it first) = 1 + index of the first non-positive pivot, or 0. */
DAPB_API int dapb_cholesky(void* A, int32_t n, int32_t f64, int32_t* info, void* stream);
/* Symmetric eigen-decomposition by block one-sided Jacobi, one cooperative launch after a small init kernel; no cuSOLVER.
 * G [n, n] fp32 / fp64 (its symmetric part is used), n % 8 == 0 and n <= 800 (fp32) / 400 (fp64) for the widest block.
 * work: 2*n*n elements + 16 bytes; evals [n] ascending; evecs [n, n] row-major, eigenvector j in column j.
 * tol <= 0 and max_sweeps <= 0 select defaults; sweeps_out (device int32) is optional. */
DAPB_API int dapb_eigh(const void* G, int32_t n, int32_t f64, void* work, void* evals, void* evecs, float tol,
                       int32_t max_sweeps, int32_t* sweeps_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DAPB200_H_ */
