/* nmf.h — C ABI of libnmf_b200.so: a B200 (sm_100a) drop-in for the update_div hot path of
 * recoord/nmf-gpu ("nmf-cuda").  Plain pointers and sizes only; no C++/torch types.
 *
 * Part 1 is the reference's documented API (its README.md:31-54) — same names, argument
 * meaning and error behaviour.  Part 2 is the session API the Python/C++ hosts, the benchmark
 * and the multi-GPU launcher use; it exposes the same iteration with device-resident data.
 * Part 3 are kernel-level entry points used only by the parity tests.
 *
 * Error behaviour (reference: cuda/error-check.hpp:9-17,55-60 and the shape checks at
 * cuda/matrix.cu:130-134,206-210,...): a failing CUDA call or a shape mismatch prints to stderr
 * and exit()s.  Hosts that prefer status codes call nmf_set_error_mode(1): every function of
 * parts 2 and 3 then returns non-zero on failure and nmf_last_error() holds the message
 * (part-1 functions are void, as in the reference; they still exit in mode 0 and set the
 * message / return early in mode 1).
 */
#ifndef NMF_B200_H
#define NMF_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------
 * Part 1 — the reference's host API
 * ------------------------------------------------------------------------------------------ */

/* The README's "matrix struct" (README.md:31-36): a host column-major float array plus its two
 * dims.  The struct itself is not defined in the reference tree (its code moved on to the C++
 * class Matrix of cuda/matrix.cuh:18-39); this is the layout of the upstream nmf-cuda lineage.
 *   mat    host, column-major, dim[0] x dim[1], leading dimension dim[0]
 *   mat_d  optional device mirror owned by the library (NULL unless the library set it)
 *   dim    {rows, cols}                                                                  */
typedef struct {
    float *mat;
    float *mat_d;
    int    dim[2];
} matrix;

/* Replaces `Matrix read_matrix(std::string file, cudaStream_t)`, cuda/nmf.cu:188-218.
 * File: int32 rows, int32 cols, float32[rows*cols] column-major (README.md:33-34).
 * Values < EPS are raised to EPS ("make sure no zero elements", nmf.cu:210-211).
 * Prints "read <file> [<rows_padded>x<cols_padded>]" like nmf.cu:215. A->mat is malloc'ed. */
void read_matrix(matrix *A, const char *file);

/* Replaces `void write_matrix(Matrix*, std::string, cudaStream_t)`, cuda/nmf.cu:220-259.
 * Same format, un-padded; prints "write <file> [<rows>x<cols>]" like nmf.cu:258. */
void write_matrix(matrix A, const char *file);

/* rows x cols matrix filled with `value` (reference: Matrix(float value, rows, cols, stream),
 * cuda/matrix.cu:69-80). */
void create_matrix(matrix *A, int rows, int cols, float value);

/* Frees A->mat (and A->mat_d if set); reference: Matrix::~Matrix, cuda/matrix.cu:82-86. */
void destroy_matrix(matrix *A);

/* README.md:38-54.  Replaces run_async + update_h + update_w, cuda/nmf.cu:76-176.
 *   W (MxK), H (KxN): initial factors, UPDATED IN PLACE (through W.mat / H.mat)
 *   X (MxN)         : target
 *   CONVERGE_THRESH : stop when (previous cost - cost) / cost < CONVERGE_THRESH, tested every
 *                     25 iterations (ITER_CHECK, nmf.cu:9); 0 => always max_iter iterations
 *   max_iter        : maximum number of iterations
 *   t               : NULL, or >= 10 doubles receiving seconds per phase:
 *                     t[0] W*H GEMM with the fused clamp+divide (+cost), t[1] 0 (fused into t[0]),
 *                     t[2] row/column sums, t[3] W'Y / YH' GEMMs with the fused divide+multiply,
 *                     t[4] 0 (fused into t[3]), t[5] W multiply/update kernel, t[6] cost read-back
 *                     and convergence test, t[7] allreduce, t[8] H2D/D2H + allocation + setup,
 *                     t[9] total wall time of the call.  Where a half-iteration runs as ONE fused
 *                     kernel (the default whenever the padded K is a multiple of 64 and <= 256) its
 *                     W*H GEMM, quotient and second GEMM are all charged to t[3] and t[0] stays 0;
 *                     on check iterations t[6] also holds the copy that saves H
 *   verbose         : 1 prints the cost every 25 iterations
 * Semantics (exact | refcompat) and GPU selection come from the environment: NMF_SEMANTICS,
 * NMF_DEVICE, NMF_ENGINE; defaults: exact, device 0, tcgen05 engine.
 * NMF_NUM_GPUS=g (default 1) runs the call on g GPUs of this box from this one process — the GPU
 * count is not part of the reference's signature: a host thread per GPU (devices NMF_DEVICE ..
 * NMF_DEVICE+g-1), X and H sharded by columns, W replicated, one NCCL allreduce per iteration. */
void update_div(matrix W, matrix H, matrix X, float CONVERGE_THRESH, int max_iter, double t[10],
                int verbose);

/* ------------------------------------------------------------------------------------------
 * Part 2 — options, sessions (device-resident data), multi-GPU
 * ------------------------------------------------------------------------------------------ */

enum { NMF_EXACT = 0, NMF_REFCOMPAT = 1 };   /* SURVEY.md §0.4 */
enum { NMF_ENGINE_TCGEN05 = 0,               /* tcgen05/TMEM/TMA 3xTF32 kernels (the product path) */
       NMF_ENGINE_SIMT = 1 };                /* plain fp32 FFMA kernels: bring-up / debugging aid  */

typedef struct {
    int semantics;      /* NMF_EXACT | NMF_REFCOMPAT                                        */
    int engine;         /* NMF_ENGINE_*                                                     */
    int device;         /* CUDA device ordinal; -1 = current device                         */
    int check_every;    /* cost / convergence period; 0 => 25 (ITER_CHECK)                  */
    int use_graph;      /* 1: replay one CUDA graph per iteration (reference: nmf.cu:100-115) */
    int tf32_passes;    /* 3 = 3xTF32 (fp32-equivalent, default); 1 = single-pass TF32      */
    int restarts;       /* sessions: R >= 1 factor pairs sharing X (batched restarts); 0 => 1 */
    int reserved[9];
} nmf_options;

typedef struct {
    int    iterations;        /* iterations executed                                          */
    int    converged;         /* 1 if stopped by CONVERGE_THRESH                              */
    double cost;              /* last evaluated KL divergence (NaN if never evaluated)        */
    double device_ms;         /* cudaEvent time of the iteration loop only                    */
    long long gpu_launches;   /* kernels of this library launched inside the loop             */
} nmf_result;

void nmf_default_options(nmf_options *o);          /* fills defaults, then applies NMF_* env    */
int  nmf_set_error_mode(int mode);                 /* 0 exit (reference), 1 return codes        */
const char *nmf_last_error(void);
const char *nmf_version(void);

/* update_div with explicit options and a result record. */
int nmf_update_div_ex(matrix W, matrix H, matrix X, float CONVERGE_THRESH, int max_iter,
                      double t[10], int verbose, const nmf_options *opt, nmf_result *res);

/* Batched random restarts (the paper's reason for wanting speed: "several random initialisations" per
 * decomposition, nmf_ismir_2009.pdf p.2 section 3.2; the reference runs exactly one, cuda/nmf.cu:42).
 * W[r] (MxK), H[r] (KxN), r = 0..restarts-1: independent initial factor pairs for the SAME X, all updated in
 * place exactly as `restarts` separate update_div calls would (each restart stops on its own convergence
 * check), but in the same kernel launches: the restarts are stacked along k and every fused half-iteration
 * kernel walks the work items of all of them, so small problems fill the GPU and X is shared.
 *   iterations, converged, cost : per-restart arrays of length `restarts` (any may be NULL); cost = KL
 *       divergence at the restart's last check (NaN if never evaluated: CONVERGE_THRESH = 0 and verbose = 0
 *       evaluate it once at the end)
 *   returns the index of the restart with the lowest cost ("keep the best"), or -1 on failure in error mode 1.
 * Shapes the stacked kernels do not cover (K > 256, fewer than 129 rows or columns, refcompat semantics, single-pass
 * TF32) run the restarts one after another through the same CUDA path. */
int nmf_update_div_batch(matrix *W, matrix *H, int restarts, matrix X, float CONVERGE_THRESH, int max_iter,
                         double t[10], int verbose, const nmf_options *opt,
                         int *iterations, int *converged, double *cost);

/* A session owns the device buffers of one rank's shard: W (M x K, replicated), the local column
 * block of H (K x n_local) and X (M x n_local).  n_offset / n_global locate the block in the
 * full problem (needed by the refcompat column mask and the synthetic generator). */
typedef struct nmf_session nmf_session;

nmf_session *nmf_session_create(int M, int n_local, int K, long long n_offset, long long n_global,
                                const nmf_options *opt);
void nmf_session_destroy(nmf_session *s);

/* Host (column-major, unpadded) -> device; values < EPS are clamped as read_matrix does.
 * Any of the three may be NULL to leave that matrix unchanged. */
int nmf_session_upload(nmf_session *s, const float *W, const float *H_local, const float *X_local);
/* Device -> host; either may be NULL. */
int nmf_session_download(nmf_session *s, float *W, float *H_local);
/* Sessions created with opt->restarts = R > 1: the factors of restart r (same layouts as above; X is shared and
 * goes up through nmf_session_upload(s, NULL, NULL, X)). */
int nmf_session_upload_restart(nmf_session *s, int r, const float *W, const float *H_local);
int nmf_session_download_restart(nmf_session *s, int r, float *W, float *H_local);
/* Per-restart outcome of the last nmf_session_run: arrays of R entries, any may be NULL. */
int nmf_session_restart_results(const nmf_session *s, int *iterations, int *converged, double *cost);
/* One half of an iteration on the resident data: which = 0 the H step (cuda/nmf.cu:118-146), 1 the W step (:148-176).
 * Parity tests use it to look at a full-size problem between the two halves. */
int nmf_session_half_step(nmf_session *s, int which);
/* `count` rows of the resident X (count x N, column-major, leading dimension count) and of W (count x K):
 * the W step is row-local, so sampled rows of a full-size run can be checked against the oracle. */
int nmf_session_download_rows(nmf_session *s, const int *rows, int count, float *Xr, float *Wr);
/* `count` columns of the resident X (M x count) and H (K x count) by GLOBAL column index (inside this
 * shard), unpadded, either output may be NULL: lets a full-size run be spot-checked against the oracle
 * on sampled columns (the H step is column-local). */
int nmf_session_download_cols(nmf_session *s, const long long *cols, int count, float *Xc, float *Hc);

/* Synthetic uniform(0,1) data generated on the device from a counter-based hash of the GLOBAL
 * (row, col) index, so any column sharding yields the same problem.  which: bit0 W, bit1 H, bit2 X */
int nmf_session_fill_uniform(nmf_session *s, unsigned long long seed, int which);
/* X <- (A*B) .* (1 + noise*(2u-1)), A (M x rank), B (rank x N), u uniform(0,1) from the same global-index
 * hash: a planted low-rank target on which the convergence test is meaningful (BASELINE config C3). */
int nmf_session_fill_planted(nmf_session *s, unsigned long long seed, int rank, float noise);

/* Join an NCCL communicator (one process per GPU).  id = 128-byte ncclUniqueId obtained from
 * nmf_nccl_unique_id() on rank 0 and broadcast by the host (torch.distributed, MPI, a file...). */
int nmf_nccl_unique_id(void *id128);
int nmf_session_join(nmf_session *s, int rank, int nranks, const void *id128);
/* A further session of the same process reuses src's communicator (ncclCommInitRank costs seconds; a
 * process makes it once).  The communicator lives until the last session that holds it is destroyed. */
int nmf_session_share_comm(nmf_session *dst, nmf_session *src);

/* Runs the iteration (the body of update_div) on the resident data. */
int nmf_session_run(nmf_session *s, float CONVERGE_THRESH, int max_iter, double t[10], int verbose,
                    nmf_result *res);

/* KL divergence of the resident factors (allreduced over ranks when joined); R values for a session with
 * R restarts. */
int nmf_session_cost(nmf_session *s, double *cost);

/* Which half-iterations this session runs as ONE kernel (the quotient never leaves the SM): bit 0 the W
 * step, bit 1 the H step; 0 for the two-kernel path (padded K not a multiple of 64 or > 256, fewer than
 * 129 rows / columns, NMF_UMMA_FUSE=0).  Informational: results are within the same tolerance either
 * way.  A session with both bits set allocates no M x N quotient buffer. */
int nmf_session_fused_steps(const nmf_session *s);

/* 1 when this session's tensor-core kernels run the two correction passes of every 3xTF32 product (lo*hi, hi*lo) as BF16
 * MMAs on bf16(x) / bf16(x - tf32(x)) operands — 8 instead of 12 tensor-pipe time units per 32-deep block, ~2^-20
 * relative error per product.  Default (NMF_CORR_BF16=7): the fused kernels at a padded K of 256 (bit 0)
 * and 128 (bit 1), the two-kernel path at K > 256, a multiple of 128 (bit 2); NMF_CORR_BF16=0: never.
 * Informational, like nmf_session_fused_steps: results stay within the same tolerances. */
int nmf_session_corr_bf16(const nmf_session *s);

/* ------------------------------------------------------------------------------------------
 * Part 3 — kernel-level entry points for the parity tests (host arrays, column-major)
 * ------------------------------------------------------------------------------------------ */

/* Y = X ./ max(W*H, EPS) (reference: matrix_multiply + set_epsilon + element_divide,
 * nmf.cu:125-131); *cost (may be NULL) = KL divergence of (X, W*H). */
int nmf_dbg_quotient(const float *W, const float *H, const float *X, float *Y, double *cost,
                     int M, int N, int K, const nmf_options *opt);
/* One H half-step (nmf.cu:118-146) / one W half-step (nmf.cu:148-176), in place. */
int nmf_dbg_h_step(const float *W, float *H, const float *X, int M, int N, int K, const nmf_options *opt);
int nmf_dbg_w_step(float *W, const float *H, const float *X, int M, int N, int K, const nmf_options *opt);
/* Column sums of W (len K) and row sums of H (len K), clamped to >= EPS
 * (sum_cols/sum_rows + set_epsilon, nmf.cu:134-135,164-165). */
int nmf_dbg_sums(const float *W, const float *H, float *sumW, float *sumH, int M, int N, int K,
                 const nmf_options *opt);
/* Row sums of H with NO allocation — no counterpart in the reference; added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The same two kernels as the session's sum_rows replacement (cuda/matrix.cu:379-503 there), but their per-chunk
 * partial sums live in a device array built into the library (4 MB per GPU that loads it, held until exit) instead of a
 * session buffer.  The call allocates nothing on host-visible CUDA APIs: no cudaMalloc / cudaFree / session.
 *   H_dev    : DEVICE pointer, K x N column-major fp32, leading dimension K, 16-byte aligned; K a multiple of 4
 *   sumH_dev : DEVICE pointer, K floats: sumH[k] = sum_n H[k,n] (not clamped)
 *   stream   : a cudaStream_t (NULL = the default stream); the call only enqueues, it does not synchronise
 * Limit: ceil(N/512) * K <= 1 048 576 (e.g. K = 256: N <= 2 097 152); beyond it the call fails with a message
 * (non-zero return in error mode 1) and launches nothing.  The pointers must belong to the CURRENT device.
 * Every call on a device shares the one array: calls that overlap on the GPU (different streams, no ordering) race.
 * Nothing in the array has to start at zero, so back-to-back calls on one stream need no clearing. */
int nmf_rowsum_noalloc(const float *H_dev, int K, int N, float *sumH_dev, void *stream);

/* Raw tensor-core contraction used by all three GEMMs, for descriptor/layout bring-up:
 * which = 0: C(MxN) = A(MxK)*B(KxN);  1: C(KxN) = A(MxK)'*B(MxN);  2: C(MxK) = A(MxN)*B(KxN)'
 * (shapes named as in the NMF use: A,B,C column-major).  passes = 1 or 3. */
int nmf_dbg_gemm(int which, const float *A, const float *B, float *C, int M, int N, int K,
                 int passes, const nmf_options *opt);

#ifdef __cplusplus
}
#endif
#endif /* NMF_B200_H */
