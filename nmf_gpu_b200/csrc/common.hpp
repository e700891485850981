// common.hpp — shared host-side declarations of libnmf_b200.so (error convention, device layout,
// kernel launch prototypes).  Product code: must never include or link anything under oracle/.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>

#include "../../include/nmf.h"

namespace nmf {

// (float)2.2204E-16 — reference cuda/matrix.cu:10 (fp32 bits 0x257ffea4, not 2^-52).
#define NMF_EPS_F ((float)(2.2204E-16))

constexpr int kCheckEveryDefault = 25;   // ITER_CHECK, reference cuda/nmf.cu:9

struct Error : std::runtime_error {
    int code;
    Error(const std::string& m, int c = 1) : std::runtime_error(m), code(c) {}
};

[[noreturn]] void fail(const char* file, int line, const std::string& msg, int code = 1);

#define NMF_FAIL(msg) ::nmf::fail(__FILE__, __LINE__, (msg))
#define NMF_CHECK(cond, msg) do { if (!(cond)) ::nmf::fail(__FILE__, __LINE__, (msg)); } while (0)
// Reference convention (cuda/error-check.hpp): print "<name> in <file>:<line>" and exit(code).
// Unlike the reference, launches are polled too (it never calls cudaGetLastError).
#define CUDA_CHECK(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) \
    ::nmf::fail(__FILE__, __LINE__, std::string(#expr) + ": " + cudaGetErrorName(e__) + " - " + \
                cudaGetErrorString(e__), (int)e__); } while (0)
#define CUDA_CHECK_LAUNCH() CUDA_CHECK(cudaGetLastError())

// Programmatic dependent launch (PDL): the kernels of an iteration are launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, so a kernel's CTAs are scheduled as the SMs of its predecessor
// drain and its prologue (barrier init, TMEM allocation, tensor-map prefetch) overlaps the predecessor's tail; every kernel
// executes griddepcontrol.wait before it touches global memory.  Thread-local switch: the session driver turns it on
// around its launches (NMF_PDL=0 turns it off), everything else launches plainly.
void set_launch_pdl(bool on);
bool launch_pdl();

// cudaLaunchKernelEx with the PDL attribute when the switch is on
template <typename... KArgs, typename... Args>
void launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    if (launch_pdl()) {
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
    }
    cudaError_t e__ = cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
    if (e__ != cudaSuccess) ::nmf::fail(__FILE__, __LINE__, std::string("cudaLaunchKernelEx: ") + cudaGetErrorName(e__) + " - " + cudaGetErrorString(e__), (int)e__);
    e__ = cudaGetLastError();
    if (e__ != cudaSuccess) ::nmf::fail(__FILE__, __LINE__, std::string("kernel launch: ") + cudaGetErrorName(e__) + " - " + cudaGetErrorString(e__), (int)e__);
}
#ifdef __CUDACC__
// first statements of every iteration kernel: wait for the predecessor grid (no-op without the launch attribute), then
// let the successor's CTAs be scheduled as this grid's SMs drain
__device__ __forceinline__ void pdl_sync() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
#endif

inline int round_up(int v, int m) { return (v + m - 1) / m * m; }
inline long long round_up_ll(long long v, long long m) { return (v + m - 1) / m * m; }

// Reference padding rule, cuda/matrix.cu:88-95 (dims equal to 1 are not padded).
inline int ref_pad32(int v) { return (v != 1 && v % 32 != 0) ? v + (32 - v % 32) : v; }

// Extents of one rank's shard as the kernels see them.
//   logical   : M, N (local columns), K
//   "valid"   : Mv, Nv, Kv — the index ranges the kernels iterate over.
//       exact     : Mv = round_up(M,32), Kv = round_up(K,32) with ZERO fill (rows of zeros are
//                   fixed points of the update and contribute nothing), Nv = N.
//       refcompat : the reference's padded dims (cuda/matrix.cu:88-95) with EPS fill, because its
//                   clamp runs over the whole padded buffer (cuda/nmf.cu:211).
//   leading dimensions: ldM = Mv (W, X, Y, P), ldK = Kv (H) — multiples of 32 floats (128 B).
//   restarts  : R >= 1 factor pairs that share X (batched random initialisations), stacked along k: W is
//               [Mv x R*Kr], H is [R*Kr x Nv]; Kr = the padded K of ONE restart and Kv = R*Kr is what every
//               elementwise / per-k kernel iterates over (restart r owns k in [r*Kr, r*Kr + K)).
struct Dims {
    int M = 0, N = 0, K = 0;
    int Mv = 0, Nv = 0, Kv = 0;
    int R = 1, Kr = 0;       // restarts and the padded K of one of them (Kv = R * Kr)
    int Nva = 0;             // columns ALLOCATED for X, Y, H (Nv rounded up to 768 = lcm of the GEMM tile widths, zero fill):
                             // lets the fused epilogues store whole tiles without per-element column predicates
    long long n_offset = 0, n_global = 0;
    int refcompat = 0;
    int w_divide = 1;        // 0 iff refcompat and rows_padded(W) > 1024 (cuda/matrix.cu:212-217)
    size_t szW() const { return (size_t)Mv * Kv; }
    size_t szH() const { return (size_t)Kv * Nv; }
    size_t szX() const { return (size_t)Mv * Nv; }
    size_t szHa() const { return (size_t)Kv * Nva; }
    size_t szXa() const { return (size_t)Mv * Nva; }
};

Dims make_dims(int M, int n_local, int K, long long n_offset, long long n_global, int semantics,
               int restarts = 1, bool pad_k64 = false);

// ---------------------------------------------------------------------------------------------
// Small kernels (kernels_small.cu).  All take the stream explicitly; none synchronises.
// ---------------------------------------------------------------------------------------------
// a[i] = max(a[i], EPS) over rows x cols of a column-major buffer with leading dimension ld.
void launch_clamp_eps(float* a, int rows, int cols, int ld, cudaStream_t st);
// Fill rows [r0,r1) x cols [c0,c1) with value (used for the refcompat EPS padding).
void launch_fill_region(float* a, int ld, int r0, int r1, int c0, int c1, float value, cudaStream_t st);
// uniform(0,1) from a hash of (seed, tag, global row, global col); rows x cols, then clampE.
void launch_fill_planted(float* x, int rows, int cols, int ld, long long col_offset, unsigned long long seed,
                         int rank, float noise, cudaStream_t st);
void launch_fill_uniform(float* a, int rows, int cols, int ld, long long col_offset,
                         unsigned long long seed, int tag, cudaStream_t st);
// hi = x with the low 13 mantissa bits cleared (exactly representable in TF32), lo = x - hi (exact).
// bf / lb (both or neither): also the BF16 forms of the correction operands, bf16(x) and bf16(x - hi)
void launch_split_tf32(const float* x, float* hi, float* lo, size_t n, cudaStream_t st,
                       unsigned short* bf = nullptr, unsigned short* lb = nullptr);

// sumW[k] = sum_m W[m,k]  (R1; reference: Matrix::sum_cols + reduce2d, cuda/matrix.cu:261-377,642-687)
// refcompat: only rows m with (m mod 64) in {0,1,2,4,8,16,32} (the reference's stale-read tail).
// Result is NOT clamped; consumers apply max(.,EPS) (reference: sumW->set_epsilon, nmf.cu:135).
void launch_colsum(const float* W, float* sumW, const Dims& d, cudaStream_t st);
// sumH[k] = sum_n H[k,n]  (R2; reference: Matrix::sum_rows + reduce2d_strided, matrix.cu:379-503,689-735)
// Two passes: per-column-block partials, then an ordered finish (deterministic, no float atomics).
// `partials` must hold rowsum_partial_count(d) * d.Kv floats.
int  rowsum_partial_count(const Dims& d);
void launch_rowsum(const float* H, float* partials, float* sumH, const Dims& d, cudaStream_t st);

// sumH_dev[k] = sum_n H_dev[k + n*K] with the partial sums in a built-in device array: no allocation of any kind
// (nmf_rowsum_noalloc).  Limit: ceil(N/512) * K <= rowsum_noalloc_max_floats().
size_t rowsum_noalloc_max_floats();
void rowsum_noalloc(const float* H_dev, int K, int N, float* sumH_dev, cudaStream_t st);

// W[m,k] <- W[m,k] * ((sum_s P[s][m,k]) / max(sumH[k],EPS))   (reference: row_divide + element_multiply,
// cuda/nmf.cu:172-175).  The divide is skipped when !d.w_divide.  Optionally writes the TF32 split.
// Small problems (the launch count matters more than the bytes): one block per column k, which can also
//   fold the `rs_blocks` partial row sums of the new H that the H update left (rs_partials != nullptr) into sumH[k] first, and
//   emit sumW[k] = the column sum of the NEW W (what the next H step divides by; refcompat: the subsample) when sumW_out != nullptr
// — then neither the row-sum finish nor the column-sum kernel runs.  Returns true when sumW_out was written.
bool launch_update_w(float* W, float* Whi, float* Wlo, const float* P, int nsplit, size_t split_stride,
                     float* sumH, const Dims& d, cudaStream_t st,
                     const float* rs_partials = nullptr, int rs_blocks = 0, float* sumW_out = nullptr,
                     unsigned short* Wbf = nullptr, unsigned short* Wlb = nullptr);
// whether launch_update_w will run the one-block-per-column kernel for this shape (the caller may then skip the row-sum finish)
bool update_w_by_column(const Dims& d, int nsplit);

// W step for the `t` (1..4) rows beyond the last whole 256-row block (row0 = 256*floor(M/256); spectrogram heights are
// 2^k + 1 = 513, 1025, 2049: one row would cost the fused kernel a whole 256-row work item per column split, 11 % of the
// W step at 2049 rows, 33 % at 513).  On the CUDA cores, plain fp32, IEEE divide:
//   P[m,k] = sum_n ( X[m,n] / max( sum_k' W[m,k'] H[k',n], EPS ) ) * H[k,n]      for m = row0 .. row0+t-1
// per-block partial sums -> `part` (w_tail_blocks(Nv) * Kv float4), folded in a fixed order into the rows of P[split 0]
// (the other splits' rows are zeroed), so that everything downstream — fold, allreduce, W update — runs unchanged.
int  w_tail_blocks(int Nv);
void launch_w_tail(const float* W, const float* H, const float* X, float* part, const Dims& d, int row0, int t,
                   float* P, int nsplit, size_t split_stride, cudaStream_t st);

// dst[j*count + i] = src[j*ld + rows[i]] for j < cols (rows = device array of `count` row indices): sampled rows of a
// column-major matrix, for the spot checks of full-size runs
void launch_gather_rows(const float* src, int ld, int cols, const int* rows, int count, float* dst, cudaStream_t st);

// out[e] = sum_s P[s][e] in split order (multi-GPU: fold the local splits before the allreduce).
void launch_fold_splits(const float* P, int nsplit, size_t split_stride, float* out, size_t n, cudaStream_t st);
// H <- H .* ((sum of the nsplit partial W'*Y) ./ max(sumW,EPS)), plus the TF32 pair of the new H
// With `rowsum_partials` (rowsum_partial_count(d) * d.Kv floats) and Kv in {64, 128, 256} the kernel also emits per-block
// partial row sums of the new H; returns the number of blocks that wrote them (0: none, run launch_rowsum instead).
int  launch_update_h(float* H, float* Hhi, float* Hlo, const float* P, int nsplit, size_t split_stride,
                     const float* sumW, const Dims& d, cudaStream_t st, float* rowsum_partials = nullptr,
                     unsigned short* Hbf = nullptr, unsigned short* Hlb = nullptr);
// sumH = ordered fold of the `nblocks` partial row sums launch_update_h left in `partials`
void launch_rowsum_finish2(const float* partials, float* sumH, int nblocks, const Dims& d, cudaStream_t st);

// out[b] = sum of the `n` double partials of batch b (partials + b*n), fixed order; nbatch blocks.
void launch_sum_partials(const double* partials, int n, int nbatch, double* out, cudaStream_t st);

// ---------------------------------------------------------------------------------------------
// SIMT engine (kernels_simt.cu): fp32 FFMA GEMMs with the same fused epilogues.  Bring-up aid.
// ---------------------------------------------------------------------------------------------
int  simt_cost_partial_count(const Dims& d);
// Y = X ./ max(W*H,EPS); if cost_partials != nullptr also per-block KL partial sums.
void simt_quotient(const float* W, const float* H, const float* X, float* Y, double* cost_partials,
                   const Dims& d, cudaStream_t st);
// H <- H .* ((W'*Y) ./ max(sumW,EPS))
void simt_update_h(const float* W, const float* Y, const float* sumW, float* H, const Dims& d, cudaStream_t st);
// P = Y*H'
void simt_numer(const float* Y, const float* H, float* P, const Dims& d, cudaStream_t st);

}  // namespace nmf
