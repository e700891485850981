// session.hpp — one rank's device-resident NMF problem and the iteration driver.
#pragma once
#include <vector>

#include "common.hpp"

namespace nmf {

namespace umma {
struct Plan;
// Tuning / A-B switches of the tcgen05 engine (NMF_UMMA_* environment variables): read ONCE when a session is created
// (read_tuning) and carried by it — no getenv on any launch path.
struct Tuning {
    int pair = 1;          // NMF_UMMA_PAIR      CTA pairs (cta_group::2) wherever the output has > 128 rows
    int chunk = 4;         // NMF_UMMA_CHUNK     32-deep blocks accumulated in TMEM between round-to-nearest promotions (1..16)
    int w3_max_k = 192;    // NMF_UMMA_W3_MAXK   whole-tile kernels: 12 epilogue warps up to this reduction length
    int hint_h = 0;        // NMF_UMMA_HINT_H    L2 eviction priority of the H operand
    int np_bn_cap = 256;   // NMF_UMMA_NP_BNCAP  width cap of one-CTA tiles of the chunked kernels
    int hints = 1;         // NMF_UMMA_HINTS     L2 eviction-priority hints on the operand / epilogue streams
    int xpf = 0;           // NMF_UMMA_XPF       G1: L2 prefetch of X two chunks ahead
    int fuse = 1;          // NMF_UMMA_FUSE      0 never fuse a half-iteration, 1/2 whenever the shape allows
    int fuse_steps = 3;    // NMF_UMMA_FUSE_STEPS bit 0 W step, bit 1 H step
    int fhsplit = 0;       // NMF_UMMA_FHSPLIT   forced row-split count of the fused H step (0: the cost model chooses)
    int fuse_nw = 0;       // NMF_UMMA_FUSE_NW   worker warps per TMEM quadrant of the fused kernels (0: 4 when Kv <= 128, else 2)
    int fh_defer = 0;      // NMF_UMMA_FH_DEFER  fused H step leaves the update to k_update_h even without a row split
    int corr_bf16 = 7;     // NMF_CORR_BF16      the two correction passes (lo*hi, hi*lo) as BF16 MMAs: fused kernels (bit 0: K = 256, bit 1: K = 128),
                           //                    two-kernel path of sessions without fused half-iterations (bit 2: K > 256)
    int weave = -1;        // NMF_UMMA_WEAVE     fused kernels, two S tiles: a- and b-stages interleaved in the ring (even L2 -> SM demand); -1: when K > 128
                           //                    (measured: C4 -3 %, C5 K=256 -3 %; at K <= 128, quotient in tensor memory, +11 % — off there)
    int prod2 = 0;         // NMF_UMMA_PROD2     fused kernels: warps 0 and 3 share the TMA loads of every ring stage
    int drain_early = 1;   // NMF_UMMA_DRAIN_EARLY  fused kernels, quotient through shared memory: drain P before the ring-slot wait of a chunk, not after its hand-off
    int fuse_alt = 0;      // NMF_UMMA_FUSE_ALT  two worker warps per quadrant (K > 128): a warp takes whole chunks in turn instead of half of every chunk
    int fuse_yt = 1;       // NMF_UMMA_FUSE_YT   K <= 128: stage the quotient in tensor memory (0: in shared memory, as K > 128 does)
    int fuse_bn = 128;     // NMF_UMMA_FUSE_BN   S tile width of the fused kernels (128 double-buffered, or 256)
    int hsplit = 0;        // NMF_UMMA_HSPLIT    forced row-split count of the two-kernel H update (0: automatic)
    int keep_quotient = 0; // NMF_KEEP_QUOTIENT  allocate the M x N quotient buffer even when both half-iterations are fused
    int variant = 0;       // NMF_UMMA_VARIANT   bring-up / timing switches that give WRONG results: honoured only by diagnostic
                           //                    builds (NMF_BUILD_DIAG=1 -> -DNMF_DIAG); a release build ignores the variable
    int w_tail = 1;        // NMF_W_TAIL         W step: rows beyond the last whole 256-row block (at most 4) on the CUDA cores
                           //                    (1: from 65536 columns, where it pays; 2: always; 0: never)
    int pdl = 1;           // NMF_PDL            programmatic dependent launch between the kernels of an iteration
    int pad64 = 1;         // NMF_PAD_K64        exact semantics: pad K to a multiple of 64 when that lets the half-iterations fuse
};
Tuning read_tuning();
}

struct Nccl;   // dlopen'ed NCCL entry points + communicator (dist.cpp)

struct nmf_session_impl {
    nmf_options opt{};
    umma::Tuning tune{};
    Dims d{};
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;            // H2D of X in column blocks (created on first use)

    // fp32 masters (column-major, ld = Mv / Kv)
    float *W = nullptr, *H = nullptr, *X = nullptr;
    // quotient Y = X ./ max(W*H,EPS), stored once as fp32 by the two-kernel path.  nullptr when both half-iterations
    // run fused (the quotient then never leaves the SM): the M x N buffer is not even allocated, which halves the
    // footprint of a session (session_need_quotient allocates it for the debug entry points that read it)
    float *Y = nullptr;
    // TF32 splits of the small operands (tcgen05 engine)
    float *Whi = nullptr, *Wlo = nullptr, *Hhi = nullptr, *Hlo = nullptr;
    // BF16 forms of the correction operands (fused kernels with K = 256): bf16(x) and bf16(x - tf32(x));
    // nullptr when the session does not use them
    unsigned short *Wbf = nullptr, *Wlb = nullptr, *Hbf = nullptr, *Hlb = nullptr;
    // sums and the packed allreduce buffer [ numerator P (Mv*Kv) | sumH (Kv) ]
    float *sumW = nullptr, *packed = nullptr, *sumH = nullptr /* = packed + Mv*Kv */;
    float *Psplit = nullptr; int nsplit = 1;      // split-N partial numerators (tcgen05 engine)
    float *Hpart = nullptr; int hsplit = 1;       // split-M partial W'*Y of the H update (small problems only)
    int fhsplit = 1;                              // the same for the fused H step (shares Hpart)
    // W step, tail rows (M = 256 q + t, 1 <= t <= 4, exact semantics): rows 256 q .. are taken out of the fused W kernel
    // (a whole 256-row work item for one spectrogram bin) and run on the CUDA cores (launch_w_tail); 0: no tail path
    int tail_t = 0, tail_row0 = 0; float* tail_part = nullptr;
    float *rowsum_partials = nullptr;
    int rowsum_blocks = 0;                        // > 0: the H update of this iteration left that many partial row sums of the new H
    bool sumW_valid = false;                      // sumW holds the column sums of the resident W (the W update emits them for the
                                                  // next iteration; anything else that writes W clears the flag)
    double *cost_partials = nullptr; int n_cost_partials = 0;
    int n_cost_partials_fused = 0;                // partials of the fused H step's cost variant (0: that step is not fused)
    float *Hbak = nullptr;                        // H before a fused H step that carries the cost (allocated on first use)
    double *cost_dev = nullptr, *cost_host = nullptr /* pinned */;   // R values each
    // restarts (d.R > 1): per-restart state of the last run; a restart that met the threshold at a check is retired — its
    // factors of that moment are kept in Wfin / Hfin and put back when the run ends
    std::vector<double> costs; std::vector<int> iters_r; std::vector<char> conv_r, retired;
    float *Wfin = nullptr, *Hfin = nullptr;

    umma::Plan* plan = nullptr;

    // multi-GPU
    Nccl* nccl = nullptr; int rank = 0, nranks = 1;

    // graphs (one iteration; and the iteration minus its leading colsum+quotient)
    cudaGraphExec_t g_full = nullptr, g_rest = nullptr;
    long long launches_full = 0, launches_rest = 0, launches_head = 0;
    long long launches = 0;    // running count of this library's kernel launches

    virtual ~nmf_session_impl();
};

// phases for t[10]
enum Phase { PH_QUOT = 0, PH_SUMS = 2, PH_GEMM2 = 3, PH_UPDW = 5, PH_COST = 6, PH_COMM = 7, PH_SETUP = 8, PH_TOTAL = 9 };

struct PhaseTimer;   // session.cu

// iteration pieces (session.cu)
void iteration_head(nmf_session_impl& s, bool with_cost, PhaseTimer* pt);   // colsum(W); Y = quotient (+cost partials)
void iteration_rest(nmf_session_impl& s, PhaseTimer* pt);                   // H update; rowsum; quotient; numerator; [allreduce]; W update
void iteration_full(nmf_session_impl& s, PhaseTimer* pt);                   // one whole iteration without a cost (fused H step when enabled)

void session_need_quotient(nmf_session_impl& s);   // allocate Y (and re-plan) if the session was created without it

// NCCL (dist.cpp)
int  nccl_unique_id(void* id128);
void nccl_join(nmf_session_impl& s, int rank, int nranks, const void* id128);
void nccl_allreduce_f32(nmf_session_impl& s, float* buf, size_t n);
void nccl_allreduce_f64(nmf_session_impl& s, double* buf, size_t n);
void nccl_destroy(Nccl* n);
void nccl_init_all(nmf_session_impl* const* sessions, int g);   // one process, a session per GPU: ncclCommInitAll
void nccl_abort(nmf_session_impl& s);                           // unblock peers of a failed rank (ncclCommAbort)
void nccl_share(nmf_session_impl& dst, nmf_session_impl& src);

namespace umma {
// tcgen05 engine (kernels_umma.cu)
Plan* plan_create(nmf_session_impl& s);
void  plan_destroy(Plan* p);
int   cost_partial_count(const Dims& d, const Tuning& tn);
int   numer_splits(const Dims& d, const Tuning& tn, int row_blocks = 0);   // row_blocks > 0: that many 256-row blocks instead of all of Mv
int   updh_splits(const Dims& d, const Tuning& tn);
void  quotient(nmf_session_impl& s, bool with_cost);
void  update_h(nmf_session_impl& s);
void  numer(nmf_session_impl& s);
// fused W step: second quotient + numerator in one kernel, the quotient never reaches HBM (umma_fused.inc)
bool  fused_w_enabled(const nmf_session_impl& s);
void  fused_w(nmf_session_impl& s);
// fused H step: first quotient + W'*Y + the H update in one kernel (needs sumW)
bool  fused_h_enabled(const nmf_session_impl& s);
int   fused_h(nmf_session_impl& s, bool with_cost = false);   // returns the kernels launched (2 when the rows of W are split)
int   fused_h_splits(const Dims& d, int passes, const Tuning& tn);
int   fused_cost_partial_count(const Dims& d, int passes, const Tuning& tn);   // per restart
bool  fused_both_wanted(const Dims& d, int passes, const Tuning& tn);
bool  fused_shape_ok(int Kr, int passes, const Tuning& tn);   // Kr (one restart's padded K) is a multiple of 64 and <= 256
unsigned int protocol_error(const nmf_session_impl& s);
bool  corr_bf16_wanted_2k(const Dims& d, int passes, const Tuning& tn);   // ... its two-kernel path does (K > 256)
bool  corr_bf16_wanted(const Dims& d, int passes, const Tuning& tn);   // the session's fused kernels run their correction passes in BF16
bool  fused_h_needs_part(const Dims& d, int passes, const Tuning& tn);   // the fused H step writes partial sums (row split or deferred update)
// raw contraction for descriptor bring-up (nmf_dbg_gemm)
void  dbg_gemm(int which, const float* dA, const float* dB, float* dC, int M, int N, int K, int passes,
               cudaStream_t st);
}  // namespace umma

}  // namespace nmf

struct nmf_session : nmf::nmf_session_impl {};
