// session.hpp — one rank's device-resident NMF problem and the iteration driver.
#pragma once
#include "common.hpp"

namespace nmf {

namespace umma { struct Plan; }

struct Nccl;   // dlopen'ed NCCL entry points + communicator (dist.cpp)

struct nmf_session_impl {
    nmf_options opt{};
    Dims d{};
    int device = 0;
    cudaStream_t stream = nullptr;

    // fp32 masters (column-major, ld = Mv / Kv)
    float *W = nullptr, *H = nullptr, *X = nullptr;
    // quotient Y = X ./ max(W*H,EPS), stored once as fp32 by the two-kernel path.  nullptr when both half-iterations
    // run fused (the quotient then never leaves the SM): the M x N buffer is not even allocated, which halves the
    // footprint of a session (session_need_quotient allocates it for the debug entry points that read it)
    float *Y = nullptr;
    // TF32 splits of the small operands (tcgen05 engine)
    float *Whi = nullptr, *Wlo = nullptr, *Hhi = nullptr, *Hlo = nullptr;
    // sums and the packed allreduce buffer [ numerator P (Mv*Kv) | sumH (Kv) ]
    float *sumW = nullptr, *packed = nullptr, *sumH = nullptr /* = packed + Mv*Kv */;
    float *Psplit = nullptr; int nsplit = 1;      // split-N partial numerators (tcgen05 engine)
    float *Hpart = nullptr; int hsplit = 1;       // split-M partial W'*Y of the H update (small problems only)
    int fhsplit = 1;                              // the same for the fused H step (shares Hpart)
    float *rowsum_partials = nullptr;
    int rowsum_blocks = 0;                        // > 0: the H update of this iteration left that many partial row sums of the new H
    double *cost_partials = nullptr; int n_cost_partials = 0;
    int n_cost_partials_fused = 0;                // partials of the fused H step's cost variant (0: that step is not fused)
    float *Hbak = nullptr;                        // H before a fused H step that carries the cost (allocated on first use)
    double *cost_dev = nullptr, *cost_host = nullptr /* pinned */;

    umma::Plan* plan = nullptr;

    // multi-GPU
    Nccl* nccl = nullptr; int rank = 0, nranks = 1;

    // graphs (one iteration; and the iteration minus its leading colsum+quotient)
    cudaGraphExec_t g_full = nullptr, g_rest = nullptr;
    long long launches_full = 0, launches_rest = 0, launches_head = 0;
    long long launches = 0;    // running count of this library's kernel launches

    ~nmf_session_impl();
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
void nccl_share(nmf_session_impl& dst, nmf_session_impl& src);

namespace umma {
// tcgen05 engine (kernels_umma.cu)
Plan* plan_create(nmf_session_impl& s);
void  plan_destroy(Plan* p);
int   cost_partial_count(const Dims& d);
int   numer_splits(const Dims& d);
int   updh_splits(const Dims& d);
void  quotient(nmf_session_impl& s, bool with_cost);
void  update_h(nmf_session_impl& s);
void  numer(nmf_session_impl& s);
// fused W step: second quotient + numerator in one kernel, the quotient never reaches HBM (umma_fused.inc)
bool  fused_w_enabled(const nmf_session_impl& s);
void  fused_w(nmf_session_impl& s);
// fused H step: first quotient + W'*Y + the H update in one kernel (needs sumW)
bool  fused_h_enabled(const nmf_session_impl& s);
int   fused_h(nmf_session_impl& s, bool with_cost = false);   // returns the kernels launched (2 when the rows of W are split)
int   fused_h_splits(const Dims& d, int passes);
int   fused_cost_partial_count(const Dims& d, int passes);
bool  fused_both_wanted(const Dims& d, int passes);
unsigned int protocol_error(const nmf_session_impl& s);
bool  fused_h_needs_part(const Dims& d, int passes);   // the fused H step writes partial sums (row split or deferred update)
// raw contraction for descriptor bring-up (nmf_dbg_gemm)
void  dbg_gemm(int which, const float* dA, const float* dB, float* dC, int M, int N, int K, int passes,
               cudaStream_t st);
}  // namespace umma

}  // namespace nmf

struct nmf_session : nmf::nmf_session_impl {};
