// session.cu — device buffers of one rank's shard and the iteration driver that replaces the
// reference's run_async / update_h / update_w (cuda/nmf.cu:76-176).
#include "session.hpp"

#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>

namespace nmf {

// ---------------------------------------------------------------------------------------------
// error convention
// ---------------------------------------------------------------------------------------------
[[noreturn]] void fail(const char* file, int line, const std::string& msg, int code) {
    const char* base = std::strrchr(file, '/');
    throw Error(std::string(base ? base + 1 : file) + ":" + std::to_string(line) + ": " + msg, code ? code : 1);
}

Dims make_dims(int M, int n_local, int K, long long n_offset, long long n_global, int semantics) {
    NMF_CHECK(M > 0 && n_local > 0 && K > 0, "matrix dimensions must be positive");
    Dims d;
    d.M = M; d.N = n_local; d.K = K;
    d.n_offset = n_offset; d.n_global = n_global > 0 ? n_global : n_local;
    d.refcompat = (semantics == NMF_REFCOMPAT);
    if (d.refcompat) {
        NMF_CHECK(M != 1 && K != 1, "refcompat semantics: dimensions equal to 1 are not padded by the "
                                    "reference (cuda/matrix.cu:88-95) and are not supported here");
        d.Mv = ref_pad32(M); d.Kv = ref_pad32(K);
        // the reference pads the GLOBAL column count; only the last shard carries that padding
        const bool last = (n_offset + n_local >= d.n_global);
        const long long gpad = (d.n_global != 1 && d.n_global % 32 != 0) ? 32 - d.n_global % 32 : 0;
        d.Nv = n_local + (last ? (int)gpad : 0);
        d.w_divide = (d.Mv <= 1024) ? 1 : 0;      // row_div<<<N, rows_padded>>>, cuda/matrix.cu:212-217
    } else {
        d.Mv = round_up(M, 32); d.Kv = round_up(K, 32); d.Nv = n_local;
        d.w_divide = 1;
    }
    d.Nva = round_up(d.Nv, 768);      // whole tiles of every width the engine uses (64, 128, 192, 256)
    return d;
}

namespace {

template <typename T>
T* dalloc(size_t n) {
    T* p = nullptr;
    CUDA_CHECK(cudaMalloc((void**)&p, n * sizeof(T)));
    return p;
}

void count_launches(nmf_session_impl& s, long long n) { s.launches += n; }

}  // namespace

nmf_session_impl::~nmf_session_impl() {
    if (g_full) cudaGraphExecDestroy(g_full);
    if (g_rest) cudaGraphExecDestroy(g_rest);
    if (plan) umma::plan_destroy(plan);
    if (nccl) nccl_destroy(nccl);
    float* fl[] = {W, H, X, Y, Whi, Wlo, Hhi, Hlo, sumW, packed, Psplit, Hpart, rowsum_partials, Hbak};
    for (float* p : fl) if (p) cudaFree(p);
    if (cost_partials) cudaFree(cost_partials);
    if (cost_dev) cudaFree(cost_dev);
    if (cost_host) cudaFreeHost(cost_host);
    if (stream) cudaStreamDestroy(stream);
}

nmf_session_impl* session_create(int M, int n_local, int K, long long n_offset, long long n_global,
                                 const nmf_options& opt) {
    int ndev = 0;
    CUDA_CHECK(cudaGetDeviceCount(&ndev));
    NMF_CHECK(ndev > 0, "no CUDA device: libnmf_b200 has no CPU fallback");
    auto* s = new nmf_session();
    try {
        s->opt = opt;
        if (s->opt.check_every <= 0) s->opt.check_every = kCheckEveryDefault;
        if (opt.device >= 0) { CUDA_CHECK(cudaSetDevice(opt.device)); }
        CUDA_CHECK(cudaGetDevice(&s->device));
        s->d = make_dims(M, n_local, K, n_offset, n_global, opt.semantics);
        const Dims& d = s->d;
        CUDA_CHECK(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
        s->W = dalloc<float>(d.szW()); s->H = dalloc<float>(d.szHa()); s->X = dalloc<float>(d.szXa());
        CUDA_CHECK(cudaMemsetAsync(s->W, 0, d.szW() * 4, s->stream));
        CUDA_CHECK(cudaMemsetAsync(s->H, 0, d.szHa() * 4, s->stream));
        CUDA_CHECK(cudaMemsetAsync(s->X, 0, d.szXa() * 4, s->stream));
        s->sumW = dalloc<float>(d.Kv);
        s->packed = dalloc<float>(d.szW() + d.Kv);
        s->sumH = s->packed + d.szW();
        s->rowsum_partials = dalloc<float>((size_t)rowsum_partial_count(d) * d.Kv);
        s->cost_dev = dalloc<double>(1);
        CUDA_CHECK(cudaMallocHost((void**)&s->cost_host, sizeof(double)));
        if (opt.engine == NMF_ENGINE_SIMT) {
            s->Y = dalloc<float>(d.szXa());
            s->n_cost_partials = simt_cost_partial_count(d);
            s->nsplit = 1;
        } else {
            // stored once as fp32; the GEMMs that consume it split it in the SM.  Not needed at all when both
            // half-iterations are fused (NMF_KEEP_QUOTIENT=1 allocates it anyway)
            const char* keep = std::getenv("NMF_KEEP_QUOTIENT");
            if (!umma::fused_both_wanted(d, opt.tf32_passes) || (keep && std::atoi(keep) != 0)) s->Y = dalloc<float>(d.szXa());
            s->Whi = dalloc<float>(d.szW()); s->Wlo = dalloc<float>(d.szW());
            s->Hhi = dalloc<float>(d.szHa()); s->Hlo = dalloc<float>(d.szHa());
            s->n_cost_partials = umma::cost_partial_count(d);
            s->n_cost_partials_fused = umma::fused_cost_partial_count(d, opt.tf32_passes);
            s->nsplit = umma::numer_splits(d);
            if (s->nsplit > 1) s->Psplit = dalloc<float>((size_t)s->nsplit * d.szW());
            s->hsplit = umma::updh_splits(d);
            s->fhsplit = umma::fused_h_splits(d, opt.tf32_passes);
            if (std::max(s->hsplit, s->fhsplit) > 1 || umma::fused_h_needs_part(d, opt.tf32_passes))
                s->Hpart = dalloc<float>((size_t)std::max(s->hsplit, s->fhsplit) * d.szHa());
            s->plan = umma::plan_create(*s);
        }
        s->cost_partials = dalloc<double>(std::max(1, std::max(s->n_cost_partials, s->n_cost_partials_fused)));
        CUDA_CHECK(cudaStreamSynchronize(s->stream));
    } catch (...) {
        delete s;
        throw;
    }
    return s;
}

void session_need_quotient(nmf_session_impl& s) {
    if (s.Y) return;
    CUDA_CHECK(cudaSetDevice(s.device));
    s.Y = dalloc<float>(s.d.szXa());
    if (s.plan) {                     // its tensor maps and kernel parameters hold the buffer's address
        umma::plan_destroy(s.plan); s.plan = nullptr;
        s.plan = umma::plan_create(s);
        if (s.g_full) { cudaGraphExecDestroy(s.g_full); s.g_full = nullptr; }
        if (s.g_rest) { cudaGraphExecDestroy(s.g_rest); s.g_rest = nullptr; }
    }
}

// After any change of W or H masters: refresh the TF32 splits the tcgen05 engine reads.
static void refresh_splits(nmf_session_impl& s, bool w, bool h) {
    if (s.opt.engine != NMF_ENGINE_TCGEN05) return;
    if (w) { launch_split_tf32(s.W, s.Whi, s.Wlo, s.d.szW(), s.stream); }
    if (h) { launch_split_tf32(s.H, s.Hhi, s.Hlo, s.d.szH(), s.stream); }
}

// Host (unpadded, column-major) -> padded device buffer + clamp, as read_matrix + Matrix(host,...)
// do (cuda/nmf.cu:208-211, cuda/matrix.cu:53-67).
static void upload_one(nmf_session_impl& s, float* dst, const float* src, int rows, int cols, int rows_v, int cols_v) {
    CUDA_CHECK(cudaMemsetAsync(dst, 0, (size_t)rows_v * cols_v * 4, s.stream));
    CUDA_CHECK(cudaMemcpy2DAsync(dst, (size_t)rows_v * 4, src, (size_t)rows * 4, (size_t)rows * 4, cols,
                                 cudaMemcpyHostToDevice, s.stream));
    if (s.d.refcompat) launch_clamp_eps(dst, rows_v, cols_v, rows_v, s.stream);   // padding becomes EPS
    else               launch_clamp_eps(dst, rows, cols, rows_v, s.stream);       // padding stays 0
}

void session_upload(nmf_session_impl& s, const float* W, const float* H, const float* X) {
    const Dims& d = s.d;
    CUDA_CHECK(cudaSetDevice(s.device));
    if (W) upload_one(s, s.W, W, d.M, d.K, d.Mv, d.Kv);
    if (H) upload_one(s, s.H, H, d.K, d.N, d.Kv, d.Nv);
    if (X) upload_one(s, s.X, X, d.M, d.N, d.Mv, d.Nv);
    refresh_splits(s, W != nullptr, H != nullptr);
    CUDA_CHECK(cudaStreamSynchronize(s.stream));
}

void session_download(nmf_session_impl& s, float* W, float* H) {
    const Dims& d = s.d;
    CUDA_CHECK(cudaSetDevice(s.device));
    if (W) CUDA_CHECK(cudaMemcpy2DAsync(W, (size_t)d.M * 4, s.W, (size_t)d.Mv * 4, (size_t)d.M * 4, d.K,
                                        cudaMemcpyDeviceToHost, s.stream));
    if (H) CUDA_CHECK(cudaMemcpy2DAsync(H, (size_t)d.K * 4, s.H, (size_t)d.Kv * 4, (size_t)d.K * 4, d.N,
                                        cudaMemcpyDeviceToHost, s.stream));
    CUDA_CHECK(cudaStreamSynchronize(s.stream));
}

void session_fill_uniform(nmf_session_impl& s, unsigned long long seed, int which) {
    const Dims& d = s.d;
    CUDA_CHECK(cudaSetDevice(s.device));
    auto fill = [&](float* dst, int rows, int cols, int rows_v, int cols_v, long long coff, int tag) {
        CUDA_CHECK(cudaMemsetAsync(dst, 0, (size_t)rows_v * cols_v * 4, s.stream));
        launch_fill_uniform(dst, rows, cols, rows_v, coff, seed, tag, s.stream);
        if (d.refcompat) launch_clamp_eps(dst, rows_v, cols_v, rows_v, s.stream);
    };
    if (which & 1) fill(s.W, d.M, d.K, d.Mv, d.Kv, 0, 1);
    if (which & 2) fill(s.H, d.K, d.N, d.Kv, d.Nv, d.n_offset, 2);
    if (which & 4) fill(s.X, d.M, d.N, d.Mv, d.Nv, d.n_offset, 3);
    refresh_splits(s, which & 1, which & 2);
    CUDA_CHECK(cudaStreamSynchronize(s.stream));
}

// X <- planted rank-`rank` product with multiplicative uniform noise (BASELINE C3's input); W, H untouched
void session_fill_planted(nmf_session_impl& s, unsigned long long seed, int rank, float noise) {
    const Dims& d = s.d;
    CUDA_CHECK(cudaSetDevice(s.device));
    NMF_CHECK(rank >= 1, "planted rank must be positive");
    CUDA_CHECK(cudaMemsetAsync(s.X, 0, d.szXa() * 4, s.stream));
    launch_fill_planted(s.X, d.M, d.N, d.Mv, d.n_offset, seed, rank, noise, s.stream);
    if (d.refcompat) launch_clamp_eps(s.X, d.Mv, d.Nv, d.Mv, s.stream);
    CUDA_CHECK(cudaStreamSynchronize(s.stream));
}

// ---------------------------------------------------------------------------------------------
// per-phase timing for t[10]: cudaEvent pairs around each phase, accumulated after a sync.
// ---------------------------------------------------------------------------------------------
struct PhaseTimer {
    cudaStream_t st;
    struct Span { int phase; cudaEvent_t a, b; };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> pool;
    double acc[10] = {0};
    explicit PhaseTimer(cudaStream_t s) : st(s) {}
    cudaEvent_t ev() {
        if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
        cudaEvent_t e; CUDA_CHECK(cudaEventCreate(&e)); return e;
    }
    void begin(int phase) { Span sp{phase, ev(), ev()}; CUDA_CHECK(cudaEventRecord(sp.a, st)); spans.push_back(sp); }
    void end() { CUDA_CHECK(cudaEventRecord(spans.back().b, st)); if (spans.size() >= 512) flush(); }
    void flush() {
        CUDA_CHECK(cudaStreamSynchronize(st));
        for (auto& sp : spans) {
            float ms = 0; CUDA_CHECK(cudaEventElapsedTime(&ms, sp.a, sp.b));
            acc[sp.phase] += ms * 1e-3;
            pool.push_back(sp.a); pool.push_back(sp.b);
        }
        spans.clear();
    }
    ~PhaseTimer() { for (auto& sp : spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); } for (auto e : pool) cudaEventDestroy(e); }
};

#define PHASE(pt, ph, body) do { if (pt) (pt)->begin(ph); body; if (pt) (pt)->end(); } while (0)

// ---------------------------------------------------------------------------------------------
// One iteration = head + rest  (reference: update_h then update_w, cuda/nmf.cu:108-109)
// ---------------------------------------------------------------------------------------------
static void do_quotient(nmf_session_impl& s, bool with_cost) {
    if (s.opt.engine == NMF_ENGINE_SIMT) {
        simt_quotient(s.W, s.H, s.X, s.Y, with_cost ? s.cost_partials : nullptr, s.d, s.stream);
        count_launches(s, 1);
    } else {
        umma::quotient(s, with_cost);
        count_launches(s, 1);
    }
    if (with_cost) { launch_sum_partials(s.cost_partials, s.n_cost_partials, s.cost_dev, s.stream); count_launches(s, 1); }
}

void iteration_head(nmf_session_impl& s, bool with_cost, PhaseTimer* pt) {
    // sum cols of W (nmf.cu:134) — first, so the head is exactly "everything that only needs W,H,X"
    PHASE(pt, PH_SUMS, { launch_colsum(s.W, s.sumW, s.d, s.stream); count_launches(s, 1); });
    // Z = X ./ max(W*H, EPS)  (nmf.cu:125-131); the KL cost of the current factors rides along
    PHASE(pt, PH_QUOT, do_quotient(s, with_cost));
}

static void w_half(nmf_session_impl& s, PhaseTimer* pt);

// Check iteration with the fused H step: the KL cost of the factors the iteration starts from is a by-product of
// that step's quotient, which also rewrites H.  H is saved first so that a run that turns out to have converged
// can hand back the factors the cost belongs to (the oracle stops before this H step).
static bool fused_check_enabled(const nmf_session_impl& s) {
    return s.opt.engine == NMF_ENGINE_TCGEN05 && umma::fused_h_enabled(s) && s.n_cost_partials_fused > 0;
}
static void iteration_head_fused_cost(nmf_session_impl& s, PhaseTimer* pt) {
    const Dims& d = s.d;
    if (!s.Hbak) s.Hbak = dalloc<float>(d.szHa());
    PHASE(pt, PH_SUMS, { launch_colsum(s.W, s.sumW, d, s.stream); count_launches(s, 1); });
    PHASE(pt, PH_COST, CUDA_CHECK(cudaMemcpyAsync(s.Hbak, s.H, d.szHa() * 4, cudaMemcpyDeviceToDevice, s.stream)));
    PHASE(pt, PH_GEMM2, {
        const int n = umma::fused_h(s, true); count_launches(s, n);
        launch_sum_partials(s.cost_partials, s.n_cost_partials_fused, s.cost_dev, s.stream); count_launches(s, 1);
    });
}
static void restore_h(nmf_session_impl& s) {
    s.rowsum_blocks = 0;              // partial row sums the abandoned H step may have left belong to the H being discarded
    CUDA_CHECK(cudaMemcpyAsync(s.H, s.Hbak, s.d.szHa() * 4, cudaMemcpyDeviceToDevice, s.stream));
    launch_split_tf32(s.H, s.Hhi, s.Hlo, s.d.szH(), s.stream); count_launches(s, 1);
}

void iteration_rest(nmf_session_impl& s, PhaseTimer* pt) {
    const Dims& d = s.d;
    // H = H .* ((W'*Z) ./ sumW)   (nmf.cu:138-145)
    PHASE(pt, PH_GEMM2, {
        if (s.opt.engine == NMF_ENGINE_SIMT) simt_update_h(s.W, s.Y, s.sumW, s.H, d, s.stream);
        else umma::update_h(s);
        count_launches(s, (s.opt.engine == NMF_ENGINE_TCGEN05 && s.hsplit > 1) ? 2 : 1);   // split: GEMM partials + fold/update kernel
    });
    w_half(s, pt);
}

// One whole iteration when no cost is wanted: with the fused H step the first quotient never reaches HBM.
void iteration_full(nmf_session_impl& s, PhaseTimer* pt) {
    if (s.opt.engine == NMF_ENGINE_TCGEN05 && umma::fused_h_enabled(s)) {
        PHASE(pt, PH_SUMS, { launch_colsum(s.W, s.sumW, s.d, s.stream); count_launches(s, 1); });
        PHASE(pt, PH_GEMM2, { const int n = umma::fused_h(s); count_launches(s, n); });
        w_half(s, pt);
    } else {
        iteration_head(s, false, pt);
        iteration_rest(s, pt);
    }
}

static void w_half(nmf_session_impl& s, PhaseTimer* pt) {
    const Dims& d = s.d;
    // sum rows of H (nmf.cu:164) into the packed buffer's tail
    // (when the H update ran as its own kernel it left per-block partial sums of the new H: only the fold remains)
    PHASE(pt, PH_SUMS, {
        if (s.rowsum_blocks > 0) { launch_rowsum_finish2(s.rowsum_partials, s.sumH, s.rowsum_blocks, d, s.stream); count_launches(s, 1); }
        else { launch_rowsum(s.H, s.rowsum_partials, s.sumH, d, s.stream); count_launches(s, 2); }
        s.rowsum_blocks = 0;
    });
    const float* P = s.packed; int nsplit = 1; size_t stride = 0;
    if (s.opt.engine == NMF_ENGINE_TCGEN05 && umma::fused_w_enabled(s)) {
        // both of the steps below in one kernel: the quotient goes from TMEM through shared memory into the
        // second MMA and never reaches HBM
        PHASE(pt, PH_GEMM2, { umma::fused_w(s); count_launches(s, 1); });
    } else {
        // Z = X ./ max(W*H, EPS) with the new H (nmf.cu:155-161)
        PHASE(pt, PH_QUOT, do_quotient(s, false));
        // numerator Z*H' (nmf.cu:168): split over column blocks, partials reduced in order by the W update
        PHASE(pt, PH_GEMM2, {
            if (s.opt.engine == NMF_ENGINE_SIMT) simt_numer(s.Y, s.H, s.packed, d, s.stream);
            else umma::numer(s);
            count_launches(s, 1);
        });
    }
    if (s.opt.engine == NMF_ENGINE_TCGEN05 && s.nsplit > 1) { P = s.Psplit; nsplit = s.nsplit; stride = d.szW(); }
    if (s.nranks > 1) {
        // One exchange per iteration: [numerator | row sums of H] summed over the column shards.
        if (nsplit > 1) {   // fold the local splits first so only Mv*Kv floats cross NVLink
            PHASE(pt, PH_UPDW, { launch_fold_splits(P, nsplit, stride, s.packed, d.szW(), s.stream); count_launches(s, 1); });
        }
        PHASE(pt, PH_COMM, nccl_allreduce_f32(s, s.packed, d.szW() + d.Kv));
        P = s.packed; nsplit = 1; stride = 0;
    }
    // W = W .* ((Z*H') ./ sumH)   (nmf.cu:172-175)
    PHASE(pt, PH_UPDW, {
        launch_update_w(s.W, s.Whi, s.Wlo, P, nsplit, stride, s.sumH, d, s.stream);
        count_launches(s, 1);
    });
}

// ---------------------------------------------------------------------------------------------
// run loop
// ---------------------------------------------------------------------------------------------
static double read_cost(nmf_session_impl& s) {
    if (s.nranks > 1) nccl_allreduce_f64(s, s.cost_dev, 1);
    CUDA_CHECK(cudaMemcpyAsync(s.cost_host, s.cost_dev, sizeof(double), cudaMemcpyDeviceToHost, s.stream));
    CUDA_CHECK(cudaStreamSynchronize(s.stream));
    return *s.cost_host;
}

static void build_graphs(nmf_session_impl& s) {
    if (s.g_full) return;
    cudaGraph_t g = nullptr;
    long long c0 = s.launches;
    CUDA_CHECK(cudaStreamBeginCapture(s.stream, cudaStreamCaptureModeThreadLocal));   // as nmf.cu:103
    iteration_full(s, nullptr);
    const long long c1 = s.launches;
    CUDA_CHECK(cudaStreamEndCapture(s.stream, &g));
    CUDA_CHECK(cudaGraphInstantiate(&s.g_full, g, 0));
    CUDA_CHECK(cudaGraphDestroy(g));
    s.launches_head = c1 - c0;
    s.launches_full = s.launches - c0;
    s.launches_rest = 0;
    if (!fused_check_enabled(s)) {      // check iterations of the two-kernel path: head (eager, with the cost) + this graph
        c0 = s.launches;
        CUDA_CHECK(cudaStreamBeginCapture(s.stream, cudaStreamCaptureModeThreadLocal));
        iteration_rest(s, nullptr);
        CUDA_CHECK(cudaStreamEndCapture(s.stream, &g));
        CUDA_CHECK(cudaGraphInstantiate(&s.g_rest, g, 0));
        CUDA_CHECK(cudaGraphDestroy(g));
        s.launches_rest = s.launches - c0;
    }
    s.launches -= s.launches_full + s.launches_rest;    // capture launched nothing
}

void session_cost(nmf_session_impl& s, double* cost) {
    CUDA_CHECK(cudaSetDevice(s.device));
    do_quotient(s, true);        // cost of the resident factors (the quotient it also writes is scratch)
    *cost = read_cost(s);
}

void session_run(nmf_session_impl& s, float thresh, int max_iter, double t[10], int verbose, nmf_result* res) {
    using clk = std::chrono::steady_clock;
    const auto wall0 = clk::now();
    CUDA_CHECK(cudaSetDevice(s.device));
    const bool need_cost = (thresh > 0.0f) || verbose;
    const int every = s.opt.check_every;
    PhaseTimer* pt = t ? new PhaseTimer(s.stream) : nullptr;
    const bool graph = s.opt.use_graph && !pt;
    s.rowsum_blocks = 0;
    if (graph) build_graphs(s);

    cudaEvent_t e0, e1;
    CUDA_CHECK(cudaEventCreate(&e0)); CUDA_CHECK(cudaEventCreate(&e1));
    const long long launches0 = s.launches;
    CUDA_CHECK(cudaEventRecord(e0, s.stream));

    double prev = std::numeric_limits<double>::infinity(), cost = std::nan("");
    int done = 0, converged = 0;
    double t_cost = 0.0;
    for (int it = 1; it <= max_iter; ++it) {
        // The cost after iteration it-1 is a by-product of iteration it's first quotient.
        const bool check = need_cost && it > 1 && ((it - 1) % every == 0);
        if (check) {
            const bool fcheck = fused_check_enabled(s);
            if (fcheck) iteration_head_fused_cost(s, pt); else iteration_head(s, true, pt);
            const auto c0 = clk::now();
            cost = read_cost(s);
            t_cost += std::chrono::duration<double>(clk::now() - c0).count();
            const double rel = (prev - cost) / cost;
            if (verbose) std::printf("iter %d  cost %.9g  rel.change %.3g\n", it - 1, cost, rel);
            if (thresh > 0.0f && rel < (double)thresh) { if (fcheck) restore_h(s); converged = 1; break; }
            prev = cost;
            if (fcheck) w_half(s, pt);
            else if (graph) { CUDA_CHECK(cudaGraphLaunch(s.g_rest, s.stream)); s.launches += s.launches_rest; }
            else iteration_rest(s, pt);
        } else if (graph) {
            CUDA_CHECK(cudaGraphLaunch(s.g_full, s.stream)); s.launches += s.launches_full;
        } else {
            iteration_full(s, pt);
        }
        done = it;
    }
    if (need_cost && !converged && done > 0 && (done % every == 0 || verbose)) {
        // final check point (or the closing line of a verbose run)
        if (pt) pt->begin(PH_QUOT);
        do_quotient(s, true);
        if (pt) pt->end();
        const auto c0 = clk::now();
        cost = read_cost(s);
        t_cost += std::chrono::duration<double>(clk::now() - c0).count();
        const double rel = (prev - cost) / cost;
        if (verbose) std::printf("iter %d  cost %.9g  rel.change %.3g\n", done, cost, rel);
        if (thresh > 0.0f && done % every == 0 && rel < (double)thresh) converged = 1;
    }
    CUDA_CHECK(cudaEventRecord(e1, s.stream));
    {
        const cudaError_t e = cudaStreamSynchronize(s.stream);
        if (e != cudaSuccess && s.opt.engine == NMF_ENGINE_TCGEN05 && umma::protocol_error(s))
            NMF_CHECK(false, std::string("kernel trapped: bounded mbarrier wait ") + std::to_string(umma::protocol_error(s)) +
                             " gave up (" + cudaGetErrorString(e) + ")");
        CUDA_CHECK(e);
    }
    float ms = 0; CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    CUDA_CHECK(cudaEventDestroy(e0)); CUDA_CHECK(cudaEventDestroy(e1));
    if (pt) {
        pt->flush();
        for (int i = 0; i < 10; ++i) t[i] = 0.0;
        for (int i = 0; i < 8; ++i) t[i] = pt->acc[i];
        t[PH_COST] += t_cost;
        t[PH_TOTAL] = std::chrono::duration<double>(clk::now() - wall0).count();
        delete pt;
    }
    if (res) {
        res->iterations = done; res->converged = converged; res->cost = cost;
        res->device_ms = ms; res->gpu_launches = s.launches - launches0;
    }
}

}  // namespace nmf
