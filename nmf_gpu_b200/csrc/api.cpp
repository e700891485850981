// api.cpp — the extern "C" boundary of libnmf_b200.so (include/nmf.h): the reference's host API
// (read_matrix / write_matrix / update_div, README.md:31-54) plus the session and test entry points.
#include <chrono>
#include <memory>
#include <thread>
#include <cmath>
#include <cstring>
#include <functional>
#include <string>
#include <vector>

#include "session.hpp"

namespace nmf {
nmf_session_impl* session_create(int M, int n_local, int K, long long n_offset, long long n_global, const nmf_options& opt);
void session_upload(nmf_session_impl& s, const float* W, const float* H, const float* X);
void session_download(nmf_session_impl& s, float* W, float* H);
void session_fill_uniform(nmf_session_impl& s, unsigned long long seed, int which);
void session_fill_planted(nmf_session_impl& s, unsigned long long seed, int rank, float noise);
void session_run(nmf_session_impl& s, float thresh, int max_iter, double t[10], int verbose, nmf_result* res);
void session_cost(nmf_session_impl& s, double* cost);
void session_upload_restart(nmf_session_impl& s, int r, const float* W, const float* H);
void session_download_restart(nmf_session_impl& s, int r, float* W, float* H);
void session_download_cols(nmf_session_impl& s, const long long* cols, int count, float* Xc, float* Hc);
void session_download_rows(nmf_session_impl& s, const int* rows, int count, float* Xr, float* Wr);
void session_half_step(nmf_session_impl& s, int which);
}  // namespace nmf

using namespace nmf;

namespace {

int g_error_mode = 0;                 // 0: print + exit (reference convention), 1: return codes
thread_local std::string g_last_error;

// Runs body; on failure follows the reference's convention (cuda/error-check.hpp:9-17: print and
// exit(code)) or, in mode 1, records the message and returns the code.
int guarded(const std::function<void()>& body) {
    try {
        g_last_error.clear();          // nmf_last_error() is empty after a call that succeeded
        body();
        return 0;
    } catch (const Error& e) {
        g_last_error = e.what();
        if (g_error_mode == 0) { std::fprintf(stderr, "nmf_b200: %s\n", e.what()); std::exit(e.code); }
        return e.code ? e.code : 1;
    } catch (const std::exception& e) {
        g_last_error = e.what();
        if (g_error_mode == 0) { std::fprintf(stderr, "nmf_b200: %s\n", e.what()); std::exit(1); }
        return 1;
    }
}

int env_int(const char* name, int dflt) {
    const char* v = std::getenv(name);
    return (v && *v) ? std::atoi(v) : dflt;
}

nmf_options resolve(const nmf_options* opt) {
    nmf_options o;
    if (opt) o = *opt; else nmf_default_options(&o);
    return o;
}

void check_shapes(const matrix& W, const matrix& H, const matrix& X) {
    NMF_CHECK(W.mat && H.mat && X.mat, "update_div: NULL matrix data");
    // the reference's wrappers exit on mismatching dims (e.g. cuda/matrix.cu:130-134)
    NMF_CHECK(W.dim[0] == X.dim[0] && H.dim[1] == X.dim[1] && W.dim[1] == H.dim[0],
              "update_div: dimensions do not agree (need W MxK, H KxN, X MxN)");
}

}  // namespace

extern "C" {

const char* nmf_version(void) { return "nmf-b200 0.1 (sm_100a)"; }
int nmf_set_error_mode(int mode) { const int old = g_error_mode; g_error_mode = mode ? 1 : 0; return old; }
const char* nmf_last_error(void) { return g_last_error.c_str(); }

void nmf_default_options(nmf_options* o) {
    std::memset(o, 0, sizeof(*o));
    o->semantics = NMF_EXACT;
    o->engine = NMF_ENGINE_TCGEN05;
    o->device = -1;
    o->check_every = kCheckEveryDefault;
    o->use_graph = 1;
    o->tf32_passes = 3;
    if (const char* v = std::getenv("NMF_SEMANTICS")) {
        if (!std::strcmp(v, "refcompat") || !std::strcmp(v, "1")) o->semantics = NMF_REFCOMPAT;
    }
    if (const char* v = std::getenv("NMF_ENGINE")) {
        if (!std::strcmp(v, "simt") || !std::strcmp(v, "1")) o->engine = NMF_ENGINE_SIMT;
    }
    o->device = env_int("NMF_DEVICE", o->device);
    o->check_every = env_int("NMF_CHECK_EVERY", o->check_every);
    o->use_graph = env_int("NMF_GRAPH", o->use_graph);
    o->tf32_passes = env_int("NMF_TF32_PASSES", o->tf32_passes);
}

// ---------------------------------------------------------------------------------------------
// Part 1: file I/O and update_div
// ---------------------------------------------------------------------------------------------
void create_matrix(matrix* A, int rows, int cols, float value) {
    guarded([&] {
        NMF_CHECK(A && rows > 0 && cols > 0, "create_matrix: bad arguments");
        A->dim[0] = rows; A->dim[1] = cols; A->mat_d = nullptr;
        A->mat = (float*)std::malloc(sizeof(float) * (size_t)rows * cols);
        NMF_CHECK(A->mat != nullptr, "create_matrix: out of memory");
        for (size_t i = 0; i < (size_t)rows * cols; ++i) A->mat[i] = value;
    });
}

void destroy_matrix(matrix* A) {
    if (!A) return;
    std::free(A->mat);
    if (A->mat_d) cudaFree(A->mat_d);
    A->mat = nullptr; A->mat_d = nullptr; A->dim[0] = A->dim[1] = 0;
}

void read_matrix(matrix* A, const char* file) {
    guarded([&] {
        NMF_CHECK(A && file, "read_matrix: bad arguments");
        FILE* fp = std::fopen(file, "rb");
        NMF_CHECK(fp != nullptr, std::string("read_matrix: cannot open ") + file);   // reference leaves this unchecked
        int32_t hdr[2] = {0, 0};
        if (std::fread(hdr, sizeof(int32_t), 2, fp) < 2) { std::fclose(fp); NMF_FAIL("read_matrix: fread error"); }
        NMF_CHECK(hdr[0] > 0 && hdr[1] > 0, "read_matrix: bad header");
        const size_t n = (size_t)hdr[0] * (size_t)hdr[1];
        float* buf = (float*)std::malloc(sizeof(float) * n);
        if (!buf) { std::fclose(fp); NMF_FAIL("read_matrix: out of memory"); }
        const size_t got = std::fread(buf, sizeof(float), n, fp);
        std::fclose(fp);
        if (got < n) { std::free(buf); NMF_FAIL("read_matrix: fread error"); }
        for (size_t i = 0; i < n; ++i) if (buf[i] < NMF_EPS_F) buf[i] = NMF_EPS_F;   // set_epsilon after load, nmf.cu:210-211
        A->mat = buf; A->mat_d = nullptr; A->dim[0] = hdr[0]; A->dim[1] = hdr[1];
        std::printf("read %s [%ix%i]\n", file, ref_pad32(hdr[0]), ref_pad32(hdr[1]));   // padded dims, nmf.cu:215
    });
}

void write_matrix(matrix A, const char* file) {
    guarded([&] {
        NMF_CHECK(A.mat && file, "write_matrix: bad arguments");
        FILE* fp = std::fopen(file, "wb");
        NMF_CHECK(fp != nullptr, std::string("write_matrix: cannot open ") + file);
        const int32_t hdr[2] = {A.dim[0], A.dim[1]};
        const size_t n = (size_t)A.dim[0] * (size_t)A.dim[1];
        const bool ok = std::fwrite(hdr, sizeof(int32_t), 2, fp) == 2 && std::fwrite(A.mat, sizeof(float), n, fp) == n;
        std::fclose(fp);
        NMF_CHECK(ok, "write_matrix: fwrite error");
        std::printf("write %s [%ix%i]\n", file, A.dim[0], A.dim[1]);                     // nmf.cu:258
    });
}

namespace {

// contiguous column blocks, dealt out in tile-wide (256-column) units as evenly as possible
// (the rule of nmf_gpu_b200/sharding.py)
void column_shard(long long N, int rank, int world, long long* first, long long* count) {
    const long long mult = 256, nblocks = (N + mult - 1) / mult, base = nblocks / world, extra = nblocks % world;
    const long long f = rank * base + std::min<long long>(rank, extra), c = base + (rank < extra ? 1 : 0);
    *first = std::min(N, f * mult);
    *count = std::min(N, (f + c) * mult) - *first;
}

// update_div on g GPUs of this box from ONE process: a host thread per GPU, X and H sharded by columns,
// W replicated, one NCCL allreduce per iteration — what `NMF_NUM_GPUS=g ./nmf` runs.
// Every rank gets the same thresh / max_iter / verbose (the collective schedule depends on them; only rank 0 prints).
// Failure handling: the shards are validated here, before any thread exists; sessions are created and filled in a first
// phase that contains no collective; the communicators are made in one ncclCommInitAll call on this thread; and a rank
// that fails inside the iteration aborts every communicator so that its peers return instead of waiting for it.
void update_div_multi(matrix W, matrix H, matrix X, float thresh, int max_iter, double t[10], int verbose,
                      const nmf_options& o, int g, nmf_result* res) {
    const int M = W.dim[0], K = W.dim[1];
    const long long N = H.dim[1];
    const int dev0 = o.device >= 0 ? o.device : 0;
    std::vector<long long> first((size_t)g), count((size_t)g);
    for (int r = 0; r < g; ++r) {
        column_shard(N, r, g, &first[(size_t)r], &count[(size_t)r]);
        NMF_CHECK(count[(size_t)r] > 0, "update_div: internal error: empty column shard");
    }
    std::vector<std::unique_ptr<nmf_session_impl>> sess((size_t)g);
    std::vector<std::string> errors((size_t)g);
    std::vector<nmf_result> results((size_t)g);
    auto run_phase = [&](const std::function<void(int)>& body, bool abort_peers) {
        std::vector<std::thread> threads;
        for (int r = 0; r < g; ++r) {
            threads.emplace_back([&, r] {
                try { body(r); }
                catch (const std::exception& e) {
                    errors[(size_t)r] = e.what();
                    if (errors[(size_t)r].empty()) errors[(size_t)r] = "unknown error";
                    if (abort_peers) for (int q = 0; q < g; ++q) if (sess[(size_t)q]) nccl_abort(*sess[(size_t)q]);
                }
            });
        }
        for (auto& th : threads) th.join();
        for (int r = 0; r < g; ++r) NMF_CHECK(errors[(size_t)r].empty(), "GPU " + std::to_string(dev0 + r) + ": " + errors[(size_t)r]);
    };
    run_phase([&](int r) {
        nmf_options orank = o;
        orank.device = dev0 + r;
        sess[(size_t)r].reset(session_create(M, (int)count[(size_t)r], K, first[(size_t)r], N, orank));
        session_upload(*sess[(size_t)r], W.mat, H.mat + (size_t)first[(size_t)r] * K, X.mat + (size_t)first[(size_t)r] * M);
    }, false);
    {
        std::vector<nmf_session_impl*> raw;
        for (auto& p : sess) raw.push_back(p.get());
        nccl_init_all(raw.data(), g);
    }
    run_phase([&](int r) {
        session_run(*sess[(size_t)r], thresh, max_iter, r == 0 ? t : nullptr, verbose, &results[(size_t)r]);
        session_download(*sess[(size_t)r], r == 0 ? W.mat : nullptr, H.mat + (size_t)first[(size_t)r] * K);
    }, true);
    if (res) *res = results[0];
}

// NMF_NUM_GPUS, limited to the 256-column blocks there are to hand out (a GPU without columns has nothing to do)
int resolve_num_gpus(long long N) {
    const char* v = std::getenv("NMF_NUM_GPUS");
    const int want = (v && *v) ? std::max(1, std::atoi(v)) : 1;
    const long long nblocks = (N + 255) / 256;
    return (int)std::min<long long>(want, std::max<long long>(1, nblocks));
}

}  // namespace

int nmf_update_div_ex(matrix W, matrix H, matrix X, float CONVERGE_THRESH, int max_iter, double t[10],
                      int verbose, const nmf_options* opt, nmf_result* res) {
    return guarded([&] {
        using clk = std::chrono::steady_clock;
        const auto w0 = clk::now();
        check_shapes(W, H, X);
        NMF_CHECK(max_iter >= 0, "update_div: max_iter must be >= 0");
        const nmf_options o = resolve(opt);
        // the GPU count is not in the reference's signature (README.md:39-47): NMF_NUM_GPUS selects it
        const int g = resolve_num_gpus(H.dim[1]);
        if (g > 1) {
            int ndev = 0;
            CUDA_CHECK(cudaGetDeviceCount(&ndev));
            NMF_CHECK((o.device >= 0 ? o.device : 0) + g <= ndev, "NMF_NUM_GPUS exceeds the visible devices");
            update_div_multi(W, H, X, CONVERGE_THRESH, max_iter, t, verbose, o, g, res);
            if (t) t[PH_TOTAL] = std::chrono::duration<double>(clk::now() - w0).count();
            return;
        }
        nmf_options o1 = o; o1.restarts = 1;
        std::unique_ptr<nmf_session_impl> s(session_create(W.dim[0], H.dim[1], W.dim[1], 0, H.dim[1], o1));
        session_upload(*s, W.mat, H.mat, X.mat);
        const double setup = std::chrono::duration<double>(clk::now() - w0).count();
        session_run(*s, CONVERGE_THRESH, max_iter, t, verbose, res);
        const auto d0 = clk::now();
        session_download(*s, W.mat, H.mat);
        if (t) {
            t[PH_SETUP] = setup + std::chrono::duration<double>(clk::now() - d0).count();
            t[PH_TOTAL] = std::chrono::duration<double>(clk::now() - w0).count();
        }
    });
}

int nmf_update_div_batch(matrix* W, matrix* H, int restarts, matrix X, float CONVERGE_THRESH, int max_iter,
                         double t[10], int verbose, const nmf_options* opt, int* iterations, int* converged, double* cost) {
    int best = -1;
    const int rc = guarded([&] {
        using clk = std::chrono::steady_clock;
        const auto w0 = clk::now();
        NMF_CHECK(W && H && restarts >= 1 && restarts <= 64, "update_div_batch: need 1..64 restarts");
        for (int r = 0; r < restarts; ++r) {
            check_shapes(W[r], H[r], X);
            NMF_CHECK(W[r].dim[1] == W[0].dim[1], "update_div_batch: every restart must have the same K");
        }
        NMF_CHECK(max_iter >= 0, "update_div: max_iter must be >= 0");
        nmf_options o = resolve(opt);
        const int M = X.dim[0], N = X.dim[1], K = W[0].dim[1];
        std::vector<int> its((size_t)restarts, 0), conv((size_t)restarts, 0);
        std::vector<double> costs((size_t)restarts, std::nan(""));
        if (t) for (int i = 0; i < 10; ++i) t[i] = 0.0;
        // the stacked path needs both half-iterations fused; anything else runs the restarts one after another
        bool stacked = restarts > 1 && o.engine == NMF_ENGINE_TCGEN05 && o.semantics == NMF_EXACT && o.tf32_passes == 3;
        if (stacked) {
            int ndev = 0; CUDA_CHECK(cudaGetDeviceCount(&ndev));
            NMF_CHECK(ndev > 0, "no CUDA device: libnmf_b200 has no CPU fallback");
            if (o.device >= 0) CUDA_CHECK(cudaSetDevice(o.device));
            const umma::Tuning tn = umma::read_tuning();
            const Dims d = make_dims(M, N, K, 0, N, o.semantics, restarts, tn.pad64 != 0);
            stacked = umma::fused_both_wanted(d, o.tf32_passes, tn);
        }
        if (stacked) {
            o.restarts = restarts;
            std::unique_ptr<nmf_session_impl> s(session_create(M, N, K, 0, N, o));
            session_upload(*s, nullptr, nullptr, X.mat);
            for (int r = 0; r < restarts; ++r) session_upload_restart(*s, r, W[r].mat, H[r].mat);
            const double setup = std::chrono::duration<double>(clk::now() - w0).count();
            nmf_result res{};
            session_run(*s, CONVERGE_THRESH, max_iter, t, verbose, &res);
            if (std::isnan(s->costs[0])) session_cost(*s, s->costs.data());     // keep-best needs the costs even when no check ran
            const auto d0 = clk::now();
            for (int r = 0; r < restarts; ++r) {
                session_download_restart(*s, r, W[r].mat, H[r].mat);
                its[(size_t)r] = s->iters_r[(size_t)r]; conv[(size_t)r] = s->conv_r[(size_t)r]; costs[(size_t)r] = s->costs[(size_t)r];
            }
            if (t) t[PH_SETUP] = setup + std::chrono::duration<double>(clk::now() - d0).count();
        } else {
            o.restarts = 1;
            std::unique_ptr<nmf_session_impl> s(session_create(M, N, K, 0, N, o));
            session_upload(*s, nullptr, nullptr, X.mat);                         // X goes up once
            for (int r = 0; r < restarts; ++r) {
                double tr[10];
                session_upload(*s, W[r].mat, H[r].mat, nullptr);
                nmf_result res{};
                session_run(*s, CONVERGE_THRESH, max_iter, t ? tr : nullptr, verbose, &res);
                if (std::isnan(res.cost)) session_cost(*s, &res.cost);
                session_download(*s, W[r].mat, H[r].mat);
                its[(size_t)r] = res.iterations; conv[(size_t)r] = res.converged; costs[(size_t)r] = res.cost;
                if (t) for (int i = 0; i < 8; ++i) t[i] += tr[i];
            }
        }
        if (t) t[PH_TOTAL] = std::chrono::duration<double>(clk::now() - w0).count();
        best = 0;
        for (int r = 1; r < restarts; ++r) if (costs[(size_t)r] < costs[(size_t)best]) best = r;
        for (int r = 0; r < restarts; ++r) {
            if (iterations) iterations[r] = its[(size_t)r];
            if (converged) converged[r] = conv[(size_t)r];
            if (cost) cost[r] = costs[(size_t)r];
        }
    });
    return rc == 0 ? best : -1;
}

void update_div(matrix W, matrix H, matrix X, float CONVERGE_THRESH, int max_iter, double t[10], int verbose) {
    nmf_update_div_ex(W, H, X, CONVERGE_THRESH, max_iter, t, verbose, nullptr, nullptr);
}

// ---------------------------------------------------------------------------------------------
// Part 2: sessions
// ---------------------------------------------------------------------------------------------
nmf_session* nmf_session_create(int M, int n_local, int K, long long n_offset, long long n_global,
                                const nmf_options* opt) {
    nmf_session* out = nullptr;
    guarded([&] { out = static_cast<nmf_session*>(session_create(M, n_local, K, n_offset, n_global, resolve(opt))); });
    return out;
}

void nmf_session_destroy(nmf_session* s) { delete s; }

int nmf_session_upload(nmf_session* s, const float* W, const float* H, const float* X) {
    return guarded([&] { NMF_CHECK(s, "NULL session"); session_upload(*s, W, H, X); });
}
int nmf_session_download(nmf_session* s, float* W, float* H) {
    return guarded([&] { NMF_CHECK(s, "NULL session"); session_download(*s, W, H); });
}
int nmf_session_upload_restart(nmf_session* s, int r, const float* W, const float* H) {
    return guarded([&] { NMF_CHECK(s, "NULL session"); session_upload_restart(*s, r, W, H); });
}
int nmf_session_download_restart(nmf_session* s, int r, float* W, float* H) {
    return guarded([&] { NMF_CHECK(s, "NULL session"); session_download_restart(*s, r, W, H); });
}
int nmf_session_restart_results(const nmf_session* s, int* iterations, int* converged, double* cost) {
    return guarded([&] {
        NMF_CHECK(s, "NULL session");
        for (int r = 0; r < s->d.R; ++r) {
            if (iterations) iterations[r] = s->iters_r[(size_t)r];
            if (converged) converged[r] = s->conv_r[(size_t)r];
            if (cost) cost[r] = s->costs[(size_t)r];
        }
    });
}
int nmf_session_download_rows(nmf_session* s, const int* rows, int count, float* Xr, float* Wr) {
    return guarded([&] { NMF_CHECK(s && rows, "bad arguments"); session_download_rows(*s, rows, count, Xr, Wr); });
}
int nmf_session_half_step(nmf_session* s, int which) {
    return guarded([&] { NMF_CHECK(s && (which == 0 || which == 1), "bad arguments"); session_half_step(*s, which); });
}
int nmf_session_download_cols(nmf_session* s, const long long* cols, int count, float* Xc, float* Hc) {
    return guarded([&] { NMF_CHECK(s && cols && count >= 0, "bad arguments"); session_download_cols(*s, cols, count, Xc, Hc); });
}
int nmf_session_fill_uniform(nmf_session* s, unsigned long long seed, int which) {
    return guarded([&] { NMF_CHECK(s, "NULL session"); session_fill_uniform(*s, seed, which); });
}
int nmf_session_fill_planted(nmf_session* s, unsigned long long seed, int rank, float noise) {
    return guarded([&] { NMF_CHECK(s, "NULL session"); session_fill_planted(*s, seed, rank, noise); });
}

int nmf_session_share_comm(nmf_session* dst, nmf_session* src) {
    return guarded([&] { NMF_CHECK(dst && src, "NULL session"); nccl_share(*dst, *src); });
}

int nmf_session_fused_steps(const nmf_session* s) {
    if (!s || s->opt.engine != NMF_ENGINE_TCGEN05) return 0;
    return (umma::fused_w_enabled(*s) ? 1 : 0) | (umma::fused_h_enabled(*s) ? 2 : 0);
}

int nmf_session_corr_bf16(const nmf_session* s) { return (s && s->Wbf != nullptr) ? 1 : 0; }

int nmf_nccl_unique_id(void* id128) { return guarded([&] { nccl_unique_id(id128); }); }
int nmf_session_join(nmf_session* s, int rank, int nranks, const void* id128) {
    return guarded([&] { NMF_CHECK(s, "NULL session"); nccl_join(*s, rank, nranks, id128); });
}
int nmf_session_run(nmf_session* s, float CONVERGE_THRESH, int max_iter, double t[10], int verbose, nmf_result* res) {
    return guarded([&] { NMF_CHECK(s, "NULL session"); session_run(*s, CONVERGE_THRESH, max_iter, t, verbose, res); });
}
int nmf_session_cost(nmf_session* s, double* cost) {
    return guarded([&] { NMF_CHECK(s && cost, "NULL argument"); session_cost(*s, cost); });
}

// ---------------------------------------------------------------------------------------------
// Part 3: kernel-level entry points for the parity tests
// ---------------------------------------------------------------------------------------------
static void with_session(const float* W, const float* H, const float* X, int M, int N, int K, const nmf_options* opt,
                         const std::function<void(nmf_session_impl&)>& body) {
    nmf_options o = resolve(opt);
    o.use_graph = 0;
    o.restarts = 1;
    std::unique_ptr<nmf_session_impl> s(session_create(M, N, K, 0, N, o));
    session_upload(*s, W, H, X);
    body(*s);
    CUDA_CHECK(cudaStreamSynchronize(s->stream));
}

int nmf_dbg_quotient(const float* W, const float* H, const float* X, float* Y, double* cost,
                     int M, int N, int K, const nmf_options* opt) {
    return guarded([&] {
        with_session(W, H, X, M, N, K, opt, [&](nmf_session_impl& s) {
            session_need_quotient(s);        // a session whose half-iterations are both fused keeps no quotient buffer
            iteration_head(s, cost != nullptr, nullptr);
            const Dims& d = s.d;
            std::vector<float> a((size_t)d.Mv * d.Nv), b;
            CUDA_CHECK(cudaMemcpyAsync(a.data(), s.Y, a.size() * 4, cudaMemcpyDeviceToHost, s.stream));
            if (cost) CUDA_CHECK(cudaMemcpyAsync(s.cost_host, s.cost_dev, 8, cudaMemcpyDeviceToHost, s.stream));
            CUDA_CHECK(cudaStreamSynchronize(s.stream));
            if (cost) *cost = *s.cost_host;
            for (int n = 0; n < N; ++n)
                for (int m = 0; m < M; ++m) {
                    const size_t i = (size_t)n * d.Mv + m;
                    Y[(size_t)n * M + m] = b.empty() ? a[i] : a[i] + b[i];
                }
        });
    });
}

int nmf_dbg_h_step(const float* W, float* H, const float* X, int M, int N, int K, const nmf_options* opt) {
    return guarded([&] {
        with_session(W, H, X, M, N, K, opt, [&](nmf_session_impl& s) {
            session_half_step(s, 0);              // exactly the H half of an iteration (column sums, quotient, W'Y, update)
            session_download(s, nullptr, H);
        });
    });
}

int nmf_dbg_w_step(float* W, const float* H, const float* X, int M, int N, int K, const nmf_options* opt) {
    return guarded([&] {
        with_session(W, H, X, M, N, K, opt, [&](nmf_session_impl& s) {
            session_half_step(s, 1);              // exactly the W half of an iteration, starting from the given H
            session_download(s, W, nullptr);
        });
    });
}

int nmf_dbg_sums(const float* W, const float* H, float* sumW, float* sumH, int M, int N, int K, const nmf_options* opt) {
    return guarded([&] {
        std::vector<float> X((size_t)M * N, 1.0f);
        with_session(W, H, X.data(), M, N, K, opt, [&](nmf_session_impl& s) {
            launch_colsum(s.W, s.sumW, s.d, s.stream);
            launch_rowsum(s.H, s.rowsum_partials, s.sumH, s.d, s.stream);
            std::vector<float> a(s.d.Kv), b(s.d.Kv);
            CUDA_CHECK(cudaMemcpyAsync(a.data(), s.sumW, a.size() * 4, cudaMemcpyDeviceToHost, s.stream));
            CUDA_CHECK(cudaMemcpyAsync(b.data(), s.sumH, b.size() * 4, cudaMemcpyDeviceToHost, s.stream));
            CUDA_CHECK(cudaStreamSynchronize(s.stream));
            for (int k = 0; k < K; ++k) {
                sumW[k] = a[k] < NMF_EPS_F ? NMF_EPS_F : a[k];   // set_epsilon, nmf.cu:135,165
                sumH[k] = b[k] < NMF_EPS_F ? NMF_EPS_F : b[k];
            }
        });
    });
}

int nmf_rowsum_noalloc(const float* H_dev, int K, int N, float* sumH_dev, void* stream) {
    return guarded([&] { rowsum_noalloc(H_dev, K, N, sumH_dev, (cudaStream_t)stream); });
}

int nmf_dbg_gemm(int which, const float* A, const float* B, float* C, int M, int N, int K, int passes,
                 const nmf_options* opt) {
    return guarded([&] {
        NMF_CHECK(which >= 0 && which <= 2, "nmf_dbg_gemm: which must be 0..2");
        NMF_CHECK(M % 32 == 0 && K % 32 == 0 && N % 4 == 0, "nmf_dbg_gemm: needs M,K multiples of 32 and N of 4");
        const nmf_options o = resolve(opt);
        if (o.device >= 0) CUDA_CHECK(cudaSetDevice(o.device));
        // operand shapes (column-major): 0: A MxK, B KxN, C MxN; 1: A MxK, B MxN, C KxN; 2: A MxN, B KxN, C MxK
        const size_t nA = (which == 2) ? (size_t)M * N : (size_t)M * K;
        const size_t nB = (which == 1) ? (size_t)M * N : (size_t)K * N;
        const size_t nC = (which == 0) ? (size_t)M * N : (which == 1 ? (size_t)K * N : (size_t)M * K);
        float *dA = nullptr, *dB = nullptr, *dC = nullptr;
        CUDA_CHECK(cudaMalloc((void**)&dA, nA * 4)); CUDA_CHECK(cudaMalloc((void**)&dB, nB * 4));
        CUDA_CHECK(cudaMalloc((void**)&dC, nC * 4));
        try {
            CUDA_CHECK(cudaMemcpy(dA, A, nA * 4, cudaMemcpyHostToDevice));
            CUDA_CHECK(cudaMemcpy(dB, B, nB * 4, cudaMemcpyHostToDevice));
            CUDA_CHECK(cudaMemset(dC, 0xFF, nC * 4));
            umma::dbg_gemm(which, dA, dB, dC, M, N, K, passes, nullptr);
            CUDA_CHECK(cudaDeviceSynchronize());
            CUDA_CHECK(cudaMemcpy(C, dC, nC * 4, cudaMemcpyDeviceToHost));
        } catch (...) { cudaFree(dA); cudaFree(dB); cudaFree(dC); throw; }
        cudaFree(dA); cudaFree(dB); cudaFree(dC);
    });
}

}  // extern "C"
