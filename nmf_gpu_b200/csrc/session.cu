// session.cu — device buffers of one rank's shard and the iteration driver that replaces the
// reference's run_async / update_h / update_w (cuda/nmf.cu:76-176).
#include "session.hpp"

#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <memory>
#include <vector>

namespace nmf {

// ---------------------------------------------------------------------------------------------
// error convention
// ---------------------------------------------------------------------------------------------
[[noreturn]] void fail(const char* file, int line, const std::string& msg, int code) {
    const char* base = std::strrchr(file, '/');
    throw Error(std::string(base ? base + 1 : file) + ":" + std::to_string(line) + ": " + msg, code ? code : 1);
}

Dims make_dims(int M, int n_local, int K, long long n_offset, long long n_global, int semantics, int restarts, bool pad_k64) {
    NMF_CHECK(M > 0 && n_local > 0 && K > 0, "matrix dimensions must be positive");
    NMF_CHECK(restarts >= 1 && restarts <= 64, "restarts must be in 1..64");
    Dims d;
    d.M = M; d.N = n_local; d.K = K; d.R = restarts;
    d.n_offset = n_offset; d.n_global = n_global > 0 ? n_global : n_local;
    d.refcompat = (semantics == NMF_REFCOMPAT);
    if (d.refcompat) {
        NMF_CHECK(M != 1 && K != 1, "refcompat semantics: dimensions equal to 1 are not padded by the "
                                    "reference (cuda/matrix.cu:88-95) and are not supported here");
        NMF_CHECK(restarts == 1, "refcompat semantics: batched restarts are not supported");
        d.Mv = ref_pad32(M); d.Kr = ref_pad32(K);
        // the reference pads the GLOBAL column count; only the last shard carries that padding
        const bool last = (n_offset + n_local >= d.n_global);
        const long long gpad = (d.n_global != 1 && d.n_global % 32 != 0) ? 32 - d.n_global % 32 : 0;
        d.Nv = n_local + (last ? (int)gpad : 0);
        d.w_divide = (d.Mv <= 1024) ? 1 : 0;      // row_div<<<N, rows_padded>>>, cuda/matrix.cu:212-217
    } else {
        // K padded to 64 (zero fill: rows/columns of zeros are fixed points of the update) when that lets the
        // half-iterations run as one fused kernel each (session_create decides)
        d.Mv = round_up(M, 32); d.Kr = round_up(K, pad_k64 ? 64 : 32); d.Nv = n_local;
        d.w_divide = 1;
    }
    d.Kv = d.R * d.Kr;
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

// zero the padding of a column-major buffer: rows [rows, rows_v) of the first cols columns, and columns [cols, cols_alloc)
void zero_padding(float* a, int rows, int cols, int rows_v, int cols_alloc, cudaStream_t st) {
    if (rows_v > rows && cols > 0)
        CUDA_CHECK(cudaMemset2DAsync(a + rows, (size_t)rows_v * 4, 0, (size_t)(rows_v - rows) * 4, (size_t)cols, st));
    if (cols_alloc > cols)
        CUDA_CHECK(cudaMemsetAsync(a + (size_t)rows_v * cols, 0, (size_t)rows_v * (cols_alloc - cols) * 4, st));
}

}  // namespace

nmf_session_impl::~nmf_session_impl() {
    if (g_full) cudaGraphExecDestroy(g_full);
    if (g_rest) cudaGraphExecDestroy(g_rest);
    if (plan) umma::plan_destroy(plan);
    if (nccl) nccl_destroy(nccl);
    float* fl[] = {W, H, X, Y, Whi, Wlo, Hhi, Hlo, sumW, packed, Psplit, Hpart, rowsum_partials, Hbak, Wfin, Hfin, tail_part};
    for (float* p : fl) if (p) cudaFree(p);
    unsigned short* bl[] = {Wbf, Wlb, Hbf, Hlb};
    for (unsigned short* p : bl) if (p) cudaFree(p);
    if (cost_partials) cudaFree(cost_partials);
    if (cost_dev) cudaFree(cost_dev);
    if (cost_host) cudaFreeHost(cost_host);
    if (copy_stream) cudaStreamDestroy(copy_stream);
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
        const int R = std::max(1, opt.restarts);
        const bool tc = (opt.engine == NMF_ENGINE_TCGEN05);
        if (tc) s->tune = umma::read_tuning();                // every NMF_UMMA_* switch, once per session
        const umma::Tuning& tn = s->tune;
        NMF_CHECK(R == 1 || tc, "restarts > 1 need the tcgen05 engine");
        // exact semantics: K padded to a multiple of 64 when that is what lets both half-iterations fuse
        const bool pad64 = tc && opt.semantics != NMF_REFCOMPAT && tn.pad64 != 0 && round_up(K, 64) != round_up(K, 32) &&
                           umma::fused_both_wanted(make_dims(M, n_local, K, n_offset, n_global, opt.semantics, R, true), opt.tf32_passes, tn);
        s->d = make_dims(M, n_local, K, n_offset, n_global, opt.semantics, R, pad64);
        const Dims& d = s->d;
        NMF_CHECK(R == 1 || umma::fused_both_wanted(d, opt.tf32_passes, tn),
                  "restarts > 1 run on the fused half-iteration kernels only: need K <= 256, more than 128 rows and columns, 3xTF32");
        CUDA_CHECK(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
        s->W = dalloc<float>(d.szW()); s->H = dalloc<float>(d.szHa()); s->X = dalloc<float>(d.szXa());
        // zero fill: the padding (rows M..Mv, k K..Kr, columns Nv..Nva) must be zero and stays zero (fixed points of the
        // update; no kernel stores there).  X: only its padding is cleared — the payload is written by upload / fill
        CUDA_CHECK(cudaMemsetAsync(s->W, 0, d.szW() * 4, s->stream));
        CUDA_CHECK(cudaMemsetAsync(s->H, 0, d.szHa() * 4, s->stream));
        zero_padding(s->X, d.M, d.Nv, d.Mv, d.Nva, s->stream);
        s->sumW = dalloc<float>(d.Kv);
        s->packed = dalloc<float>(d.szW() + d.Kv);
        s->sumH = s->packed + d.szW();
        s->rowsum_partials = dalloc<float>((size_t)rowsum_partial_count(d) * d.Kv);
        s->cost_dev = dalloc<double>(R);
        CUDA_CHECK(cudaMallocHost((void**)&s->cost_host, sizeof(double) * R));
        s->costs.assign(R, std::nan("")); s->iters_r.assign(R, 0); s->conv_r.assign(R, 0); s->retired.assign(R, 0);
        if (opt.engine == NMF_ENGINE_SIMT) {
            s->Y = dalloc<float>(d.szXa());
            s->n_cost_partials = simt_cost_partial_count(d);
            s->nsplit = 1;
        } else {
            // stored once as fp32; the GEMMs that consume it split it in the SM.  Not needed at all when both
            // half-iterations are fused (NMF_KEEP_QUOTIENT=1 allocates it anyway)
            if (!umma::fused_both_wanted(d, opt.tf32_passes, tn) || (tn.keep_quotient && R == 1)) s->Y = dalloc<float>(d.szXa());
            s->Whi = dalloc<float>(d.szW()); s->Wlo = dalloc<float>(d.szW());
            s->Hhi = dalloc<float>(d.szHa()); s->Hlo = dalloc<float>(d.szHa());
            if (umma::corr_bf16_wanted(d, opt.tf32_passes, tn) || umma::corr_bf16_wanted_2k(d, opt.tf32_passes, tn)) {
                s->Wbf = dalloc<unsigned short>(d.szW()); s->Wlb = dalloc<unsigned short>(d.szW());
                s->Hbf = dalloc<unsigned short>(d.szHa()); s->Hlb = dalloc<unsigned short>(d.szHa());
                CUDA_CHECK(cudaMemsetAsync(s->Hbf, 0, d.szHa() * 2, s->stream));      // the padding columns stay zero
                CUDA_CHECK(cudaMemsetAsync(s->Hlb, 0, d.szHa() * 2, s->stream));
            }
            s->n_cost_partials = umma::cost_partial_count(d, tn);
            s->n_cost_partials_fused = umma::fused_cost_partial_count(d, opt.tf32_passes, tn);
            // W step: the (at most 4) rows beyond the last whole 256-row block leave the fused kernel (common.hpp, launch_w_tail)
            {
                const int q = d.M / 256, t = d.M - 256 * q;
                // measured (profiles/r2_w_tail_ab.txt): pays at C3 (10^6 columns: the fused kernel's wasted row block costs more than
                // the extra kernel), loses at C2 (20000 columns: two more launches and a latency-bound kernel on 40 blocks):
                // on from 65536 columns (NMF_W_TAIL=2 forces it, =0 never)
                const bool pays = tn.w_tail >= 2 || d.Nv >= 65536;
                if (tn.w_tail != 0 && pays && !d.refcompat && q >= 1 && t >= 1 && t <= 4 && umma::fused_both_wanted(d, opt.tf32_passes, tn)) {
                    s->tail_t = t; s->tail_row0 = 256 * q;
                    s->tail_part = dalloc<float>((size_t)w_tail_blocks(d.Nv) * d.Kv * 4);
                }
            }
            s->nsplit = umma::numer_splits(d, tn, s->tail_t ? s->tail_row0 / 256 : 0);
            if (s->nsplit > 1) s->Psplit = dalloc<float>((size_t)s->nsplit * d.szW());
            s->hsplit = umma::updh_splits(d, tn);
            s->fhsplit = umma::fused_h_splits(d, opt.tf32_passes, tn);
            if (std::max(s->hsplit, s->fhsplit) > 1 || umma::fused_h_needs_part(d, opt.tf32_passes, tn))
                s->Hpart = dalloc<float>((size_t)std::max(s->hsplit, s->fhsplit) * d.szHa());
            s->plan = umma::plan_create(*s);
        }
        s->cost_partials = dalloc<double>(std::max(1, std::max(s->n_cost_partials, s->n_cost_partials_fused * R)));
        CUDA_CHECK(cudaStreamSynchronize(s->stream));
    } catch (...) {
        delete s;
        throw;
    }
    return s;
}

void session_need_quotient(nmf_session_impl& s) {
    if (s.Y) return;
    NMF_CHECK(s.d.R == 1, "the quotient buffer is not defined for restarts > 1");
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
    if (w) s.sumW_valid = false;          // every path that rewrites W from outside the iteration comes through here
    if (s.opt.engine != NMF_ENGINE_TCGEN05) return;
    if (w) { launch_split_tf32(s.W, s.Whi, s.Wlo, s.d.szW(), s.stream, s.Wbf, s.Wlb); }
    if (h) { launch_split_tf32(s.H, s.Hhi, s.Hlo, s.d.szH(), s.stream, s.Hbf, s.Hlb); }
}

// Host (unpadded, column-major, rows x cols) -> a block of a padded device buffer (leading dimension ld) + clamp, as
// read_matrix + Matrix(host,...) do (cuda/nmf.cu:208-211, cuda/matrix.cu:53-67).  exact: the padding is zero since
// session_create and no kernel writes it, so nothing but the payload moves.  refcompat: the reference's clamp runs over
// the whole padded buffer (nmf.cu:211) and its padding evolves with the iteration, so the padding is reset to EPS.
static void upload_block(nmf_session_impl& s, float* dst, int ld, const float* src, int rows, int cols) {
    CUDA_CHECK(cudaMemcpy2DAsync(dst, (size_t)ld * 4, src, (size_t)rows * 4, (size_t)rows * 4, cols,
                                 cudaMemcpyHostToDevice, s.stream));
    launch_clamp_eps(dst, rows, cols, ld, s.stream);
}
static void upload_one(nmf_session_impl& s, float* dst, const float* src, int rows, int cols, int rows_v, int cols_v) {
    if (s.d.refcompat) {
        CUDA_CHECK(cudaMemsetAsync(dst, 0, (size_t)rows_v * cols_v * 4, s.stream));
        CUDA_CHECK(cudaMemcpy2DAsync(dst, (size_t)rows_v * 4, src, (size_t)rows * 4, (size_t)rows * 4, cols,
                                     cudaMemcpyHostToDevice, s.stream));
        launch_clamp_eps(dst, rows_v, cols_v, rows_v, s.stream);                // padding becomes EPS
    } else upload_block(s, dst, rows_v, src, rows, cols);
}

// X goes up in column blocks on a copy stream, each followed by its clamp on the compute stream: the clamp of block b
// runs while block b + 1 is on the wire (one cudaMemcpy2DAsync per block)
static void upload_x(nmf_session_impl& s, const float* X) {
    const Dims& d = s.d;
    if (d.refcompat) { upload_one(s, s.X, X, d.M, d.N, d.Mv, d.Nv); return; }
    const size_t col_bytes = (size_t)d.M * 4;
    const int blk = (int)std::max<size_t>(256, std::min<size_t>((size_t)d.N, ((size_t)256 << 20) / col_bytes / 256 * 256));   // ~256 MB
    if (!s.copy_stream) CUDA_CHECK(cudaStreamCreateWithFlags(&s.copy_stream, cudaStreamNonBlocking));
    std::vector<cudaEvent_t> evs;
    try {
        for (int c0 = 0; c0 < d.N; c0 += blk) {
            const int nc = std::min(blk, d.N - c0);
            float* dst = s.X + (size_t)c0 * d.Mv;
            CUDA_CHECK(cudaMemcpy2DAsync(dst, (size_t)d.Mv * 4, X + (size_t)c0 * d.M, col_bytes, col_bytes, nc,
                                         cudaMemcpyHostToDevice, s.copy_stream));
            cudaEvent_t e; CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); evs.push_back(e);
            CUDA_CHECK(cudaEventRecord(e, s.copy_stream));
            CUDA_CHECK(cudaStreamWaitEvent(s.stream, e, 0));
            launch_clamp_eps(dst, d.M, nc, d.Mv, s.stream);
        }
        CUDA_CHECK(cudaStreamSynchronize(s.stream));
    } catch (...) { for (auto e : evs) cudaEventDestroy(e); throw; }
    for (auto e : evs) cudaEventDestroy(e);
}

void session_upload(nmf_session_impl& s, const float* W, const float* H, const float* X) {
    const Dims& d = s.d;
    NMF_CHECK(d.R == 1 || (W == nullptr && H == nullptr), "restarts > 1: upload the factors with session_upload_restart");
    CUDA_CHECK(cudaSetDevice(s.device));
    if (W) upload_one(s, s.W, W, d.M, d.K, d.Mv, d.Kv);
    if (H) upload_one(s, s.H, H, d.K, d.N, d.Kv, d.Nv);
    refresh_splits(s, W != nullptr, H != nullptr);
    if (X) upload_x(s, X);
    CUDA_CHECK(cudaStreamSynchronize(s.stream));
}

// factors of restart r (W: M x K, H: K x N, unpadded column-major) into their block of the stacked arrays
void session_upload_restart(nmf_session_impl& s, int r, const float* W, const float* H) {
    const Dims& d = s.d;
    NMF_CHECK(r >= 0 && r < d.R, "restart index out of range");
    NMF_CHECK(!d.refcompat, "refcompat semantics: batched restarts are not supported");
    CUDA_CHECK(cudaSetDevice(s.device));
    if (W) upload_block(s, s.W + (size_t)r * d.Kr * d.Mv, d.Mv, W, d.M, d.K);
    if (H) upload_block(s, s.H + (size_t)r * d.Kr, d.Kv, H, d.K, d.N);
    refresh_splits(s, W != nullptr, H != nullptr);
    CUDA_CHECK(cudaStreamSynchronize(s.stream));
}

void session_download_restart(nmf_session_impl& s, int r, float* W, float* H) {
    const Dims& d = s.d;
    NMF_CHECK(r >= 0 && r < d.R, "restart index out of range");
    CUDA_CHECK(cudaSetDevice(s.device));
    if (W) CUDA_CHECK(cudaMemcpy2DAsync(W, (size_t)d.M * 4, s.W + (size_t)r * d.Kr * d.Mv, (size_t)d.Mv * 4, (size_t)d.M * 4, d.K,
                                        cudaMemcpyDeviceToHost, s.stream));
    if (H) CUDA_CHECK(cudaMemcpy2DAsync(H, (size_t)d.K * 4, s.H + (size_t)r * d.Kr, (size_t)d.Kv * 4, (size_t)d.K * 4, d.N,
                                        cudaMemcpyDeviceToHost, s.stream));
    CUDA_CHECK(cudaStreamSynchronize(s.stream));
}

void session_download(nmf_session_impl& s, float* W, float* H) {
    NMF_CHECK(s.d.R == 1, "restarts > 1: download the factors with session_download_restart");
    session_download_restart(s, 0, W, H);
}

// `count` columns of the resident X and H (global column indices inside this shard's range, any order) -> host,
// unpadded: Xc is M x count, Hc is K x count.  What a full-size spot check compares against the oracle.
void session_download_cols(nmf_session_impl& s, const long long* cols, int count, float* Xc, float* Hc) {
    const Dims& d = s.d;
    NMF_CHECK(d.R == 1, "restarts > 1: not supported");
    CUDA_CHECK(cudaSetDevice(s.device));
    for (int i = 0; i < count; ++i) {
        const long long c = cols[i] - d.n_offset;
        NMF_CHECK(c >= 0 && c < d.N, "column outside this shard");
        if (Xc) CUDA_CHECK(cudaMemcpyAsync(Xc + (size_t)i * d.M, s.X + (size_t)c * d.Mv, (size_t)d.M * 4, cudaMemcpyDeviceToHost, s.stream));
        if (Hc) CUDA_CHECK(cudaMemcpyAsync(Hc + (size_t)i * d.K, s.H + (size_t)c * d.Kv, (size_t)d.K * 4, cudaMemcpyDeviceToHost, s.stream));
    }
    CUDA_CHECK(cudaStreamSynchronize(s.stream));
}

void session_download_rows(nmf_session_impl& s, const int* rows, int count, float* Xr, float* Wr) {
    const Dims& d = s.d;
    NMF_CHECK(d.R == 1, "restarts > 1: not supported");
    NMF_CHECK(count > 0, "no rows");
    for (int i = 0; i < count; ++i) NMF_CHECK(rows[i] >= 0 && rows[i] < d.M, "row out of range");
    CUDA_CHECK(cudaSetDevice(s.device));
    int* drows = dalloc<int>((size_t)count);
    float* tmp = dalloc<float>((size_t)count * std::max(d.N, d.K));
    try {
        CUDA_CHECK(cudaMemcpyAsync(drows, rows, sizeof(int) * count, cudaMemcpyHostToDevice, s.stream));
        if (Xr) {
            launch_gather_rows(s.X, d.Mv, d.N, drows, count, tmp, s.stream);
            CUDA_CHECK(cudaMemcpyAsync(Xr, tmp, (size_t)count * d.N * 4, cudaMemcpyDeviceToHost, s.stream));
        }
        if (Wr) {
            launch_gather_rows(s.W, d.Mv, d.K, drows, count, tmp, s.stream);
            CUDA_CHECK(cudaMemcpyAsync(Wr, tmp, (size_t)count * d.K * 4, cudaMemcpyDeviceToHost, s.stream));
        }
        CUDA_CHECK(cudaStreamSynchronize(s.stream));
    } catch (...) { cudaFree(drows); cudaFree(tmp); throw; }
    cudaFree(drows); cudaFree(tmp);
}

void session_fill_uniform(nmf_session_impl& s, unsigned long long seed, int which) {
    const Dims& d = s.d;
    CUDA_CHECK(cudaSetDevice(s.device));
    // restart r draws from seed + 7919 r, by the (row, col) index inside the restart: restart 0 of a batched session
    // starts from the factors a single session of the same seed does
    if (which & 1) {
        CUDA_CHECK(cudaMemsetAsync(s.W, 0, d.szW() * 4, s.stream));
        for (int r = 0; r < d.R; ++r)
            launch_fill_uniform(s.W + (size_t)r * d.Kr * d.Mv, d.M, d.K, d.Mv, 0, seed + 7919ull * r, 1, s.stream);
        if (d.refcompat) launch_clamp_eps(s.W, d.Mv, d.Kv, d.Mv, s.stream);
    }
    if (which & 2) {
        CUDA_CHECK(cudaMemsetAsync(s.H, 0, d.szHa() * 4, s.stream));
        for (int r = 0; r < d.R; ++r)
            launch_fill_uniform(s.H + (size_t)r * d.Kr, d.K, d.N, d.Kv, d.n_offset, seed + 7919ull * r, 2, s.stream);
        if (d.refcompat) launch_clamp_eps(s.H, d.Kv, d.Nv, d.Kv, s.stream);
    }
    if (which & 4) {
        zero_padding(s.X, d.M, d.Nv, d.Mv, d.Nva, s.stream);
        launch_fill_uniform(s.X, d.M, d.N, d.Mv, d.n_offset, seed, 3, s.stream);
        if (d.refcompat) launch_clamp_eps(s.X, d.Mv, d.Nv, d.Mv, s.stream);
    }
    refresh_splits(s, which & 1, which & 2);
    CUDA_CHECK(cudaStreamSynchronize(s.stream));
}

// X <- planted rank-`rank` product with multiplicative uniform noise (BASELINE C3's input); W, H untouched
void session_fill_planted(nmf_session_impl& s, unsigned long long seed, int rank, float noise) {
    const Dims& d = s.d;
    CUDA_CHECK(cudaSetDevice(s.device));
    NMF_CHECK(rank >= 1, "planted rank must be positive");
    zero_padding(s.X, d.M, d.Nv, d.Mv, d.Nva, s.stream);
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
    if (with_cost) { launch_sum_partials(s.cost_partials, s.n_cost_partials, 1, s.cost_dev, s.stream); count_launches(s, 1); }
}

// sum cols of W (Matrix::sum_cols, nmf.cu:134) at the head of an iteration.  Small problems (w_update_emits_colsum: a property
// of the shape, fixed for the session): the W update of the previous iteration left them, and the kernel runs only when W
// came from somewhere else (upload, generator, a restored restart) — session_run makes that call eagerly before it captures,
// so a captured iteration never contains it.  Everything else: one k_colsum launch per iteration, as the reference.
static int nsplit_at_update(const nmf_session_impl& s) {
    return (s.nranks == 1 && s.opt.engine == NMF_ENGINE_TCGEN05 && s.nsplit > 1) ? s.nsplit : 1;
}
static bool w_update_emits_colsum(const nmf_session_impl& s) { return update_w_by_column(s.d, nsplit_at_update(s)); }
static void ensure_sumw(nmf_session_impl& s, PhaseTimer* pt) {
    if (w_update_emits_colsum(s) && s.sumW_valid) return;
    PHASE(pt, PH_SUMS, { launch_colsum(s.W, s.sumW, s.d, s.stream); count_launches(s, 1); });
    s.sumW_valid = true;
}

void iteration_head(nmf_session_impl& s, bool with_cost, PhaseTimer* pt) {
    // sum cols of W (nmf.cu:134): the W update of the previous iteration left them (ensure_sumw otherwise)
    ensure_sumw(s, pt);
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
    ensure_sumw(s, pt);
    PHASE(pt, PH_COST, CUDA_CHECK(cudaMemcpyAsync(s.Hbak, s.H, d.szHa() * 4, cudaMemcpyDeviceToDevice, s.stream)));
    PHASE(pt, PH_GEMM2, {
        const int n = umma::fused_h(s, true); count_launches(s, n);
        launch_sum_partials(s.cost_partials, s.n_cost_partials_fused, d.R, s.cost_dev, s.stream); count_launches(s, 1);   // one cost per restart
    });
}
static void restore_h(nmf_session_impl& s) {
    s.rowsum_blocks = 0;              // partial row sums the abandoned H step may have left belong to the H being discarded
    CUDA_CHECK(cudaMemcpyAsync(s.H, s.Hbak, s.d.szHa() * 4, cudaMemcpyDeviceToDevice, s.stream));
    launch_split_tf32(s.H, s.Hhi, s.Hlo, s.d.szH(), s.stream, s.Hbf, s.Hlb); count_launches(s, 1);
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
        ensure_sumw(s, pt);
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
    // Small single-GPU problems: the W update kernel folds those partial sums itself (one launch fewer).
    const bool fold_in_update = s.rowsum_blocks > 0 && s.nranks == 1 && w_update_emits_colsum(s);
    const int rs_blocks = s.rowsum_blocks;
    if (!fold_in_update) {
        PHASE(pt, PH_SUMS, {
            if (s.rowsum_blocks > 0) { launch_rowsum_finish2(s.rowsum_partials, s.sumH, s.rowsum_blocks, d, s.stream); count_launches(s, 1); }
            else { launch_rowsum(s.H, s.rowsum_partials, s.sumH, d, s.stream); count_launches(s, 2); }
        });
    }
    s.rowsum_blocks = 0;
    const float* P = s.packed; int nsplit = 1; size_t stride = 0;
    if (s.opt.engine == NMF_ENGINE_TCGEN05 && umma::fused_w_enabled(s)) {
        // both of the steps below in one kernel: the quotient goes from TMEM through shared memory into the
        // second MMA and never reaches HBM
        PHASE(pt, PH_GEMM2, {
            umma::fused_w(s); count_launches(s, 1);
            if (s.tail_t) {
                launch_w_tail(s.W, s.H, s.X, s.tail_part, d, s.tail_row0, s.tail_t, s.nsplit > 1 ? s.Psplit : s.packed, s.nsplit, d.szW(), s.stream);
                count_launches(s, 2);
            }
        });
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
    // (small problems: the same kernel leaves the column sums of the new W for the next H step, nmf.cu:134)
    PHASE(pt, PH_UPDW, {
        s.sumW_valid = launch_update_w(s.W, s.Whi, s.Wlo, P, nsplit, stride, s.sumH, d, s.stream,
                                       fold_in_update ? s.rowsum_partials : nullptr, rs_blocks, s.sumW, s.Wbf, s.Wlb);
        count_launches(s, 1);
    });
}

// one half of an iteration, eagerly (parity tests look at a full-size problem between the halves)
void session_half_step(nmf_session_impl& s, int which) {
    CUDA_CHECK(cudaSetDevice(s.device));
    if (which == 0) {
        if (s.opt.engine == NMF_ENGINE_TCGEN05 && umma::fused_h_enabled(s)) {
            ensure_sumw(s, nullptr);
            umma::fused_h(s);
        } else {
            iteration_head(s, false, nullptr);
            if (s.opt.engine == NMF_ENGINE_SIMT) simt_update_h(s.W, s.Y, s.sumW, s.H, s.d, s.stream);
            else umma::update_h(s);
        }
    } else w_half(s, nullptr);
    CUDA_CHECK(cudaStreamSynchronize(s.stream));
}

// ---------------------------------------------------------------------------------------------
// run loop
// ---------------------------------------------------------------------------------------------
// the R costs the last cost kernel left in cost_dev -> s.costs (summed over the column shards when joined)
static void read_costs(nmf_session_impl& s) {
    const int R = s.d.R;
    if (s.nranks > 1) nccl_allreduce_f64(s, s.cost_dev, (size_t)R);
    CUDA_CHECK(cudaMemcpyAsync(s.cost_host, s.cost_dev, sizeof(double) * R, cudaMemcpyDeviceToHost, s.stream));
    CUDA_CHECK(cudaStreamSynchronize(s.stream));
    for (int r = 0; r < R; ++r) s.costs[r] = s.cost_host[r];
}

// KL divergence(s) of the resident factors, outside a check iteration.  One restart: the quotient kernel run for its
// cost.  Restarts > 1 (that kernel would multiply the stacked factors): the cost variant of the fused H step on a saved
// H, which is then put back.
static void eval_costs(nmf_session_impl& s, PhaseTimer* pt) {
    if (s.d.R == 1) {
        PHASE(pt, PH_QUOT, do_quotient(s, true));
    } else {
        iteration_head_fused_cost(s, pt);
        PHASE(pt, PH_COST, restore_h(s));
    }
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
    eval_costs(s, nullptr);      // cost of the resident factors (a quotient the kernel also writes is scratch)
    read_costs(s);
    for (int r = 0; r < s.d.R; ++r) cost[r] = s.costs[r];
}

// restart r of a batched run met the threshold at a check: keep the factors it had then (W is still untouched by this
// iteration; H as saved before the H step) — the other restarts go on, and so, harmlessly, does r's block
static void retire_restart(nmf_session_impl& s, int r) {
    const Dims& d = s.d;
    if (!s.Wfin) { s.Wfin = dalloc<float>(d.szW()); s.Hfin = dalloc<float>(d.szHa()); }
    s.retired[r] = 1;
    CUDA_CHECK(cudaMemcpyAsync(s.Wfin + (size_t)r * d.Kr * d.Mv, s.W + (size_t)r * d.Kr * d.Mv, (size_t)d.Kr * d.Mv * 4,
                               cudaMemcpyDeviceToDevice, s.stream));
    CUDA_CHECK(cudaMemcpy2DAsync(s.Hfin + (size_t)r * d.Kr, (size_t)d.Kv * 4, s.Hbak + (size_t)r * d.Kr, (size_t)d.Kv * 4,
                                 (size_t)d.Kr * 4, d.Nv, cudaMemcpyDeviceToDevice, s.stream));
}
static void put_back_retired(nmf_session_impl& s) {
    const Dims& d = s.d;
    bool any = false;
    for (int r = 0; r < d.R; ++r) {
        if (!s.retired[r]) continue;
        any = true;
        CUDA_CHECK(cudaMemcpyAsync(s.W + (size_t)r * d.Kr * d.Mv, s.Wfin + (size_t)r * d.Kr * d.Mv, (size_t)d.Kr * d.Mv * 4,
                                   cudaMemcpyDeviceToDevice, s.stream));
        CUDA_CHECK(cudaMemcpy2DAsync(s.H + (size_t)r * d.Kr, (size_t)d.Kv * 4, s.Hfin + (size_t)r * d.Kr, (size_t)d.Kv * 4,
                                     (size_t)d.Kr * 4, d.Nv, cudaMemcpyDeviceToDevice, s.stream));
    }
    if (any) { refresh_splits(s, true, true); count_launches(s, 2); }
}

namespace {
struct EventPair {          // RAII: an exception between create and destroy must not leak the events
    cudaEvent_t a = nullptr, b = nullptr;
    EventPair() { CUDA_CHECK(cudaEventCreate(&a)); CUDA_CHECK(cudaEventCreate(&b)); }
    ~EventPair() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
};
}  // namespace

// The collective schedule (which iterations evaluate / allreduce the cost) depends only on thresh, verbose, max_iter
// and check_every — the same on every rank of a sharded run; only rank 0 prints.
void session_run(nmf_session_impl& s, float thresh, int max_iter, double t[10], int verbose, nmf_result* res) {
    using clk = std::chrono::steady_clock;
    const auto wall0 = clk::now();
    CUDA_CHECK(cudaSetDevice(s.device));
    const bool need_cost = (thresh > 0.0f) || verbose;
    const bool print = verbose && s.rank == 0;
    const int every = s.opt.check_every, R = s.d.R;
    std::unique_ptr<PhaseTimer> pt_own(t ? new PhaseTimer(s.stream) : nullptr);
    PhaseTimer* pt = pt_own.get();
    const bool graph = s.opt.use_graph && !pt;
    s.rowsum_blocks = 0;
    struct PdlGuard { explicit PdlGuard(bool on) { set_launch_pdl(on); } ~PdlGuard() { set_launch_pdl(false); } } pdl_guard(s.tune.pdl != 0);
    if (max_iter > 0 && w_update_emits_colsum(s)) ensure_sumw(s, pt);   // a captured iteration of a small problem never contains the column-sum kernel
    if (graph) build_graphs(s);

    EventPair ev;
    const long long launches0 = s.launches;
    CUDA_CHECK(cudaEventRecord(ev.a, s.stream));

    std::vector<double> prev((size_t)R, std::numeric_limits<double>::infinity());
    for (int r = 0; r < R; ++r) { s.costs[r] = std::nan(""); s.iters_r[r] = 0; s.conv_r[r] = 0; s.retired[r] = 0; }
    int done = 0, converged = 0, live = R;
    double t_cost = 0.0;
    auto report = [&](int it, int r, double rel) {
        if (!print) return;
        if (R == 1) std::printf("iter %d  cost %.9g  rel.change %.3g\n", it, s.costs[0], rel);
        else std::printf("iter %d  restart %d  cost %.9g  rel.change %.3g%s\n", it, r, s.costs[r], rel, s.conv_r[r] ? "  (converged)" : "");
    };
    for (int it = 1; it <= max_iter; ++it) {
        // The cost after iteration it-1 is a by-product of iteration it's first quotient.
        const bool check = need_cost && it > 1 && ((it - 1) % every == 0);
        if (check) {
            const bool fcheck = fused_check_enabled(s);
            if (fcheck) iteration_head_fused_cost(s, pt); else iteration_head(s, true, pt);
            const auto c0 = clk::now();
            const std::vector<double> before = s.costs;
            read_costs(s);
            t_cost += std::chrono::duration<double>(clk::now() - c0).count();
            for (int r = 0; r < R; ++r) {
                if (s.conv_r[r]) { s.costs[r] = before[r]; continue; }   // retired: its cost stays the one it stopped with
                const double rel = (prev[r] - s.costs[r]) / s.costs[r];
                if (thresh > 0.0f && rel < (double)thresh) {
                    s.conv_r[r] = 1; s.iters_r[r] = it - 1; --live;
                    if (R > 1) retire_restart(s, r);
                }
                report(it - 1, r, rel);
                prev[r] = s.costs[r];
            }
            if (live == 0) { if (fcheck) restore_h(s); converged = 1; break; }
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
        eval_costs(s, pt);
        const auto c0 = clk::now();
        const std::vector<double> before = s.costs;
        read_costs(s);
        t_cost += std::chrono::duration<double>(clk::now() - c0).count();
        for (int r = 0; r < R; ++r) {
            if (s.conv_r[r]) { s.costs[r] = before[r]; continue; }
            const double rel = (prev[r] - s.costs[r]) / s.costs[r];
            if (thresh > 0.0f && done % every == 0 && rel < (double)thresh) { s.conv_r[r] = 1; s.iters_r[r] = done; --live; }
            report(done, r, rel);
        }
        if (live == 0) converged = 1;
    }
    for (int r = 0; r < R; ++r) if (!s.conv_r[r]) s.iters_r[r] = done;
    if (R > 1) put_back_retired(s);
    CUDA_CHECK(cudaEventRecord(ev.b, s.stream));
    {
        const cudaError_t e = cudaStreamSynchronize(s.stream);
        if (e != cudaSuccess && s.opt.engine == NMF_ENGINE_TCGEN05 && umma::protocol_error(s))
            NMF_CHECK(false, std::string("kernel trapped: bounded mbarrier wait ") + std::to_string(umma::protocol_error(s)) +
                             " gave up (" + cudaGetErrorString(e) + ")");
        CUDA_CHECK(e);
    }
    float ms = 0; CUDA_CHECK(cudaEventElapsedTime(&ms, ev.a, ev.b));
    if (pt) {
        pt->flush();
        for (int i = 0; i < 10; ++i) t[i] = 0.0;
        for (int i = 0; i < 8; ++i) t[i] = pt->acc[i];
        t[PH_COST] += t_cost;
        t[PH_TOTAL] = std::chrono::duration<double>(clk::now() - wall0).count();
    }
    if (res) {
        int best = 0;
        for (int r = 1; r < R; ++r) if (s.costs[r] < s.costs[best]) best = r;     // NaN (never evaluated) compares false
        res->iterations = done; res->converged = converged; res->cost = s.costs[best];
        res->device_ms = ms; res->gpu_launches = s.launches - launches0;
    }
}

}  // namespace nmf
