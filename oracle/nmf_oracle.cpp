// nmf_oracle.cpp — CPU restatement (C++/OpenMP) of the reference's KL-divergence NMF iteration.
//
// THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load it.  The product path
// (nmf_gpu_b200/csrc, libnmf_b200.so) never links, loads or calls anything in oracle/.
//
// The reference (recoord/nmf-gpu) has no CPU path: every operation is a CUDA kernel or a cuBLAS
// call.  This file restates, on the host, the arithmetic that reference executes per iteration:
//
//   cuda/nmf.cu:118-146  update_h : Z = X ./ clampE(W*H); sW = clampE(colsum(W));
//                                   H = H .* ((W'*Z) ./ sW)          (divide, then multiply)
//   cuda/nmf.cu:148-176  update_w : Z = X ./ clampE(W*H); sH = clampE(rowsum(H));
//                                   W = W .* ((Z*H') ./ sH)
//   cuda/matrix.cu:182-188  kernel_set_epsilon : a = (a < EPS) ? EPS : a   (a clamp, not "+EPS")
//   cuda/matrix.cu:146-152  vec_div  : IEEE divide
//   cuda/matrix.cu:88-95    add_padding : every dim != 1 rounded up to a multiple of 32
//   cuda/nmf.cu:188-218     read_matrix : column-major payload, clamp to >= EPS after load
//   cuda/matrix.cu:578-640  reduce1d_div : cost term x*(log x - log y) - x + y  (dead code in the
//                                   reference, but the only in-tree definition of the cost)
//   cuda/matrix.cu:642-735  reduce2d / reduce2d_strided : the column / row sums.
//
// Two semantics modes:
//   mode 0 "exact"     : what README / paper eq.(2) / the MATLAB comment at nmf.cu:104-107 specify:
//                        true sums, always divide, zero padding (i.e. no padding effects).
//   mode 1 "refcompat" : what the reference binary actually computes, including its two latent
//                        defects, because the bundled golden Wtest.bin/Htest.bin encode them:
//                        (a) the reduction tail at matrix.cu:676-683 / :724-731 reads stale shared
//                            memory, so each "sum" only covers indices i with
//                            (i mod 64) in {0,1,2,4,8,16,32};
//                        (b) row_divide (matrix.cu:203-218) launches rows_padded threads per block,
//                            an invalid configuration when rows_padded > 1024, and no launch is
//                            error-checked: the W normalisation silently does not happen;
//                        (c) matrices are padded to multiples of 32 and the clamp runs over the
//                            whole padded buffer, so padding holds EPS, not 0.
//
// Parity pinning: mode 1 is pinned against the reference's own golden vectors
// (/root/reference/Wtest.bin, Htest.bin -> tests/golden/) by tests/test_oracle_golden.py.
// Mode 0 shares every line with mode 1 except the two flags above; there is no golden for it in
// the reference tree — it is pinned by the paper's update rule (monotone KL
// decrease, fixed points) and by the known-answer scalars recorded in BASELINE.md.
//
// The GEMMs of the reference are cuBLAS calls (closed source, cuda/matrix.cu:101,111,121); their
// summation order is not specified, so parity against them is tolerance-level by construction.
//
// Build: g++ -O3 -mavx2 -mfma -fopenmp -shared -fPIC nmf_oracle.cpp -o libnmf_oracle.so
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// (float)2.2204E-16, cuda/matrix.cu:10 — NOT 2^-52; fp32 bit pattern 0x257ffea4.
const float kEps = (float)(2.2204E-16);

inline float clampE(float v) { return v < kEps ? kEps : v; }   // matrix.cu:185-187 (NaN passes through)

inline int pad32(int v) {                                      // matrix.cu:88-95
    if (v != 1 && v % 32 != 0) return v + (32 - v % 32);
    return v;
}

// ---------------------------------------------------------------------------------------------
// Column-major padded matrix on the host.
// ---------------------------------------------------------------------------------------------
struct Mat {
    int rows = 0, cols = 0, ld = 0, cols_p = 0;   // logical dims, leading dimension (= padded rows)
    std::vector<float> v;
    float*       col(int j)       { return v.data() + (size_t)j * ld; }
    const float* col(int j) const { return v.data() + (size_t)j * ld; }
};

// Load a host column-major rows x cols array the way read_matrix + Matrix(host,...) do
// (nmf.cu:208-211, matrix.cu:53-67): pitch copy into a zeroed padded buffer, then clamp.
// refcompat: padded to 32 and the clamp covers the padding (padding == EPS).
// exact    : no padding, clamp covers the data only.
Mat load(const float* src, int rows, int cols, bool refcompat) {
    Mat m;
    m.rows = rows; m.cols = cols;
    m.ld = refcompat ? pad32(rows) : rows;
    m.cols_p = refcompat ? pad32(cols) : cols;
    m.v.assign((size_t)m.ld * m.cols_p, 0.0f);
    for (int j = 0; j < cols; ++j)
        std::memcpy(m.col(j), src + (size_t)j * rows, sizeof(float) * rows);
    for (auto& x : m.v) x = clampE(x);
    return m;
}

void store(const Mat& m, float* dst) {                          // write_matrix's un-padding copy, nmf.cu:228-231
    for (int j = 0; j < m.cols; ++j)
        std::memcpy(dst + (size_t)j * m.rows, m.col(j), sizeof(float) * m.rows);
}

// ---------------------------------------------------------------------------------------------
// The reference's sum kernels as compiled (refcompat), and the true sums (exact).
// ---------------------------------------------------------------------------------------------
// Launch parameters chosen by init_params, nmf.cu:53-74: returns block size (bs).
int ref_block_size(int padded) {
    unsigned x = (unsigned)(padded / 128 + ((padded % 128) ? 1 : 0));
    unsigned p = 1; while (p < x) p <<= 1;                      // nextpow2, nmf.cu:178-186
    if (p <= 128) return 128;
    return (int)p;                                              // <= 512 or the reference exits
}

// One block of reduce2d<bs> / reduce2d_strided<bs> (matrix.cu:642-735) over n values with the
// given stride, in the kernel's own accumulation order, with the stale-read tail: the loads of
// the last seven steps are all issued before the single store, so thread 0 ends with
// s[0]+s[32]+s[16]+s[8]+s[4]+s[2]+s[1] where s[] is the state after the last __syncthreads level.
float ref_block_sum(const float* g, int n, size_t stride, int bs) {
    std::vector<float> sd(bs, 0.0f);
    if (n < bs) {
        // n - bs underflows in the kernel (uint32) and it would read out of bounds; none of the
        // supported configurations get here.  Keep the closed form so the oracle stays defined.
        for (int i = 0; i < n; ++i) sd[i % bs] += g[(size_t)i * stride];
    } else {
        const int lim = n - bs;
        for (int tid = 0; tid < bs; ++tid) {
            float acc = 0.0f;
            int i = tid;
            while (i < lim) {                                   // grid.x == 1 (else reference exits)
                acc += g[(size_t)i * stride] + g[(size_t)(i + bs) * stride];
                i += 2 * bs;
            }
            if (i < n) acc += g[(size_t)i * stride];
            sd[tid] = acc;
        }
    }
    for (int half = bs / 2; half >= 64; half /= 2)              // the __syncthreads-separated levels
        for (int t = 0; t < half; ++t) sd[t] += sd[t + half];
    float r = sd[0];
    r += sd[32]; r += sd[16]; r += sd[8]; r += sd[4]; r += sd[2]; r += sd[1];
    return r;
}

template <typename Acc>
void col_sums(const Mat& W, bool refcompat, std::vector<float>& out) {   // sum over rows, per column
    out.assign(W.cols_p, 0.0f);
    const int bs = refcompat ? ref_block_size(W.ld) : 0;
    #pragma omp parallel for schedule(static)
    for (int k = 0; k < W.cols_p; ++k) {
        if (refcompat) {
            out[k] = ref_block_sum(W.col(k), W.ld, 1, bs);
        } else {
            Acc s = 0;
            const float* c = W.col(k);
            for (int m = 0; m < W.ld; ++m) s += (Acc)c[m];
            out[k] = (float)s;
        }
        out[k] = clampE(out[k]);                                // sumW->set_epsilon, nmf.cu:135
    }
}

template <typename Acc>
void row_sums(const Mat& H, bool refcompat, std::vector<float>& out) {   // sum over columns, per row
    out.assign(H.ld, 0.0f);
    const int bs = refcompat ? ref_block_size(H.cols_p) : 0;
    #pragma omp parallel for schedule(static)
    for (int k = 0; k < H.ld; ++k) {
        if (refcompat) {
            out[k] = ref_block_sum(H.v.data() + k, H.cols_p, (size_t)H.ld, bs);
        } else {
            Acc s = 0;
            for (int n = 0; n < H.cols_p; ++n) s += (Acc)H.v[(size_t)n * H.ld + k];
            out[k] = (float)s;
        }
        out[k] = clampE(out[k]);                                // sumH2->set_epsilon, nmf.cu:165
    }
}

// ---------------------------------------------------------------------------------------------
// The three contractions (cuBLAS SGEMM in the reference, matrix.cu:97-125), on padded dims.
// Acc = float reproduces an fp32 SGEMM (sequential-k order); Acc = double is the "truth" run.
// ---------------------------------------------------------------------------------------------
constexpr int NB = 4;   // columns processed together (register / cache blocking only)

// Z = X ./ clampE(W*H)      nmf.cu:125-131 / :155-161
template <typename Acc>
void quotient(const Mat& W, const Mat& H, const Mat& X, Mat& Z) {
    const int Mp = W.ld, Kp = W.cols_p, Np = H.cols_p;
    #pragma omp parallel
    {
        std::vector<Acc> acc((size_t)NB * Mp);
        #pragma omp for schedule(dynamic, 4)
        for (int n0 = 0; n0 < Np; n0 += NB) {
            const int nb = std::min(NB, Np - n0);
            std::fill(acc.begin(), acc.end(), (Acc)0);
            for (int k = 0; k < Kp; ++k) {
                const float* w = W.col(k);
                for (int j = 0; j < nb; ++j) {
                    const Acc h = (Acc)H.col(n0 + j)[k];
                    Acc* a = acc.data() + (size_t)j * Mp;
                    #pragma omp simd
                    for (int m = 0; m < Mp; ++m) a[m] += (Acc)w[m] * h;
                }
            }
            for (int j = 0; j < nb; ++j) {
                const float* x = X.col(n0 + j);
                float* z = Z.col(n0 + j);
                const Acc* a = acc.data() + (size_t)j * Mp;
                for (int m = 0; m < Mp; ++m) z[m] = x[m] / clampE((float)a[m]);
            }
        }
    }
}

// H = H .* ((W'*Z) ./ sW)    nmf.cu:138-145
template <typename Acc>
void update_h(const Mat& W, const Mat& Z, const std::vector<float>& sW, Mat& H) {
    const int Mp = W.ld, Kp = W.cols_p, Np = H.cols_p;
    #pragma omp parallel for schedule(dynamic, 4)
    for (int n = 0; n < Np; ++n) {
        const float* z = Z.col(n);
        float* h = H.col(n);
        for (int k = 0; k < Kp; ++k) {
            const float* w = W.col(k);
            Acc s = 0;
            #pragma omp simd reduction(+ : s)
            for (int m = 0; m < Mp; ++m) s += (Acc)w[m] * (Acc)z[m];
            const float q = (float)s / sW[k];                   // col_div, matrix.cu:244-250
            h[k] = h[k] * q;                                    // vec_mul, matrix.cu:174-180
        }
    }
}

// W = W .* ((Z*H') ./ sH)    nmf.cu:168-175 ; divide skipped iff !do_divide (refcompat, Mp > 1024)
template <typename Acc>
void update_w(const Mat& Z, const Mat& H, const std::vector<float>& sH, bool do_divide, Mat& W) {
    const int Mp = W.ld, Kp = W.cols_p, Np = H.cols_p;
    #pragma omp parallel
    {
        std::vector<Acc> acc((size_t)NB * Mp);
        #pragma omp for schedule(dynamic, 1)
        for (int k0 = 0; k0 < Kp; k0 += NB) {
            const int kb = std::min(NB, Kp - k0);
            std::fill(acc.begin(), acc.end(), (Acc)0);
            for (int n = 0; n < Np; ++n) {
                const float* z = Z.col(n);
                const float* hc = H.col(n);
                for (int j = 0; j < kb; ++j) {
                    const Acc h = (Acc)hc[k0 + j];
                    Acc* a = acc.data() + (size_t)j * Mp;
                    #pragma omp simd
                    for (int m = 0; m < Mp; ++m) a[m] += (Acc)z[m] * h;
                }
            }
            for (int j = 0; j < kb; ++j) {
                float* w = W.col(k0 + j);
                const Acc* a = acc.data() + (size_t)j * Mp;
                const float d = sH[k0 + j];
                for (int m = 0; m < Mp; ++m) {
                    float q = (float)a[m];
                    if (do_divide) q = q / d;                   // row_div, matrix.cu:220-224
                    w[m] = w[m] * q;                            // vec_mul
                }
            }
        }
    }
}

// KL divergence  sum x*(log x - log y) - x + y,  y = clampE(W*H), over the UNPADDED block.
// Term from reduce1d_div (matrix.cu:590-593); evaluated in fp64 (this is the yard-stick).
double kl_cost(const Mat& W, const Mat& H, const Mat& X) {
    const int M = X.rows, N = X.cols, Kp = W.cols_p;
    double total = 0.0;
    #pragma omp parallel
    {
        std::vector<double> acc(M);
        #pragma omp for schedule(dynamic, 8) reduction(+ : total)
        for (int n = 0; n < N; ++n) {
            std::fill(acc.begin(), acc.end(), 0.0);
            for (int k = 0; k < Kp; ++k) {
                const float* w = W.col(k);
                const double h = (double)H.col(n)[k];
                #pragma omp simd
                for (int m = 0; m < M; ++m) acc[m] += (double)w[m] * h;
            }
            const float* x = X.col(n);
            double s = 0.0;
            for (int m = 0; m < M; ++m) {
                const double xv = (double)x[m];
                const double yv = std::max(acc[m], (double)kEps);
                s += xv * (std::log(xv) - std::log(yv)) - xv + yv;
            }
            total += s;
        }
    }
    return total;
}

struct Problem {
    Mat W, H, X, Z;
    bool refcompat;
    bool w_divide;
    Problem(const float* w, const float* h, const float* x, int M, int N, int K, int mode) {
        refcompat = (mode == 1);
        W = load(w, M, K, refcompat);
        H = load(h, K, N, refcompat);
        X = load(x, M, N, refcompat);
        Z = X;
        // row_div launches rows_padded threads per block (matrix.cu:212-217): invalid above 1024.
        w_divide = refcompat ? (W.ld <= 1024) : true;
    }
};

template <typename Acc>
void h_half(Problem& p) {
    std::vector<float> sW;
    quotient<Acc>(p.W, p.H, p.X, p.Z);
    col_sums<Acc>(p.W, p.refcompat, sW);
    update_h<Acc>(p.W, p.Z, sW, p.H);
}

template <typename Acc>
void w_half(Problem& p) {
    std::vector<float> sH;
    quotient<Acc>(p.W, p.H, p.X, p.Z);
    row_sums<Acc>(p.H, p.refcompat, sH);
    update_w<Acc>(p.Z, p.H, sH, p.w_divide, p.W);
}

}  // namespace

extern "C" {

int nmf_oracle_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// OpenMP threads for every later call (torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU baseline wants them all).
void nmf_oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

float nmf_oracle_eps(void) { return kEps; }

// Full update_div on host arrays (column-major, unpadded: W MxK, H KxN, X MxN), in place on W, H.
// Iteration contract: H half-step then W half-step per iteration; when
// thresh > 0 or verbose, the KL cost is evaluated every check_every (=25, ITER_CHECK nmf.cu:9)
// iterations and the loop stops when (prev - cur) / cur < thresh (first check never stops).
// Returns the number of iterations executed; *final_cost (if non-NULL) gets the fp64 KL cost of
// the returned factors.
int nmf_oracle_update_div(float* W, float* H, const float* X, int M, int N, int K,
                          float thresh, int max_iter, int mode, int acc64, int check_every,
                          int verbose, double* final_cost) {
    Problem p(W, H, X, M, N, K, mode);
    if (check_every <= 0) check_every = 25;
    double prev = std::numeric_limits<double>::infinity();
    int it = 0;
    for (it = 1; it <= max_iter; ++it) {
        if (acc64) { h_half<double>(p); w_half<double>(p); }
        else       { h_half<float>(p);  w_half<float>(p);  }
        if ((thresh > 0.0f || verbose) && it % check_every == 0) {
            const double cur = kl_cost(p.W, p.H, p.X);
            const double rel = (prev - cur) / cur;
            if (verbose) std::printf("oracle iter %d  cost %.9g  rel.change %.3g\n", it, cur, rel);
            if (thresh > 0.0f && rel < (double)thresh) { ++it; break; }
            prev = cur;
        }
    }
    --it;
    if (final_cost) *final_cost = kl_cost(p.W, p.H, p.X);
    store(p.W, W);
    store(p.H, H);
    return it;
}

// fp64 KL cost of given factors (inputs clamped to >= EPS as read_matrix does).
double nmf_oracle_kl(const float* W, const float* H, const float* X, int M, int N, int K) {
    Problem p(W, H, X, M, N, K, 0);
    return kl_cost(p.W, p.H, p.X);
}

// Kernel-level pieces for the CUDA parity tests (exact semantics, unpadded).
// Y (MxN) = X ./ clampE(W*H)
void nmf_oracle_quotient(const float* W, const float* H, const float* X, float* Y,
                         int M, int N, int K, int acc64) {
    Problem p(W, H, X, M, N, K, 0);
    if (acc64) quotient<double>(p.W, p.H, p.X, p.Z); else quotient<float>(p.W, p.H, p.X, p.Z);
    store(p.Z, Y);
}

// One H half-step / one W half-step, in place.
void nmf_oracle_h_step(const float* W, float* H, const float* X, int M, int N, int K, int mode, int acc64) {
    Problem p(W, H, X, M, N, K, mode);
    if (acc64) h_half<double>(p); else h_half<float>(p);
    store(p.H, H);
}

void nmf_oracle_w_step(float* W, const float* H, const float* X, int M, int N, int K, int mode, int acc64) {
    Problem p(W, H, X, M, N, K, mode);
    if (acc64) w_half<double>(p); else w_half<float>(p);
    store(p.W, W);
}

// Timed loop for the CPU baseline: runs `iters` full iterations (exact, fp32) and returns seconds
// (omp_get_wtime around the loop only — load/store excluded, like the GPU timing).
double nmf_oracle_time_iterations(float* W, float* H, const float* X, int M, int N, int K, int iters) {
    Problem p(W, H, X, M, N, K, 0);
#ifdef _OPENMP
    const double t0 = omp_get_wtime();
#else
    const double t0 = 0;
#endif
    for (int i = 0; i < iters; ++i) { h_half<float>(p); w_half<float>(p); }
#ifdef _OPENMP
    const double t1 = omp_get_wtime();
#else
    const double t1 = 0;
#endif
    store(p.W, W);
    store(p.H, H);
    return t1 - t0;
}

}  // extern "C"
