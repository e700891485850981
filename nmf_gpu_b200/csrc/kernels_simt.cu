// kernels_simt.cu — the SIMT engine: plain fp32 FFMA tiled GEMMs with the same fused epilogues as
// the tcgen05 engine.  It exists for bring-up and as an on-device cross-check at sizes the CPU
// oracle cannot reach; it is NOT the product path (NMF_ENGINE=simt selects it explicitly).
//
// One kernel template, three contractions (reference: the three cublasSgemm wrappers,
// cuda/matrix.cu:97-125) with generic strides:
//   MODE 0  C[m,n] = sum_k W[m,k] H[k,n]   epilogue  Y = X ./ max(C,EPS)  (+ KL partial sums)
//   MODE 1  C[k,n] = sum_m W[m,k] Y[m,n]   epilogue  H = H .* (C ./ max(sumW[k],EPS))
//   MODE 2  C[m,k] = sum_n Y[m,n] H[k,n]   epilogue  P = C
#include "common.hpp"

namespace nmf {
namespace {

constexpr int TS = 64;    // output tile TS x TS
constexpr int TR = 16;    // reduction slice
constexpr int NT = 256;   // threads; each computes 4 x 4

struct SimtArgs {
    const float* A; long long sAi, sAr;      // A(i,r) = A[i*sAi + r*sAr]
    const float* B; long long sBr, sBj;      // B(r,j) = B[r*sBr + j*sBj]
    int I, J, R;                             // output I x J, reduction length R
    // epilogue operands
    const float* X; float* Y; long long ldO; // MODE 0: X, Y (ld ldO); MODE 1: H in Y (ld ldO); MODE 2: P in Y
    const float* vec;                        // MODE 1: sumW
    double* cost_partials; int Mlog, Nlog;   // MODE 0: optional KL partials over the logical block
};

__device__ __forceinline__ float clampE(float v) { return v < NMF_EPS_F ? NMF_EPS_F : v; }

template <int MODE, bool A_I_CONTIG, bool B_J_CONTIG>
__global__ void __launch_bounds__(NT) k_simt_gemm(SimtArgs a) {
    __shared__ float As[TR][TS + 4];
    __shared__ float Bs[TR][TS + 4];
    const int i0 = blockIdx.x * TS, j0 = blockIdx.y * TS;
    const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;   // thread computes rows tx*4.., cols ty*4..
    float acc[4][4];
    #pragma unroll
    for (int p = 0; p < 4; ++p)
        #pragma unroll
        for (int q = 0; q < 4; ++q) acc[p][q] = 0.0f;

    for (int r0 = 0; r0 < a.R; r0 += TR) {
        #pragma unroll
        for (int e = threadIdx.x; e < TS * TR; e += NT) {
            int ii, rr;
            if (A_I_CONTIG) { ii = e % TS; rr = e / TS; } else { rr = e % TR; ii = e / TR; }
            const int gi = i0 + ii, gr = r0 + rr;
            As[rr][ii] = (gi < a.I && gr < a.R) ? a.A[(long long)gi * a.sAi + (long long)gr * a.sAr] : 0.0f;
            int jj, r2;
            if (B_J_CONTIG) { jj = e % TS; r2 = e / TS; } else { r2 = e % TR; jj = e / TR; }
            const int gj = j0 + jj, gr2 = r0 + r2;
            Bs[r2][jj] = (gj < a.J && gr2 < a.R) ? a.B[(long long)gr2 * a.sBr + (long long)gj * a.sBj] : 0.0f;
        }
        __syncthreads();
        float part[4][4];
        #pragma unroll
        for (int p2 = 0; p2 < 4; ++p2)
            #pragma unroll
            for (int q2 = 0; q2 < 4; ++q2) part[p2][q2] = 0.0f;
        #pragma unroll
        for (int r = 0; r < TR; ++r) {
            const float4 av = *reinterpret_cast<const float4*>(&As[r][tx * 4]);
            const float4 bv = *reinterpret_cast<const float4*>(&Bs[r][ty * 4]);
            const float ar[4] = {av.x, av.y, av.z, av.w}, br[4] = {bv.x, bv.y, bv.z, bv.w};
            #pragma unroll
            for (int p = 0; p < 4; ++p)
                #pragma unroll
                for (int q = 0; q < 4; ++q) part[p][q] = fmaf(ar[p], br[q], part[p][q]);
        }
        // two-level summation: 16-term partial sums, then one add into the running total
        #pragma unroll
        for (int p2 = 0; p2 < 4; ++p2)
            #pragma unroll
            for (int q2 = 0; q2 < 4; ++q2) acc[p2][q2] += part[p2][q2];
        __syncthreads();
    }

    double cost = 0.0;
    #pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int j = j0 + ty * 4 + q;
        if (j >= a.J) continue;
        #pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int i = i0 + tx * 4 + p;
            if (i >= a.I) continue;
            const long long o = (long long)j * a.ldO + i;
            if (MODE == 0) {
                const float y = clampE(acc[p][q]);               // set_epsilon, cuda/matrix.cu:182-188
                const float x = a.X[o];
                a.Y[o] = x / y;                                  // vec_div, cuda/matrix.cu:146-152
                if (a.cost_partials != nullptr && i < a.Mlog && j < a.Nlog)
                    cost += (double)(x * (logf(x) - logf(y)) - x + y);   // reduce1d_div term, matrix.cu:593
            } else if (MODE == 1) {
                const float q1 = acc[p][q] / clampE(a.vec[i]);   // col_div, cuda/matrix.cu:244-250
                a.Y[o] = a.Y[o] * q1;                            // vec_mul
            } else {
                a.Y[o] = acc[p][q];
            }
        }
    }
    if (MODE == 0 && a.cost_partials != nullptr) {
        __shared__ double red[NT / 32];
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) cost += __shfl_xor_sync(0xffffffffu, cost, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = cost;
        __syncthreads();
        if (threadIdx.x == 0) {
            double s = 0.0;
            for (int w = 0; w < NT / 32; ++w) s += red[w];
            a.cost_partials[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = s;
        }
    }
}

inline dim3 tiles(int I, int J) { return dim3((I + TS - 1) / TS, (J + TS - 1) / TS); }

}  // namespace

int simt_cost_partial_count(const Dims& d) {
    const dim3 g = tiles(d.Mv, d.Nv);
    return (int)(g.x * g.y);
}

void simt_quotient(const float* W, const float* H, const float* X, float* Y, double* cost_partials,
                   const Dims& d, cudaStream_t st) {
    SimtArgs a{};
    a.A = W; a.sAi = 1; a.sAr = d.Mv;
    a.B = H; a.sBr = 1; a.sBj = d.Kv;
    a.I = d.Mv; a.J = d.Nv; a.R = d.Kv;
    a.X = X; a.Y = Y; a.ldO = d.Mv;
    a.cost_partials = cost_partials; a.Mlog = d.M; a.Nlog = d.N;
    k_simt_gemm<0, true, false><<<tiles(a.I, a.J), NT, 0, st>>>(a);
    CUDA_CHECK_LAUNCH();
}

void simt_update_h(const float* W, const float* Y, const float* sumW, float* H, const Dims& d, cudaStream_t st) {
    SimtArgs a{};
    a.A = W; a.sAi = d.Mv; a.sAr = 1;          // A(i=k, r=m) = W[m + k*ld]
    a.B = Y; a.sBr = 1; a.sBj = d.Mv;          // B(r=m, j=n) = Y[m + n*ld]
    a.I = d.Kv; a.J = d.Nv; a.R = d.Mv;
    a.Y = H; a.ldO = d.Kv; a.vec = sumW;
    k_simt_gemm<1, false, false><<<tiles(a.I, a.J), NT, 0, st>>>(a);
    CUDA_CHECK_LAUNCH();
}

void simt_numer(const float* Y, const float* H, float* P, const Dims& d, cudaStream_t st) {
    SimtArgs a{};
    a.A = Y; a.sAi = 1; a.sAr = d.Mv;          // A(i=m, r=n) = Y[m + n*ld]
    a.B = H; a.sBr = d.Kv; a.sBj = 1;          // B(r=n, j=k) = H[k + n*ldK]
    a.I = d.Mv; a.J = d.Kv; a.R = d.Nv;
    a.Y = P; a.ldO = d.Mv;
    k_simt_gemm<2, true, true><<<tiles(a.I, a.J), NT, 0, st>>>(a);
    CUDA_CHECK_LAUNCH();
}

}  // namespace nmf
