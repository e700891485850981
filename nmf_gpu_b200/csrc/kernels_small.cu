// kernels_small.cu — HBM-bound helper kernels: clamp, fills, TF32 split, the W column sums (R1),
// the H row sums (R2), the W multiplicative update, ordered partial-sum finishes.
// All reductions are warp-shuffle trees with a fixed order (no float atomics): run-to-run
// bit-reproducible, which the 1e-4 @ 100 iterations parity budget needs (SURVEY.md §7 "Determinism").
#include <mutex>

#include "common.hpp"
#include "tf32.cuh"

namespace nmf {

namespace { thread_local bool t_pdl = false; }
void set_launch_pdl(bool on) { t_pdl = on; }
bool launch_pdl() { return t_pdl; }

namespace {

__device__ __forceinline__ float clampE(float v) { return v < NMF_EPS_F ? NMF_EPS_F : v; }

__device__ __forceinline__ float warp_sum(float v) {
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// The reference's reduce2d tail (cuda/matrix.cu:676-683) reads stale shared memory; what its
// "sum" covers is exactly the indices with (i mod 64) in {0,1,2,4,8,16,32}  (SURVEY.md §0.2).
__device__ __forceinline__ bool ref_sum_mask(long long i) {
    const unsigned r = (unsigned)(i & 63);
    return r == 0 || (r & (r - 1)) == 0;     // 0 or a power of two < 64
}

__global__ void k_clamp_eps(float* __restrict__ a, int rows, int cols, int ld) {
    const size_t total = (size_t)rows * cols;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (size_t)gridDim.x * blockDim.x) {
        const size_t c = i / rows, r = i - c * rows;
        float* p = a + c * (size_t)ld + r;
        *p = clampE(*p);
    }
}

__global__ void k_fill_region(float* __restrict__ a, int ld, int r0, int r1, int c0, int c1, float value) {
    const int nr = r1 - r0;
    const size_t total = (size_t)nr * (c1 - c0);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (size_t)gridDim.x * blockDim.x) {
        const size_t c = i / nr, r = i - c * nr;
        a[(c0 + c) * (size_t)ld + r0 + r] = value;
    }
}

__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {   // splitmix64 finaliser
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__global__ void k_fill_uniform(float* __restrict__ a, int rows, int cols, int ld, long long col_offset,
                               unsigned long long seed, int tag) {
    const size_t total = (size_t)rows * cols;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (size_t)gridDim.x * blockDim.x) {
        const size_t c = i / rows, r = i - c * rows;
        const unsigned long long gc = (unsigned long long)(col_offset + (long long)c);
        unsigned long long h = mix64(seed * 0x9E3779B97F4A7C15ull + (unsigned long long)tag);
        h = mix64(h ^ (gc * 0xD1B54A32D192ED03ull + (unsigned long long)r));
        const float u = (float)(h >> 40) * (1.0f / 16777216.0f);   // 24 random bits -> [0,1)
        a[c * (size_t)ld + r] = clampE(u);
    }
}

// Planted low-rank target (BASELINE config C3): X[m,n] = (sum_k A[m,k] B[k,n]) * (1 + noise*(2u-1)) with
// A, B, u uniform(0,1) drawn from the same counter-based hash of the GLOBAL indices (tags 11, 12, 13),
// so any column sharding generates the same matrix.  One thread per element, rank `rank` dot product.
__device__ __forceinline__ float hash_u01(unsigned long long seed, int tag, unsigned long long a, unsigned long long b) {
    unsigned long long h = mix64(seed * 0x9E3779B97F4A7C15ull + (unsigned long long)tag);
    h = mix64(h ^ (b * 0xD1B54A32D192ED03ull + a));
    return (float)(h >> 40) * (1.0f / 16777216.0f);
}

__global__ void k_fill_planted(float* __restrict__ x, int rows, int cols, int ld, long long col_offset,
                               unsigned long long seed, int rank, float noise) {
    const size_t total = (size_t)rows * cols;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (size_t)gridDim.x * blockDim.x) {
        const size_t c = i / rows, r = i - c * rows;
        const unsigned long long gc = (unsigned long long)(col_offset + (long long)c);
        float acc = 0.0f;
        for (int k = 0; k < rank; ++k)
            acc = fmaf(hash_u01(seed, 11, r, (unsigned long long)k), hash_u01(seed, 12, (unsigned long long)k, gc), acc);
        const float u = hash_u01(seed, 13, r, gc);
        x[c * (size_t)ld + r] = clampE(acc * (1.0f + noise * (2.0f * u - 1.0f)));
    }
}

// The BF16 forms of the correction operands: bf = bf16(x), lb = bf16(x - tf32(x)), round to nearest.
// Four consecutive elements as one 8-byte store each.
__device__ __forceinline__ void store_bf16_pair4(unsigned short* __restrict__ bf, unsigned short* __restrict__ lb, size_t e,
                                                 const float4& v, const float4& h) {
    *reinterpret_cast<uint2*>(bf + e) = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
    *reinterpret_cast<uint2*>(lb + e) = make_uint2(pack_bf16x2(v.x - h.x, v.y - h.y), pack_bf16x2(v.z - h.z, v.w - h.w));
}

__global__ void k_split_tf32(const float* __restrict__ x, float* __restrict__ hi, float* __restrict__ lo, size_t n,
                             unsigned short* __restrict__ bf, unsigned short* __restrict__ lb) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float h, l;
        tf32_split(x[i], h, l);
        hi[i] = h;
        lo[i] = l;
        if (bf != nullptr) {
            bf[i] = (unsigned short)(pack_bf16x2(x[i], x[i]) & 0xffffu);
            lb[i] = (unsigned short)(pack_bf16x2(x[i] - h, x[i] - h) & 0xffffu);
        }
    }
}

// R1: one block per column of W (contiguous, coalesced), 256 threads, shuffle tree.
__global__ void __launch_bounds__(256) k_colsum(const float* __restrict__ W, float* __restrict__ sumW,
                                                int Mv, int ld, int refcompat) {
    pdl_sync();
    const float* col = W + (size_t)blockIdx.x * ld;
    float s = 0.0f;
    if (refcompat) {
        for (int m = threadIdx.x; m < Mv; m += 256) if (ref_sum_mask(m)) s += col[m];
    } else {
        for (int m = threadIdx.x; m < Mv; m += 256) s += col[m];
    }
    __shared__ float part[8];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        float v = threadIdx.x < 8 ? part[threadIdx.x] : 0.0f;
        v = warp_sum(v);
        if (threadIdx.x == 0) sumW[blockIdx.x] = v;
    }
}

// R2 pass 1: a block is KX threads along k (each owns 4 consecutive rows k: one 16-byte load per column, the KX threads of
// a column read KX*16 contiguous bytes) x NY column lanes, and covers kRowsumChunk columns; every thread keeps 4 independent
// float4 accumulators, i.e. 64 bytes in flight (with one scalar load in flight per thread the kernel ran at 54 % of the
// HBM peak: latency-, not bandwidth-bound).  partial[chunk][k]; fixed summation order.
constexpr int kRowsumChunk = 512;
__global__ void __launch_bounds__(256) k_rowsum_partial(const float* __restrict__ H, float* __restrict__ partials,
                                                        int Kv, int Nv, long long n_offset, int refcompat) {
    pdl_sync();
    const int KX = blockDim.x, NY = blockDim.y;
    const int k4 = blockIdx.x * KX + threadIdx.x;              // float4 index along k
    const bool k_ok = 4 * k4 < Kv;
    const int n0 = blockIdx.y * kRowsumChunk;
    const int n1 = min(Nv, n0 + kRowsumChunk);
    float4 a[4];
    #pragma unroll
    for (int u = 0; u < 4; ++u) a[u] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (k_ok) {
        const float4* Hp = reinterpret_cast<const float4*>(H) + k4;
        const size_t ld4 = (size_t)(Kv >> 2);
        for (int n = n0 + threadIdx.y; n < n1; n += 4 * NY) {
            float4 v[4];
            #pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int nn = n + u * NY;
                const bool take = nn < n1 && (!refcompat || ref_sum_mask(n_offset + nn));
                v[u] = take ? Hp[(size_t)nn * ld4] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            }
            #pragma unroll
            for (int u = 0; u < 4; ++u) { a[u].x += v[u].x; a[u].y += v[u].y; a[u].z += v[u].z; a[u].w += v[u].w; }
        }
    }
    __shared__ float4 sm[256];
    float4 t;
    t.x = (a[0].x + a[1].x) + (a[2].x + a[3].x); t.y = (a[0].y + a[1].y) + (a[2].y + a[3].y);
    t.z = (a[0].z + a[1].z) + (a[2].z + a[3].z); t.w = (a[0].w + a[1].w) + (a[2].w + a[3].w);
    sm[threadIdx.y * KX + threadIdx.x] = t;
    __syncthreads();
    if (threadIdx.y == 0 && k_ok) {
        for (int j = 1; j < NY; ++j) { const float4 v = sm[j * KX + threadIdx.x]; t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w; }
        *reinterpret_cast<float4*>(partials + (size_t)blockIdx.y * Kv + 4 * k4) = t;
    }
}

// W update: one thread per 4 consecutive rows of one column (float4, ld is a multiple of 32).
__global__ void __launch_bounds__(256) k_update_w(float* __restrict__ W, float* __restrict__ Whi, float* __restrict__ Wlo,
                                                  const float* __restrict__ P, int nsplit, size_t split_stride,
                                                  const float* __restrict__ sumH, int Mv, int Kv, int divide,
                                                  unsigned short* __restrict__ Wbf, unsigned short* __restrict__ Wlb) {
    pdl_sync();
    const size_t total4 = (size_t)Mv * Kv / 4;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (size_t)gridDim.x * blockDim.x) {
        const size_t e = i * 4;
        const int k = (int)(e / Mv);
        float4 acc = *reinterpret_cast<const float4*>(P + e);
        for (int s = 1; s < nsplit; ++s) {                       // ordered: split 0,1,2,...
            const float4 p = *reinterpret_cast<const float4*>(P + (size_t)s * split_stride + e);
            acc.x += p.x; acc.y += p.y; acc.z += p.z; acc.w += p.w;
        }
        if (divide) {                                            // row_div, cuda/matrix.cu:220-224
            const float d = clampE(sumH[k]);
            acc.x = acc.x / d; acc.y = acc.y / d; acc.z = acc.z / d; acc.w = acc.w / d;
        }
        float4 w = *reinterpret_cast<float4*>(W + e);
        w.x *= acc.x; w.y *= acc.y; w.z *= acc.z; w.w *= acc.w;  // vec_mul, cuda/matrix.cu:174-180
        *reinterpret_cast<float4*>(W + e) = w;
        if (Whi != nullptr) {
            float4 h, l;
            tf32_split(w.x, h.x, l.x); tf32_split(w.y, h.y, l.y);
            tf32_split(w.z, h.z, l.z); tf32_split(w.w, h.w, l.w);
            *reinterpret_cast<float4*>(Whi + e) = h;
            *reinterpret_cast<float4*>(Wlo + e) = l;
            if (Wbf != nullptr) store_bf16_pair4(Wbf, Wlb, e, w, h);
        }
    }
}

// W update, one block per column k (small problems: fewer launches).  Deterministic: thread-strided partial sums, a fixed
// shuffle tree per warp, then warp 0 over the warp sums.
__device__ __forceinline__ float block_sum_512(float v, float* red) {      // every thread returns the block's sum
    v = warp_sum(v);
    __syncthreads();                                     // `red` may still be read from a previous call
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = (threadIdx.x & 31) < 16 ? red[threadIdx.x & 31] : 0.0f;
    return warp_sum(t);
}
__global__ void __launch_bounds__(512) k_update_w_col(float* __restrict__ W, float* __restrict__ Whi, float* __restrict__ Wlo,
                                                      const float* __restrict__ P, int nsplit, size_t split_stride,
                                                      float* __restrict__ sumH, const float* __restrict__ rs_partials, int rs_blocks,
                                                      int Mv, int Kv, int divide, float* __restrict__ sumW, int refcompat,
                                                      unsigned short* __restrict__ Wbf, unsigned short* __restrict__ Wlb) {
    pdl_sync();
    __shared__ float red[16];
    const int k = blockIdx.x;
    float d;
    if (rs_partials != nullptr) {                        // sum_rows of the new H (nmf.cu:164): ordered fold of the H update's partials
        float s = 0.0f;
        for (int c = threadIdx.x; c < rs_blocks; c += 512) s += rs_partials[(size_t)c * Kv + k];
        s = block_sum_512(s, red);
        if (threadIdx.x == 0) sumH[k] = s;
        d = clampE(s);
    } else d = clampE(sumH[k]);
    float cs = 0.0f;
    const size_t col = (size_t)k * Mv;
    for (int m = 4 * threadIdx.x; m < Mv; m += 4 * 512) {
        const size_t e = col + m;
        float4 acc = *reinterpret_cast<const float4*>(P + e);
        for (int s = 1; s < nsplit; ++s) {                       // ordered: split 0,1,2,...
            const float4 p = *reinterpret_cast<const float4*>(P + (size_t)s * split_stride + e);
            acc.x += p.x; acc.y += p.y; acc.z += p.z; acc.w += p.w;
        }
        if (divide) { acc.x = acc.x / d; acc.y = acc.y / d; acc.z = acc.z / d; acc.w = acc.w / d; }   // row_div, cuda/matrix.cu:220-224
        float4 w = *reinterpret_cast<float4*>(W + e);
        w.x *= acc.x; w.y *= acc.y; w.z *= acc.z; w.w *= acc.w;  // vec_mul, cuda/matrix.cu:174-180
        *reinterpret_cast<float4*>(W + e) = w;
        if (Whi != nullptr) {
            float4 h, l;
            tf32_split(w.x, h.x, l.x); tf32_split(w.y, h.y, l.y);
            tf32_split(w.z, h.z, l.z); tf32_split(w.w, h.w, l.w);
            *reinterpret_cast<float4*>(Whi + e) = h;
            *reinterpret_cast<float4*>(Wlo + e) = l;
            if (Wbf != nullptr) store_bf16_pair4(Wbf, Wlb, e, w, h);
        }
        if (refcompat) {                                         // the reference's stale-read tail: rows m mod 64 in {0,1,2,4,8,16,32}
            cs += (ref_sum_mask(m) ? w.x : 0.0f) + (ref_sum_mask(m + 1) ? w.y : 0.0f) + (ref_sum_mask(m + 2) ? w.z : 0.0f) +
                  (ref_sum_mask(m + 3) ? w.w : 0.0f);
        } else cs += (w.x + w.y) + (w.z + w.w);
    }
    if (sumW != nullptr) {                                       // sum_cols of the new W (nmf.cu:134 of the NEXT iteration)
        cs = block_sum_512(cs, red);
        if (threadIdx.x == 0) sumW[k] = cs;
    }
}

// H update from split partial sums of W'*Y (small problems: the H GEMM is split over the rows of W so
// that it fills the SMs): H <- H .* ((sum_s P[s]) ./ max(sumW,EPS)), ordered sum, emits the TF32 pair.
// One thread per 4 consecutive k of one column (float4; Kv is a multiple of 32).
// ROWSUM (Kv divides 1024: every thread keeps the same 4 rows k over its grid-stride loop): the block also emits
// its partial row sums of the NEW H, rs_partials[block][k] (refcompat: only the columns of the reference's subsample),
// so that the W half needs no pass of its own over H: k_rowsum_finish2 folds the blocks in a fixed order.
template <bool ROWSUM>
__global__ void __launch_bounds__(256) k_update_h(float* __restrict__ H, float* __restrict__ Hhi, float* __restrict__ Hlo,
                                                  const float* __restrict__ P, int nsplit, size_t split_stride,
                                                  const float* __restrict__ sumW, int Kv, size_t total4,
                                                  float* __restrict__ rs_partials, long long n_offset, int refcompat,
                                                  unsigned short* __restrict__ Hbf, unsigned short* __restrict__ Hlb) {
    pdl_sync();
    float4 rs = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (size_t)gridDim.x * blockDim.x) {
        const size_t e = i * 4;
        const int k = (int)(e % (size_t)Kv);
        float4 acc = *reinterpret_cast<const float4*>(P + e);
        for (int s = 1; s < nsplit; ++s) {
            const float4 p = *reinterpret_cast<const float4*>(P + (size_t)s * split_stride + e);
            acc.x += p.x; acc.y += p.y; acc.z += p.z; acc.w += p.w;
        }
        const float4 sw = *reinterpret_cast<const float4*>(sumW + k);
        acc.x = acc.x / clampE(sw.x); acc.y = acc.y / clampE(sw.y);      // col_div, cuda/matrix.cu:244-250
        acc.z = acc.z / clampE(sw.z); acc.w = acc.w / clampE(sw.w);
        float4 h = *reinterpret_cast<float4*>(H + e);
        h.x *= acc.x; h.y *= acc.y; h.z *= acc.z; h.w *= acc.w;          // vec_mul, cuda/matrix.cu:174-180
        *reinterpret_cast<float4*>(H + e) = h;
        if (Hhi != nullptr) {
            float4 hi, lo;
            tf32_split(h.x, hi.x, lo.x); tf32_split(h.y, hi.y, lo.y);
            tf32_split(h.z, hi.z, lo.z); tf32_split(h.w, hi.w, lo.w);
            *reinterpret_cast<float4*>(Hhi + e) = hi;
            *reinterpret_cast<float4*>(Hlo + e) = lo;
            if (Hbf != nullptr) store_bf16_pair4(Hbf, Hlb, e, h, hi);
        }
        if (ROWSUM && (!refcompat || ref_sum_mask(n_offset + (long long)(e / (size_t)Kv)))) {
            rs.x += h.x; rs.y += h.y; rs.z += h.z; rs.w += h.w;
        }
    }
    if (ROWSUM) {
        // threads tid, tid + Kv/4, tid + 2 Kv/4, .. own the same rows (256 and the grid stride are multiples of Kv/4)
        __shared__ float4 sm[256];
        sm[threadIdx.x] = rs;
        __syncthreads();
        const int k4n = Kv >> 2;
        if ((int)threadIdx.x < k4n) {
            float4 t = sm[threadIdx.x];
            for (int j = (int)threadIdx.x + k4n; j < 256; j += k4n) { const float4 v = sm[j]; t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w; }
            *reinterpret_cast<float4*>(rs_partials + (size_t)blockIdx.x * Kv + 4 * threadIdx.x) = t;
        }
    }
}

// ordered fold of per-block partial row sums: block = 32 rows k x 32 groups; group g sums blocks g, g + 32, ..
__global__ void __launch_bounds__(1024) k_rowsum_finish2(const float* __restrict__ partials, float* __restrict__ sumH, int Kv, int nblocks) {
    pdl_sync();
    __shared__ float sm[32][33];
    const int k = blockIdx.x * 32 + threadIdx.x;
    float s = 0.0f;
    if (k < Kv) for (int c = threadIdx.y; c < nblocks; c += 32) s += partials[(size_t)c * Kv + k];
    sm[threadIdx.y][threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.y == 0 && k < Kv) {
        float t = 0.0f;
        #pragma unroll
        for (int j = 0; j < 32; ++j) t += sm[j][threadIdx.x];
        sumH[k] = t;
    }
}

__global__ void __launch_bounds__(256) k_fold_splits(const float* __restrict__ P, int nsplit, size_t split_stride,
                                                     float* __restrict__ out, size_t n4) {
    pdl_sync();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 acc = *reinterpret_cast<const float4*>(P + i * 4);
        for (int s = 1; s < nsplit; ++s) {
            const float4 p = *reinterpret_cast<const float4*>(P + (size_t)s * split_stride + i * 4);
            acc.x += p.x; acc.y += p.y; acc.z += p.z; acc.w += p.w;
        }
        *reinterpret_cast<float4*>(out + i * 4) = acc;
    }
}

__global__ void __launch_bounds__(1024) k_sum_partials(const double* __restrict__ partials, int n, double* __restrict__ out) {
    pdl_sync();
    // one block; thread-strided ordered sums, a fixed shuffle tree per warp, then warp 0 over the 32 warp sums:
    // deterministic (no atomics), and wide enough for the ~1e5 per-warp partials of the fused H step
    __shared__ double wsum[32];
    partials += (size_t)blockIdx.x * n;                       // one block per batch entry (restart)
    out += blockIdx.x;
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += 1024) s += partials[i];
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        s = wsum[threadIdx.x];
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (threadIdx.x == 0) *out = s;
    }
}

// ---- W step of the tail rows (see common.hpp) ------------------------------------------------------------------------
constexpr int kTailCols = 512;            // columns per block (8 warps x 64 columns)
// KJ = Kr / 32 values of k per lane.  A warp walks columns; its lanes hold W[tail rows, k] and the running sums for their k.
template <int KJ>
__global__ void __launch_bounds__(256) k_w_tail(const float* __restrict__ W, const float* __restrict__ H, const float* __restrict__ X,
                                                float4* __restrict__ part, int Mv, int ldH, int Kr, int Nv, int row0, int t) {
    pdl_sync();
    const int koff = blockIdx.y * Kr;                       // restart
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float w[4][KJ], acc[4][KJ];
    #pragma unroll
    for (int i = 0; i < 4; ++i)
        #pragma unroll
        for (int j = 0; j < KJ; ++j) {
            w[i][j] = i < t ? W[(size_t)(koff + lane + 32 * j) * Mv + row0 + i] : 0.0f;
            acc[i][j] = 0.0f;
        }
    const int n0 = blockIdx.x * kTailCols, n1 = min(Nv, n0 + kTailCols);
    // two columns per trip (n and n + 8): two independent load -> reduce -> divide chains in flight per warp
    for (int n = n0 + warp; n < n1; n += 16) {
        const bool two = n + 8 < n1;
        const int nb = two ? n + 8 : n;
        float ha[KJ], hb[KJ];
        const float* hca = H + (size_t)n * ldH + koff + lane;
        const float* hcb = H + (size_t)nb * ldH + koff + lane;
        #pragma unroll
        for (int j = 0; j < KJ; ++j) { ha[j] = hca[32 * j]; hb[j] = hcb[32 * j]; }
        const float xa = lane < t ? X[(size_t)n * Mv + row0 + lane] : 0.0f;      // lane i holds x of tail row i
        const float xb = (two && lane < t) ? X[(size_t)nb * Mv + row0 + lane] : 0.0f;   // (0 for a missing second column: adds nothing)
        #pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (i < t) {                                    // uniform
                float sa = 0.0f, sb = 0.0f;
                #pragma unroll
                for (int j = 0; j < KJ; ++j) { sa = fmaf(w[i][j], ha[j], sa); sb = fmaf(w[i][j], hb[j], sb); }
                sa = warp_sum(sa); sb = warp_sum(sb);       // (W*H)[m,n], every lane
                const float ya = __shfl_sync(0xffffffffu, xa, i) / clampE(sa);  // set_epsilon + vec_div, nmf.cu:158-161
                const float yb = __shfl_sync(0xffffffffu, xb, i) / clampE(sb);
                #pragma unroll
                for (int j = 0; j < KJ; ++j) acc[i][j] = fmaf(yb, hb[j], fmaf(ya, ha[j], acc[i][j]));
            }
        }
    }
    // the 8 warps of the block, in order
    __shared__ float sm[8][4][KJ * 32];
    #pragma unroll
    for (int i = 0; i < 4; ++i)
        #pragma unroll
        for (int j = 0; j < KJ; ++j) sm[warp][i][lane + 32 * j] = acc[i][j];
    __syncthreads();
    for (int k = threadIdx.x; k < Kr; k += 256) {
        float4 v = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        for (int wv = 0; wv < 8; ++wv) { v.x += sm[wv][0][k]; v.y += sm[wv][1][k]; v.z += sm[wv][2][k]; v.w += sm[wv][3][k]; }
        part[(size_t)blockIdx.x * ((size_t)gridDim.y * Kr) + koff + k] = v;
    }
}

// one block per stacked k: ordered fold of the per-block partial sums (double), then the tail rows of P: split 0 = the sums,
// every other split = 0 (row0 is a multiple of 4 and ld a multiple of 32: one aligned float4 per column)
__global__ void __launch_bounds__(256) k_w_tail_fold(const float4* __restrict__ part, int nblocks, int Kv, float* __restrict__ P, int nsplit,
                                                     size_t split_stride, int Mv, int row0) {
    pdl_sync();
    const int k = blockIdx.x;
    double a[4] = {0.0, 0.0, 0.0, 0.0};
    for (int b = threadIdx.x; b < nblocks; b += 256) {
        const float4 v = part[(size_t)b * Kv + k];
        a[0] += v.x; a[1] += v.y; a[2] += v.z; a[3] += v.w;
    }
    __shared__ double red[8][4];
    #pragma unroll
    for (int i = 0; i < 4; ++i) {
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) a[i] += __shfl_xor_sync(0xffffffffu, a[i], o);
    }
    if ((threadIdx.x & 31) == 0) { for (int i = 0; i < 4; ++i) red[threadIdx.x >> 5][i] = a[i]; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double s[4] = {0.0, 0.0, 0.0, 0.0};
        for (int wv = 0; wv < 8; ++wv) for (int i = 0; i < 4; ++i) s[i] += red[wv][i];
        float* p0 = P + (size_t)k * Mv + row0;
        *reinterpret_cast<float4*>(p0) = make_float4((float)s[0], (float)s[1], (float)s[2], (float)s[3]);
        for (int sp = 1; sp < nsplit; ++sp) *reinterpret_cast<float4*>(p0 + (size_t)sp * split_stride) = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
}

__global__ void k_gather_rows(const float* __restrict__ src, int ld, int cols, const int* __restrict__ rows, int count,
                              float* __restrict__ dst) {
    const size_t total = (size_t)cols * count;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const size_t j = e / count; const int i = (int)(e - j * count);
        dst[e] = src[j * (size_t)ld + rows[i]];
    }
}

inline int grid_for(size_t total, int block, int max_blocks = 148 * 16) {
    size_t g = (total + block - 1) / block;
    if (g < 1) g = 1;
    if (g > (size_t)max_blocks) g = max_blocks;
    return (int)g;
}

}  // namespace

void launch_clamp_eps(float* a, int rows, int cols, int ld, cudaStream_t st) {
    if ((size_t)rows * cols == 0) return;
    k_clamp_eps<<<grid_for((size_t)rows * cols, 256), 256, 0, st>>>(a, rows, cols, ld);
    CUDA_CHECK_LAUNCH();
}

void launch_fill_region(float* a, int ld, int r0, int r1, int c0, int c1, float value, cudaStream_t st) {
    if (r1 <= r0 || c1 <= c0) return;
    k_fill_region<<<grid_for((size_t)(r1 - r0) * (c1 - c0), 256), 256, 0, st>>>(a, ld, r0, r1, c0, c1, value);
    CUDA_CHECK_LAUNCH();
}

void launch_fill_uniform(float* a, int rows, int cols, int ld, long long col_offset,
                         unsigned long long seed, int tag, cudaStream_t st) {
    if ((size_t)rows * cols == 0) return;
    k_fill_uniform<<<grid_for((size_t)rows * cols, 256), 256, 0, st>>>(a, rows, cols, ld, col_offset, seed, tag);
    CUDA_CHECK_LAUNCH();
}

void launch_fill_planted(float* x, int rows, int cols, int ld, long long col_offset, unsigned long long seed,
                         int rank, float noise, cudaStream_t st) {
    if ((size_t)rows * cols == 0) return;
    k_fill_planted<<<grid_for((size_t)rows * cols, 256), 256, 0, st>>>(x, rows, cols, ld, col_offset, seed, rank, noise);
    CUDA_CHECK_LAUNCH();
}

void launch_split_tf32(const float* x, float* hi, float* lo, size_t n, cudaStream_t st, unsigned short* bf, unsigned short* lb) {
    if (n == 0) return;
    k_split_tf32<<<grid_for(n, 256), 256, 0, st>>>(x, hi, lo, n, bf, lb);
    CUDA_CHECK_LAUNCH();
}

void launch_colsum(const float* W, float* sumW, const Dims& d, cudaStream_t st) {
    launch_k(k_colsum, dim3(d.Kv), dim3(256), 0, st, W, sumW, d.Mv, d.Mv, d.refcompat);
}

int rowsum_partial_count(const Dims& d) {
    // enough for either producer: 512-column chunks (k_rowsum_partial) or the blocks of k_update_h<ROWSUM>
    const int chunks = (d.Nv + kRowsumChunk - 1) / kRowsumChunk;
    return std::max(chunks, grid_for((size_t)d.Kv * d.Nv / 4, 256));
}

void launch_rowsum(const float* H, float* partials, float* sumH, const Dims& d, cudaStream_t st) {
    const int nch = (d.Nv + kRowsumChunk - 1) / kRowsumChunk;
    const int KX = std::min(d.Kv / 4, 64), NY = 256 / KX;
    dim3 grid((d.Kv / 4 + KX - 1) / KX, nch), block(KX, NY);
    launch_k(k_rowsum_partial, grid, block, 0, st, H, partials, d.Kv, d.Nv, d.n_offset, d.refcompat);
    // ordered fold of the chunks: 32 groups of chunks in parallel per k (one thread per k walking all chunks was a
    // chain of ~2000 dependent loads at 10^6 columns: 100+ us, more than the pass over H itself)
    launch_k(k_rowsum_finish2, dim3((d.Kv + 31) / 32), dim3(32, 32), 0, st, (const float*)partials, sumH, d.Kv, nch);
}

// ---------------------------------------------------------------------------------------------
// Row sums of H without any allocation (nmf_rowsum_noalloc, include/nmf.h): the same two kernels as launch_rowsum, with
// their per-chunk partial sums in a file-scope device array instead of a session's cudaMalloc'ed buffer — for callers
// that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid it).  4 MB per GPU that loads the
// library, held until exit.  Every call on a device shares the array: calls that overlap on the GPU (different streams)
// race on it.  Nothing in it has to start at zero — pass 1 overwrites every partial that pass 2 reads — so the previous
// call's contents are harmless.
// ---------------------------------------------------------------------------------------------
constexpr size_t kNoAllocScratchFloats = (size_t)1 << 20;
static __device__ __align__(256) float g_rowsum_scratch[kNoAllocScratchFloats];

size_t rowsum_noalloc_max_floats() { return kNoAllocScratchFloats; }

void rowsum_noalloc(const float* H_dev, int K, int N, float* sumH_dev, cudaStream_t st) {
    NMF_CHECK(H_dev != nullptr && sumH_dev != nullptr, "nmf_rowsum_noalloc: NULL pointer");
    NMF_CHECK(K > 0 && N > 0, "nmf_rowsum_noalloc: K and N must be positive");
    NMF_CHECK(K % 4 == 0 && ((uintptr_t)H_dev & 15) == 0, "nmf_rowsum_noalloc: K must be a multiple of 4 and H 16-byte aligned");
    const int nch = (N + kRowsumChunk - 1) / kRowsumChunk;
    NMF_CHECK((size_t)nch * (size_t)K <= kNoAllocScratchFloats && nch <= 65535,
              "nmf_rowsum_noalloc: ceil(N/512) * K = " + std::to_string((size_t)nch * (size_t)K) + " partial sums exceed the built-in scratch of " +
              std::to_string(kNoAllocScratchFloats) + " floats (4 MB): use a session, which allocates its own");
    // In host code `g_rowsum_scratch` is only a stand-in for the symbol, not a device address: resolve it once per device,
    // with that device current (the caller's current device is the one the pointers belong to)
    static float* scratch_of[64] = {nullptr};
    static std::mutex mu;
    int dev = 0;
    CUDA_CHECK(cudaGetDevice(&dev));
    NMF_CHECK(dev >= 0 && dev < 64, "nmf_rowsum_noalloc: device ordinal out of range");
    float* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!scratch_of[dev]) {
            float* p = nullptr;
            CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_rowsum_scratch));
            NMF_CHECK(p != nullptr, "nmf_rowsum_noalloc: cudaGetSymbolAddress returned NULL");
            scratch_of[dev] = p;
        }
        scratch = scratch_of[dev];
    }
    const int KX = std::min(K / 4, 64), NY = 256 / KX;
    dim3 grid((K / 4 + KX - 1) / KX, nch), block(KX, NY);
    launch_k(k_rowsum_partial, grid, block, 0, st, H_dev, scratch, K, N, 0ll, 0);
    launch_k(k_rowsum_finish2, dim3((K + 31) / 32), dim3(32, 32), 0, st, (const float*)scratch, sumH_dev, K, nch);
}

// the one-block-per-column kernel pays when the launch count matters more than the bytes: up to 64 MB of partials + W
bool update_w_by_column(const Dims& d, int nsplit) { return d.szW() * (size_t)(nsplit + 4) * 4 <= ((size_t)64 << 20); }

bool launch_update_w(float* W, float* Whi, float* Wlo, const float* P, int nsplit, size_t split_stride,
                     float* sumH, const Dims& d, cudaStream_t st, const float* rs_partials, int rs_blocks, float* sumW_out,
                     unsigned short* Wbf, unsigned short* Wlb) {
    if (update_w_by_column(d, nsplit)) {
        launch_k(k_update_w_col, dim3(d.Kv), dim3(512), 0, st, W, Whi, Wlo, P, nsplit, split_stride, sumH, rs_partials, rs_blocks,
                 d.Mv, d.Kv, d.w_divide, sumW_out, d.refcompat, Wbf, Wlb);
        return sumW_out != nullptr;
    }
    NMF_CHECK(rs_partials == nullptr, "launch_update_w: fold the row-sum partials first for this shape");
    launch_k(k_update_w, dim3(grid_for(d.szW() / 4, 256)), dim3(256), 0, st, W, Whi, Wlo, P, nsplit, split_stride, (const float*)sumH,
             d.Mv, d.Kv, d.w_divide, Wbf, Wlb);
    return false;
}

int launch_update_h(float* H, float* Hhi, float* Hlo, const float* P, int nsplit, size_t split_stride,
                    const float* sumW, const Dims& d, cudaStream_t st, float* rowsum_partials, unsigned short* Hbf, unsigned short* Hlb) {
    const size_t total4 = (size_t)d.Kv * d.Nv / 4;
    const int grid = grid_for(total4, 256);
    const bool rowsum = rowsum_partials != nullptr && d.Kv <= 1024 && 1024 % d.Kv == 0;   // every thread keeps the same 4 rows k over its loop
    if (rowsum) launch_k(k_update_h<true>, dim3(grid), dim3(256), 0, st, H, Hhi, Hlo, P, nsplit, split_stride, sumW, d.Kv, total4, rowsum_partials, d.n_offset, d.refcompat, Hbf, Hlb);
    else        launch_k(k_update_h<false>, dim3(grid), dim3(256), 0, st, H, Hhi, Hlo, P, nsplit, split_stride, sumW, d.Kv, total4, (float*)nullptr, 0ll, 0, Hbf, Hlb);
    return rowsum ? grid : 0;
}

void launch_rowsum_finish2(const float* partials, float* sumH, int nblocks, const Dims& d, cudaStream_t st) {
    launch_k(k_rowsum_finish2, dim3((d.Kv + 31) / 32), dim3(32, 32), 0, st, partials, sumH, d.Kv, nblocks);
}

int w_tail_blocks(int Nv) { return (Nv + kTailCols - 1) / kTailCols; }

void launch_w_tail(const float* W, const float* H, const float* X, float* part, const Dims& d, int row0, int t,
                   float* P, int nsplit, size_t split_stride, cudaStream_t st) {
    NMF_CHECK(t >= 1 && t <= 4 && row0 % 4 == 0 && d.Kr % 64 == 0 && d.Kr <= 256, "launch_w_tail: unsupported shape");
    const dim3 grid(w_tail_blocks(d.Nv), d.R), block(256);
    float4* part4 = reinterpret_cast<float4*>(part);
    switch (d.Kr / 32) {
        case 2: launch_k(k_w_tail<2>, grid, block, 0, st, W, H, X, part4, d.Mv, d.Kv, d.Kr, d.Nv, row0, t); break;
        case 4: launch_k(k_w_tail<4>, grid, block, 0, st, W, H, X, part4, d.Mv, d.Kv, d.Kr, d.Nv, row0, t); break;
        case 6: launch_k(k_w_tail<6>, grid, block, 0, st, W, H, X, part4, d.Mv, d.Kv, d.Kr, d.Nv, row0, t); break;
        default: launch_k(k_w_tail<8>, grid, block, 0, st, W, H, X, part4, d.Mv, d.Kv, d.Kr, d.Nv, row0, t); break;
    }
    launch_k(k_w_tail_fold, dim3(d.Kv), dim3(256), 0, st, (const float4*)part4, (int)grid.x, d.Kv, P, nsplit, split_stride, d.Mv, row0);
}

void launch_gather_rows(const float* src, int ld, int cols, const int* rows, int count, float* dst, cudaStream_t st) {
    if ((size_t)cols * count == 0) return;
    k_gather_rows<<<grid_for((size_t)cols * count, 256), 256, 0, st>>>(src, ld, cols, rows, count, dst);
    CUDA_CHECK_LAUNCH();
}

void launch_fold_splits(const float* P, int nsplit, size_t split_stride, float* out, size_t n, cudaStream_t st) {
    launch_k(k_fold_splits, dim3(grid_for(n / 4, 256)), dim3(256), 0, st, P, nsplit, split_stride, out, n / 4);
}

void launch_sum_partials(const double* partials, int n, int nbatch, double* out, cudaStream_t st) {
    launch_k(k_sum_partials, dim3(nbatch), dim3(1024), 0, st, partials, n, out);
}

}  // namespace nmf
