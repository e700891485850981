// umma_fused.cuh — parameters and launcher of the fused half-iteration kernels (kernels_fused.cu), shared with the
// launch plan (umma_host.inc).
#pragma once
#include "umma_common.cuh"

namespace nmf { namespace umma {

constexpr int F_STAGES = 3;
constexpr int F_STAGE_BYTES = 64 * 1024;           // A hi, A lo (16 KB each) + B hi, B lo (16 KB each)
constexpr int F_B_TILE = 16 * 1024;
constexpr int F_BN = 256;                          // S tile width (one S tile + P fill the 512 TMEM columns when K = 256)
constexpr int F_BN_DB = 128;                       // two 128-wide S tiles (double-buffered) + P
constexpr int F_NCH = F_BN / 32;                   // 32-column chunks per tile (upper bound)
// NW = worker warps per TMEM lane quadrant (each takes 32 / NW columns of every S chunk and every NW-th chunk of P):
//   NW = 2 : 384 threads, setmaxnreg 48 / 224 (launch gives 168: 120 x 128 released >= 56 x 256 acquired); P up to 256 columns
//   NW = 4 : 640 threads, setmaxnreg 40 / 104 (launch gives 96: 56 x 128 released >= 8 x 512 acquired); P up to 128 columns —
//            with K <= 128 the MMAs of a chunk take half as long and the quotient is the bound, so twice the warps work on it
template <int NW> struct FCfg {
    static constexpr int NTHREADS = 128 + 128 * NW;
    static constexpr int REGS_CTRL = NW == 2 ? 48 : 40, REGS_WORK = NW == 2 ? 224 : 104;
    static constexpr int CW = 32 / NW;                 // S / X / Y columns per thread and chunk
    static constexpr int ACC = NW == 2 ? 4 : 1;        // 32-column chunks of P a thread accumulates (chunks h, h + NW, ..)
    static constexpr int NARR = 4 * NW;                // worker warps per CTA (arrival counts)
};
constexpr int F_WORK_WARP0 = 4;
constexpr int F_YRING = 4;                         // yfull barriers (> ring depth, see the workers)
constexpr int F_XBUF_BYTES = BM * 32 * 4;          // one X chunk of this CTA: 128 lane rows x 32 columns fp32
constexpr int F_XBUFS = 2;
constexpr int F_EMPTY_WAYS = 4;                    // `empty` barriers per ring slot, taken in turn by the slot's uses (see the kernel)
constexpr int F_SMEM_BYTES = F_STAGES * F_STAGE_BYTES + F_XBUFS * F_XBUF_BYTES + 1024 /*align*/ + 512 /*barriers*/;

// Restarts (nbatch > 1, the batched random initialisations of the paper's section 3.2): R independent factor pairs that
// share X, stacked along k — W_all [Mv x R*Kv] (restart r = columns r*Kv ..), H_all [R*Kv x Nv] with leading dimension
// ldH = R*Kv (restart r = rows r*Kv ..).  A work item belongs to one restart; its two GEMMs only touch that restart's
// k range (the block-diagonal structure), so the kernel sees one more item index and a k offset on its TMA coordinates
// and output addresses.  Every other kernel of the iteration is elementwise or a per-k sum and runs on the stacked
// arrays as they are.
struct alignas(64) FusedParams {
    CUtensorMap mapWhi, mapWlo;      // W pair [Mv x R*Kv], MN-major boxes 32 x 32    (a-stages: A in the W step, B in the H step)
    CUtensorMap mapHhi, mapHlo;      // H pair [R*Kv x Nv], K-major boxes 32 x 128    (a-stages: B in the W step, A in the H step)
    CUtensorMap mapGhi, mapGlo;      // b-stage B: W step H pair, MN-major boxes 32 x 32; H step W pair, K-major boxes 32 x Kv/2
    // corr (BF16 correction passes, K = 256): the same three operands as BF16 arrays — bf = bf16(x), lb = bf16(x - tf32(x)):
    // MN-major boxes 64 x 32 (128-byte swizzle), K-major boxes 32 x rows (64-byte swizzle)
    CUtensorMap mapWbf, mapWlb, mapHbf, mapHlb, mapGbf, mapGlb;
    CUtensorMap mapX;                // X [Mv x Nv]: W step boxes 128 m x 32 n (no swizzle), H step boxes 32 m x 128 n (128B swizzle)
    int Ldim, Cdim, Kv;              // extents of the lane-row and tile-column dimensions (W step Mv, Nv; H step Nv, Mv); Kv of ONE restart
    int tiles_l, nsplit, ntiles_c;   // 256-row blocks of L, splits of the C range, 256-wide tiles of C
    int nbatch;                      // restarts (>= 1): work items = tiles_l * nsplit * nbatch, the restart index slowest
    int ldH;                         // leading dimension of H = nbatch * Kv
    int kblocks;                     // Kv / 32
    int period;                      // chunks accumulated in TMEM between promotions of P
    int hintW, hintH, hintS;
    // W step output: partial numerators P[split][m + k*ldP]; H step with nsplit > 1: partial W'*Y sums P[split][k + n*ldH]
    float* P; unsigned long long split_stride; int ldP;
    // H step with nsplit == 1: the update itself
    float* H; float* Hhi; float* Hlo; const float* sumW;
    // COST instantiation (H step of a check iteration): the KL divergence of the factors the step starts from
    // (reduce1d_div, cuda/matrix.cu:578-640) over the logical Llog x Clog block; one partial per worker warp and work item
    double* cost_partials; int Llog, Clog;
    int bn;                          // S tile width the launch runs with (256, or 128 double-buffered)
    int nw;                          // worker warps per TMEM lane quadrant (2 or 4)
    int ytmem;                       // the quotient chunk is staged in tensor memory (A operand of the second GEMM from TMEM); nw == 4 only
    int weave;                       // two S tiles: the next tile's a-stages are spread between this tile's b-stages in the ring
    int prod2;                       // two TMA producer warps (0 and 3) share the loads of every ring stage
    int drain_early;                 // quotient through shared memory: drain a pending period of P before waiting for the b-stage's ring slot
    int alt;                         // nw == 2: a worker warp takes all 32 columns of every other chunk (else 16 columns of every chunk)
    int corr;                        // 1: correction passes lo*hi and hi*lo run as kind::f16 BF16 MMAs (nw == 2, Kv == 256 only)
    unsigned short* Hbf; unsigned short* Hlb;   // corr, H step with nsplit == 1: the BF16 forms of the new H
    int defer;                       // H step: store W'*Y sums to P even when nsplit == 1, k_update_h applies the update
    unsigned int* err;
};

// One launch of k_fused<hstep, p.nw, cost, p.bn> on CTA pairs (kernels_fused.cu).
void launch_fused(bool hstep, bool cost, const FusedParams& p, cudaStream_t st);

}}  // namespace nmf::umma
