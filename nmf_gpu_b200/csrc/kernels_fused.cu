// kernels_fused.cu — the fused half-iteration kernels of the tcgen05 engine.
//
//   W step (HSTEP = false)   P[m,k] = sum_n ( X[m,n] / max( (W*H)[m,n], EPS ) ) * H[k,n]         nmf.cu:155-168
//   H step (HSTEP = true )   H[k,n] *= ( sum_m W[m,k] * X[m,n] / max( (W*H)[m,n], EPS ) ) / sumW[k]   nmf.cu:125-145
//
// i.e. a quotient of the iteration and the GEMM that consumes it in ONE kernel: the M x N quotient never
// exists in HBM.  Measured reason: at the power cap the 17 GB write + 17 GB read of the
// quotient per half-iteration cost more than the flops they feed.
//
// One CTA pair (cta_group::2) owns 256 "lane rows" L (rows m of W in the W step, columns n of H in the H
// step) and walks over 256-wide tiles of the other dimension C (n resp. m).  Per tile the tensor core
// alternates between two contractions that share one smem ring, in issue order:
//   a-stages (K/32 of them)  S[256 L x 256 C]  = sum_k ...                      -> TMEM columns   0..255
//        W step: A = W[L, kblock] (MN-major), B = H[kblock, C] (K-major)
//        H step: A = H[kblock, L] (K-major),  B = W[C, kblock] (MN-major)         (S is the transpose of W*H)
//   b-stages (8 of them)     P[256 L x K]     += Y[L, 32 C] * G[32 C, K]          -> TMEM columns 256..511
//        W step: G = H[:, 32 n]' (MN-major)      H step: G = W[32 m, :] (K-major)
// where the A operand of a b-stage is NOT loaded: the worker warps read the S chunk from TMEM, divide X by
// it, split the quotient into its TF32 (hi, lo) pair and write the pair straight into the stage's A slots
// in the K-major SWIZZLE_128B layout the MMA expects (what TMA would have produced from a stored Y).
// Every `period` chunks (default 4 = 128 columns, as the chunked kernels) the workers drain P into fp32
// registers with round-to-nearest adds (the TMEM accumulator truncates).
//
// Warp roles (384 threads, registers re-balanced with setmaxnreg 48 / 224 / 224):
//   warp 0       TMA producer   a-stage operands and the b-stage G operand (the ring)
//   warp 2       X loader       X chunks into a 2-deep buffer, each as soon as the workers release its slot
//   warp 1       MMA issuer     (leader CTA)
//   warps 4..11  workers        warp w owns TMEM lanes 32*(w%4)..+31 (lane rows) and half h = (w-4)/4 of the columns:
//                               quotient: columns 16h..16h+15 of every S chunk -> Y chunk (hi, lo) in smem;
//                               accumulation: P chunks h, h+2, .. TMEM -> registers every `period` chunks (one
//                               chunk late, so that the drain overlaps the MMAs fed by the next chunk's quotient);
//                               end of the work item: W step: partial sums -> P[split]; H step: the H update
//
// BN = 128 (K <= 128, where P leaves TMEM columns free): TWO S tiles of 128 columns.  The MMA warp issues the
// a-stages of tile g+1 BEFORE the b-stages of tile g, so S of the next tile is complete while the workers are still
// dividing the current one: they never wait for the S GEMM (with one S tile that wait was 26 % of their time at
// K = 128, ncu source page of C3) and the drain of a tile's P is taken after the first quotient chunk of the next tile.
#include <atomic>

#include "umma_fused.cuh"

namespace nmf { namespace umma {

// 16 consecutive fp32 columns of this warp's 32 TMEM lanes
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}

// 8 consecutive fp32 columns of this warp's 32 TMEM lanes
__device__ __forceinline__ void tmem_ld_32x8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, uint32_t (&r)[16]) { tmem_ld_32x16(taddr, r); }
__device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, uint32_t (&r)[8]) { tmem_ld_32x8(taddr, r); }

// registers -> TMEM: 8 / 16 consecutive 32-bit columns of this warp's 32 lanes (the mirror of the loads above)
__device__ __forceinline__ void tmem_st_cols(uint32_t taddr, const uint32_t (&r)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}
__device__ __forceinline__ void tmem_st_cols(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
                 :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
                    "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void tmem_st_x4(uint32_t taddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" :: "r"(taddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void stg128u(unsigned short* ptr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" :: "l"(ptr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// kind::f16 (BF16) 256-row MMA with the A operand in TENSOR MEMORY: 16 K elements = 8 columns (two BF16 per 32-bit cell, the
// lower k in the low half)
__device__ __forceinline__ void mma_bf16_cg2_ts(uint32_t d_tmem, uint32_t a_tmem, uint32_t blo, uint32_t bhi, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], db, %4, p;\n\t}"
        :: "r"(d_tmem), "r"(a_tmem), "r"(blo), "r"(bhi), "r"(idesc), "r"(accumulate) : "memory");
}

// 256-row MMA over the CTA pair with the A operand in TENSOR MEMORY (each CTA's own 128 rows x 8 tf32 columns at `a_tmem`;
// K-major by construction) and B in shared memory
__device__ __forceinline__ void mma_tf32_cg2_ts(uint32_t d_tmem, uint32_t a_tmem, uint32_t blo, uint32_t bhi, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], db, %4, p;\n\t}"
        :: "r"(d_tmem), "r"(a_tmem), "r"(blo), "r"(bhi), "r"(idesc), "r"(accumulate) : "memory");
}

// 32-byte global accesses (sm_100: LDG/STG.256): the H-step epilogue moves 128 contiguous bytes per lane row, and a
// warp instruction touches 32 lines either way, so twice the bytes per instruction halve its trips through the LSU
__device__ __forceinline__ void ldg256(const float* ptr, float (&r)[8]) {
    asm volatile("ld.global.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]), "=f"(r[7]) : "l"(ptr));
}
__device__ __forceinline__ void stg256(float* ptr, const float (&r)[8]) {
    asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 :: "l"(ptr), "f"(r[0]), "f"(r[1]), "f"(r[2]), "f"(r[3]), "f"(r[4]), "f"(r[5]), "f"(r[6]), "f"(r[7]) : "memory");
}

// YT (K <= 128, where P leaves 128 TMEM columns free): the quotient chunk goes to TENSOR MEMORY instead of shared memory
// and the second GEMM takes its A operand from there (tcgen05.mma with A in TMEM, the arrangement FlashAttention uses
// for P*V).  The hi half OVERWRITES the S chunk it was computed from (each worker warp reads and rewrites its own lanes
// and columns; the tensor pipe runs the MMAs of one issuer in order, so the a-stages that later reuse the S buffer cannot
// overtake the b-stages that read it), the lo half goes to a ring of four 32-column slots at columns 384..511, released
// by the b-stage's own commit (`yempty`): the hand-off is four chunks deep instead of the three ring slots.  Measured reason (ncu source page of the
// K = 64 kernels, profiles/r2_ncu_source_c5k64_fused_before.txt): 15-18 % of all warp samples sat on the
// fence.proxy.async that publishes the workers' shared-memory stores to the tensor core, and the swizzled store
// addressing is pure issue-slot cost in a loop whose CUDA-core work (~800 issue cycles per chunk and scheduler) already
// exceeds the chunk's MMA time (768 cycles at K = 64).  A TMEM store needs no proxy fence and no address arithmetic, and
// the b-stage MMAs stop reading 4 KB of A per dispatch from shared memory.
// CB (K = 256, two worker warps per quadrant): the two correction passes of every product, lo*hi and hi*lo, run
// as kind::f16 BF16 MMAs (K = 16 per instruction: half the tensor time of a TF32 pass each) on BF16 forms of the operands —
// bf = bf16(x) and lb = bf16(x - tf32(x)), which sit where the TF32 `lo` tile was (8 KB each: bf at +0, lb at +8 KB):
//   D += A_lb * B_bf + A_bf * B_lb (2 + 2 BF16 MMAs)  +  A_hi * B_hi (4 TF32 MMAs)      — 8 instead of 12 MMA time units.
// The correction terms are 2^-11 of the product and BF16 keeps 8 bits of each factor: ~2^-20 relative per product, the
// level of the dropped lo*lo term.  BF16 tiles: MN-major = TMA boxes of 64 x 32 with the 128-byte swizzle (descriptor
// layout SWIZZLE_128B, LBO 4096 between 64-wide boxes, SBO 1024 between 8-deep groups, 2048 B per K = 16 step);
// K-major = 64-byte rows with the 64-byte swizzle (SWIZZLE_64B, SBO 512, 32 B per K = 16 step).
template <bool HSTEP, int NW, bool COST, int BN, bool YT, bool CB>
__global__ void __launch_bounds__(FCfg<NW>::NTHREADS, 1) k_fused(const __grid_constant__ FusedParams p) {
    static_assert(!CB || (NW == 2 && !YT) || (NW == 4 && YT), "BF16 correction passes: K = 256 (two worker warps, quotient in shared memory) or K = 128 (four, quotient in tensor memory)");
    using FC = FCfg<NW>;
    constexpr int CW = FC::CW;
    constexpr int NCH = BN / 32;                       // 32-column chunks per S tile
    constexpr int SBUF = (BN == F_BN_DB) ? 2 : 1;      // S tiles in TMEM
    static_assert(SBUF * BN == F_BN, "S occupies TMEM columns 0..255, P starts at 256");
    constexpr uint32_t F_YT_COL = 384, F_YT_SLOTS = 4; // YT: ring of lo slots in TMEM (P then holds <= 128 columns: 256..383)
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* xbuf = smem + F_STAGES * F_STAGE_BYTES;                          // [2] X chunks
    uint64_t* bars = (uint64_t*)(xbuf + F_XBUFS * F_XBUF_BYTES);
    uint64_t* full = bars;               // [3]  TMA -> MMA                      (leader's is used)
    uint64_t* yfull = bars + 4;          // [4]  workers -> MMA: Y chunk written (both CTAs arrive at the leader's)
    uint64_t* sfull = bars + 8;          // [2]  MMA -> workers: S tile complete (multicast)
    uint64_t* sempty = bars + 10;        // [2]  workers -> MMA: S tile read     (leader's)
    uint64_t* pfull = bars + 12;         //      MMA -> workers: a period of P complete (multicast)
    uint64_t* pempty = bars + 13;        //      workers -> MMA: P drained       (leader's)
    uint64_t* xfull = bars + 14;         // [2]  TMA -> workers: X chunk landed  (CTA-local)
    uint64_t* xempty = bars + 16;        // [2]  workers -> TMA: X chunk read    (CTA-local)
    // [3][4] MMA -> TMA and workers: ring slot consumed (multicast to both CTAs).  FOUR barriers per slot, taken in turn
    // by the slot's successive uses: a parity wait can only tell a barrier's current phase from the one before, and the
    // workers do not take part in the a-stages — with two S tiles they step over a later tile's a-stages, so one barrier
    // per slot could be two phases away from the use they wait for (behind: a false "free"; ahead: a wait that never
    // ends).  With four, that takes a drift of 24 ring positions; the roles are never more than KB + 6 <= 14 apart.
    uint64_t* empty = bars + 20;
    uint64_t* yempty = bars + 20 + F_STAGES * F_EMPTY_WAYS;      // [4]  YT: MMA -> workers: lo slot read (multicast)
    uint32_t* tmem_slot = (uint32_t*)(bars + 24 + F_STAGES * F_EMPTY_WAYS);

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int worker = (int)(blockIdx.x >> 1), nworkers = (int)(gridDim.x >> 1);
    const int per_batch = p.tiles_l * p.nsplit;                    // work items of one restart
    const int items = per_batch * p.nbatch;
    const int KB = p.kblocks;
    const int bnh2 = p.Kv >> 1;                                   // P columns this CTA stages per b-stage
    // WEAVE (p.weave, two S tiles): the ring carries the next tile's a-stages BETWEEN this tile's b-stages (chunk c is
    // preceded by k blocks [c KB / nch, (c + 1) KB / nch)) instead of all of them first.  With the BF16 correction passes an
    // a-stage's MMAs take 512 cycles for 48 KB of operands (94 B/clk, what one SM can draw from L2 at best) and a b-stage's
    // 1024 cycles for 32 KB; woven, the demand is even (128 KB per 2048 cycles) and the quotient chunks are due one per
    // 2048 cycles instead of four in a row at 1024.
    const bool weave = SBUF == 2 && p.weave != 0;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&p.mapWhi); prefetch_tmap(&p.mapWlo); prefetch_tmap(&p.mapHhi); prefetch_tmap(&p.mapHlo);
        prefetch_tmap(&p.mapGhi); prefetch_tmap(&p.mapGlo); prefetch_tmap(&p.mapX);
        if (CB) {
            prefetch_tmap(&p.mapWbf); prefetch_tmap(&p.mapWlb); prefetch_tmap(&p.mapHbf); prefetch_tmap(&p.mapHlb);
            prefetch_tmap(&p.mapGbf); prefetch_tmap(&p.mapGlb);
        }
        for (int s = 0; s < F_XBUFS; ++s) { mbar_init(&xfull[s], 1); mbar_init(&xempty[s], FC::NARR); }
        for (int s = 0; s < F_STAGES; ++s) mbar_init(&full[s], 1);
        for (int s = 0; s < F_STAGES * F_EMPTY_WAYS; ++s) mbar_init(&empty[s], 1);
        for (int s = 0; s < F_YRING; ++s) mbar_init(&yfull[s], 2 * FC::NARR);   // the worker warps of both CTAs
        for (int s = 0; s < 2; ++s) { mbar_init(&sfull[s], 1); mbar_init(&sempty[s], 2 * FC::NARR); }
        mbar_init(pfull, 1); mbar_init(pempty, 2 * FC::NARR);
        for (int s = 0; s < 4; ++s) mbar_init(&yempty[s], 1);
        fence_barrier_init();
    }
    if (warp == 1) { tmem_alloc_cg2(tmem_slot, TMEM_COLS); tmem_relinquish_cg2(); }
    tc_fence_before();
    __syncwarp(); cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_S = tmem_base, tmem_P = tmem_base + (uint32_t)F_BN;
    pdl_sync();            // everything above overlapped the predecessor's tail; global memory is touched only from here on
    // one opaque 32-bit shared-space base; every ring slot and barrier below is this plus a constant (see umma_ptx.cuh)
    uint32_t sb = smem_u32(smem);
    asm volatile("" : "+r"(sb));
    const uint32_t xbuf_a = sb + (uint32_t)(F_STAGES * F_STAGE_BYTES);
    const uint32_t bar_a = xbuf_a + (uint32_t)(F_XBUFS * F_XBUF_BYTES);
    const uint32_t full_a = bar_a, yfull_a = bar_a + 4 * 8, sfull_a = bar_a + 8 * 8, sempty_a = bar_a + 10 * 8,
                   pfull_a = bar_a + 12 * 8, pempty_a = bar_a + 13 * 8, xfull_a = bar_a + 14 * 8, xempty_a = bar_a + 16 * 8,
                   empty_a = bar_a + 20 * 8, yempty_a = bar_a + (20 + F_STAGES * F_EMPTY_WAYS) * 8;
    // ring position = (stage, use): use counts the slot's uses.  Use u of a slot is released on empty[stage][u % 4];
    // whoever fills the slot for use u first waits for use u - 1 (u = 0: a fresh barrier's "preceding phase", passes)
    auto empty_prev = [&](uint32_t stage, uint32_t use, uint32_t& par) {
        const uint32_t u = use - 1u;
        par = (u >> 2) & 1u;
        return empty_a + (stage * (uint32_t)F_EMPTY_WAYS + (u & 3u)) * 8u;
    };
    auto empty_of = [&](uint32_t stage, uint32_t use) { return empty_a + (stage * (uint32_t)F_EMPTY_WAYS + (use & 3u)) * 8u; };

    // tile range (along C) of work item t = (lane-row block, split sp)
    auto tile_range = [&](int t, int& a, int& b) {
        const int sp = (t / p.tiles_l) % p.nsplit;
        a = (int)((long long)sp * p.ntiles_c / p.nsplit); b = (int)((long long)(sp + 1) * p.ntiles_c / p.nsplit);
    };
    auto chunks_of = [&](int tj) { const int left = p.Cdim - tj * BN; return left >= BN ? NCH : (left + 31) >> 5; };
    // this CTA pair's tiles in order, across its work items (items whose tile range is empty are stepped over; with
    // SBUF == 2 the host keeps nsplit <= ntiles_c, so every item owns a tile).  The producer and the MMA warp run one
    // tile ahead with the a-stages when SBUF == 2.
    struct Cur { int t, tj, nt1; };
    auto cur_seek = [&](Cur& c) { while (c.t < items) { tile_range(c.t, c.tj, c.nt1); if (c.tj < c.nt1) break; c.t += nworkers; } };
    auto cur_init = [&](Cur& c) { c.t = worker; c.tj = 0; c.nt1 = 0; cur_seek(c); };
    auto cur_valid = [&](const Cur& c) { return c.t < items; };
    auto cur_next = [&](Cur& c) { if (++c.tj >= c.nt1) { c.t += nworkers; cur_seek(c); } };

    if (warp < F_WORK_WARP0) {
      setmaxnreg_dec<FC::REGS_CTRL>();
      if (warp == 0 || (warp == 3 && p.prod2)) {
        // ===================== TMA producer =====================
        // p.prod2: warps 0 and 3 share the loads of every stage (part 0: the TF32 hi tiles and the barrier's expect_tx;
        // part 1: the lo / BF16 tiles) — a stage's 7-11 TMA instructions are issued by one thread each, and that issue
        // time sits on the path slot free -> data landed -> MMAs issued that bounds the a-stages with a 3-slot ring.
        const bool part0 = warp == 0, part1 = (warp == 3) || !p.prod2;
        const uint64_t polW = l2_policy(p.hintW), polH = l2_policy(p.hintH);
        const uint32_t bytes_a = 2u * (2u * (uint32_t)A_TILE_BYTES + 2u * (uint32_t)((BN / 2) * BK * 4));   // both CTAs' W and H pairs
        const uint32_t bytes_b = 2u * 2u * (uint32_t)(bnh2 * BK * 4);                               // both CTAs' G chunk pairs
        const uint32_t full_r = mapa_u32(full_a, 0u);                                               // the leader's `full` barriers
        constexpr int NBOX_W = HSTEP ? (BN / 2) / 32 : BM / 32;     // 32-row MN-major boxes of W (A: lane rows; B: this CTA's half of the tile)
        uint32_t stage = 0, use = 0;
        auto load_a = [&](const Cur& c, int s0, int s1) {           // a-stages [s0, s1) of a tile: W and H pairs of k block s
            const int l0 = (c.t % p.tiles_l) * 2 * BM + (int)rank * BM;      // this CTA's lane rows
            const int c0 = c.tj * BN;
            const int koff = (c.t / per_batch) * p.Kv;                        // this restart's k range
            for (int s = s0; s < s1; ++s) {
                { uint32_t par; const uint32_t eb = empty_prev(stage, use, par); mbar_wait_a(eb, par, p.err, 11); }
                if (elect_one()) {
                    const uint32_t fullc = full_r + stage * 8;
                    const uint32_t sA = sb + stage * F_STAGE_BYTES;
                    const uint32_t sB = sA + 2 * A_TILE_BYTES;
                    if (rank == 0 && part0) mbar_arrive_expect_tx_a(full_a + stage * 8, bytes_a);
                    const int r0 = koff + s * BK;
                    const uint32_t sW = HSTEP ? sB : sA;
                    const uint32_t sH = HSTEP ? sA : sB;           // H: one K-major box (W step: BN/2 columns, H step: 128 lane rows)
                    const int w0 = HSTEP ? c0 + (int)rank * (BN / 2) : l0;
                    const int h0 = HSTEP ? l0 : c0 + (int)rank * (BN / 2);
                    if (CB) {
                        // hi tiles as TF32; the BF16 pair (bf, lb) in the two halves of the `lo` slot.  Same byte count.
                        if (part0) {
                        for (int j = 0; j < NBOX_W; ++j) tma_load_2d_cg2_a(sW + j * 4096, &p.mapWhi, fullc, w0 + 32 * j, r0, polW);
                        tma_load_2d_cg2_a(sH, &p.mapHhi, fullc, r0, h0, polH);
                        }
                        if (part1) {
                        for (int j = 0; j < NBOX_W / 2; ++j) {
                            tma_load_2d_cg2_a(sW + A_TILE_BYTES + j * 4096, &p.mapWbf, fullc, w0 + 64 * j, r0, polW);
                            tma_load_2d_cg2_a(sW + A_TILE_BYTES + 8192 + j * 4096, &p.mapWlb, fullc, w0 + 64 * j, r0, polW);
                        }
                        tma_load_2d_cg2_a(sH + F_B_TILE, &p.mapHbf, fullc, r0, h0, polH);
                        tma_load_2d_cg2_a(sH + F_B_TILE + 8192, &p.mapHlb, fullc, r0, h0, polH);
                        }
                    } else
                    for (int h = 0; h < 2; ++h) {
                        if (!(h ? part1 : part0)) continue;
                        const CUtensorMap* mW = h ? &p.mapWlo : &p.mapWhi;
                        for (int j = 0; j < NBOX_W; ++j)
                            tma_load_2d_cg2_a(sW + h * A_TILE_BYTES + j * 4096, mW, fullc, w0 + 32 * j, r0, polW);
                        tma_load_2d_cg2_a(sH + h * F_B_TILE, h ? &p.mapHlo : &p.mapHhi, fullc, r0, h0, polH);
                    }
                }
                __syncwarp();
                if (++stage == (uint32_t)F_STAGES) { stage = 0; ++use; }
            }
        };
        auto load_b = [&](const Cur& c, int s0, int s1) {           // b-stages [s0, s1) of a tile: G chunk (the Y chunk is written by the workers)
            const int c0 = c.tj * BN;
            const int koff = (c.t / per_batch) * p.Kv + (int)rank * bnh2;    // this restart's k range, this CTA's half of it
            for (int s = s0; s < s1; ++s) {
                { uint32_t par; const uint32_t eb = empty_prev(stage, use, par); mbar_wait_a(eb, par, p.err, 11); }
                if (elect_one()) {
                    const uint32_t fullc = full_r + stage * 8;
                    const uint32_t sB = sb + stage * F_STAGE_BYTES + 2 * A_TILE_BYTES;
                    if (rank == 0 && part0) mbar_arrive_expect_tx_a(full_a + stage * 8, bytes_b);
                    const int n0 = c0 + s * 32;
                    if (CB) {
                        if (HSTEP) {
                            if (part0) tma_load_2d_cg2_a(sB, &p.mapGhi, fullc, n0, koff, polW);
                            if (part1) {
                            tma_load_2d_cg2_a(sB + F_B_TILE, &p.mapGbf, fullc, n0, koff, polW);
                            tma_load_2d_cg2_a(sB + F_B_TILE + 8192, &p.mapGlb, fullc, n0, koff, polW);
                            }
                        } else {
                            if (part0)
                            for (int j = 0; j < bnh2 / 32; ++j) tma_load_2d_cg2_a(sB + j * 4096, &p.mapGhi, fullc, koff + 32 * j, n0, polH);
                            if (part1)
                            for (int j = 0; j < bnh2 / 64; ++j) {
                                tma_load_2d_cg2_a(sB + F_B_TILE + j * 4096, &p.mapGbf, fullc, koff + 64 * j, n0, polH);
                                tma_load_2d_cg2_a(sB + F_B_TILE + 8192 + j * 4096, &p.mapGlb, fullc, koff + 64 * j, n0, polH);
                            }
                        }
                    } else
                    for (int h = 0; h < 2; ++h) {
                        if (!(h ? part1 : part0)) continue;
                        const CUtensorMap* mG = h ? &p.mapGlo : &p.mapGhi;
                        if (HSTEP) tma_load_2d_cg2_a(sB + h * F_B_TILE, mG, fullc, n0, koff, polW);
                        else {
                            for (int j = 0; j < bnh2 / 32; ++j)
                                tma_load_2d_cg2_a(sB + h * F_B_TILE + j * 4096, mG, fullc, koff + 32 * j, n0, polH);
                        }
                    }
                }
                __syncwarp();
                if (++stage == (uint32_t)F_STAGES) { stage = 0; ++use; }
            }
        };
        Cur cur, nxt; cur_init(cur); nxt = cur;
        if (SBUF == 2 && cur_valid(nxt)) { load_a(nxt, 0, KB); cur_next(nxt); }
        while (cur_valid(cur)) {
            const int nch = chunks_of(cur.tj);
            if (weave) {                                            // next tile's a-stages spread between this tile's b-stages
                const bool more = cur_valid(nxt);
                for (int c = 0; c < nch; ++c) {
                    if (more) load_a(nxt, c * KB / nch, (c + 1) * KB / nch);
                    load_b(cur, c, c + 1);
                }
                if (more) cur_next(nxt);
            } else {
                if (SBUF == 2) { if (cur_valid(nxt)) { load_a(nxt, 0, KB); cur_next(nxt); } }
                else load_a(cur, 0, KB);
                load_b(cur, 0, nch);
            }
            cur_next(cur);
        }
      } else if (warp == 2) {
        // ===================== X loader: chunk by chunk into the 2-deep buffer, as soon as the workers release a slot =====================
        // (its own warp: behind the ring producer it waited for ring slots it does not need)
        const uint64_t polS = l2_policy(p.hintS);
        uint32_t xi = 0;
        for (int t = worker; t < items; t += nworkers) {
            const int l0 = (t % p.tiles_l) * 2 * BM + (int)rank * BM;
            int nt0, nt1; tile_range(t, nt0, nt1);
            for (int tj = nt0; tj < nt1; ++tj) {
                const int c0 = tj * BN, nch = chunks_of(tj);
                for (int c = 0; c < nch; ++c, ++xi) {
                    const uint32_t xs = xi & 1u;
                    mbar_wait_a(xempty_a + xs * 8, ((xi >> 1) & 1u) ^ 1u, p.err, 20);
                    if (elect_one()) {
                        mbar_arrive_expect_tx_a(xfull_a + xs * 8, (uint32_t)F_XBUF_BYTES);
                        if (HSTEP) tma_load_2d_hint_a(xbuf_a + xs * F_XBUF_BYTES, &p.mapX, xfull_a + xs * 8, c0 + c * 32, l0, polS);
                        else       tma_load_2d_hint_a(xbuf_a + xs * F_XBUF_BYTES, &p.mapX, xfull_a + xs * 8, l0, c0 + c * 32, polS);
                    }
                    __syncwarp();
                }
            }
        }
      } else if (warp == 1 && rank == 0) {
        // ===================== MMA issuer (leader CTA) =====================
        const uint32_t sbase = sb;
        // operand layouts: MN-major tiles (SWIZZLE_128B_BASE32B: LBO 4096, SBO 512, 1024 B per K step) and
        // K-major tiles (SWIZZLE_128B: SBO 1024, 32 B per K step)
        const uint64_t dA_mn = smem_desc(sbase, 4096u, 512u, 1u), dA_k = smem_desc(sbase, 16u, 1024u, 2u);
        const uint64_t dB_mn = smem_desc(sbase + 2 * A_TILE_BYTES, 4096u, 512u, 1u), dB_k = smem_desc(sbase + 2 * A_TILE_BYTES, 16u, 1024u, 2u);
        constexpr uint32_t STEP_MN = 1024u >> 4, STEP_K = (uint32_t)(UK * 4) >> 4;
        // S:  W step A = W MN-major, B = H K-major;   H step A = H K-major, B = W MN-major
        // P:  W step A = Y K-major,  B = H MN-major;  H step A = Y K-major, B = W K-major
        const uint32_t idesc_a = HSTEP ? idesc_tf32(2 * BM, BN, 0, 1) : idesc_tf32(2 * BM, BN, 1, 0);
        const uint32_t idesc_b = HSTEP ? idesc_tf32(2 * BM, p.Kv, 0, 0) : idesc_tf32(2 * BM, p.Kv, 0, 1);
        const uint64_t dAa = HSTEP ? dA_k : dA_mn, dBa = HSTEP ? dB_mn : dB_k;
        const uint64_t dAb = dA_k,                 dBb = HSTEP ? dB_k : dB_mn;
        const uint32_t sAa = HSTEP ? STEP_K : STEP_MN, sBa = HSTEP ? STEP_MN : STEP_K;
        const uint32_t sAb = STEP_K,                   sBb = HSTEP ? STEP_K : STEP_MN;
        const uint32_t aHiW_a = (uint32_t)(dAa >> 32), bHiW_a = (uint32_t)(dBa >> 32), aLo_a = (uint32_t)dAa, bLo_a = (uint32_t)dBa;
        const uint32_t aHiW_b = (uint32_t)(dAb >> 32), bHiW_b = (uint32_t)(dBb >> 32), aLo_b = (uint32_t)dAb, bLo_b = (uint32_t)dBb;
        // CB: the BF16 tiles (bf at the `lo` slot, lb 8 KB further): MN-major SWIZZLE_128B / K-major SWIZZLE_64B descriptors
        const uint64_t eA_mn = smem_desc(sbase + A_TILE_BYTES, 4096u, 1024u, 2u), eA_k = smem_desc(sbase + A_TILE_BYTES, 16u, 512u, 4u);
        const uint64_t eB_mn = smem_desc(sbase + 2 * A_TILE_BYTES + F_B_TILE, 4096u, 1024u, 2u), eB_k = smem_desc(sbase + 2 * A_TILE_BYTES + F_B_TILE, 16u, 512u, 4u);
        constexpr uint32_t STEP16_MN = 2048u >> 4, STEP16_K = 32u >> 4, OFF_LB = 8192u >> 4;
        const uint64_t eAa = HSTEP ? eA_k : eA_mn, eBa = HSTEP ? eB_mn : eB_k;
        const uint64_t eAb = eA_k,                 eBb = HSTEP ? eB_k : eB_mn;
        const uint32_t tAa = HSTEP ? STEP16_K : STEP16_MN, tBa = HSTEP ? STEP16_MN : STEP16_K;
        const uint32_t tAb = STEP16_K,                     tBb = HSTEP ? STEP16_K : STEP16_MN;
        const uint32_t idesc16_a = HSTEP ? idesc_bf16(2 * BM, BN, 0, 1) : idesc_bf16(2 * BM, BN, 1, 0);
        const uint32_t idesc16_b = HSTEP ? idesc_bf16(2 * BM, p.Kv, 0, 0) : idesc_bf16(2 * BM, p.Kv, 0, 1);
        // CB: the 8 MMA time units of one 32-deep block: lb*bf, bf*lb (BF16, K = 16), hi*hi (TF32, K = 8); small terms first
        auto block_cb = [&](uint32_t d, uint32_t aH, uint32_t bH, uint32_t aW, uint32_t bW, uint32_t sa, uint32_t sb, uint32_t idesc,
                            uint32_t a16, uint32_t b16, uint32_t a16W, uint32_t b16W, uint32_t ta, uint32_t tb, uint32_t idesc16, uint32_t acc) {
            #pragma unroll
            for (int ks = 0; ks < 2; ++ks) mma_bf16_cg2_w(d, a16 + OFF_LB + ks * ta, a16W, b16 + ks * tb, b16W, idesc16, ks == 0 ? acc : 1u);
            #pragma unroll
            for (int ks = 0; ks < 2; ++ks) mma_bf16_cg2_w(d, a16 + ks * ta, a16W, b16 + OFF_LB + ks * tb, b16W, idesc16, 1u);
            #pragma unroll
            for (int ks = 0; ks < BK / UK; ++ks) mma_tf32_cg2_w(d, aH + ks * sa, aW, bH + ks * sb, bW, idesc, 1u);
        };
        // the 12 MMAs of one 32-deep block: lo*hi, hi*lo, hi*hi (small terms first)
        auto block = [&](uint32_t d, uint32_t aH, uint32_t aL, uint32_t bH, uint32_t bL, uint32_t aW, uint32_t bW,
                         uint32_t sa, uint32_t sb, uint32_t idesc, uint32_t acc) {
            #pragma unroll
            for (int ks = 0; ks < BK / UK; ++ks) mma_tf32_cg2_w(d, aL + ks * sa, aW, bH + ks * sb, bW, idesc, ks == 0 ? acc : 1u);
            #pragma unroll
            for (int ks = 0; ks < BK / UK; ++ks) mma_tf32_cg2_w(d, aH + ks * sa, aW, bL + ks * sb, bW, idesc, 1u);
            #pragma unroll
            for (int ks = 0; ks < BK / UK; ++ks) mma_tf32_cg2_w(d, aH + ks * sa, aW, bH + ks * sb, bW, idesc, 1u);
        };
        uint32_t stage = 0, use = 0;
        uint32_t ga = 0, gb = 0, pctr = 0, yctr = 0;
        auto issue_a = [&](int kb0, int kb1) {                   // S tile ga (TMEM columns (ga % SBUF) * BN ..) += k blocks [kb0, kb1)
            if (kb0 >= kb1) return;
            const uint32_t b = ga % (uint32_t)SBUF;
            if (kb0 == 0) {
                mbar_wait_a(sempty_a + b * 8, ((ga / (uint32_t)SBUF) & 1u) ^ 1u, p.err, 12);   // the tile that last used this S buffer has been read
                tc_fence_after();
            }
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait_a(full_a + stage * 8, use & 1u, p.err, 13);
                tc_fence_after();
                const uint32_t so = (uint32_t)(stage * F_STAGE_BYTES) >> 4;
                const uint32_t aH = aLo_a + so, aL = aH + (A_TILE_BYTES >> 4), bH = bLo_a + so, bL = bH + (F_B_TILE >> 4);
                if (elect_one()) {
                    if (CB) block_cb(tmem_S + b * (uint32_t)BN, aH, bH, aHiW_a, bHiW_a, sAa, sBa, idesc_a, (uint32_t)eAa + so, (uint32_t)eBa + so,
                                     (uint32_t)(eAa >> 32), (uint32_t)(eBa >> 32), tAa, tBa, idesc16_a, kb == 0 ? 0u : 1u);
                    else
                    block(tmem_S + b * (uint32_t)BN, aH, aL, bH, bL, aHiW_a, bHiW_a, sAa, sBa, idesc_a, kb == 0 ? 0u : 1u);
                    mma_commit_cg2_a(empty_of(stage, use), (uint16_t)3);
                    if (kb == KB - 1) mma_commit_cg2_a(sfull_a + b * 8, (uint16_t)3);
                }
                __syncwarp();
                if (++stage == (uint32_t)F_STAGES) { stage = 0; ++use; }
            }
            if (kb1 == KB) ++ga;
        };
        int in_period = 0;                                       // chunks of the tile accumulated into P since its last hand-over
        auto issue_b = [&](int cb0, int cb1, int nch) {          // P += Y chunk * G chunk for chunks [cb0, cb1) of the tile
            if (cb0 == 0) in_period = 0;
            for (int c = cb0; c < cb1; ++c) {
                if (in_period == 0) { mbar_wait_a(pempty_a, (pctr & 1u) ^ 1u, p.err, 14); }     // P drained by the workers
                mbar_wait_a(full_a + stage * 8, use & 1u, p.err, 15);                               // G chunk landed
                mbar_wait_a(yfull_a + (yctr & (F_YRING - 1)) * 8, (yctr / F_YRING) & 1u, p.err, 16);  // Y chunk written (both CTAs)
                tc_fence_after();
                const uint32_t so = (uint32_t)(stage * F_STAGE_BYTES) >> 4;
                const uint32_t aH = aLo_b + so, aL = aH + (A_TILE_BYTES >> 4), bH = bLo_b + so, bL = bH + (F_B_TILE >> 4);
                const bool last = (++in_period == p.period) || (c == nch - 1);
                if (elect_one()) {
                    if (YT) {
                        // A = the quotient chunk in tensor memory: hi where the S chunk was, lo in ring slot yctr & 3 (8 columns per K = 8 step)
                        const uint32_t yH = tmem_S + (gb % (uint32_t)SBUF) * (uint32_t)BN + (uint32_t)(c * 32);
                        const uint32_t yL = tmem_base + F_YT_COL + (yctr & (F_YT_SLOTS - 1u)) * 32u;
                        const uint32_t acc0 = in_period == 1 ? 0u : 1u;
                        if (CB) {
                            // the ring slot holds the BF16 forms, packed two per cell: bf in columns 0..15, lb in columns 16..31
                            const uint32_t g16 = (uint32_t)eBb + so, g16W = (uint32_t)(eBb >> 32);
                            #pragma unroll
                            for (int ks = 0; ks < 2; ++ks) mma_bf16_cg2_ts(tmem_P, yL + 16u + ks * 8u, g16 + ks * tBb, g16W, idesc16_b, ks == 0 ? acc0 : 1u);
                            #pragma unroll
                            for (int ks = 0; ks < 2; ++ks) mma_bf16_cg2_ts(tmem_P, yL + ks * 8u, g16 + OFF_LB + ks * tBb, g16W, idesc16_b, 1u);
                            #pragma unroll
                            for (int ks = 0; ks < BK / UK; ++ks) mma_tf32_cg2_ts(tmem_P, yH + ks * UK, bH + ks * sBb, bHiW_b, idesc_b, 1u);
                        } else {
                        #pragma unroll
                        for (int ks = 0; ks < BK / UK; ++ks) mma_tf32_cg2_ts(tmem_P, yL + ks * UK, bH + ks * sBb, bHiW_b, idesc_b, ks == 0 ? acc0 : 1u);
                        #pragma unroll
                        for (int ks = 0; ks < BK / UK; ++ks) mma_tf32_cg2_ts(tmem_P, yH + ks * UK, bL + ks * sBb, bHiW_b, idesc_b, 1u);
                        #pragma unroll
                        for (int ks = 0; ks < BK / UK; ++ks) mma_tf32_cg2_ts(tmem_P, yH + ks * UK, bH + ks * sBb, bHiW_b, idesc_b, 1u);
                        }
                    } else if (CB) {
                        block_cb(tmem_P, aH, bH, aHiW_b, bHiW_b, sAb, sBb, idesc_b, (uint32_t)eAb + so, (uint32_t)eBb + so,
                                 (uint32_t)(eAb >> 32), (uint32_t)(eBb >> 32), tAb, tBb, idesc16_b, in_period == 1 ? 0u : 1u);
                    } else
                    block(tmem_P, aH, aL, bH, bL, aHiW_b, bHiW_b, sAb, sBb, idesc_b, in_period == 1 ? 0u : 1u);
                    mma_commit_cg2_a(empty_of(stage, use), (uint16_t)3);
                    if (YT) mma_commit_cg2_a(yempty_a + (yctr & (F_YT_SLOTS - 1u)) * 8, (uint16_t)3);
                    if (last) mma_commit_cg2_a(pfull_a, (uint16_t)3);
                }
                __syncwarp();
                if (last) { ++pctr; in_period = 0; }
                ++yctr;
                if (++stage == (uint32_t)F_STAGES) { stage = 0; ++use; }
            }
            if (cb1 == nch) ++gb;
        };
        Cur cur, nxt; cur_init(cur); nxt = cur;
        if (SBUF == 2 && cur_valid(nxt)) { issue_a(0, KB); cur_next(nxt); }
        while (cur_valid(cur)) {
            const int nch = chunks_of(cur.tj);
            if (weave) {
                const bool more = cur_valid(nxt);
                for (int c = 0; c < nch; ++c) {
                    if (more) issue_a(c * KB / nch, (c + 1) * KB / nch);
                    issue_b(c, c + 1, nch);
                }
                if (more) cur_next(nxt);
            } else {
                if (SBUF == 2) { if (cur_valid(nxt)) { issue_a(0, KB); cur_next(nxt); } }
                else issue_a(0, KB);
                issue_b(0, nch, nch);
            }
            cur_next(cur);
        }
      }
    } else {
        setmaxnreg_inc<FC::REGS_WORK>();
        // ===================== worker warps: quotient (S -> Y chunk in the ring) and accumulation (P -> registers) =====================
        const int q = warp & 3, h = (warp - F_WORK_WARP0) >> 2;
        const int nchP = p.Kv >> 5;
        const int rowoff = (int)rank * BM + q * 32 + lane;
        const uint32_t row = (uint32_t)(q * 32 + lane);                       // lane row inside this CTA's 128 x 32 A tile
        const uint32_t row_off = (row >> 3) * 1024u + (row & 7u) * 128u;      // K-major SWIZZLE_128B: 8-row groups of 1 KB,
        const uint32_t sw = row & 7u;                                         // 16-byte chunk index XOR (row mod 8)
        const uint32_t laneS0 = tmem_S + ((uint32_t)(q * 32) << 16);         // + the column half (hx * CW) per access
        const uint32_t laneP = tmem_P + ((uint32_t)(q * 32) << 16);
        const uint32_t laneY = tmem_base + F_YT_COL + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * CW);   // YT: this warp's columns of a lo slot
        // this thread's X elements of a chunk.  W step: [column][lane row], no swizzle: 16 scalars 512 B apart.
        // H step: [lane row][32 columns] with the 128-byte swizzle: four 16-byte pieces of its own 128-byte row.
        // (the column half hx * CW is added per access)
        const uint32_t xoff = HSTEP ? xbuf_a + row * 128u : xbuf_a + row * 4u;
        // ALT (p.alt, two worker warps per quadrant, quotient through shared memory): instead of half the columns of EVERY
        // chunk, a warp takes ALL 32 columns of every other chunk (chunk parity = h) — the per-chunk costs that do not depend
        // on the column count (the proxy fence that publishes the stores, the barrier waits and arrivals) are paid half as
        // often per warp.  A handling warp arrives twice where the two halves' warps arrived once each.
        const bool alt = (NW == 2 && !YT) && p.alt != 0;
        // the leader CTA's barriers these warps arrive on (shared::cluster addresses, computed once: mapa and the
        // special-register reads behind it sat on the worker -> MMA critical path of every chunk)
        const uint32_t yfull_r = mapa_u32(yfull_a, 0u), sempty_r = mapa_u32(sempty_a, 0u), pempty_r = mapa_u32(pempty_a, 0u);
        bool pending = false;                            // a completed period of P not yet drained
        float acc[FC::ACC][32];
        uint32_t pctr = 0;
        auto drain = [&]() {                             // P (TMEM, truncating accumulation) -> fp32 registers, round-to-nearest
            mbar_wait_a(pfull_a, pctr & 1u, p.err, 17);
            tc_fence_after();
            #pragma unroll
            for (int cc = 0; cc < FC::ACC; ++cc) {
                const int c = h + NW * cc;
                if (c < nchP) {
                    #pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        uint32_t r[16];
                        tmem_ld_32x16(laneP + (uint32_t)(c * 32 + hh * 16), r);
                        tmem_ld_wait();
                        #pragma unroll
                        for (int jj = 0; jj < 16; ++jj) acc[cc][hh * 16 + jj] += __uint_as_float(r[jj]);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (elect_one()) mbar_arrive_cluster(pempty_r);
            ++pctr;
        };
        // ring position of the next b-stage (a- and b-stages of every tile, in issue order)
        const uint32_t kb_mod = (uint32_t)KB % (uint32_t)F_STAGES, kb_div = (uint32_t)KB / (uint32_t)F_STAGES;
        uint32_t stage = 0, use = 0;
        uint32_t gw = 0, yctr = 0, xi = 0;
        // step over the KB a-stages of one tile (SBUF == 1: this tile's own; SBUF == 2: a later tile's, which may not even
        // be loaded yet — see the four-way `empty` barriers above)
        auto skip_a = [&]() {
            stage += kb_mod; use += kb_div;
            if (stage >= (uint32_t)F_STAGES) { stage -= (uint32_t)F_STAGES; ++use; }
        };
        auto skip_n = [&](uint32_t n) {                  // weave: the a-stages that precede one b-stage
            stage += n % (uint32_t)F_STAGES; use += n / (uint32_t)F_STAGES;
            if (stage >= (uint32_t)F_STAGES) { stage -= (uint32_t)F_STAGES; ++use; }
        };
        if (SBUF == 2 && worker < items) skip_a();       // the first tile's a-stages lead the ring
        for (int t = worker; t < items; t += nworkers) {
            const int l = (t % p.tiles_l) * 2 * BM + rowoff;             // this thread's lane row (W step: m, H step: n)
            const bool row_ok = l < p.Ldim;
            const bool row_log = l < p.Llog;             // COST: the lane row lies inside the logical matrix
            const int koff = (t / per_batch) * p.Kv;     // this restart's k range
            const int sp = (t / p.tiles_l) % p.nsplit;
            double cost_item = 0.0;
            int nt0, nt1; tile_range(t, nt0, nt1);
            if (HSTEP && p.nsplit == 1 && !p.defer && row_ok) {
                // the item's epilogue reads this thread's 128 bytes of H per accumulator chunk; H was last touched an iteration
                // ago (DRAM by now): pull the lines into L2 while the item's tiles are computed
                #pragma unroll
                for (int cc = 0; cc < FC::ACC; ++cc)
                    if (h + NW * cc < nchP) prefetch_l2(p.H + (size_t)l * (uint32_t)p.ldH + koff + (h + NW * cc) * 32);
            }
            #pragma unroll
            for (int cc = 0; cc < FC::ACC; ++cc)
                #pragma unroll
                for (int jj = 0; jj < 32; ++jj) acc[cc][jj] = 0.0f;
            for (int tj = nt0; tj < nt1; ++tj) {
                const int nch = chunks_of(tj);
                // ring order: SBUF == 1: this tile's a-stages, then its b-stages; SBUF == 2: the NEXT tile's a-stages
                // (if there is a next tile) come before this tile's b-stages
                const bool more = SBUF == 1 || (tj + 1 < nt1) || (t + nworkers < items);
                if (more && !weave) skip_a();
                int in_period = 0;
                float cost_tile = 0.0f;                  // <= 8 x CW terms in fp32, then into the item's double
                const uint32_t sbuf = gw % (uint32_t)SBUF;
                const uint32_t laneS = laneS0 + sbuf * (uint32_t)BN;
                // the chunk after whose S read this warp releases the S tile: its last one (ALT: of its parity; none: chunk 0)
                int rel_c = nch - 1;
                if (alt && ((nch - 1) & 1) != h) rel_c = nch >= 2 ? nch - 2 : 0;
                const int hh1 = alt ? 2 : h + 1;            // one past the last column half this warp reads of a chunk
                mbar_wait_a(sfull_a + sbuf * 8, (gw / (uint32_t)SBUF) & 1u, p.err, 18);
                tc_fence_after();
                #pragma unroll 1
                for (int c = 0; c < nch; ++c, ++yctr, ++xi) {
                    const uint32_t xs = xi & 1u;
                    const bool mine = !alt || ((c & 1) == h);
                    if (weave && more) skip_n((uint32_t)((c + 1) * KB / nch - c * KB / nch));
                    if (mine) {
                    mbar_wait_a(xfull_a + xs * 8, (xi >> 1) & 1u, p.err, 21);
                    // phase 1: the quotient of this warp's columns of the chunk into registers — its own half, or (ALT) both
                    // halves one after the other: all of it happens BEFORE the wait for the ring slot, which for a tile's first
                    // chunks lasts through the next tile's a-stages
                    constexpr int NH = (NW == 2 && !YT) ? 2 : 1;
                    float z[NH][CW];
                    #pragma unroll
                    for (int hi_ = 0; hi_ < NH; ++hi_) {
                    if (hi_ < (alt ? 2 : 1)) {
                    const int hx = alt ? hi_ : h;
                    uint32_t r[CW];
                    tmem_ld_cols(laneS + (uint32_t)(c * 32 + hx * CW), r);
                    float x[CW];
                    if (HSTEP) {
                        #pragma unroll
                        for (int v = 0; v < CW / 4; ++v) {
                            const float4 xv = lds128(xoff + xs * (uint32_t)F_XBUF_BYTES + ((((uint32_t)((CW / 4) * hx + v)) ^ sw) << 4));
                            x[4 * v] = xv.x; x[4 * v + 1] = xv.y; x[4 * v + 2] = xv.z; x[4 * v + 3] = xv.w;
                        }
                    } else {
                        const uint32_t xa = xoff + xs * (uint32_t)F_XBUF_BYTES + (uint32_t)(hx * CW) * (uint32_t)(BM * 4);
                        #pragma unroll
                        for (int jj = 0; jj < CW; ++jj)
                            asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x[jj]) : "r"(xa + (uint32_t)(jj * BM * 4)));
                    }
                    tmem_ld_wait();
                    if (c == rel_c && hx == hh1 - 1) {   // S is free for the next tile's a-stages
                        tc_fence_before();
                        __syncwarp();
                        if (elect_one()) mbar_arrive_cluster(sempty_r + sbuf * 8);
                    }
                    // quotient (set_epsilon + vec_div, nmf.cu:128-131 / 158-161), branch-free: reciprocal, quotient, one
                    // residual correction (correctly rounded for normal quotients; x = 0 — zero-filled rows and columns
                    // beyond the matrix — gives exactly 0 and a quotient that underflows goes to 0 as x / y does); one check
                    // per thread and chunk sends huge or non-finite divisors and quotients (NaN fails the comparison) to the exact divide
                    float qmax = 0.0f;
                    #pragma unroll
                    for (int jj = 0; jj < CW; ++jj) {
                        const float y = clampE(__uint_as_float(r[jj]));
                        float rc;
                        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(y));
                        const float q0 = x[jj] * rc;
                        qmax = fmaxf(qmax, fmaxf(q0, y));
                        z[hi_][jj] = fmaf(fmaf(-y, q0, x[jj]), rc, q0);
                    }
                    if (!(qmax <= 1.0e30f)) {
                        #pragma unroll
                        for (int jj = 0; jj < CW; ++jj) z[hi_][jj] = x[jj] / clampE(__uint_as_float(r[jj]));
                    }
                    if (COST) {                          // x (log x - log y) - x + y, the term of reduce1d_div (matrix.cu:593)
                        const int cb = tj * BN + c * 32 + hx * CW;
                        #pragma unroll
                        for (int jj = 0; jj < CW; ++jj) {
                            if (row_log && cb + jj < p.Clog) {
                                const float y = clampE(__uint_as_float(r[jj]));
                                cost_tile += x[jj] * (logf(x[jj]) - logf(y)) - x[jj] + y;
                            }
                        }
                    }
                    }
                    }
                    // phase 2: hand the chunk to the second GEMM
                    if (YT) {
                        uint32_t zh[CW], zl[CW];
                        #pragma unroll
                        for (int jj = 0; jj < CW; ++jj) {
                            float hv, lv;
                            tf32_split_fast(z[0][jj], hv, lv);
                            zh[jj] = __float_as_uint(hv); zl[jj] = __float_as_uint(lv);
                        }
                        // hi: in place, over the S values this thread has just read (nobody else touches these lanes and columns)
                        tmem_st_cols(laneS + (uint32_t)(c * 32 + h * CW), zh);
                        // lo: ring slot yctr & 3, free once the b-stage that read it four chunks ago has retired (first round: passes)
                        const uint32_t ys = yctr & (F_YT_SLOTS - 1u);
                        mbar_wait_a(yempty_a + ys * 8, ((yctr >> 2) & 1u) ^ 1u, p.err, 19);
                        tc_fence_after();
                        if (CB) {        // BF16 forms of this thread's 8 elements, packed: bf -> cells 4h.., lb -> cells 16 + 4h.. of the slot
                            const uint32_t cell = laneY - (uint32_t)(h * CW) + ys * 32u + (uint32_t)(h * (CW / 2));
                            tmem_st_x4(cell, pack_bf16x2(z[0][0], z[0][1]), pack_bf16x2(z[0][2], z[0][3]), pack_bf16x2(z[0][4], z[0][5]), pack_bf16x2(z[0][6], z[0][7]));
                            tmem_st_x4(cell + 16u, pack_bf16x2(__uint_as_float(zl[0]), __uint_as_float(zl[1])), pack_bf16x2(__uint_as_float(zl[2]), __uint_as_float(zl[3])),
                                       pack_bf16x2(__uint_as_float(zl[4]), __uint_as_float(zl[5])), pack_bf16x2(__uint_as_float(zl[6]), __uint_as_float(zl[7])));
                        } else
                        tmem_st_cols(laneY + ys * 32u, zl);
                        tmem_st_wait();
                        tc_fence_before();
                    } else {
                    // EARLY DRAIN (p.drain_early): the ring slot of a tile's first b-stage follows the NEXT tile's a-stages, so the
                    // workers sit here through most of an a-phase with this chunk's quotient in registers.  A period of P that
                    // ended with the previous tile completed before those a-stages were even issued: draining it now, instead
                    // of after this chunk's hand-off, takes it off the path slot free -> Y written -> b-stage issued (the MMA
                    // warp waits for `pempty` before the tile's first b-stage).
                    if (p.drain_early && pending) { drain(); pending = false; }
                    // the ring slot of this b-stage must have been released by the MMAs that last read it
                    { uint32_t par; const uint32_t eb = empty_prev(stage, use, par); mbar_wait_a(eb, par, p.err, 19); }
                    const uint32_t hi_base = sb + stage * (uint32_t)F_STAGE_BYTES + row_off;
                    #pragma unroll
                    for (int hi_ = 0; hi_ < NH; ++hi_) {
                    if (hi_ < (alt ? 2 : 1)) {
                    const int hx = alt ? hi_ : h;
                    if (CB) {
                        // hi as a TF32 tile (K-major, 128-byte swizzle); bf = bf16(z) and lb = bf16(z - hi) as BF16 tiles where the lo
                        // tile was (K-major: 64-byte rows, 8-row groups of 512 B, 16-byte chunk index XOR (row / 2) mod 4)
                        const uint32_t b16 = sb + stage * (uint32_t)F_STAGE_BYTES + (uint32_t)A_TILE_BYTES + (row >> 3) * 512u + (row & 7u) * 64u;
                        const uint32_t sw64 = (row >> 1) & 3u;
                        #pragma unroll
                        for (int v = 0; v < CW / 8; ++v) {
                            float hv[8], lv[8];
                            #pragma unroll
                            for (int e = 0; e < 8; ++e) tf32_split_fast(z[hi_][8 * v + e], hv[e], lv[e]);
                            const uint32_t a0 = hi_base + ((((uint32_t)((CW / 4) * hx + 2 * v)) ^ sw) << 4);
                            const uint32_t a1 = hi_base + ((((uint32_t)((CW / 4) * hx + 2 * v + 1)) ^ sw) << 4);
                            sts128(a0, make_float4(hv[0], hv[1], hv[2], hv[3]));
                            sts128(a1, make_float4(hv[4], hv[5], hv[6], hv[7]));
                            const uint32_t c16 = b16 + ((((uint32_t)((CW / 8) * hx + v)) ^ sw64) << 4);
                            sts128u(c16, pack_bf16x2(z[hi_][8 * v], z[hi_][8 * v + 1]), pack_bf16x2(z[hi_][8 * v + 2], z[hi_][8 * v + 3]),
                                    pack_bf16x2(z[hi_][8 * v + 4], z[hi_][8 * v + 5]), pack_bf16x2(z[hi_][8 * v + 6], z[hi_][8 * v + 7]));
                            sts128u(c16 + 8192u, pack_bf16x2(lv[0], lv[1]), pack_bf16x2(lv[2], lv[3]), pack_bf16x2(lv[4], lv[5]), pack_bf16x2(lv[6], lv[7]));
                        }
                    } else {
                    #pragma unroll
                    for (int v = 0; v < CW / 4; ++v) {
                        float4 hv, lv;
                        tf32_split_fast(z[hi_][4 * v + 0], hv.x, lv.x); tf32_split_fast(z[hi_][4 * v + 1], hv.y, lv.y);
                        tf32_split_fast(z[hi_][4 * v + 2], hv.z, lv.z); tf32_split_fast(z[hi_][4 * v + 3], hv.w, lv.w);
                        const uint32_t a = hi_base + ((((uint32_t)((CW / 4) * hx + v)) ^ sw) << 4);
                        sts128(a, hv);
                        sts128(a + (uint32_t)A_TILE_BYTES, lv);
                    }
                    }
                    }
                    }
                    }
                    if (!YT) fence_proxy_async_smem();
                    __syncwarp();
                    if (elect_one()) {
                        const uint32_t yb = yfull_r + ((yctr & (uint32_t)(F_YRING - 1)) << 3);
                        mbar_arrive_cluster(yb);
                        mbar_arrive_a(xempty_a + xs * 8);   // the X values were consumed by the math above
                        if (alt) { mbar_arrive_cluster(yb); mbar_arrive_a(xempty_a + xs * 8); }   // for the warp that sat this chunk out
                    }
                    } else if (c == rel_c) {             // ALT, a tile with no chunk of this warp's parity: nothing of S to read
                        tc_fence_before();
                        __syncwarp();
                        if (elect_one()) mbar_arrive_cluster(sempty_r + sbuf * 8);
                    }
                    if (++stage == (uint32_t)F_STAGES) { stage = 0; ++use; }
                    // a period that ended one chunk ago (ALT: one or two): its MMAs ran while this chunk's quotient was
                    // computed.  A warp drains after a chunk it worked on; a second period never ends on an undrained one
                    if (pending && mine) { drain(); pending = false; }
                    if (++in_period == p.period) { in_period = 0; if (pending) drain(); pending = true; }
                }
                if (SBUF == 2) {
                    // the tile's last period is drained after the first quotient chunk of the next tile (or below, at the
                    // end of the work item): its MMAs are queued behind the next tile's a-stages
                    if (in_period != 0) { if (pending) drain(); pending = true; }
                } else {                                 // overlaps the next tile's a-stages
                    if (pending) { drain(); pending = false; }
                    if (in_period != 0) drain();
                }
                ++gw;
                if (COST) cost_item += (double)cost_tile;
            }
            if (pending) { drain(); pending = false; }
            if (COST) {                                  // every worker warp of every work item writes its slot (no zero fill needed)
                #pragma unroll
                for (int o = 16; o > 0; o >>= 1) cost_item += __shfl_xor_sync(0xffffffffu, cost_item, o);
                if (lane == 0) p.cost_partials[((size_t)t * 2 + rank) * FC::NARR + (uint32_t)(warp - F_WORK_WARP0)] = cost_item;
            }
            if (row_ok) {
                if (HSTEP && (p.nsplit > 1 || p.defer)) {
                    // the rows of W were split over work items (or the update is left to k_update_h): partial sums
                    float* out = p.P + (size_t)sp * p.split_stride + (size_t)l * (uint32_t)p.ldH + koff;
                    #pragma unroll
                    for (int cc = 0; cc < FC::ACC; ++cc) {
                        const int c = h + NW * cc;
                        if (c < nchP) {
                            #pragma unroll
                            for (int v = 0; v < 4; ++v) {
                                float o8[8];
                                #pragma unroll
                                for (int e = 0; e < 8; ++e) o8[e] = acc[cc][8 * v + e];
                                stg256(out + c * 32 + 8 * v, o8);
                            }
                        }
                    }
                } else if (HSTEP) {
                    // H[:, n] <- H[:, n] .* (acc ./ max(sumW, EPS))  (col_divide + element_multiply, nmf.cu:142-145) + its TF32 pair
                    const size_t base = (size_t)l * (uint32_t)p.ldH + koff;
                    const float* sumW = p.sumW + koff;
                    #pragma unroll
                    for (int cc = 0; cc < FC::ACC; ++cc) {
                        const int c = h + NW * cc;
                        if (c < nchP) {
                            // 16 columns at a time (the 4-worker-warp variant has 104 registers): their loads first (one HBM
                            // latency per half chunk, not per 32 bytes).  col_div is an IEEE divide; as in the quotient:
                            // reciprocal, quotient, one residual correction (correctly rounded for normal quotients, a = 0
                            // gives exactly 0), one range check per 16
                            #pragma unroll
                            for (int hf = 0; hf < 2; ++hf) {
                                float hv[2][8];
                                #pragma unroll
                                for (int v = 0; v < 2; ++v) ldg256(p.H + base + (size_t)(c * 32 + hf * 16 + 8 * v), hv[v]);
                                float qv[16];
                                float qmax = 0.0f;
                                #pragma unroll
                                for (int jj = 0; jj < 16; ++jj) {
                                    const float a = acc[cc][hf * 16 + jj];
                                    const float y = clampE(__ldg(sumW + c * 32 + hf * 16 + jj));    // 1 KB, L1-resident
                                    float rc;
                                    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(y));
                                    const float q0 = a * rc;
                                    qmax = fmaxf(qmax, fmaxf(q0, y));
                                    qv[jj] = fmaf(fmaf(-y, q0, a), rc, q0);
                                }
                                if (!(qmax <= 1.0e30f)) {
                                    #pragma unroll
                                    for (int jj = 0; jj < 16; ++jj) qv[jj] = acc[cc][hf * 16 + jj] / clampE(__ldg(sumW + c * 32 + hf * 16 + jj));
                                }
                                #pragma unroll
                                for (int v = 0; v < 2; ++v) {
                                    const size_t o = base + (size_t)(c * 32 + hf * 16 + 8 * v);
                                    float hn[8], hh[8], hl[8];
                                    #pragma unroll
                                    for (int e = 0; e < 8; ++e) {
                                        hn[e] = hv[v][e] * qv[8 * v + e];
                                        tf32_split(hn[e], hh[e], hl[e]);
                                    }
                                    stg256(p.H + o, hn); stg256(p.Hhi + o, hh); stg256(p.Hlo + o, hl);
                                    if (CB) {
                                        stg128u(p.Hbf + o, pack_bf16x2(hn[0], hn[1]), pack_bf16x2(hn[2], hn[3]), pack_bf16x2(hn[4], hn[5]), pack_bf16x2(hn[6], hn[7]));
                                        stg128u(p.Hlb + o, pack_bf16x2(hn[0] - hh[0], hn[1] - hh[1]), pack_bf16x2(hn[2] - hh[2], hn[3] - hh[3]),
                                                pack_bf16x2(hn[4] - hh[4], hn[5] - hh[5]), pack_bf16x2(hn[6] - hh[6], hn[7] - hh[7]));
                                    }
                                }
                            }
                        }
                    }
                } else {
                    float* out = p.P + (size_t)sp * p.split_stride + (size_t)koff * (uint32_t)p.ldP + l;
                    #pragma unroll
                    for (int cc = 0; cc < FC::ACC; ++cc) {
                        const int c = h + NW * cc;
                        if (c < nchP) {
                            #pragma unroll
                            for (int jj = 0; jj < 32; ++jj) out[(size_t)(c * 32 + jj) * (uint32_t)p.ldP] = acc[cc][jj];
                        }
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncwarp();
    cluster_sync_all();
    if (warp == 1) { tc_fence_after(); tmem_dealloc_cg2(tmem_base, TMEM_COLS); }
}

// ---------------------------------------------------------------------------------------------
// host side: one launch on CTA pairs, at most one pair per TPC, persistent over the work items
// ---------------------------------------------------------------------------------------------
namespace {

int num_sms_f() {
    int dev = 0, n = 0;
    CUDA_CHECK(cudaGetDevice(&dev));
    CUDA_CHECK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    return n;
}

template <bool HSTEP, int NW, bool COST, int BN, bool YT, bool CB = false>
void launch_fused_nw(const FusedParams& p, cudaStream_t st) {
    // function attributes are per device: one host thread per GPU may launch the same instantiation
    static std::atomic<unsigned long long> configured{0};            // bit d: done on device d
    int dev = 0; CUDA_CHECK(cudaGetDevice(&dev));
    if (!(configured.load(std::memory_order_acquire) >> (dev & 63) & 1ull)) {
        CUDA_CHECK(cudaFuncSetAttribute(k_fused<HSTEP, NW, COST, BN, YT, CB>, cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM_BYTES));
        configured.fetch_or(1ull << (dev & 63), std::memory_order_release);
    }
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(FCfg<NW>::NTHREADS, 1, 1);
    cfg.dynamicSmemBytes = F_SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (launch_pdl()) {
        attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[1].val.programmaticStreamSerializationAllowed = 1;
        cfg.numAttrs = 2;
    }
    static std::atomic<int> max_pairs{0};            // co-resident CTA pairs (one per TPC): 74 on B200 (same on every GPU of the box)
    if (!max_pairs.load()) {
        const int sms = num_sms_f();
        cfg.gridDim = dim3(2u * (unsigned)(sms / 2), 1, 1);
        int n = 0;
        CUDA_CHECK(cudaOccupancyMaxActiveClusters(&n, k_fused<HSTEP, NW, COST, BN, YT, CB>, &cfg));
        NMF_CHECK(n > 0, "no CTA pair of the fused half-iteration kernel fits on this device");
        max_pairs.store(std::min(n, sms / 2));
    }
    const long long items = (long long)p.tiles_l * p.nsplit * p.nbatch;
    cfg.gridDim = dim3(2u * (unsigned)std::min<long long>(items, max_pairs.load()), 1, 1);
    CUDA_CHECK(cudaLaunchKernelEx(&cfg, k_fused<HSTEP, NW, COST, BN, YT, CB>, p));
    CUDA_CHECK_LAUNCH();
}

template <bool HSTEP, bool COST>
void launch_fused_t(const FusedParams& p, cudaStream_t st) {
    // the quotient in tensor memory (ytmem) is instantiated for the K <= 128 default (4 worker warps per quadrant) only
    if (p.nw == 4) {
        if (p.ytmem && p.corr) {
            if (p.bn == F_BN_DB) launch_fused_nw<HSTEP, 4, COST, F_BN_DB, true, true>(p, st); else launch_fused_nw<HSTEP, 4, COST, F_BN, true, true>(p, st);
        } else if (p.ytmem) {
            if (p.bn == F_BN_DB) launch_fused_nw<HSTEP, 4, COST, F_BN_DB, true>(p, st); else launch_fused_nw<HSTEP, 4, COST, F_BN, true>(p, st);
        } else {
            if (p.bn == F_BN_DB) launch_fused_nw<HSTEP, 4, COST, F_BN_DB, false>(p, st); else launch_fused_nw<HSTEP, 4, COST, F_BN, false>(p, st);
        }
    } else if (p.corr) {
        if (p.bn == F_BN_DB) launch_fused_nw<HSTEP, 2, COST, F_BN_DB, false, true>(p, st); else launch_fused_nw<HSTEP, 2, COST, F_BN, false, true>(p, st);
    } else {
        if (p.bn == F_BN_DB) launch_fused_nw<HSTEP, 2, COST, F_BN_DB, false>(p, st); else launch_fused_nw<HSTEP, 2, COST, F_BN, false>(p, st);
    }
}

}  // namespace

void launch_fused(bool hstep, bool cost, const FusedParams& p, cudaStream_t st) {
    NMF_CHECK(p.nw == 2 || (p.nw == 4 && p.Kv <= 128), "fused kernel: 4 worker warps per quadrant need a P tile of <= 128 columns");
    NMF_CHECK(!cost || hstep, "fused kernel: only the H step carries the cost");
    NMF_CHECK(!p.corr || (p.nw == 2 && p.Kv == 256) || (p.nw == 4 && p.ytmem && p.Kv == 128),
              "fused kernel: BF16 correction passes need K = 256, or K = 128 with the quotient in tensor memory");
    NMF_CHECK(!p.ytmem || (p.nw == 4 && p.Kv <= 128), "fused kernel: the quotient fits in tensor memory only beside a P tile of <= 128 columns");
    if (hstep) { if (cost) launch_fused_t<true, true>(p, st); else launch_fused_t<true, false>(p, st); }
    else launch_fused_t<false, false>(p, st);
}

}}  // namespace nmf::umma
