// umma_common.cuh — what the two translation units of the tcgen05 engine share (kernels_umma.cu: the three GEMM
// kernels with fused epilogues + the launch plan; kernels_fused.cu: the fused half-iteration kernels).
#pragma once
#include <cuda.h>

#include "session.hpp"
#include "tf32.cuh"
#include "umma_ptx.cuh"

namespace nmf { namespace umma {

using namespace ptx;

constexpr int BM = 128;              // UMMA rows per CTA (TMEM lanes)
constexpr int BN_MAX = 256;          // UMMA N upper bound
constexpr int BK = 32;               // reduction elements per stage (= one 128-byte swizzle row)
constexpr int UK = 8;                // reduction elements per tcgen05.mma.kind::tf32
constexpr int A_TILE_BYTES = BM * BK * 4;          // 16 KB
constexpr int TMEM_COLS = 512;

__device__ __forceinline__ float clampE(float v) { return v < NMF_EPS_F ? NMF_EPS_F : v; }

}}  // namespace nmf::umma
