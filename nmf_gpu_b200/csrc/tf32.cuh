// tf32.cuh — the fp32 -> (hi, lo) split behind the 3xTF32 products.
//   hi = x rounded to nearest TF32 (10 explicit mantissa bits; cvt.rna), exactly representable;
//   lo = (x - hi) rounded to nearest TF32.  x - hi is exact in fp32 and |lo| <= 2^-11 |x|, signed,
//   so neither the dropped lo*lo term nor the rounding of lo carries a systematic sign
//   (a truncating split would make lo >= 0 for x > 0: a one-sided bias that a multiplicative
//   update integrates over iterations).  hi + lo == x to 2^-22 |x|.
#pragma once
#include <cuda_runtime.h>

namespace nmf {

__device__ __forceinline__ float tf32_rn(float v) {
    unsigned r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return __uint_as_float(r);
}

__device__ __forceinline__ void tf32_split(float v, float& hi, float& lo) {
    hi = tf32_rn(v);
    lo = tf32_rn(v - hi);
}

// The in-SM splitter's form (kernels_umma.cu): cvt.rna.tf32.f32 is a 4-instruction sequence on sm_100a
// (Inf/NaN guard, add half an ulp, select, mask) and the splitter warps sit on the TMA -> MMA critical
// path, so hi is rounded by hand (add 0x1000 to the magnitude, clear 13 bits: the same round-to-
// nearest, ties away; Inf stays Inf) and lo = x - hi is left UNROUNDED: it is exact in fp32, and the
// tensor core reads only the upper 19 bits of a TF32 operand, i.e. truncates lo towards zero — lo is
// sign-symmetric, so that truncation is unbiased.  3 instructions per element instead of 9.
__device__ __forceinline__ void tf32_split_fast(float v, float& hi, float& lo) {
    hi = __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xffffe000u);
    lo = v - hi;
}

// two floats -> packed bf16x2, round to nearest even: the LOW half holds `lo_elem` (the element at the lower address)
__device__ __forceinline__ unsigned pack_bf16x2(float lo_elem, float hi_elem) {
    unsigned r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi_elem), "f"(lo_elem));
    return r;
}

}  // namespace nmf
