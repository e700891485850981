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

}  // namespace nmf
