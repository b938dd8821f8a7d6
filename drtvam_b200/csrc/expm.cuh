// 1 - 2^-x without the cancellation of  E(u_in) - E(u_out).
//
// The deposit of a voxel is  e^{-s t} (1 - e^{-s dt})  (sensor.py:400-404).  Formed as a difference of
// two ex2.approx values of the absolute optical depth it carries the rounding of both (an ulp of the
// argument and of the result, ~1e-7 each) divided by the per-voxel extinction s dt ~ 2.5e-3, i.e. ~1e-4
// relative noise per deposit - the whole parity budget.  Here the small factor is evaluated directly.
#pragma once

namespace drtvam {

// x >= 0 in units of log2: returns 1 - 2^-x, accurate to a few fp32 ulp
__device__ __forceinline__ float one_minus_ex2_neg(float x) {
    const float y = x * 0.6931471805599453f;
    float r = fmaf(-y, 1.f / 120.f, 1.f / 24.f);
    r = fmaf(-y, r, 1.f / 6.f);
    r = fmaf(-y, r, 0.5f);
    r = fmaf(-y, r, 1.f);
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-x));
    return y < 0.05f ? y * r : 1.f - e;
}

}  // namespace drtvam
