// Shared by the forward and backward blend kernels (both compiled with -fmad=false so that the per-pair
// tests -- power > 0, power < pmin, alpha < 1/255 -- take bit-identical decisions in the two passes).
#pragma once
#include "common.cuh"

namespace amav_blend {

// exp(x), x <= 0: same operation sequence as amav_oracle_exp().
__device__ __forceinline__ float amav_exp(float x)
{
    x = fmaxf(x, -80.0f);
    const float L2E_HI = 1.44269502162933349609375f;
    const float L2E_LO = 1.925963033500011e-8f;
    const float MAGIC = 12582912.0f;
    const float t = x * L2E_HI;
    const float r = t + MAGIC;
    const float n = r - MAGIC;
    float f = fmaf(x, L2E_HI, -n);
    f = fmaf(x, L2E_LO, f);
    float p = 1.5252733804059840e-5f;
    p = fmaf(p, f, 1.5403530393381610e-4f);
    p = fmaf(p, f, 1.3333558146428443e-3f);
    p = fmaf(p, f, 9.6181291076284772e-3f);
    p = fmaf(p, f, 5.5504108664821580e-2f);
    p = fmaf(p, f, 2.4022650695910071e-1f);
    p = fmaf(p, f, 6.9314718055994531e-1f);
    p = fmaf(p, f, 1.0f);
    // r = MAGIC + n exactly, so the low mantissa bits of r hold n (two's complement): p * 2^n
    const int ni = __float_as_int(r) - __float_as_int(MAGIC);
    return __int_as_float(__float_as_int(p) + (ni << 23));
}

constexpr int BLEND_THREADS = 256;
constexpr int SUB_W = 8, SUB_H = 4;        // pixels of one warp: 8 wide x 4 tall

// Upper bound of power(p) = Ah*dx^2 + Bn*dx*dy + Ch*dy^2 over the warp's pixel rectangle (dx = gx - px in
// [dxl, dxh], dy likewise).  power is a concave quadratic with its maximum (0) at the Gaussian centre, so over a
// box that does not contain the centre the maximum lies on the edge(s) facing it, where it is a clamped 1-D
// quadratic optimum.  A Gaussian whose bound is below pmin (minus a rounding margin) cannot reach alpha >= 1/255
// at any pixel of the rectangle: the exact per-pixel test would reject it 32 times, so dropping it from the
// warp's list changes nothing in the result.
__device__ __forceinline__ bool warp_rect_keep(const float4 a, const float4 b, float x0, float y0)
{
    const float Ah = a.z, Bn = a.w, Ch = b.x, pmin = b.z;
    const float dxl = a.x - (x0 + (float)(SUB_W - 1)), dxh = a.x - x0;
    const float dyl = a.y - (y0 + (float)(SUB_H - 1)), dyh = a.y - y0;
    if (!(Ah < 0.0f && Ch < 0.0f)) return true;                       // degenerate conic (or NaN): exact path decides
    const bool out_x = dxl > 0.0f || dxh < 0.0f, out_y = dyl > 0.0f || dyh < 0.0f;
    if (!out_x && !out_y) return true;                                // centre inside the rectangle
    const float ex = dxl > 0.0f ? dxl : dxh, ey = dyl > 0.0f ? dyl : dyh;
    float m = -3.0e38f;
    if (out_x) {
        const float dy = fminf(fmaxf(__fdividef(-0.5f * Bn * ex, Ch), dyl), dyh);
        m = fmaf(fmaf(Bn, dy, Ah * ex), ex, (Ch * dy) * dy);
    }
    if (out_y) {
        const float dx = fminf(fmaxf(__fdividef(-0.5f * Bn * ey, Ah), dxl), dxh);
        m = fmaxf(m, fmaf(fmaf(Bn, ey, Ah * dx), dx, (Ch * ey) * ey));
    }
    const float X = fmaxf(fabsf(dxl), fabsf(dxh)), Y = fmaxf(fabsf(dyl), fabsf(dyh));
    const float mag = -(Ah * X * X + Ch * Y * Y);
    return !(m + 0.01f + 8.0e-6f * mag < pmin);
}

}  // namespace amav_blend
