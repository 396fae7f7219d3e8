// Shared by the forward and backward blend kernels (both compiled with -fmad=false so that the per-pair
// tests -- power > 0, power < pmin, alpha < 1/255 -- take bit-identical decisions in the two passes).
#pragma once
#include "common.cuh"

namespace amav_blend {

// exp(x) for -87 < x <= 0: same operation sequence as amav_oracle_exp() (which also clamps x to >= -80; the blend
// kernels only call this with power >= pmin, and preprocess stores pmin >= -80).  12 instructions: range reduction
// by one FMA against the 1.5*2^23 magic constant, degree-5 minimax polynomial of 2^f on [-0.5, 0.5] (1.5 ulp
// over the whole range; CUDA's expf, which upstream calls, is 2 ulp), exponent inserted by an integer add.
__device__ __forceinline__ float amav_exp(float x)
{
    const float L2E_HI = 1.44269502162933349609375f;
    const float L2E_LO = 1.925963033500011e-8f;
    const float MAGIC = 12582912.0f;
    const float r = fmaf(x, L2E_HI, MAGIC);
    const float n = r - MAGIC;
    float f = fmaf(x, L2E_HI, -n);
    f = fmaf(x, L2E_LO, f);
    float p = 0.001326472731307149f;
    p = fmaf(p, f, 0.009671512991189957f);
    p = fmaf(p, f, 0.05550733581185341f);
    p = fmaf(p, f, 0.24022242426872253f);
    p = fmaf(p, f, 0.6931470036506653f);
    p = fmaf(p, f, 1.0f);
    // r = MAGIC + n exactly, so the low mantissa bits of r hold n (two's complement): p * 2^n
    const int ni = __float_as_int(r) - __float_as_int(MAGIC);
    return __int_as_float(__float_as_int(p) + (ni << 23));
}

// amav_exp has two FFMAs with two constant operands each (x * log2e + MAGIC, c5 * f + c4); an FFMA takes one
// immediate, and ptxas rematerialises the other constant (immediate move, parameter load or uniform-register copy)
// inside the list walk: 2 of ~44 instructions per pair.  A value that comes out of a volatile shared-memory load at
// a per-thread address cannot be rematerialised and is not promoted to a uniform register, so it stays in a plain
// register.  Call once per kernel with a 4-float shared scratch; contains a barrier.  amav_exp(x, ec) is
// bit-identical to amav_exp(x).
struct ExpConsts { float magic, c5; };
__device__ __forceinline__ ExpConsts exp_consts(float *s_scratch)
{
    if (threadIdx.x == 0) {
        s_scratch[0] = s_scratch[1] = 12582912.0f;
        s_scratch[2] = s_scratch[3] = 0.001326472731307149f;
    }
    __syncthreads();
    ExpConsts ec;
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(s_scratch) + 4u * (threadIdx.x & 1u);
    asm volatile("ld.volatile.shared.f32 %0, [%1];" : "=f"(ec.magic) : "r"(a));
    asm volatile("ld.volatile.shared.f32 %0, [%1+8];" : "=f"(ec.c5) : "r"(a));
    return ec;
}
__device__ __forceinline__ float amav_exp(float x, const ExpConsts ec)
{
    const float L2E_HI = 1.44269502162933349609375f;
    const float L2E_LO = 1.925963033500011e-8f;
    const float MAGIC = 12582912.0f;
    const float r = fmaf(x, L2E_HI, ec.magic);
    const float n = r - MAGIC;
    float f = fmaf(x, L2E_HI, -n);
    f = fmaf(x, L2E_LO, f);
    float p = fmaf(ec.c5, f, 0.009671512991189957f);
    p = fmaf(p, f, 0.05550733581185341f);
    p = fmaf(p, f, 0.24022242426872253f);
    p = fmaf(p, f, 0.6931470036506653f);
    p = fmaf(p, f, 1.0f);
    const int ni = __float_as_int(r) - __float_as_int(MAGIC);
    return __int_as_float(__float_as_int(p) + (ni << 23));
}

// Two independent amav_exp evaluations with their operations alternating, so that in-order issue overlaps the two
// dependent chains (each result is bit-identical to amav_exp(x)).
__device__ __forceinline__ void amav_exp_x2(float x0, float x1, const ExpConsts ec, float &e0, float &e1)
{
    const float L2E_HI = 1.44269502162933349609375f;
    const float L2E_LO = 1.925963033500011e-8f;
    const float MAGIC = 12582912.0f;
    const float r0 = fmaf(x0, L2E_HI, ec.magic), r1 = fmaf(x1, L2E_HI, ec.magic);
    const float n0 = r0 - MAGIC, n1 = r1 - MAGIC;
    float f0 = fmaf(x0, L2E_HI, -n0), f1 = fmaf(x1, L2E_HI, -n1);
    f0 = fmaf(x0, L2E_LO, f0); f1 = fmaf(x1, L2E_LO, f1);
    float p0 = fmaf(ec.c5, f0, 0.009671512991189957f), p1 = fmaf(ec.c5, f1, 0.009671512991189957f);
    p0 = fmaf(p0, f0, 0.05550733581185341f); p1 = fmaf(p1, f1, 0.05550733581185341f);
    p0 = fmaf(p0, f0, 0.24022242426872253f); p1 = fmaf(p1, f1, 0.24022242426872253f);
    p0 = fmaf(p0, f0, 0.6931470036506653f); p1 = fmaf(p1, f1, 0.6931470036506653f);
    p0 = fmaf(p0, f0, 1.0f); p1 = fmaf(p1, f1, 1.0f);
    e0 = __int_as_float(__float_as_int(p0) + ((__float_as_int(r0) - __float_as_int(MAGIC)) << 23));
    e1 = __int_as_float(__float_as_int(p1) + ((__float_as_int(r1) - __float_as_int(MAGIC)) << 23));
}

// Two amav_exp evaluations in ONE instruction stream of sm_100 packed-FP32 operations (fma.rn.f32x2 / add.rn.f32x2:
// SASS FFMA2 / FADD2, each half IEEE round-to-nearest, so e.x == amav_exp(x.x) and e.y == amav_exp(x.y) bit for bit).
// The range reduction is written on the NEGATED sum: rn = fma(x, -log2e, -MAGIC) = -r exactly (round-to-nearest is
// symmetric), so -n = MAGIC - r = rn + MAGIC needs no negated operand (the packed instructions have no negate
// modifier in PTX), and the exponent insertion only uses the low mantissa bits of r, which -r shares:
// (bits(r) - bits(MAGIC)) << 23 == bits(rn) << 23 (mod 2^32).  9 packed FP instructions + 2 integer ones per half.
__device__ __forceinline__ float2 amav_exp_f32x2(const float2 x)
{
    const float L2E_HI = 1.44269502162933349609375f;
    const float L2E_LO = 1.925963033500011e-8f;
    const float MAGIC = 12582912.0f;
    const float2 rn = __ffma2_rn(x, make_float2(-L2E_HI, -L2E_HI), make_float2(-MAGIC, -MAGIC));
    const float2 mn = __fadd2_rn(rn, make_float2(MAGIC, MAGIC));
    float2 f = __ffma2_rn(x, make_float2(L2E_HI, L2E_HI), mn);
    f = __ffma2_rn(x, make_float2(L2E_LO, L2E_LO), f);
    float2 p = __ffma2_rn(make_float2(0.001326472731307149f, 0.001326472731307149f), f,
                          make_float2(0.009671512991189957f, 0.009671512991189957f));
    p = __ffma2_rn(p, f, make_float2(0.05550733581185341f, 0.05550733581185341f));
    p = __ffma2_rn(p, f, make_float2(0.24022242426872253f, 0.24022242426872253f));
    p = __ffma2_rn(p, f, make_float2(0.6931470036506653f, 0.6931470036506653f));
    p = __ffma2_rn(p, f, make_float2(1.0f, 1.0f));
    float2 e;
    e.x = __int_as_float(__float_as_int(p.x) + (__float_as_int(rn.x) << 23));
    e.y = __int_as_float(__float_as_int(p.y) + (__float_as_int(rn.y) << 23));
    return e;
}

// 16-byte shared-memory load at a 32-bit shared address + immediate offset.  The list walk addresses its records
// as base + 48 j with the base held in a register; through a generic pointer ptxas rebuilds the shared window
// base (S2R SR_CgaCtaId + LEA) for every entry.
template <int OFF> __device__ __forceinline__ float4 lds128(uint32_t addr)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+%5];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr), "n"(OFF));
    return v;
}
template <int OFF> __device__ __forceinline__ float2 lds64(uint32_t addr)
{
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2+%3];" : "=f"(v.x), "=f"(v.y) : "r"(addr), "n"(OFF));
    return v;
}

template <int OFF> __device__ __forceinline__ uint32_t lds_u16(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u16 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(OFF));
    return v;
}

// Tile queue of the blend kernels.  A tile is one CTA's sequential walk and list lengths differ by orders of
// magnitude, so a deep tile handed out late leaves the whole GPU waiting for one CTA (a 4-frame training launch
// lasts exactly as long as its deepest tile).  amav_splat_scan lists the deep tiles (`order`: [0] = number with
// >= AMAV_DEEP0 entries, stored from the front of the list; [1] = number with >= AMAV_DEEP1, stored from its back;
// list of num_tiles words from [2]); the queue hands those out first, then every other tile in frame order.
// Returns false when the queue is empty, else the tile in `t` or t = 0xffffffff for a position to skip (a deep
// tile met again in the plain sweep).  Uniform across the CTA; `q` is the popped queue position.
__device__ __forceinline__ bool queue_tile(unsigned int q, const uint32_t *__restrict__ order,
                                           const uint32_t *__restrict__ tile_start, unsigned int total_tiles,
                                           unsigned int &t)
{
    unsigned int nA = 0, nB = 0;
    if (order) { nA = order[0]; nB = order[1]; }
    if (q < nA) { t = order[2 + q]; return true; }
    if (q < nA + nB) { t = order[2 + (total_tiles - 1 - (q - nA))]; return true; }
    t = q - nA - nB;
    if (t >= total_tiles) return false;
    if (order && tile_start[t + 1] - tile_start[t] >= (uint32_t)AMAV_DEEP1) t = 0xffffffffu;
    return true;
}

constexpr int BLEND_THREADS = 256;
constexpr int SUB_W = 8, SUB_H = 4;        // pixels of one warp: 8 wide x 4 tall

// Upper bound of power(p) = Ah*dx^2 + Bn*dx*dy + Ch*dy^2 over the warp's pixel rectangle (dx = gx - px in
// [dxl, dxh], dy likewise).  power is a concave quadratic with its maximum (0) at the Gaussian centre, so over a
// box that does not contain the centre the maximum lies on the edge(s) facing it, where it is a clamped 1-D
// quadratic optimum.  A Gaussian whose bound is below pmin (minus a rounding margin) cannot reach alpha >= 1/255
// at any pixel of the rectangle: the exact per-pixel test would reject it 32 times, so dropping it from the
// warp's list changes nothing in the result.
template <int RW = SUB_W, int RH = SUB_H>
__device__ __forceinline__ bool warp_rect_keep(const float4 a, const float4 b, float x0, float y0)
{
    const float Ah = a.z, Bn = a.w, Ch = b.x, pmin = b.z;
    const float dxl = a.x - (x0 + (float)(RW - 1)), dxh = a.x - x0;
    const float dyl = a.y - (y0 + (float)(RH - 1)), dyh = a.y - y0;
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
