// Stage 4c: front-to-back alpha blending of every tile of every frame (upstream renderCUDA,
// SURVEY.md Appendix A.3; reference call site src/models/renderer.py:555-568).
//
// Compiled with -fmad=false; the per-(pixel, Gaussian) arithmetic is written with explicit
// fmaf() exactly as in oracle/raster_oracle.c (amav_oracle_blend), including the polynomial
// exp, so colours, final_T and n_contrib are bit-identical to the oracle.
//
// Persistent CTAs (256 threads = one 16x16 tile) pull tiles from an atomic work counter, stage
// the tile's sorted Gaussian records through shared memory in batches of 256, and stop a tile as
// soon as every pixel is saturated (T < 1e-4).
#include "common.cuh"

namespace {

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

struct BlendParams {
    int F, N, W, H, gx, gy, hwc, clamp01;
    const float4 *geom;
    const uint32_t *tile_start;
    const int32_t *point_list;
    float bg0, bg1, bg2;
    float *out_color, *final_T, *out_invdepth;
    int32_t *n_contrib;
    unsigned int *work_counter;
};

constexpr int BLEND_THREADS = 256;

__global__ void __launch_bounds__(BLEND_THREADS) blend_kernel(const BlendParams p, const float *__restrict__ bg)
{
    __shared__ float4 s_g0[BLEND_THREADS];
    __shared__ float4 s_g1[BLEND_THREADS];
    __shared__ float2 s_g2[BLEND_THREADS];
    const int tiles = p.gx * p.gy;
    const int64_t total_tiles = (int64_t)p.F * tiles;
    const float bg0 = bg[0], bg1 = bg[1], bg2 = bg[2];
    const int lx = threadIdx.x & 15, ly = threadIdx.x >> 4;
    const float alpha_min = 1.0f / 255.0f;

    for (int64_t t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int f = (int)(t / tiles);
        const int tile = (int)(t - (int64_t)f * tiles);
        const int ty = tile / p.gx, tx = tile - ty * p.gx;
        const int pxi = tx * 16 + lx, pyi = ty * 16 + ly;
        const bool inside = pxi < p.W && pyi < p.H;
        const float pxf = (float)pxi, pyf = (float)pyi;
        const uint32_t r0 = p.tile_start[t], r1 = p.tile_start[t + 1];
        const float4 *geom = p.geom + (int64_t)f * p.N * 3;

        float T = 1.0f, C0 = 0.0f, C1 = 0.0f, C2 = 0.0f, Dv = 0.0f;
        int contributor = 0, last = 0;
        bool done = !inside;
        int todo = (int)(r1 - r0);
        for (uint32_t base = r0; base < r1; base += BLEND_THREADS, todo -= BLEND_THREADS) {
            if (__syncthreads_count(done) == BLEND_THREADS) break;
            if (base + threadIdx.x < r1) {
                const int g = p.point_list[base + threadIdx.x];
                const float4 a = geom[3 * (int64_t)g], b = geom[3 * (int64_t)g + 1], c = geom[3 * (int64_t)g + 2];
                s_g0[threadIdx.x] = a;
                s_g1[threadIdx.x] = b;
                s_g2[threadIdx.x] = make_float2(c.x, c.y);
            }
            __syncthreads();
            const int nb = todo < BLEND_THREADS ? todo : BLEND_THREADS;
            for (int j = 0; !done && j < nb; ++j) {
                ++contributor;
                const float4 a = s_g0[j];
                const float dx = a.x - pxf, dy = a.y - pyf;
                const float4 b = s_g1[j];
                const float power = fmaf(fmaf(a.w, dy, a.z * dx), dx, (b.x * dy) * dy);
                if (power > 0.0f) continue;
                const float alpha = fminf(0.99f, b.y * amav_exp(power));
                if (alpha < alpha_min) continue;
                const float Tn = T * (1.0f - alpha);
                if (Tn < 0.0001f) { done = true; continue; }
                const float w = alpha * T;
                const float2 c = s_g2[j];
                C0 = fmaf(b.z, w, C0);
                C1 = fmaf(b.w, w, C1);
                C2 = fmaf(c.x, w, C2);
                if (p.out_invdepth) Dv = fmaf(1.0f / c.y, w, Dv);
                T = Tn;
                last = contributor;
            }
        }
        if (inside) {
            float o0 = fmaf(T, bg0, C0), o1 = fmaf(T, bg1, C1), o2 = fmaf(T, bg2, C2);
            if (p.clamp01) {
                o0 = fminf(fmaxf(o0, 0.0f), 1.0f);
                o1 = fminf(fmaxf(o1, 0.0f), 1.0f);
                o2 = fminf(fmaxf(o2, 0.0f), 1.0f);
            }
            const int64_t pix = (int64_t)pyi * p.W + pxi;
            const int64_t hw = (int64_t)p.H * p.W;
            if (p.hwc) {
                float *o = p.out_color + ((int64_t)f * hw + pix) * 3;
                o[0] = o0; o[1] = o1; o[2] = o2;
            } else {
                float *o = p.out_color + (int64_t)f * 3 * hw + pix;
                o[0] = o0; o[hw] = o1; o[2 * hw] = o2;
            }
            if (p.final_T) p.final_T[(int64_t)f * hw + pix] = T;
            if (p.n_contrib) p.n_contrib[(int64_t)f * hw + pix] = last;
            if (p.out_invdepth) p.out_invdepth[(int64_t)f * hw + pix] = Dv;
        }
        __syncthreads();
    }
}

}  // namespace

extern "C" int amav_splat_blend(const AmavSplatDims *d, const float *geom, const uint32_t *tile_start,
                                const int32_t *point_list, const float *bg, int hwc, int clamp01,
                                float *out_color, float *final_T, int32_t *n_contrib, float *out_invdepth,
                                void *stream)
{
    AMAV_CHECK_ARG(d && d->frames >= 0 && d->width > 0 && d->height > 0, "bad dims");
    if (d->frames == 0) return AMAV_OK;
    AMAV_CHECK_ARG(tile_start && bg && out_color, "null buffer");
    BlendParams p;
    p.F = d->frames; p.N = d->num_gaussians; p.W = d->width; p.H = d->height;
    p.gx = (d->width + 15) / 16; p.gy = (d->height + 15) / 16; p.hwc = hwc; p.clamp01 = clamp01;
    p.geom = reinterpret_cast<const float4 *>(geom); p.tile_start = tile_start; p.point_list = point_list;
    p.out_color = out_color; p.final_T = final_T; p.n_contrib = n_contrib; p.out_invdepth = out_invdepth;
    p.work_counter = nullptr; p.bg0 = p.bg1 = p.bg2 = 0.f;
    const int64_t total_tiles = (int64_t)d->frames * p.gx * p.gy;
    const int grid = (int)(total_tiles < AMAV_SM_COUNT * 8 ? total_tiles : AMAV_SM_COUNT * 8);
    blend_kernel<<<grid, BLEND_THREADS, 0, (cudaStream_t)stream>>>(p, bg);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
