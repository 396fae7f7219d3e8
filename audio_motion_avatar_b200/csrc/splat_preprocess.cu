// Stage 4a: per-Gaussian projection for all frames of a call + per-tile counting.
//
// Replaces upstream preprocessCUDA of diff-gaussian-rasterization as called by the reference at
// src/models/renderer.py:516-566.  Compiled with -fmad=false: an FMA
// happens only where fmaf() is written, every expression is evaluated in the order written, so the
// results (depth bits, radius, tile rectangle, conic) are bit-identical to oracle/raster_oracle.c.
//
// One thread per (frame, Gaussian); inputs are read once, coalesced (12+12+16+4+12 B per Gaussian),
// the 48-byte geometry record is written as three float4 stores.  HBM-bound: 56 B in + 60 B out.
#include "common.cuh"

namespace {

struct PreprocessParams {
    int F, N, vps, W, H, gx, gy, activate;
    float scale_modifier;
    const float *means3d, *scales, *rots, *opac, *colors, *view, *proj, *tanfov;
    float4 *geom;
    ushort4 *rect;
    int32_t *tiles_touched;
    uint32_t *tile_count;
};

// CTA = PRE_CHUNK consecutive Gaussians of ONE frame (PRE_CHUNK / 256 per thread, strided so loads stay
// coalesced).  Per-tile counts are accumulated in a shared-memory histogram of the frame's tiles and flushed with
// one global atomicAdd per non-empty (CTA, tile): ~20x fewer global atomics than one per (Gaussian, tile) -- the
// body covers a few hundred tiles, a chunk emits ~16k entries.  `use_smem == 0` (more tiles than fit in shared
// memory) falls back to direct global atomics.
#ifndef AMAV_PRE_CHUNK
#define AMAV_PRE_CHUNK 2048
#endif
constexpr int PRE_CHUNK = AMAV_PRE_CHUNK;

#ifndef AMAV_PRE_CTAS
#define AMAV_PRE_CTAS 4
#endif
__global__ void __launch_bounds__(256, AMAV_PRE_CTAS) preprocess_kernel(const PreprocessParams p, const int use_smem)
{
    extern __shared__ uint32_t s_hist[];
    const int f = blockIdx.y;
    const int tiles = p.gx * p.gy;
    if (use_smem) {
        for (int t = threadIdx.x; t < tiles; t += 256) s_hist[t] = 0u;
        __syncthreads();
    }
    uint32_t *gcnt = p.tile_count + (int64_t)f * tiles;
    uint32_t *cnt = use_smem ? s_hist : gcnt;
    const float *view = p.view + 16 * f;
    const float *proj = p.proj + 16 * f;
    const float tanfovx = p.tanfov[2 * f], tanfovy = p.tanfov[2 * f + 1];
    for (int k = 0; k < PRE_CHUNK / 256; ++k) {
    const int i = blockIdx.x * PRE_CHUNK + k * 256 + threadIdx.x;
    if (i >= p.N) break;
    const int64_t gid = (int64_t)f * p.N + i;
    const int64_t src = (int64_t)(f / p.vps) * p.N + i;

    float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0, g2 = g0;
    ushort4 rc = make_ushort4(0, 0, 0, 0);
    int touched = 0;

    const float px = p.means3d[3 * src], py = p.means3d[3 * src + 1], pz = p.means3d[3 * src + 2];
    // step 2: view-space point, near cull
    const float pvx = view[0] * px + view[4] * py + view[8] * pz + view[12];
    const float pvy = view[1] * px + view[5] * py + view[9] * pz + view[13];
    const float pvz = view[2] * px + view[6] * py + view[10] * pz + view[14];
    bool ok = pvz > 0.2f;
    if (ok) {
        // step 3: clip space
        const float phx = proj[0] * px + proj[4] * py + proj[8] * pz + proj[12];
        const float phy = proj[1] * px + proj[5] * py + proj[9] * pz + proj[13];
        const float phw = proj[3] * px + proj[7] * py + proj[11] * pz + proj[15];
        const float pw = 1.0f / (phw + 0.0000001f);
        const float ppx = phx * pw, ppy = phy * pw;

        float s0 = p.scales[3 * src], s1 = p.scales[3 * src + 1], s2 = p.scales[3 * src + 2];
        float op = p.opac[src];
        float cr = p.colors[3 * src], cg = p.colors[3 * src + 1], cb = p.colors[3 * src + 2];
        if (p.activate) {
            // render_one, renderer.py:532-533,547
            s0 = fminf(expf(s0 - 3.9f), 0.1f);
            s1 = fminf(expf(s1 - 3.9f), 0.1f);
            s2 = fminf(expf(s2 - 3.9f), 0.1f);
            op = 1.0f / (1.0f + expf(-op));
            cr = fminf(fmaxf(cr, 0.0f), 1.0f);
            cg = fminf(fmaxf(cg, 0.0f), 1.0f);
            cb = fminf(fmaxf(cb, 0.0f), 1.0f);
        }
        // step 4: 3D covariance, quaternion (w,x,y,z) NOT normalised
        const float sx = p.scale_modifier * s0, sy = p.scale_modifier * s1, sz = p.scale_modifier * s2;
        const float4 q = reinterpret_cast<const float4 *>(p.rots)[src];
        const float qr = q.x, qx = q.y, qy = q.z, qz = q.w;
        const float R00 = 1.0f - 2.0f * (qy * qy + qz * qz);
        const float R01 = 2.0f * (qx * qy - qr * qz);
        const float R02 = 2.0f * (qx * qz + qr * qy);
        const float R10 = 2.0f * (qx * qy + qr * qz);
        const float R11 = 1.0f - 2.0f * (qx * qx + qz * qz);
        const float R12 = 2.0f * (qy * qz - qr * qx);
        const float R20 = 2.0f * (qx * qz - qr * qy);
        const float R21 = 2.0f * (qy * qz + qr * qx);
        const float R22 = 1.0f - 2.0f * (qx * qx + qy * qy);
        const float m00 = R00 * sx, m01 = R01 * sy, m02 = R02 * sz;
        const float m10 = R10 * sx, m11 = R11 * sy, m12 = R12 * sz;
        const float m20 = R20 * sx, m21 = R21 * sy, m22 = R22 * sz;
        const float S00 = m00 * m00 + m01 * m01 + m02 * m02;
        const float S01 = m00 * m10 + m01 * m11 + m02 * m12;
        const float S02 = m00 * m20 + m01 * m21 + m02 * m22;
        const float S11 = m10 * m10 + m11 * m11 + m12 * m12;
        const float S12 = m10 * m20 + m11 * m21 + m12 * m22;
        const float S22 = m20 * m20 + m21 * m21 + m22 * m22;

        // step 5: EWA 2D covariance
        const float focal_x = (float)p.W / (2.0f * tanfovx);
        const float focal_y = (float)p.H / (2.0f * tanfovy);
        const float limx = 1.3f * tanfovx, limy = 1.3f * tanfovy;
        const float txc = fminf(limx, fmaxf(-limx, pvx / pvz)) * pvz;
        const float tyc = fminf(limy, fmaxf(-limy, pvy / pvz)) * pvz;
        const float J00 = focal_x / pvz;
        const float J02 = -(focal_x * txc) / (pvz * pvz);
        const float J11 = focal_y / pvz;
        const float J12 = -(focal_y * tyc) / (pvz * pvz);
        const float A00 = J00 * view[0] + J02 * view[2];
        const float A01 = J00 * view[4] + J02 * view[6];
        const float A02 = J00 * view[8] + J02 * view[10];
        const float A10 = J11 * view[1] + J12 * view[2];
        const float A11 = J11 * view[5] + J12 * view[6];
        const float A12 = J11 * view[9] + J12 * view[10];
        const float u0 = A00 * S00 + A01 * S01 + A02 * S02;
        const float u1 = A00 * S01 + A01 * S11 + A02 * S12;
        const float u2 = A00 * S02 + A01 * S12 + A02 * S22;
        const float v0 = A10 * S00 + A11 * S01 + A12 * S02;
        const float v1 = A10 * S01 + A11 * S11 + A12 * S12;
        const float v2 = A10 * S02 + A11 * S12 + A12 * S22;
        float ca = u0 * A00 + u1 * A01 + u2 * A02;
        const float cb2 = u0 * A10 + u1 * A11 + u2 * A12;
        float cc = v0 * A10 + v1 * A11 + v2 * A12;

        // step 6/7: dilation + conic
        ca = ca + 0.3f;
        cc = cc + 0.3f;
        const float det = ca * cc - cb2 * cb2;
        ok = det != 0.0f;
        if (ok) {
            const float det_inv = 1.0f / det;
            const float conx = cc * det_inv, cony = -cb2 * det_inv, conz = ca * det_inv;
            // step 8: radius
            const float mid = 0.5f * (ca + cc);
            const float sq = sqrtf(fmaxf(0.1f, mid * mid - det));
            const float lambda1 = mid + sq, lambda2 = mid - sq;
            const float radf = ceilf(3.0f * sqrtf(fmaxf(lambda1, lambda2)));
            // step 9: pixel centre
            const float pix = ((ppx + 1.0f) * (float)p.W - 1.0f) * 0.5f;
            const float piy = ((ppy + 1.0f) * (float)p.H - 1.0f) * 0.5f;
            ok = (fabsf(pix) < 1.0e9f) && (fabsf(piy) < 1.0e9f) && (radf < 1.0e9f);
            if (ok) {
                const int rad = (int)radf;
                const int minx = min(p.gx, max(0, (int)((pix - (float)rad) / 16.0f)));
                const int miny = min(p.gy, max(0, (int)((piy - (float)rad) / 16.0f)));
                const int maxx = min(p.gx, max(0, (int)((pix + (float)rad + 15.0f) / 16.0f)));
                const int maxy = min(p.gy, max(0, (int)((piy + (float)rad + 15.0f) / 16.0f)));
                touched = (maxx - minx) * (maxy - miny);
                if (touched > 0) {
                    // blend skips a pair before evaluating exp() when power < pmin: alpha = op*exp(power) is then
                    // below 1/255 with a 1e-4 relative margin, far above the rounding of logf/exp, so the exact
                    // alpha < 1/255 test that follows sees every borderline pair (bit-exact lists are kept).
                    const float pmin = fmaxf(logf((1.0f / 255.0f) / op) - 1.0e-4f, -80.0f);   // >= -80 (also for NaN): amav_exp's domain
                    g0 = make_float4(pix, piy, -0.5f * conx, -cony);
                    g1 = make_float4(-0.5f * conz, op, pmin, cr);
                    g2 = make_float4(cg, cb, pvz, (float)rad);
                    rc = make_ushort4((unsigned short)minx, (unsigned short)miny, (unsigned short)maxx,
                                      (unsigned short)maxy);
                    {   // one flat loop over the rectangle (lanes diverge on area only, not on shape)
                        const int w = maxx - minx, n = w * (maxy - miny);
                        int t = miny * p.gx + minx, x = 0;
                        for (int e = 0; e < n; ++e) {
                            atomicAdd(cnt + t, 1u);
                            ++t;
                            if (++x == w) { x = 0; t += p.gx - w; }
                        }
                    }
                }
            }
        }
    }
    p.geom[3 * gid] = g0;
    p.geom[3 * gid + 1] = g1;
    p.geom[3 * gid + 2] = g2;
    p.rect[gid] = rc;
    p.tiles_touched[gid] = touched;
    }
    if (use_smem) {
        __syncthreads();
        for (int t = threadIdx.x; t < tiles; t += 256) {
            const uint32_t c = s_hist[t];
            if (c) atomicAdd(gcnt + t, c);
        }
    }
}

}  // namespace

extern "C" int amav_splat_preprocess(const AmavSplatDims *d, const float *means3d, const float *scales,
                                     const float *rots, const float *opacities, const float *colors,
                                     const float *view, const float *proj, const float *tanfov, float *geom,
                                     uint16_t *rect, int32_t *tiles_touched, uint32_t *tile_count, void *stream)
{
    AMAV_CHECK_ARG(d && d->frames >= 0 && d->num_gaussians >= 0 && d->views_per_set >= 1, "bad dims");
    AMAV_CHECK_ARG(d->width > 0 && d->height > 0 && d->width <= 65535 * 16 && d->height <= 65535 * 16, "bad image size");
    const int64_t total = (int64_t)d->frames * d->num_gaussians;
    if (total == 0) return AMAV_OK;
    AMAV_CHECK_ARG(means3d && scales && rots && opacities && colors && view && proj && tanfov, "null input");
    AMAV_CHECK_ARG(geom && rect && tiles_touched && tile_count, "null output");
    PreprocessParams p;
    p.F = d->frames; p.N = d->num_gaussians; p.vps = d->views_per_set; p.W = d->width; p.H = d->height;
    p.gx = (d->width + 15) / 16; p.gy = (d->height + 15) / 16; p.activate = d->activate;
    p.scale_modifier = d->scale_modifier;
    p.means3d = means3d; p.scales = scales; p.rots = rots; p.opac = opacities; p.colors = colors;
    p.view = view; p.proj = proj; p.tanfov = tanfov;
    p.geom = reinterpret_cast<float4 *>(geom); p.rect = reinterpret_cast<ushort4 *>(rect);
    p.tiles_touched = tiles_touched; p.tile_count = tile_count;
    AMAV_CHECK_ARG(d->frames <= 65535, "too many frames for one call");
    const size_t hist = (size_t)p.gx * p.gy * sizeof(uint32_t);
    const int use_smem = hist <= 160 * 1024;
    if (use_smem) amav_ensure_dyn_smem((const void *)preprocess_kernel, (int)hist);
    dim3 grid(amav_div_up(d->num_gaussians, PRE_CHUNK), d->frames);
    preprocess_kernel<<<grid, 256, use_smem ? hist : 0, (cudaStream_t)stream>>>(p, use_smem);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
