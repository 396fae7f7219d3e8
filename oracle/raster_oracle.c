/*
 * oracle/raster_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Single-threaded (optionally OpenMP-over-tiles) CPU restatement, in plain C,
 * of the tile-based 3D Gaussian splatting rasteriser that the reference calls
 * at  src/models/renderer.py:516-566  (GaussianRasterizationSettings /
 * GaussianRasterizer from graphdeco-inria/diff-gaussian-rasterization, `main`,
 * un-pinned `git clone`, README.md:117-121).  That package is NOT vendored in
 * /root/reference and cannot be fetched offline, so this file restates its
 * published algorithm:
 *     preprocess -> inclusive scan -> duplicateWithKeys -> stable sort on
 *     (tile<<32 | depth bits) -> identifyTileRanges -> 16x16 blend,
 * plus the backward pass.
 *
 * PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures
 * for this path, and the upstream CUDA binary cannot be
 * built here.  The pin is: the invariants in tests/test_oracle_raster.py
 * (analytic single-Gaussian splat, sum of weights = 1 - final_T, finite
 * differences for the backward) and the committed tests/golden vectors.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.
 *
 * Floating-point contract (so the CUDA kernels can match bit for bit):
 *   - compiled with -ffp-contract=off: no implicit FMA contraction;
 *   - an FMA happens only where fmaf() is written;
 *   - expressions evaluate left to right exactly as written;
 *   - exp() in the blend is amav_exp() below (degree-7 polynomial, <= 1 ulp
 *     of expf over the range used).  Upstream uses CUDA expf (<= 2 ulp), which
 *     no CPU can reproduce bit for bit; define AMAV_ORACLE_LIBM_EXP to use
 *     libm expf instead (tests check the two agree to ~1e-7 relative).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define BLOCK_X 16
#define BLOCK_Y 16

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }

/* exp(x) for x <= 0 (blend only evaluates it for power <= 0). */
float amav_oracle_exp(float x)
{
#ifdef AMAV_ORACLE_LIBM_EXP
    return expf(x);
#else
    if (x < -80.0f) x = -80.0f;
    const float L2E_HI = 1.44269502162933349609375f;   /* fp32(log2 e)            */
    const float L2E_LO = 1.925963033500011e-8f;        /* log2 e - fp32(log2 e)   */
    const float MAGIC = 12582912.0f;                   /* 1.5 * 2^23              */
    float t = x * L2E_HI;
    float r = t + MAGIC;
    float n = r - MAGIC;                               /* rint(t), exact          */
    float f = fmaf(x, L2E_HI, -n);
    f = fmaf(x, L2E_LO, f);                            /* f in about [-0.5, 0.5]  */
    float p = 1.5252733804059840e-5f;                  /* ln2^7/7!                */
    p = fmaf(p, f, 1.5403530393381610e-4f);            /* ln2^6/6!                */
    p = fmaf(p, f, 1.3333558146428443e-3f);            /* ln2^5/5!                */
    p = fmaf(p, f, 9.6181291076284772e-3f);            /* ln2^4/4!                */
    p = fmaf(p, f, 5.5504108664821580e-2f);            /* ln2^3/3!                */
    p = fmaf(p, f, 2.4022650695910071e-1f);            /* ln2^2/2!                */
    p = fmaf(p, f, 6.9314718055994531e-1f);            /* ln2                     */
    p = fmaf(p, f, 1.0f);
    int32_t ni = (int32_t)n;
    return u2f((uint32_t)((int32_t)f2u(p) + ni * (1 << 23)));
#endif
}

/* ------------------------------------------------------------------ */
/* A.1  per-Gaussian preprocess                                        */
/* ------------------------------------------------------------------ */
/*
 * view / proj: 16 floats each, exactly the tensors the reference passes as
 * viewmatrix / projmatrix (renderer.py:523-524), i.e. already transposed, so
 * element (row r, col c) of the mathematical matrix is m[c*4 + r].
 *
 * Outputs (all caller-allocated, length N unless noted):
 *   depth[N], radii[N], xy[2N], conic_opacity[4N], cov3d[6N],
 *   rect[4N] = (minx, miny, maxx, maxy), tiles_touched[N].
 */
void amav_oracle_preprocess(
    int N, const float *means3d, const float *scales, const float *rots,
    const float *opacities, const float *view, const float *proj,
    int W, int H, float tanfovx, float tanfovy, float scale_modifier,
    float *depth, int32_t *radii, float *xy, float *conic_opacity,
    float *cov3d, int32_t *rect, int32_t *tiles_touched)
{
    const int gx = (W + BLOCK_X - 1) / BLOCK_X;
    const int gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    const float focal_x = (float)W / (2.0f * tanfovx);
    const float focal_y = (float)H / (2.0f * tanfovy);
    const float limx = 1.3f * tanfovx;
    const float limy = 1.3f * tanfovy;

    for (int i = 0; i < N; ++i) {
        radii[i] = 0;
        tiles_touched[i] = 0;
        depth[i] = 0.0f;
        xy[2 * i] = xy[2 * i + 1] = 0.0f;
        for (int k = 0; k < 4; ++k) { conic_opacity[4 * i + k] = 0.0f; rect[4 * i + k] = 0; }
        for (int k = 0; k < 6; ++k) cov3d[6 * i + k] = 0.0f;

        const float px = means3d[3 * i], py = means3d[3 * i + 1], pz = means3d[3 * i + 2];
        /* step 2: view-space point, near cull */
        const float pvx = view[0] * px + view[4] * py + view[8] * pz + view[12];
        const float pvy = view[1] * px + view[5] * py + view[9] * pz + view[13];
        const float pvz = view[2] * px + view[6] * py + view[10] * pz + view[14];
        if (!(pvz > 0.2f)) continue;

        /* step 3: clip space */
        const float phx = proj[0] * px + proj[4] * py + proj[8] * pz + proj[12];
        const float phy = proj[1] * px + proj[5] * py + proj[9] * pz + proj[13];
        const float phw = proj[3] * px + proj[7] * py + proj[11] * pz + proj[15];
        const float pw = 1.0f / (phw + 0.0000001f);
        const float ppx = phx * pw, ppy = phy * pw;

        /* step 4: 3D covariance, Sigma = Rot S S Rot^T, quaternion NOT normalised */
        const float sx = scale_modifier * scales[3 * i];
        const float sy = scale_modifier * scales[3 * i + 1];
        const float sz = scale_modifier * scales[3 * i + 2];
        const float qr = rots[4 * i], qx = rots[4 * i + 1], qy = rots[4 * i + 2], qz = rots[4 * i + 3];
        const float R00 = 1.0f - 2.0f * (qy * qy + qz * qz);
        const float R01 = 2.0f * (qx * qy - qr * qz);
        const float R02 = 2.0f * (qx * qz + qr * qy);
        const float R10 = 2.0f * (qx * qy + qr * qz);
        const float R11 = 1.0f - 2.0f * (qx * qx + qz * qz);
        const float R12 = 2.0f * (qy * qz - qr * qx);
        const float R20 = 2.0f * (qx * qz - qr * qy);
        const float R21 = 2.0f * (qy * qz + qr * qx);
        const float R22 = 1.0f - 2.0f * (qx * qx + qy * qy);
        /* m[a][i] = Rot[a][i] * s_i */
        const float m00 = R00 * sx, m01 = R01 * sy, m02 = R02 * sz;
        const float m10 = R10 * sx, m11 = R11 * sy, m12 = R12 * sz;
        const float m20 = R20 * sx, m21 = R21 * sy, m22 = R22 * sz;
        const float S00 = m00 * m00 + m01 * m01 + m02 * m02;
        const float S01 = m00 * m10 + m01 * m11 + m02 * m12;
        const float S02 = m00 * m20 + m01 * m21 + m02 * m22;
        const float S11 = m10 * m10 + m11 * m11 + m12 * m12;
        const float S12 = m10 * m20 + m11 * m21 + m12 * m22;
        const float S22 = m20 * m20 + m21 * m21 + m22 * m22;

        /* step 5: EWA 2D covariance  cov = A Sigma A^T,  A = Jac(2x3) * Rv(3x3) */
        const float txc = fminf(limx, fmaxf(-limx, pvx / pvz)) * pvz;
        const float tyc = fminf(limy, fmaxf(-limy, pvy / pvz)) * pvz;
        const float J00 = focal_x / pvz;
        const float J02 = -(focal_x * txc) / (pvz * pvz);
        const float J11 = focal_y / pvz;
        const float J12 = -(focal_y * tyc) / (pvz * pvz);
        /* Rv[r][c] = view[c*4 + r] */
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
        const float cb = u0 * A10 + u1 * A11 + u2 * A12;
        float cc = v0 * A10 + v1 * A11 + v2 * A12;

        /* step 6/7: dilation, conic (antialiasing=False => opacity factor 1) */
        ca = ca + 0.3f;
        cc = cc + 0.3f;
        const float det = ca * cc - cb * cb;
        if (det == 0.0f) continue;
        const float det_inv = 1.0f / det;
        const float conx = cc * det_inv, cony = -cb * det_inv, conz = ca * det_inv;

        /* step 8: radius */
        const float mid = 0.5f * (ca + cc);
        const float sq = sqrtf(fmaxf(0.1f, mid * mid - det));
        const float lambda1 = mid + sq, lambda2 = mid - sq;
        const float radf = ceilf(3.0f * sqrtf(fmaxf(lambda1, lambda2)));

        /* step 9: pixel centre */
        const float pix = ((ppx + 1.0f) * (float)W - 1.0f) * 0.5f;
        const float piy = ((ppy + 1.0f) * (float)H - 1.0f) * 0.5f;
        /* non-finite values are undefined behaviour upstream ((int) of NaN);
           both this oracle and the kernels cull them. */
        if (!(fabsf(pix) < 1.0e9f) || !(fabsf(piy) < 1.0e9f) || !(radf < 1.0e9f)) continue;
        const int rad = (int)radf;

        /* step 10: tile rectangle */
        const int minx = imin(gx, imax(0, (int)((pix - (float)rad) / (float)BLOCK_X)));
        const int miny = imin(gy, imax(0, (int)((piy - (float)rad) / (float)BLOCK_Y)));
        const int maxx = imin(gx, imax(0, (int)((pix + (float)rad + (float)(BLOCK_X - 1)) / (float)BLOCK_X)));
        const int maxy = imin(gy, imax(0, (int)((piy + (float)rad + (float)(BLOCK_Y - 1)) / (float)BLOCK_Y)));
        if ((maxx - minx) * (maxy - miny) == 0) continue;

        /* step 11 */
        depth[i] = pvz;
        radii[i] = rad;
        xy[2 * i] = pix; xy[2 * i + 1] = piy;
        conic_opacity[4 * i] = conx; conic_opacity[4 * i + 1] = cony;
        conic_opacity[4 * i + 2] = conz; conic_opacity[4 * i + 3] = opacities[i];
        cov3d[6 * i] = S00; cov3d[6 * i + 1] = S01; cov3d[6 * i + 2] = S02;
        cov3d[6 * i + 3] = S11; cov3d[6 * i + 4] = S12; cov3d[6 * i + 5] = S22;
        rect[4 * i] = minx; rect[4 * i + 1] = miny; rect[4 * i + 2] = maxx; rect[4 * i + 3] = maxy;
        tiles_touched[i] = (maxy - miny) * (maxx - minx);
    }
}

/* ------------------------------------------------------------------ */
/* A.2  binning: scan -> key emit -> stable sort -> tile ranges         */
/* ------------------------------------------------------------------ */
/* Inclusive scan of tiles_touched; returns D = offsets[N-1] (0 for N==0). */
int64_t amav_oracle_scan(int N, const int32_t *tiles_touched, int64_t *offsets)
{
    int64_t acc = 0;
    for (int i = 0; i < N; ++i) { acc += tiles_touched[i]; offsets[i] = acc; }
    return acc;
}

/* keys[D] = (tile << 32) | bits(depth); values[D] = Gaussian index. Emission
 * order: Gaussian index ascending, tile rows outer, tile columns inner. */
void amav_oracle_emit_keys(int N, const int32_t *radii, const int32_t *rect,
                           const float *depth, const int64_t *offsets, int W,
                           uint64_t *keys, int32_t *values)
{
    const int gx = (W + BLOCK_X - 1) / BLOCK_X;
    for (int i = 0; i < N; ++i) {
        if (radii[i] <= 0) continue;
        int64_t off = (i == 0) ? 0 : offsets[i - 1];
        for (int y = rect[4 * i + 1]; y < rect[4 * i + 3]; ++y)
            for (int x = rect[4 * i]; x < rect[4 * i + 2]; ++x) {
                uint64_t key = (uint64_t)(uint32_t)(y * gx + x);
                key = (key << 32) | (uint64_t)f2u(depth[i]);
                keys[off] = key;
                values[off] = i;
                ++off;
            }
    }
}

/* Stable sort of (key, value) pairs by key: bottom-up merge sort.  Same result
 * as a stable LSD radix sort over all key bits. */
void amav_oracle_sort_pairs(int64_t D, uint64_t *keys, int32_t *values)
{
    if (D <= 1) return;
    uint64_t *kt = (uint64_t *)malloc((size_t)D * sizeof(uint64_t));
    int32_t *vt = (int32_t *)malloc((size_t)D * sizeof(int32_t));
    uint64_t *ka = keys, *kb = kt;
    int32_t *va = values, *vb = vt;
    for (int64_t w = 1; w < D; w *= 2) {
        for (int64_t lo = 0; lo < D; lo += 2 * w) {
            int64_t mid = lo + w < D ? lo + w : D;
            int64_t hi = lo + 2 * w < D ? lo + 2 * w : D;
            int64_t i = lo, j = mid, o = lo;
            while (i < mid && j < hi) {
                if (ka[j] < ka[i]) { kb[o] = ka[j]; vb[o] = va[j]; ++j; }
                else               { kb[o] = ka[i]; vb[o] = va[i]; ++i; }
                ++o;
            }
            while (i < mid) { kb[o] = ka[i]; vb[o] = va[i]; ++i; ++o; }
            while (j < hi)  { kb[o] = ka[j]; vb[o] = va[j]; ++j; ++o; }
        }
        uint64_t *ks = ka; ka = kb; kb = ks;
        int32_t *vs = va; va = vb; vb = vs;
    }
    if (ka != keys) {
        memcpy(keys, ka, (size_t)D * sizeof(uint64_t));
        memcpy(values, va, (size_t)D * sizeof(int32_t));
    }
    free(kt); free(vt);
}

/* ranges[2*tile] = first entry, ranges[2*tile+1] = one past last; empty tiles (0,0). */
void amav_oracle_tile_ranges(int64_t D, const uint64_t *keys, int num_tiles, int32_t *ranges)
{
    memset(ranges, 0, (size_t)num_tiles * 2 * sizeof(int32_t));
    for (int64_t k = 0; k < D; ++k) {
        uint32_t t = (uint32_t)(keys[k] >> 32);
        if (k == 0) ranges[2 * t] = 0;
        else {
            uint32_t p = (uint32_t)(keys[k - 1] >> 32);
            if (p != t) { ranges[2 * p + 1] = (int32_t)k; ranges[2 * t] = (int32_t)k; }
        }
        if (k == D - 1) ranges[2 * t + 1] = (int32_t)D;
    }
}

/* ------------------------------------------------------------------ */
/* A.3  blend (forward)                                                */
/* ------------------------------------------------------------------ */
/*
 * point_list[D]: sorted Gaussian indices; ranges[2*tiles]; per-Gaussian xy,
 * conic_opacity, rgb[3N], depth.  bg[3].  Outputs: out_color[3*H*W] (CHW),
 * final_T[H*W], n_contrib[H*W], out_invdepth[H*W] (may be NULL).
 *
 * Arithmetic contract per (pixel, Gaussian):
 *   Ah = -0.5f*conic.x; Bn = -conic.y; Ch = -0.5f*conic.z      (exact scalings)
 *   dx = xy.x - px; dy = xy.y - py
 *   power = fmaf(fmaf(Bn, dy, Ah*dx), dx, (Ch*dy)*dy)
 *   alpha = fminf(0.99f, opac * exp(power));  w = alpha*T;  C = fmaf(rgb, w, C)
 */
void amav_oracle_blend(int W, int H, const int32_t *ranges, const int32_t *point_list,
                       const float *xy, const float *conic_opacity, const float *rgb,
                       const float *depth, const float *bg,
                       float *out_color, float *final_T, int32_t *n_contrib,
                       float *out_invdepth)
{
    const int gx = (W + BLOCK_X - 1) / BLOCK_X;
    const int gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    const float alpha_min = 1.0f / 255.0f;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1)
#endif
    for (int tile = 0; tile < gx * gy; ++tile) {
        const int tx = tile % gx, ty = tile / gx;
        const int r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
        for (int ly = 0; ly < BLOCK_Y; ++ly)
            for (int lx = 0; lx < BLOCK_X; ++lx) {
                const int pxi = tx * BLOCK_X + lx, pyi = ty * BLOCK_Y + ly;
                if (pxi >= W || pyi >= H) continue;
                const float pxf = (float)pxi, pyf = (float)pyi;
                float T = 1.0f, C0 = 0.0f, C1 = 0.0f, C2 = 0.0f, Dv = 0.0f;
                int contributor = 0, last = 0;
                for (int k = r0; k < r1; ++k) {
                    const int g = point_list[k];
                    ++contributor;
                    const float Ah = -0.5f * conic_opacity[4 * g];
                    const float Bn = -conic_opacity[4 * g + 1];
                    const float Ch = -0.5f * conic_opacity[4 * g + 2];
                    const float dx = xy[2 * g] - pxf;
                    const float dy = xy[2 * g + 1] - pyf;
                    const float power = fmaf(fmaf(Bn, dy, Ah * dx), dx, (Ch * dy) * dy);
                    if (power > 0.0f) continue;
                    const float alpha = fminf(0.99f, conic_opacity[4 * g + 3] * amav_oracle_exp(power));
                    if (alpha < alpha_min) continue;
                    const float Tn = T * (1.0f - alpha);
                    if (Tn < 0.0001f) break;
                    const float w = alpha * T;
                    C0 = fmaf(rgb[3 * g], w, C0);
                    C1 = fmaf(rgb[3 * g + 1], w, C1);
                    C2 = fmaf(rgb[3 * g + 2], w, C2);
                    Dv = fmaf(1.0f / depth[g], w, Dv);
                    T = Tn;
                    last = contributor;
                }
                const size_t pix = (size_t)pyi * W + pxi;
                final_T[pix] = T;
                n_contrib[pix] = last;
                out_color[pix] = fmaf(T, bg[0], C0);
                out_color[(size_t)H * W + pix] = fmaf(T, bg[1], C1);
                out_color[(size_t)2 * H * W + pix] = fmaf(T, bg[2], C2);
                if (out_invdepth) out_invdepth[pix] = Dv;
            }
    }
}
