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
 * built here.  The pin is: the invariants in tests/test_oracle_invariants.py and tests/test_oracle_backward.py
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
 *   - exp() in the blend is amav_exp() below (degree-5 minimax polynomial,
 *     <= 1.5 ulp of exp over the range used).  Upstream uses CUDA expf (<= 2 ulp), which
 *     no CPU can reproduce bit for bit; define AMAV_ORACLE_LIBM_EXP to use
 *     libm expf instead (tests check the two agree to ~1e-7 relative).
 *   - AMAV_ORACLE_UPSTREAM_ARITH (libamav_oracle_upstream.so, `make upstream`): the blend's
 *     per-pair arithmetic as upstream WRITES it -- libm expf and the literal
 *     power = -0.5f*(con.x*dx*dx + con.z*dy*dy) - con.y*dx*dy, no FMA -- instead of the
 *     FMA nesting and polynomial shared with the kernels.  tests/test_oracle_upstream.py and
 *     bench.py count how many alpha >= 1/255 / T < 1e-4 decisions (n_contrib) and pixels
 *     differ between the two builds: the distance between "our oracle" and "upstream's
 *     arithmetic".  Preprocess and binning are identical in both builds.
 *   - the loops over Gaussians / merge runs carry OpenMP pragmas; every iteration is
 *     independent, so results do not depend on the thread count.
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
#if defined(AMAV_ORACLE_LIBM_EXP) || defined(AMAV_ORACLE_UPSTREAM_ARITH)
    return expf(x);
#else
    if (x < -80.0f) x = -80.0f;
    const float L2E_HI = 1.44269502162933349609375f;   /* fp32(log2 e)            */
    const float L2E_LO = 1.925963033500011e-8f;        /* log2 e - fp32(log2 e)   */
    const float MAGIC = 12582912.0f;                   /* 1.5 * 2^23              */
    float r = fmaf(x, L2E_HI, MAGIC);
    float n = r - MAGIC;                               /* rint(x log2 e), exact   */
    float f = fmaf(x, L2E_HI, -n);
    f = fmaf(x, L2E_LO, f);                            /* f in about [-0.5, 0.5]  */
    float p = 0.001326472731307149f;                   /* degree-5 minimax of     */
    p = fmaf(p, f, 0.009671512991189957f);             /* (2^f - 1)/f on          */
    p = fmaf(p, f, 0.05550733581185341f);              /* [-0.5, 0.5]             */
    p = fmaf(p, f, 0.24022242426872253f);
    p = fmaf(p, f, 0.6931470036506653f);
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

#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
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
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
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
#ifdef _OPENMP
#pragma omp parallel for schedule(static) if (D / (2 * w) >= 4)
#endif
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
#ifdef AMAV_ORACLE_UPSTREAM_ARITH
                    (void)Ah; (void)Bn; (void)Ch;
                    const float power = -0.5f * (conic_opacity[4 * g] * dx * dx + conic_opacity[4 * g + 2] * dy * dy)
                                        - conic_opacity[4 * g + 1] * dx * dy;
#else
                    const float power = fmaf(fmaf(Bn, dy, Ah * dx), dx, (Ch * dy) * dy);
#endif
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

/* ------------------------------------------------------------------ */
/* B.1  blend backward (upstream renderCUDA of  */
/*      backward.cu, reached through autograd from renderer.py:555-568) */
/* ------------------------------------------------------------------ */
/*
 * Per pixel, back to front over the tile's list, entries with 1-based position
 * <= n_contrib only, same skips as the forward.  Per-pair arithmetic in float as
 * upstream; the per-Gaussian sums are kept in double (upstream uses float
 * atomicAdd in arbitrary order, so its own result is only defined up to that
 * rounding; gradient parity is a tolerance check).
 *   dL_dpix [3,H,W] (CHW, gradient w.r.t. the un-clamped colour output)
 *   dL_dmean2D [N,2] (already scaled by 0.5*W / 0.5*H as upstream), dL_dconic [N,3]
 *   = (A, B, C) with the B entry holding HALF the true derivative (upstream
 *   convention, undone in B.2), dL_dopacity [N], dL_dcolor [N,3].
 */
void amav_oracle_blend_backward(int W, int H, int N, const int32_t *ranges, const int32_t *point_list,
                                const float *xy, const float *conic_opacity, const float *rgb,
                                const float *bg, const float *final_T, const int32_t *n_contrib,
                                const float *dL_dpix, float *dL_dmean2D, float *dL_dconic,
                                float *dL_dopacity, float *dL_dcolor)
{
    const int gx = (W + BLOCK_X - 1) / BLOCK_X;
    const int gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    const float alpha_min = 1.0f / 255.0f;
    const float ddelx_dx = 0.5f * (float)W, ddely_dy = 0.5f * (float)H;
    double *acc = (double *)calloc((size_t)(N > 0 ? N : 1) * 9, sizeof(double));
    for (int tile = 0; tile < gx * gy; ++tile) {
        const int tx = tile % gx, ty = tile / gx;
        const int r0 = ranges[2 * tile];
        for (int ly = 0; ly < BLOCK_Y; ++ly)
            for (int lx = 0; lx < BLOCK_X; ++lx) {
                const int pxi = tx * BLOCK_X + lx, pyi = ty * BLOCK_Y + ly;
                if (pxi >= W || pyi >= H) continue;
                const size_t pix = (size_t)pyi * W + pxi;
                const float pxf = (float)pxi, pyf = (float)pyi;
                const float T_final = final_T[pix];
                float T = T_final;
                const int last_contributor = n_contrib[pix];
                const float g0 = dL_dpix[pix], g1 = dL_dpix[(size_t)H * W + pix], g2 = dL_dpix[(size_t)2 * H * W + pix];
                const float bg_dot = bg[0] * g0 + bg[1] * g1 + bg[2] * g2;
                float accum0 = 0.0f, accum1 = 0.0f, accum2 = 0.0f, last_alpha = 0.0f;
                float lc0 = 0.0f, lc1 = 0.0f, lc2 = 0.0f;
                for (int k = r0 + last_contributor - 1; k >= r0; --k) {
                    const int g = point_list[k];
                    const float conx = conic_opacity[4 * g], cony = conic_opacity[4 * g + 1];
                    const float conz = conic_opacity[4 * g + 2], op = conic_opacity[4 * g + 3];
                    const float Ah = -0.5f * conx, Bn = -cony, Ch = -0.5f * conz;
                    const float dx = xy[2 * g] - pxf, dy = xy[2 * g + 1] - pyf;
                    const float power = fmaf(fmaf(Bn, dy, Ah * dx), dx, (Ch * dy) * dy);
                    if (power > 0.0f) continue;
                    const float G = amav_oracle_exp(power);
                    const float alpha = fminf(0.99f, op * G);
                    if (alpha < alpha_min) continue;
                    T = T / (1.0f - alpha);
                    const float w = alpha * T;
                    float dL_dalpha = 0.0f;
                    accum0 = last_alpha * lc0 + (1.0f - last_alpha) * accum0; lc0 = rgb[3 * g];
                    accum1 = last_alpha * lc1 + (1.0f - last_alpha) * accum1; lc1 = rgb[3 * g + 1];
                    accum2 = last_alpha * lc2 + (1.0f - last_alpha) * accum2; lc2 = rgb[3 * g + 2];
                    dL_dalpha += (lc0 - accum0) * g0;
                    dL_dalpha += (lc1 - accum1) * g1;
                    dL_dalpha += (lc2 - accum2) * g2;
                    dL_dalpha *= T;
                    last_alpha = alpha;
                    dL_dalpha += (-T_final / (1.0f - alpha)) * bg_dot;
                    const float dL_dG = op * dL_dalpha;
                    const float gdx = G * dx, gdy = G * dy;
                    const float dG_ddelx = -gdx * conx - gdy * cony;
                    const float dG_ddely = -gdy * conz - gdx * cony;
                    double *a = acc + (size_t)g * 9;
                    a[0] += (double)(dL_dG * dG_ddelx * ddelx_dx);
                    a[1] += (double)(dL_dG * dG_ddely * ddely_dy);
                    a[2] += (double)(-0.5f * gdx * dx * dL_dG);
                    a[3] += (double)(-0.5f * gdx * dy * dL_dG);
                    a[4] += (double)(-0.5f * gdy * dy * dL_dG);
                    a[5] += (double)(G * dL_dalpha);
                    a[6] += (double)(w * g0);
                    a[7] += (double)(w * g1);
                    a[8] += (double)(w * g2);
                }
            }
    }
    for (int i = 0; i < N; ++i) {
        const double *a = acc + (size_t)i * 9;
        dL_dmean2D[2 * i] = (float)a[0]; dL_dmean2D[2 * i + 1] = (float)a[1];
        dL_dconic[3 * i] = (float)a[2]; dL_dconic[3 * i + 1] = (float)a[3]; dL_dconic[3 * i + 2] = (float)a[4];
        dL_dopacity[i] = (float)a[5];
        dL_dcolor[3 * i] = (float)a[6]; dL_dcolor[3 * i + 1] = (float)a[7]; dL_dcolor[3 * i + 2] = (float)a[8];
    }
    free(acc);
}

/* ------------------------------------------------------------------ */
/* B.2-B.4  per-Gaussian backward: conic -> cov2D -> (cov3D, view-space */
/*          point) -> (scale, rotation, mean3D); mean2D -> mean3D       */
/* ------------------------------------------------------------------ */
/*
 * Follows upstream computeCov2DCUDA / preprocessCUDA / computeCov3D of backward.cu:
 *   - denom2inv = 1/(det^2 + 1e-7);
 *   - where t.x/t.z (t.y/t.z) was clamped to +-1.3 tanfov the x (y) gradient is masked and the
 *     clamped value is treated as independent of t.z (upstream's approximation);
 *   - the quaternion is not normalised (gradient of the raw quaternion -> matrix map).
 * Inputs are the forward's inputs plus radii (> 0 selects the Gaussians that were rendered).
 * Outputs dL_dmeans3d [N,3], dL_dscales [N,3], dL_drots [N,4] are overwritten.
 */
void amav_oracle_preprocess_backward(
    int N, const float *means3d, const float *scales, const float *rots,
    const float *view, const float *proj, int W, int H, float tanfovx, float tanfovy,
    float scale_modifier, const int32_t *radii, const float *dL_dmean2D, const float *dL_dconic,
    float *dL_dmeans3d, float *dL_dscales, float *dL_drots)
{
    const float focal_x = (float)W / (2.0f * tanfovx);
    const float focal_y = (float)H / (2.0f * tanfovy);
    const float limx = 1.3f * tanfovx;
    const float limy = 1.3f * tanfovy;
    for (int i = 0; i < N; ++i) {
        for (int k = 0; k < 3; ++k) { dL_dmeans3d[3 * i + k] = 0.0f; dL_dscales[3 * i + k] = 0.0f; }
        for (int k = 0; k < 4; ++k) dL_drots[4 * i + k] = 0.0f;
        if (radii[i] <= 0) continue;
        const float px = means3d[3 * i], py = means3d[3 * i + 1], pz = means3d[3 * i + 2];
        const float pvx = view[0] * px + view[4] * py + view[8] * pz + view[12];
        const float pvy = view[1] * px + view[5] * py + view[9] * pz + view[13];
        const float pvz = view[2] * px + view[6] * py + view[10] * pz + view[14];
        /* forward quantities again (same expressions as amav_oracle_preprocess) */
        const float s[3] = {scale_modifier * scales[3 * i], scale_modifier * scales[3 * i + 1], scale_modifier * scales[3 * i + 2]};
        const float qr = rots[4 * i], qx = rots[4 * i + 1], qy = rots[4 * i + 2], qz = rots[4 * i + 3];
        float Rm[3][3];
        Rm[0][0] = 1.0f - 2.0f * (qy * qy + qz * qz); Rm[0][1] = 2.0f * (qx * qy - qr * qz); Rm[0][2] = 2.0f * (qx * qz + qr * qy);
        Rm[1][0] = 2.0f * (qx * qy + qr * qz); Rm[1][1] = 1.0f - 2.0f * (qx * qx + qz * qz); Rm[1][2] = 2.0f * (qy * qz - qr * qx);
        Rm[2][0] = 2.0f * (qx * qz - qr * qy); Rm[2][1] = 2.0f * (qy * qz + qr * qx); Rm[2][2] = 1.0f - 2.0f * (qx * qx + qy * qy);
        float m[3][3], S[3][3];
        for (int a = 0; a < 3; ++a) for (int k = 0; k < 3; ++k) m[a][k] = Rm[a][k] * s[k];
        for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) S[a][b] = m[a][0] * m[b][0] + m[a][1] * m[b][1] + m[a][2] * m[b][2];
        const float txtz = pvx / pvz, tytz = pvy / pvz;
        const float txc = fminf(limx, fmaxf(-limx, txtz)) * pvz;
        const float tyc = fminf(limy, fmaxf(-limy, tytz)) * pvz;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.0f : 1.0f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.0f : 1.0f;
        const float J00 = focal_x / pvz, J02 = -(focal_x * txc) / (pvz * pvz);
        const float J11 = focal_y / pvz, J12 = -(focal_y * tyc) / (pvz * pvz);
        float A0[3], A1[3], u[3], v[3];
        for (int k = 0; k < 3; ++k) {              /* Rv[r][k] = view[k*4 + r] */
            A0[k] = J00 * view[k * 4] + J02 * view[k * 4 + 2];
            A1[k] = J11 * view[k * 4 + 1] + J12 * view[k * 4 + 2];
        }
        for (int k = 0; k < 3; ++k) {
            u[k] = A0[0] * S[0][k] + A0[1] * S[1][k] + A0[2] * S[2][k];
            v[k] = A1[0] * S[0][k] + A1[1] * S[1][k] + A1[2] * S[2][k];
        }
        const float a = (u[0] * A0[0] + u[1] * A0[1] + u[2] * A0[2]) + 0.3f;
        const float b = u[0] * A1[0] + u[1] * A1[1] + u[2] * A1[2];
        const float c = (v[0] * A1[0] + v[1] * A1[1] + v[2] * A1[2]) + 0.3f;
        /* B.2: conic -> (a, b, c) */
        const float X = dL_dconic[3 * i], Y = dL_dconic[3 * i + 1], Z = dL_dconic[3 * i + 2];
        const float denom = a * c - b * b;
        const float denom2inv = 1.0f / (denom * denom + 0.0000001f);
        float dL_da = 0.0f, dL_db = 0.0f, dL_dc = 0.0f;
        if (denom2inv != 0.0f) {
            dL_da = denom2inv * (-c * c * X + 2.0f * b * c * Y + (denom - a * c) * Z);
            dL_dc = denom2inv * (-a * a * Z + 2.0f * a * b * Y + (denom - a * c) * X);
            dL_db = denom2inv * 2.0f * (b * c * X - (denom + 2.0f * b * b) * Y + a * b * Z);
        }
        /* (a, b, c) -> Sigma (full symmetric gradient Gs) and -> A */
        float Gs[3][3];
        for (int k = 0; k < 3; ++k)
            for (int l = 0; l < 3; ++l)
                Gs[k][l] = A0[k] * A0[l] * dL_da + 0.5f * (A0[k] * A1[l] + A0[l] * A1[k]) * dL_db + A1[k] * A1[l] * dL_dc;
        float dA0[3], dA1[3];
        for (int k = 0; k < 3; ++k) {
            dA0[k] = 2.0f * u[k] * dL_da + v[k] * dL_db;
            dA1[k] = 2.0f * v[k] * dL_dc + u[k] * dL_db;
        }
        float dJ00 = 0.0f, dJ02 = 0.0f, dJ11 = 0.0f, dJ12 = 0.0f;
        for (int k = 0; k < 3; ++k) {
            dJ00 += view[k * 4] * dA0[k];
            dJ02 += view[k * 4 + 2] * dA0[k];
            dJ11 += view[k * 4 + 1] * dA1[k];
            dJ12 += view[k * 4 + 2] * dA1[k];
        }
        const float tz = 1.0f / pvz, tz2 = tz * tz, tz3 = tz2 * tz;
        const float dL_dtx = x_grad_mul * -focal_x * tz2 * dJ02;
        const float dL_dty = y_grad_mul * -focal_y * tz2 * dJ12;
        const float dL_dtz = -focal_x * tz2 * dJ00 - focal_y * tz2 * dJ11 + (2.0f * focal_x * txc) * tz3 * dJ02 +
                             (2.0f * focal_y * tyc) * tz3 * dJ12;
        float dmean[3];
        for (int k = 0; k < 3; ++k) dmean[k] = view[k * 4] * dL_dtx + view[k * 4 + 1] * dL_dty + view[k * 4 + 2] * dL_dtz;
        /* B.3: mean2D (NDC-scaled) -> mean3D */
        const float phx = proj[0] * px + proj[4] * py + proj[8] * pz + proj[12];
        const float phy = proj[1] * px + proj[5] * py + proj[9] * pz + proj[13];
        const float phw = proj[3] * px + proj[7] * py + proj[11] * pz + proj[15];
        const float m_w = 1.0f / (phw + 0.0000001f);
        const float mul1 = phx * m_w * m_w, mul2 = phy * m_w * m_w;
        const float d2x = dL_dmean2D[2 * i], d2y = dL_dmean2D[2 * i + 1];
        for (int k = 0; k < 3; ++k)
            dmean[k] += (proj[k * 4] * m_w - proj[k * 4 + 3] * mul1) * d2x + (proj[k * 4 + 1] * m_w - proj[k * 4 + 3] * mul2) * d2y;
        for (int k = 0; k < 3; ++k) dL_dmeans3d[3 * i + k] = dmean[k];
        /* B.4: Sigma = m m^T, m[a][k] = Rot[a][k] s_k */
        float dm[3][3], D[3][3];
        for (int a2 = 0; a2 < 3; ++a2)
            for (int k = 0; k < 3; ++k)
                dm[a2][k] = 2.0f * (Gs[a2][0] * m[0][k] + Gs[a2][1] * m[1][k] + Gs[a2][2] * m[2][k]);
        for (int k = 0; k < 3; ++k)
            dL_dscales[3 * i + k] = (dm[0][k] * Rm[0][k] + dm[1][k] * Rm[1][k] + dm[2][k] * Rm[2][k]) * scale_modifier;
        for (int a2 = 0; a2 < 3; ++a2) for (int k = 0; k < 3; ++k) D[a2][k] = dm[a2][k] * s[k];
        dL_drots[4 * i] = 2.0f * (-qz * D[0][1] + qy * D[0][2] + qz * D[1][0] - qx * D[1][2] - qy * D[2][0] + qx * D[2][1]);
        dL_drots[4 * i + 1] = 2.0f * (qy * D[0][1] + qz * D[0][2] + qy * D[1][0] - 2.0f * qx * D[1][1] - qr * D[1][2] +
                                      qz * D[2][0] + qr * D[2][1] - 2.0f * qx * D[2][2]);
        dL_drots[4 * i + 2] = 2.0f * (-2.0f * qy * D[0][0] + qx * D[0][1] + qr * D[0][2] + qx * D[1][0] + qz * D[1][2] -
                                      qr * D[2][0] + qz * D[2][1] - 2.0f * qy * D[2][2]);
        dL_drots[4 * i + 3] = 2.0f * (-2.0f * qz * D[0][0] - qr * D[0][1] + qx * D[0][2] + qr * D[1][0] - 2.0f * qz * D[1][1] +
                                      qy * D[1][2] + qx * D[2][0] + qy * D[2][1]);
    }
}
