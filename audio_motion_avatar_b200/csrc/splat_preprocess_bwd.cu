// Stage 4a backward: per-Gaussian chain from the screen-space sums of the blend backward to the rasteriser's
// inputs (upstream computeCov2DCUDA / preprocessCUDA / computeCov3D of backward.cu)
// and on through render_one's activations (src/models/renderer.py:532-533,547).
//
// One thread per (Gaussian set, Gaussian); the views that share a set (render_multi_view, renderer.py:431-445)
// are summed in registers, so no atomics and no T-fold expanded copies of the attributes.  The forward
// intermediates are recomputed from the inputs (cheaper than storing cov3D / cov2D: 56 B in, 56 B out per Gaussian
// plus 36 B of sums per view).
#include "common.cuh"

namespace {

struct PreBwdParams {
    int G, N, vps, W, H, activate;
    float scale_modifier;
    const float *means3d, *scales, *rots, *opac, *colors, *view, *proj, *tanfov;
    const float4 *geom;      // [F,N,3] float4: radius in g2.w
    const float *dsum;       // [F,N,12]
    float *d_means, *d_scales, *d_rots, *d_opac, *d_colors;
};

__global__ void __launch_bounds__(256) preprocess_bwd_kernel(const PreBwdParams p)
{
    const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (int64_t)p.G * p.N) return;
    const int set = (int)(gid / p.N);
    const int i = (int)(gid - (int64_t)set * p.N);

    const float px = p.means3d[3 * gid], py = p.means3d[3 * gid + 1], pz = p.means3d[3 * gid + 2];
    const float raw_s[3] = {p.scales[3 * gid], p.scales[3 * gid + 1], p.scales[3 * gid + 2]};
    float act_s[3], s[3], ds_act[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        if (p.activate) {
            const float e = expf(raw_s[k] - 3.9f);
            act_s[k] = fminf(e, 0.1f);
            ds_act[k] = e < 0.1f ? e : 0.0f;          // d min(exp(r - 3.9), 0.1) / dr
        } else {
            act_s[k] = raw_s[k];
            ds_act[k] = 1.0f;
        }
        s[k] = p.scale_modifier * act_s[k];
    }
    const float raw_o = p.opac[gid];
    const float op = p.activate ? 1.0f / (1.0f + expf(-raw_o)) : raw_o;
    const float dop_act = p.activate ? op * (1.0f - op) : 1.0f;
    float col_pass[3] = {1.0f, 1.0f, 1.0f};
    if (p.activate) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float c = p.colors[3 * gid + k];
            col_pass[k] = (c >= 0.0f && c <= 1.0f) ? 1.0f : 0.0f;
        }
    }
    const float4 q = reinterpret_cast<const float4 *>(p.rots)[gid];
    const float qr = q.x, qx = q.y, qy = q.z, qz = q.w;
    float Rm[3][3];
    Rm[0][0] = 1.0f - 2.0f * (qy * qy + qz * qz); Rm[0][1] = 2.0f * (qx * qy - qr * qz); Rm[0][2] = 2.0f * (qx * qz + qr * qy);
    Rm[1][0] = 2.0f * (qx * qy + qr * qz); Rm[1][1] = 1.0f - 2.0f * (qx * qx + qz * qz); Rm[1][2] = 2.0f * (qy * qz - qr * qx);
    Rm[2][0] = 2.0f * (qx * qz - qr * qy); Rm[2][1] = 2.0f * (qy * qz + qr * qx); Rm[2][2] = 1.0f - 2.0f * (qx * qx + qy * qy);
    float m[3][3], S[3][3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int k = 0; k < 3; ++k) m[a][k] = Rm[a][k] * s[k];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) S[a][b] = m[a][0] * m[b][0] + m[a][1] * m[b][1] + m[a][2] * m[b][2];

    float dmean[3] = {0.f, 0.f, 0.f}, dop = 0.0f, dcol[3] = {0.f, 0.f, 0.f};
    float Gs[3][3];                                   // dL/dSigma summed over views (Sigma does not depend on the view)
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) Gs[a][b] = 0.0f;

    for (int vw = 0; vw < p.vps; ++vw) {
        const int f = set * p.vps + vw;
        const int64_t fi = (int64_t)f * p.N + i;
        if (!(p.geom[3 * fi + 2].w > 0.0f)) continue;                  // culled in this view
        const float4 d0 = reinterpret_cast<const float4 *>(p.dsum)[3 * fi];
        const float4 d1 = reinterpret_cast<const float4 *>(p.dsum)[3 * fi + 1];
        const float d2 = p.dsum[12 * fi + 8];
        const float *view = p.view + 16 * f, *proj = p.proj + 16 * f;
        const float tanfovx = p.tanfov[2 * f], tanfovy = p.tanfov[2 * f + 1];
        dop += d1.y;
        dcol[0] += d1.z; dcol[1] += d1.w; dcol[2] += d2;
        const float pvx = view[0] * px + view[4] * py + view[8] * pz + view[12];
        const float pvy = view[1] * px + view[5] * py + view[9] * pz + view[13];
        const float pvz = view[2] * px + view[6] * py + view[10] * pz + view[14];
        const float focal_x = (float)p.W / (2.0f * tanfovx), focal_y = (float)p.H / (2.0f * tanfovy);
        const float limx = 1.3f * tanfovx, limy = 1.3f * tanfovy;
        const float txtz = pvx / pvz, tytz = pvy / pvz;
        const float txc = fminf(limx, fmaxf(-limx, txtz)) * pvz;
        const float tyc = fminf(limy, fmaxf(-limy, tytz)) * pvz;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.0f : 1.0f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.0f : 1.0f;
        const float J00 = focal_x / pvz, J02 = -(focal_x * txc) / (pvz * pvz);
        const float J11 = focal_y / pvz, J12 = -(focal_y * tyc) / (pvz * pvz);
        float A0[3], A1[3], u[3], v[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            A0[k] = J00 * view[k * 4] + J02 * view[k * 4 + 2];
            A1[k] = J11 * view[k * 4 + 1] + J12 * view[k * 4 + 2];
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            u[k] = A0[0] * S[0][k] + A0[1] * S[1][k] + A0[2] * S[2][k];
            v[k] = A1[0] * S[0][k] + A1[1] * S[1][k] + A1[2] * S[2][k];
        }
        const float a = (u[0] * A0[0] + u[1] * A0[1] + u[2] * A0[2]) + 0.3f;
        const float b = u[0] * A1[0] + u[1] * A1[1] + u[2] * A1[2];
        const float c = (v[0] * A1[0] + v[1] * A1[1] + v[2] * A1[2]) + 0.3f;
        const float X = d0.z, Y = d0.w, Z = d1.x;
        const float denom = a * c - b * b;
        const float denom2inv = 1.0f / (denom * denom + 0.0000001f);
        float dL_da = 0.0f, dL_db = 0.0f, dL_dc = 0.0f;
        if (denom2inv != 0.0f) {
            dL_da = denom2inv * (-c * c * X + 2.0f * b * c * Y + (denom - a * c) * Z);
            dL_dc = denom2inv * (-a * a * Z + 2.0f * a * b * Y + (denom - a * c) * X);
            dL_db = denom2inv * 2.0f * (b * c * X - (denom + 2.0f * b * b) * Y + a * b * Z);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k)
#pragma unroll
            for (int l = 0; l < 3; ++l)
                Gs[k][l] += A0[k] * A0[l] * dL_da + 0.5f * (A0[k] * A1[l] + A0[l] * A1[k]) * dL_db + A1[k] * A1[l] * dL_dc;
        float dJ00 = 0.0f, dJ02 = 0.0f, dJ11 = 0.0f, dJ12 = 0.0f;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float dA0 = 2.0f * u[k] * dL_da + v[k] * dL_db;
            const float dA1 = 2.0f * v[k] * dL_dc + u[k] * dL_db;
            dJ00 += view[k * 4] * dA0;
            dJ02 += view[k * 4 + 2] * dA0;
            dJ11 += view[k * 4 + 1] * dA1;
            dJ12 += view[k * 4 + 2] * dA1;
        }
        const float tz = 1.0f / pvz, tz2 = tz * tz, tz3 = tz2 * tz;
        const float dL_dtx = x_grad_mul * -focal_x * tz2 * dJ02;
        const float dL_dty = y_grad_mul * -focal_y * tz2 * dJ12;
        const float dL_dtz = -focal_x * tz2 * dJ00 - focal_y * tz2 * dJ11 + (2.0f * focal_x * txc) * tz3 * dJ02 +
                             (2.0f * focal_y * tyc) * tz3 * dJ12;
        const float phx = proj[0] * px + proj[4] * py + proj[8] * pz + proj[12];
        const float phy = proj[1] * px + proj[5] * py + proj[9] * pz + proj[13];
        const float phw = proj[3] * px + proj[7] * py + proj[11] * pz + proj[15];
        const float m_w = 1.0f / (phw + 0.0000001f);
        const float mul1 = phx * m_w * m_w, mul2 = phy * m_w * m_w;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            dmean[k] += view[k * 4] * dL_dtx + view[k * 4 + 1] * dL_dty + view[k * 4 + 2] * dL_dtz;
            dmean[k] += (proj[k * 4] * m_w - proj[k * 4 + 3] * mul1) * d0.x + (proj[k * 4 + 1] * m_w - proj[k * 4 + 3] * mul2) * d0.y;
        }
    }
    // Sigma = m m^T with m[a][k] = Rot[a][k] s_k
    float D[3][3], dscale[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float acc = 0.0f;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float dm = 2.0f * (Gs[a][0] * m[0][k] + Gs[a][1] * m[1][k] + Gs[a][2] * m[2][k]);
            acc += dm * Rm[a][k];
            D[a][k] = dm * s[k];
        }
        dscale[k] = acc * p.scale_modifier * ds_act[k];
    }
    const float dq_r = 2.0f * (-qz * D[0][1] + qy * D[0][2] + qz * D[1][0] - qx * D[1][2] - qy * D[2][0] + qx * D[2][1]);
    const float dq_x = 2.0f * (qy * D[0][1] + qz * D[0][2] + qy * D[1][0] - 2.0f * qx * D[1][1] - qr * D[1][2] + qz * D[2][0] +
                               qr * D[2][1] - 2.0f * qx * D[2][2]);
    const float dq_y = 2.0f * (-2.0f * qy * D[0][0] + qx * D[0][1] + qr * D[0][2] + qx * D[1][0] + qz * D[1][2] - qr * D[2][0] +
                               qz * D[2][1] - 2.0f * qy * D[2][2]);
    const float dq_z = 2.0f * (-2.0f * qz * D[0][0] - qr * D[0][1] + qx * D[0][2] + qr * D[1][0] - 2.0f * qz * D[1][1] +
                               qy * D[1][2] + qx * D[2][0] + qy * D[2][1]);
    p.d_means[3 * gid] = dmean[0]; p.d_means[3 * gid + 1] = dmean[1]; p.d_means[3 * gid + 2] = dmean[2];
    p.d_scales[3 * gid] = dscale[0]; p.d_scales[3 * gid + 1] = dscale[1]; p.d_scales[3 * gid + 2] = dscale[2];
    reinterpret_cast<float4 *>(p.d_rots)[gid] = make_float4(dq_r, dq_x, dq_y, dq_z);
    p.d_opac[gid] = dop * dop_act;
    p.d_colors[3 * gid] = dcol[0] * col_pass[0]; p.d_colors[3 * gid + 1] = dcol[1] * col_pass[1];
    p.d_colors[3 * gid + 2] = dcol[2] * col_pass[2];
}

}  // namespace

extern "C" int amav_splat_preprocess_backward(const AmavSplatDims *d, const float *means3d, const float *scales,
                                              const float *rots, const float *opacities, const float *colors,
                                              const float *view, const float *proj, const float *tanfov,
                                              const float *geom, const float *dsum, float *d_means3d, float *d_scales,
                                              float *d_rots, float *d_opacities, float *d_colors, void *stream)
{
    AMAV_CHECK_ARG(d && d->frames >= 0 && d->num_gaussians >= 0 && d->views_per_set >= 1, "bad dims");
    AMAV_CHECK_ARG(d->frames % d->views_per_set == 0, "frames must be a multiple of views_per_set");
    const int sets = d->frames / d->views_per_set;
    const int64_t total = (int64_t)sets * d->num_gaussians;
    if (total == 0) return AMAV_OK;
    AMAV_CHECK_ARG(means3d && scales && rots && opacities && colors && view && proj && tanfov && geom && dsum, "null input");
    AMAV_CHECK_ARG(d_means3d && d_scales && d_rots && d_opacities && d_colors, "null output");
    AMAV_CHECK_ARG((((uintptr_t)dsum | (uintptr_t)geom | (uintptr_t)rots | (uintptr_t)d_rots) & 15) == 0, "16-byte alignment");
    PreBwdParams p;
    p.G = sets; p.N = d->num_gaussians; p.vps = d->views_per_set; p.W = d->width; p.H = d->height;
    p.activate = d->activate; p.scale_modifier = d->scale_modifier;
    p.means3d = means3d; p.scales = scales; p.rots = rots; p.opac = opacities; p.colors = colors;
    p.view = view; p.proj = proj; p.tanfov = tanfov;
    p.geom = reinterpret_cast<const float4 *>(geom); p.dsum = dsum;
    p.d_means = d_means3d; p.d_scales = d_scales; p.d_rots = d_rots; p.d_opac = d_opacities; p.d_colors = d_colors;
    preprocess_bwd_kernel<<<amav_div_up(total, 256), 256, 0, (cudaStream_t)stream>>>(p);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
