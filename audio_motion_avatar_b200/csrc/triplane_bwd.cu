// Stages 1+2 backward: gradients of the triplane query + linear decode (autograd of F.grid_sample, the five
// nn.Linear heads, normalize and sigmoid in the reference: src/models/renderer.py:292-317, :165-171, :319-346).
//
// Mirrors the forward's project-then-sample factorisation:
//   sample_decode_bwd   per site: dL/d(gaussians) -> dL/d(raw head outputs) -> scatter into the PROJECTED planes'
//                       gradient (12 taps x 64 B, float4 reductions into an L2-resident 0.8 MB/frame buffer),
//                       dL/dpoints (direct + xyz columns of the heads + bilinear position derivative), dL/dtransl,
//                       and the xyz-column / bias part of dL/dW;
//   project_bwd_tokens  dL/dtokens = W_feat^T . dL/dproj: streams 25 MB/frame OUT once (HBM-bound, the mirror of
//                       project_kernel);
//   project_bwd_weight  dL/dW_feat = dL/dproj . tokens^T: streams the tokens IN once more.
// The reference's backward instead re-reads / writes the 614 MB/frame sampled-feature tensor several times and
// scatters 1536 channels per tap with scalar atomics (grid_sampler_2d_backward).
#include "common.cuh"

namespace {

__constant__ int c_slot_row_b[16] = {3, 4, 5, 6, 0, 1, 2, 10, 7, 8, 9, -1, 11, 12, 13, -1};

struct SampleBwdParams {
    int F, N, R;
    float radius;
    const float4 *proj;
    const float *points, *weight, *bias;
    int ldw;
    const float *g_xyz, *g_scale, *g_rot, *g_opacity, *g_color;   // any may be NULL (= zero)
    float4 *d_proj;        // [F,3,R*R,4] float4, zeroed by the caller
    float *d_points;       // [F,N,3]
    float *d_transl;       // [F,3], zeroed by the caller
    float *d_wb;           // [14,4]: columns x,y,z of dL/dW and dL/dbias, zeroed by the caller
};

__device__ __forceinline__ void tap_coords_b(float g, int R, int &i0, float &w1)
{
    const float i = ((g + 1.0f) * (float)R - 1.0f) * 0.5f;
    const float fl = floorf(i);
    i0 = (int)fl;
    w1 = i - fl;
}

__global__ void __launch_bounds__(256) sample_decode_bwd_kernel(const SampleBwdParams p)
{
    __shared__ float4 s_wb[16];
    __shared__ float s_dwb[64];        // [slot][x,y,z,bias]
    __shared__ float s_dt[3];
    if (threadIdx.x < 16) {
        const int row = c_slot_row_b[threadIdx.x];
        s_wb[threadIdx.x] = row < 0 ? make_float4(0.f, 0.f, 0.f, 0.f)
                                    : make_float4(p.weight[(int64_t)row * p.ldw], p.weight[(int64_t)row * p.ldw + 1],
                                                  p.weight[(int64_t)row * p.ldw + 2], p.bias[row]);
    }
    if (threadIdx.x < 64) s_dwb[threadIdx.x] = 0.0f;
    if (threadIdx.x < 3) s_dt[threadIdx.x] = 0.0f;
    __syncthreads();
    const int f = blockIdx.y;
    const int site = blockIdx.x * 64 + (threadIdx.x >> 2);
    const int lane4 = threadIdx.x & 3;
    const bool live = site < p.N;
    const int64_t gid = (int64_t)f * p.N + (live ? site : 0);
    const float px = p.points[3 * gid], py = p.points[3 * gid + 1], pz = p.points[3 * gid + 2];
    const float qx = px / p.radius, qy = py / p.radius, qz = pz / p.radius;
    const float gx = fminf(fmaxf(qx, -1.0f), 1.0f), gy = fminf(fmaxf(qy, -1.0f), 1.0f), gz = fminf(fmaxf(qz, -1.0f), 1.0f);
    const int R = p.R, R2 = R * R;
    const float ua[3] = {gx, gx, gy}, va[3] = {gy, gz, gz};
    const float4 *base = p.proj + ((int64_t)f * 3 * R2) * 4 + lane4;
    float4 *dbase = p.d_proj + ((int64_t)f * 3 * R2) * 4 + lane4;

    // pass 1: the forward value of this lane's slot group (needed by normalize / sigmoid backward)
    int x0[3], y0[3];
    float wx1[3], wy1[3];
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 tex[3][4];
#pragma unroll
    for (int pl = 0; pl < 3; ++pl) {
        tap_coords_b(ua[pl], R, x0[pl], wx1[pl]);
        tap_coords_b(va[pl], R, y0[pl], wy1[pl]);
        const float4 *plane = base + (int64_t)pl * R2 * 4;
        float4 pacc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int dx = t & 1, dy = t >> 1;
            const int xi = x0[pl] + dx, yi = y0[pl] + dy;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (xi >= 0 && xi < R && yi >= 0 && yi < R) v = __ldg(plane + (int64_t)(yi * R + xi) * 4);
            tex[pl][t] = v;
            const float w = (dx ? wx1[pl] : 1.0f - wx1[pl]) * (dy ? wy1[pl] : 1.0f - wy1[pl]);
            pacc.x = fmaf(v.x, w, pacc.x); pacc.y = fmaf(v.y, w, pacc.y);
            pacc.z = fmaf(v.z, w, pacc.z); pacc.w = fmaf(v.w, w, pacc.w);
        }
        acc.x += pacc.x; acc.y += pacc.y; acc.z += pacc.z; acc.w += pacc.w;
    }
    float v[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float4 wb = s_wb[lane4 * 4 + k];
        v[k] = v[k] + (fmaf(wb.x, px, fmaf(wb.y, py, wb.z * pz)) + wb.w);
    }
    // dL/d(raw head outputs) of this lane's slots
    float dv[4] = {0.f, 0.f, 0.f, 0.f};
    float dxyz[3] = {0.f, 0.f, 0.f};
    if (live) {
        if (lane4 == 0) {
            if (p.g_rot) {
                const float4 g = reinterpret_cast<const float4 *>(p.g_rot)[gid];
                const float nrm = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3]);
                const float n = fmaxf(nrm, 1e-12f);
                const float y0n = v[0] / n, y1n = v[1] / n, y2n = v[2] / n, y3n = v[3] / n;
                // y = v / max(|v|, eps): below eps the denominator is constant
                const float dot = nrm > 1e-12f ? (y0n * g.x + y1n * g.y + y2n * g.z + y3n * g.w) : 0.0f;
                dv[0] = (g.x - y0n * dot) / n; dv[1] = (g.y - y1n * dot) / n;
                dv[2] = (g.z - y2n * dot) / n; dv[3] = (g.w - y3n * dot) / n;
            }
        } else if (lane4 == 1) {
            if (p.g_xyz) {
                dxyz[0] = p.g_xyz[3 * gid]; dxyz[1] = p.g_xyz[3 * gid + 1]; dxyz[2] = p.g_xyz[3 * gid + 2];
                dv[0] = dxyz[0]; dv[1] = dxyz[1]; dv[2] = dxyz[2];
            }
            if (p.g_opacity) dv[3] = p.g_opacity[gid];
        } else if (lane4 == 2) {
            if (p.g_scale) { dv[0] = p.g_scale[3 * gid]; dv[1] = p.g_scale[3 * gid + 1]; dv[2] = p.g_scale[3 * gid + 2]; }
        } else if (p.g_color) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const float c = 1.0f / (1.0f + expf(-v[k]));
                dv[k] = p.g_color[3 * gid + k] * c * (1.0f - c);
            }
        }
    }
    // pass 2: scatter into the projected planes' gradient + position derivative
    float dpx = dxyz[0], dpy = dxyz[1], dpz = dxyz[2];      // xyz = points + offset + transl
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float4 wb = s_wb[lane4 * 4 + k];
        dpx = fmaf(wb.x, dv[k], dpx); dpy = fmaf(wb.y, dv[k], dpy); dpz = fmaf(wb.z, dv[k], dpz);
    }
    const float half_r = 0.5f * (float)R / p.radius;
    const float mx = (qx > -1.0f && qx < 1.0f) ? half_r : 0.0f;      // d clamp(p/radius) / dp
    const float my = (qy > -1.0f && qy < 1.0f) ? half_r : 0.0f;
    const float mz = (qz > -1.0f && qz < 1.0f) ? half_r : 0.0f;
    const bool any_dv = dv[0] != 0.0f || dv[1] != 0.0f || dv[2] != 0.0f || dv[3] != 0.0f;
#pragma unroll
    for (int pl = 0; pl < 3; ++pl) {
        float4 *dplane = dbase + (int64_t)pl * R2 * 4;
        float s[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int dx = t & 1, dy = t >> 1;
            const int xi = x0[pl] + dx, yi = y0[pl] + dy;
            const float4 tv = tex[pl][t];
            s[t] = dv[0] * tv.x + dv[1] * tv.y + dv[2] * tv.z + dv[3] * tv.w;
            if (any_dv && xi >= 0 && xi < R && yi >= 0 && yi < R) {
                const float w = (dx ? wx1[pl] : 1.0f - wx1[pl]) * (dy ? wy1[pl] : 1.0f - wy1[pl]);
                atomicAdd(dplane + (int64_t)(yi * R + xi) * 4, make_float4(w * dv[0], w * dv[1], w * dv[2], w * dv[3]));
            }
        }
        const float du = (1.0f - wy1[pl]) * (s[1] - s[0]) + wy1[pl] * (s[3] - s[2]);     // d/d(ix)
        const float dw = (1.0f - wx1[pl]) * (s[2] - s[0]) + wx1[pl] * (s[3] - s[1]);     // d/d(iy)
        if (pl == 0) { dpx = fmaf(du, mx, dpx); dpy = fmaf(dw, my, dpy); }
        else if (pl == 1) { dpx = fmaf(du, mx, dpx); dpz = fmaf(dw, mz, dpz); }
        else { dpy = fmaf(du, my, dpy); dpz = fmaf(dw, mz, dpz); }
    }
    // the four lanes of a site hold partial position gradients
    dpx += __shfl_xor_sync(0xffffffffu, dpx, 1); dpx += __shfl_xor_sync(0xffffffffu, dpx, 2);
    dpy += __shfl_xor_sync(0xffffffffu, dpy, 1); dpy += __shfl_xor_sync(0xffffffffu, dpy, 2);
    dpz += __shfl_xor_sync(0xffffffffu, dpz, 1); dpz += __shfl_xor_sync(0xffffffffu, dpz, 2);
    if (live && lane4 == 0 && p.d_points) {
        p.d_points[3 * gid] = dpx; p.d_points[3 * gid + 1] = dpy; p.d_points[3 * gid + 2] = dpz;
    }
    // dL/dW[:, 0:3], dL/dbias and dL/dtransl: reduce over the 8 sites of the warp, then the CTA, then one atomic each
    float r[16];
#pragma unroll
    for (int k = 0; k < 4; ++k) { r[4 * k] = dv[k] * px; r[4 * k + 1] = dv[k] * py; r[4 * k + 2] = dv[k] * pz; r[4 * k + 3] = dv[k]; }
#pragma unroll
    for (int d = 4; d <= 16; d <<= 1) {
#pragma unroll
        for (int k = 0; k < 16; ++k) r[k] += __shfl_xor_sync(0xffffffffu, r[k], d);
#pragma unroll
        for (int k = 0; k < 3; ++k) dxyz[k] += __shfl_xor_sync(0xffffffffu, dxyz[k], d);
    }
    if ((threadIdx.x & 31) < 4) {
#pragma unroll
        for (int k = 0; k < 16; ++k) atomicAdd(&s_dwb[lane4 * 16 + k], r[k]);
        if (lane4 == 1) { atomicAdd(&s_dt[0], dxyz[0]); atomicAdd(&s_dt[1], dxyz[1]); atomicAdd(&s_dt[2], dxyz[2]); }
    }
    __syncthreads();
    if (threadIdx.x < 64) {
        const int row = c_slot_row_b[threadIdx.x >> 2];
        if (row >= 0 && s_dwb[threadIdx.x] != 0.0f) atomicAdd(p.d_wb + row * 4 + (threadIdx.x & 3), s_dwb[threadIdx.x]);
    } else if (threadIdx.x < 67 && p.d_transl) {
        atomicAdd(p.d_transl + 3 * f + (threadIdx.x - 64), s_dt[threadIdx.x - 64]);
    }
}

// ---------------------------------------------------------------------------------------------
// dL/dtokens[f][c][p*R2 + t] = sum_slot W[row(slot)][3 + p*C + c] * dproj[f][p][t][slot]
constexpr int PB_THREADS_T = 256;
constexpr int PB_TEX = 128;
constexpr int PB_KGROUPS = 8;

__global__ void __launch_bounds__(PB_THREADS_T, 2)
project_bwd_tokens_kernel(int C, int R2, const float *__restrict__ d_proj, const float *__restrict__ weight, int ldw,
                          float *__restrict__ d_tokens, int64_t frame_stride, int accumulate)
{
    extern __shared__ float4 s_dyn_b[];
    float *s_w = reinterpret_cast<float *>(s_dyn_b);                 // [C][16]
    const int tiles_per_plane = R2 / PB_TEX;
    const int tile = blockIdx.x % tiles_per_plane;
    const int pl = (blockIdx.x / tiles_per_plane) % 3;
    const int f = blockIdx.x / (3 * tiles_per_plane);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < C * 16; i += PB_THREADS_T) {
        const int c = i >> 4, s = i & 15;
        const int row = c_slot_row_b[s];
        s_w[i] = row < 0 ? 0.0f : weight[(int64_t)row * ldw + 3 + pl * C + c];
    }
    // this lane's 4 texels x 16 slots
    float dp[4][16];
    const float4 *src = reinterpret_cast<const float4 *>(d_proj + (((int64_t)f * 3 + pl) * R2 + tile * PB_TEX + lane * 4) * 16);
#pragma unroll
    for (int t = 0; t < 4; ++t)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float4 v = __ldg(src + t * 4 + q);
            dp[t][4 * q] = v.x; dp[t][4 * q + 1] = v.y; dp[t][4 * q + 2] = v.z; dp[t][4 * q + 3] = v.w;
        }
    __syncthreads();
    const int kper = C / PB_KGROUPS;
    const int c0 = warp * kper;
    float *dst = d_tokens + (int64_t)f * frame_stride + (int64_t)c0 * (3 * R2) + (int64_t)pl * R2 + tile * PB_TEX + lane * 4;
    for (int c = 0; c < kper; ++c) {
        const float4 *wp = reinterpret_cast<const float4 *>(s_w + (size_t)(c0 + c) * 16);
        const float4 w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3];
        const float w[14] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x, w2.y, w2.z, w3.x, w3.y, w3.z};
        const int sl[14] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 13, 14};
        float o[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < 14; ++k)
#pragma unroll
            for (int t = 0; t < 4; ++t) o[t] = fmaf(w[k], dp[t][sl[k]], o[t]);
        float4 *d4 = reinterpret_cast<float4 *>(dst + (int64_t)c * (3 * R2));
        if (accumulate) {
            const float4 old = *d4;
            o[0] += old.x; o[1] += old.y; o[2] += old.z; o[3] += old.w;
        }
        __stcs(d4, make_float4(o[0], o[1], o[2], o[3]));
    }
}

// dL/dW[row(slot)][3 + p*C + c] += sum_t dproj[f][p][t][slot] * tokens[f][c][p*R2 + t]
__global__ void __launch_bounds__(PB_THREADS_T, 2)
project_bwd_weight_kernel(int C, int R2, const float *__restrict__ d_proj, const float *__restrict__ tokens,
                          int64_t frame_stride, float *__restrict__ d_weight, int ldw)
{
    const int tiles_per_plane = R2 / PB_TEX;
    const int tile = blockIdx.x % tiles_per_plane;
    const int pl = (blockIdx.x / tiles_per_plane) % 3;
    const int f = blockIdx.x / (3 * tiles_per_plane);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float dp[4][16];
    const float4 *src = reinterpret_cast<const float4 *>(d_proj + (((int64_t)f * 3 + pl) * R2 + tile * PB_TEX + lane * 4) * 16);
#pragma unroll
    for (int t = 0; t < 4; ++t)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float4 v = __ldg(src + t * 4 + q);
            dp[t][4 * q] = v.x; dp[t][4 * q + 1] = v.y; dp[t][4 * q + 2] = v.z; dp[t][4 * q + 3] = v.w;
        }
    const int kper = C / PB_KGROUPS;
    const int c0 = warp * kper;
    const float *tsrc = tokens + (int64_t)f * frame_stride + (int64_t)c0 * (3 * R2) + (int64_t)pl * R2 + tile * PB_TEX + lane * 4;
    const bool up16 = lane & 16, up8 = lane & 8, up4 = lane & 4, up2 = lane & 2;
    // after the split butterfly lane L holds slot ((L>>4)&1)*8 + ((L>>3)&1)*4 + ((L>>2)&1)*2 + ((L>>1)&1)
    const int my_slot = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
    const int out_row = c_slot_row_b[my_slot];
    constexpr int U = 4;
    for (int cb = 0; cb < kper; cb += U) {
        float4 tv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) tv[u] = __ldcs(reinterpret_cast<const float4 *>(tsrc + (int64_t)(cb + u) * (3 * R2)));
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float a[16];
#pragma unroll
            for (int s = 0; s < 16; ++s)
                a[s] = fmaf(tv[u].x, dp[0][s], fmaf(tv[u].y, dp[1][s], fmaf(tv[u].z, dp[2][s], tv[u].w * dp[3][s])));
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const float send = up16 ? a[q] : a[q + 8];
                const float keep = up16 ? a[q + 8] : a[q];
                a[q] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float send = up8 ? a[q] : a[q + 4];
                const float keep = up8 ? a[q + 4] : a[q];
                a[q] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
            }
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const float send = up4 ? a[q] : a[q + 2];
                const float keep = up4 ? a[q + 2] : a[q];
                a[q] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
            }
            {
                const float send = up2 ? a[0] : a[1];
                const float keep = up2 ? a[1] : a[0];
                a[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
            }
            a[0] += __shfl_xor_sync(0xffffffffu, a[0], 1);
            if ((lane & 1) == 0 && out_row >= 0)
                atomicAdd(d_weight + (int64_t)out_row * ldw + 3 + pl * C + c0 + cb + u, a[0]);
        }
    }
}

}  // namespace

extern "C" int amav_sample_decode_backward(int F, int N, int C, int R, float radius, const float *proj,
                                           const float *points, const float *weight, const float *bias,
                                           const float *g_xyz, const float *g_scale, const float *g_rot,
                                           const float *g_opacity, const float *g_color, float *d_proj,
                                           float *d_points, float *d_transl, float *d_wb, void *stream)
{
    AMAV_CHECK_ARG(F >= 0 && F <= 65535 && N >= 0 && C > 0 && R > 0 && radius > 0.0f, "bad dims");
    if ((int64_t)F * N == 0) return AMAV_OK;
    AMAV_CHECK_ARG(proj && points && weight && bias && d_proj && d_wb, "null buffer");
    AMAV_CHECK_ARG(((uintptr_t)d_proj & 15) == 0 && (!g_rot || ((uintptr_t)g_rot & 15) == 0), "16-byte alignment");
    SampleBwdParams p;
    p.F = F; p.N = N; p.R = R; p.radius = radius;
    p.proj = reinterpret_cast<const float4 *>(proj); p.points = points; p.weight = weight; p.bias = bias;
    p.ldw = 3 + 3 * C;
    p.g_xyz = g_xyz; p.g_scale = g_scale; p.g_rot = g_rot; p.g_opacity = g_opacity; p.g_color = g_color;
    p.d_proj = reinterpret_cast<float4 *>(d_proj); p.d_points = d_points; p.d_transl = d_transl; p.d_wb = d_wb;
    dim3 grid(amav_div_up(N, 64), F);
    sample_decode_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}

extern "C" int amav_triplane_project_backward(int F, int C, int R, const float *d_proj, const float *tokens,
                                              int64_t token_frame_stride, const float *weight, float *d_tokens,
                                              int64_t d_token_frame_stride, int accumulate_tokens, float *d_weight,
                                              void *stream_)
{
    AMAV_CHECK_ARG(F >= 0 && C > 0 && R > 0, "bad dims");
    AMAV_CHECK_ARG((R * R) % PB_TEX == 0, "R*R must be a multiple of 128");
    AMAV_CHECK_ARG(C % (PB_KGROUPS * 8) == 0, "C must be a multiple of 64");
    if (F == 0) return AMAV_OK;
    AMAV_CHECK_ARG(d_proj && weight, "null buffer");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int64_t grid = (int64_t)F * 3 * (R * R / PB_TEX);
    AMAV_CHECK_ARG(grid < ((int64_t)1 << 31), "too many tiles");
    if (d_tokens) {
        AMAV_CHECK_ARG(d_token_frame_stride >= (int64_t)C * 3 * R * R && (d_token_frame_stride & 3) == 0 &&
                       ((uintptr_t)d_tokens & 15) == 0, "bad d_tokens stride / alignment");
        const size_t smem = (size_t)C * 16 * sizeof(float);
        AMAV_CHECK_ARG(smem <= 100 * 1024, "C too large for the shared-memory weight tile");
        amav_ensure_dyn_smem((const void *)project_bwd_tokens_kernel, (int)smem);
        project_bwd_tokens_kernel<<<(unsigned)grid, PB_THREADS_T, smem, stream>>>(C, R * R, d_proj, weight, 3 + 3 * C, d_tokens,
                                                                                 d_token_frame_stride, accumulate_tokens);
        AMAV_CHECK_LAUNCH();
    }
    if (d_weight) {
        AMAV_CHECK_ARG(tokens && token_frame_stride >= (int64_t)C * 3 * R * R && (token_frame_stride & 3) == 0 &&
                       ((uintptr_t)tokens & 15) == 0, "bad tokens stride / alignment");
        project_bwd_weight_kernel<<<(unsigned)grid, PB_THREADS_T, 0, stream>>>(C, R * R, d_proj, tokens, token_frame_stride,
                                                                              d_weight, 3 + 3 * C);
        AMAV_CHECK_LAUNCH();
    }
    return AMAV_OK;
}
