// Stages 1+2: triplane query and linear decode (reference src/models/renderer.py:292-317 sample_from_triplane,
// :165-171 the five nn.Linear heads, :319-346 construct_gaussians).
//
// The decode is linear in the sampled features, so decode(bilinear(T)) == bilinear(decode(T)):
//   1. triplane_project: every texel of every plane is pushed through the decoder heads once
//      (a [R*R x C] . [C x 14] product per plane, packed fp32 FMA, HBM-bound: it reads the 25 MB/frame
//      triplane exactly once and writes 0.8 MB/frame of 16-wide projected texels; persistent CTAs per plane);
//   2. sample_decode: per Gaussian site, 3 planes x 4 bilinear taps of 64-byte projected texels
//      (L2-resident), the xyz/bias terms and construct_gaussians' activations.
// The reference instead materialises [F,N,1536] sampled features (614 MB/frame) and reads them 5x.
//
// Projected texel layout (16 floats, one float4 per lane of a 4-lane group in sample_decode):
//   slot 0..3  rotation (w,x,y,z)      decoder rows 3..6
//   slot 4..6  xyz offset, slot 7 opacity   rows 0..2, 10
//   slot 8..10 scaling, slot 11 = 0         rows 7..9
//   slot 12..14 shs,    slot 15 = 0         rows 11..13
#include "common.cuh"

namespace {

__constant__ int c_slot_row[16] = {3, 4, 5, 6, 0, 1, 2, 10, 7, 8, 9, -1, 11, 12, 13, -1};

constexpr int PJ_THREADS = 256;
constexpr int PJ_TEX = 128;            // texels per CTA (32 lanes x float4)
constexpr int PJ_KGROUPS = 8;          // one warp per K slice

// smem: w[C][16] (slot-major per channel) then, after the main loop, red[8][128][16] reuses/extends it
__global__ void __launch_bounds__(PJ_THREADS, 2)
project_kernel(int F, int C, int R2, const float *__restrict__ tokens, int64_t frame_stride,
               const float *__restrict__ weight, int ldw, float *__restrict__ proj)
{
    extern __shared__ float4 s_dyn[];
    float *s_w = reinterpret_cast<float *>(s_dyn);                   // [C][16]
    float *s_red = s_w + (size_t)C * 16;                             // [8][14][128]
    // persistent CTAs: blockIdx.y = plane (its decoder columns stay in shared memory), blockIdx.x strides over the
    // (frame, 128-texel tile) items of that plane
    const int tiles_per_plane = R2 / PJ_TEX;
    const int p = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    for (int i = threadIdx.x; i < C * 16; i += PJ_THREADS) {
        const int c = i >> 4, s = i & 15;
        const int row = c_slot_row[s];
        s_w[i] = row < 0 ? 0.0f : weight[(int64_t)row * ldw + 3 + p * C + c];
    }
    __syncthreads();

    const int kper = C / PJ_KGROUPS;
    const int c0 = warp * kper;
    for (int item = blockIdx.x; item < F * tiles_per_plane; item += gridDim.x) {
    const int f = item / tiles_per_plane, tile = item - f * tiles_per_plane;
    const float *src = tokens + (int64_t)f * frame_stride + (int64_t)c0 * (3 * R2) + (int64_t)p * R2 + tile * PJ_TEX + lane * 4;
    // packed FP32 (sm_100 FFMA2): one instruction per (texel pair, output); each half is an IEEE fma, so the bits are
    // those of the scalar loop
    float2 acc[2][14];
#pragma unroll
    for (int t = 0; t < 2; ++t)
#pragma unroll
        for (int o = 0; o < 14; ++o) acc[t][o] = make_float2(0.0f, 0.0f);

    constexpr int U = 8;
    for (int cb = 0; cb < kper; cb += U) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            v[u] = __ldcs(reinterpret_cast<const float4 *>(src + (int64_t)(cb + u) * (3 * R2)));
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const float4 *wp = reinterpret_cast<const float4 *>(s_w + (size_t)(c0 + cb + u) * 16);
            const float4 w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3];
            const float w[14] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x, w2.y, w2.z, w3.x, w3.y, w3.z};
            const float2 t01 = make_float2(v[u].x, v[u].y), t23 = make_float2(v[u].z, v[u].w);
#pragma unroll
            for (int o = 0; o < 14; ++o) {
                const float2 ww = make_float2(w[o], w[o]);
                acc[0][o] = __ffma2_rn(t01, ww, acc[0][o]);
                acc[1][o] = __ffma2_rn(t23, ww, acc[1][o]);
            }
        }
    }
    // cross-warp (K slice) reduction through shared memory: red[warp][o][texel]
#pragma unroll
    for (int o = 0; o < 14; ++o)
        *reinterpret_cast<float4 *>(s_red + ((size_t)warp * 14 + o) * PJ_TEX + lane * 4) =
            make_float4(acc[0][o].x, acc[0][o].y, acc[1][o].x, acc[1][o].y);
    __syncthreads();
    // thread -> (texel, half of the 14 outputs)
    const int tex = threadIdx.x & (PJ_TEX - 1), half = threadIdx.x >> 7;
    float r[7];
#pragma unroll
    for (int o = 0; o < 7; ++o) r[o] = 0.0f;
    for (int g = 0; g < PJ_KGROUPS; ++g) {
#pragma unroll
        for (int o = 0; o < 7; ++o) r[o] += s_red[((size_t)g * 14 + half * 7 + o) * PJ_TEX + tex];
    }
    // compact index 0..13 -> slots (skip 11 and 15)
    float *out = proj + (((int64_t)f * 3 + p) * R2 + tile * PJ_TEX + tex) * 16;
    if (half == 0) {
        reinterpret_cast<float4 *>(out)[0] = make_float4(r[0], r[1], r[2], r[3]);
        out[4] = r[4]; out[5] = r[5]; out[6] = r[6];
    } else {
        out[7] = r[0];
        reinterpret_cast<float4 *>(out)[2] = make_float4(r[1], r[2], r[3], 0.0f);
        reinterpret_cast<float4 *>(out)[3] = make_float4(r[4], r[5], r[6], 0.0f);
    }
    __syncthreads();                                     // s_red is free for the next item
    }
}

// ---------------------------------------------------------------------------------------------
// sample_decode: 4 lanes per Gaussian site, lane l owns float4 slot group l of the projected texel.
struct SampleParams {
    int F, N, R;
    float radius;
    const float4 *proj;       // [F,3,R*R,4] float4
    const float *points;      // [F,N,3]
    const float *transl;      // [F,3]
    const float *weight, *bias;   // decoder [14,3+3C], [14]
    int ldw;
    float *xyz, *scale, *rot, *opacity, *color;
    const int32_t *order;     // processing order of the sites (a locality order fixed at set-up), or NULL
};


__device__ __forceinline__ void tap_coords(float g, int R, int &i0, float &w1)
{
    // grid_sample, align_corners=False: i = ((g+1)*R-1)/2
    const float i = ((g + 1.0f) * (float)R - 1.0f) * 0.5f;
    const float fl = floorf(i);
    i0 = (int)fl;
    w1 = i - fl;
}

#ifndef AMAV_SD_CTAS
#define AMAV_SD_CTAS 6            // 40 registers, 48 warps per SM: the kernel is latency-bound (0.93 -> 0.88 ms; 8 CTAs spill: 0.92)
#endif
__global__ void __launch_bounds__(256, AMAV_SD_CTAS) sample_decode_kernel(const SampleParams p)
{
    // per slot: (w_x, w_y, w_z, bias) of the decoder's xyz columns / bias
    __shared__ float4 s_wb[16];
    if (threadIdx.x < 16) {
        const int row = c_slot_row[threadIdx.x];
        s_wb[threadIdx.x] = row < 0 ? make_float4(0.f, 0.f, 0.f, 0.f)
                                    : make_float4(p.weight[(int64_t)row * p.ldw], p.weight[(int64_t)row * p.ldw + 1],
                                                  p.weight[(int64_t)row * p.ldw + 2], p.bias[row]);
    }
    __syncthreads();
    // grid = (site groups, frames): no 64-bit division to find the frame
    const int f = blockIdx.y;
    const int slot = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 2);
    const int lane4 = (int)(threadIdx.x & 3);
    if (slot >= p.N) return;
    // neighbouring thread groups take sites that are neighbours in space (not in the reference's randperm order), so
    // their taps share texels in L1; every result is written at the site's own index
    const int site = p.order ? __ldg(p.order + slot) : slot;
    const int64_t gid = (int64_t)f * p.N + site;
    const float px = p.points[3 * gid], py = p.points[3 * gid + 1], pz = p.points[3 * gid + 2];
    const float gx = fminf(fmaxf(px / p.radius, -1.0f), 1.0f);
    const float gy = fminf(fmaxf(py / p.radius, -1.0f), 1.0f);
    const float gz = fminf(fmaxf(pz / p.radius, -1.0f), 1.0f);
    const int R = p.R, R2 = R * R;
    // planes (x,y) (x,z) (y,z): first coordinate indexes W (columns), second H (rows); renderer.py:300-303
    const float ua[3] = {gx, gx, gy}, va[3] = {gy, gz, gz};
    const float4 *base = p.proj + ((int64_t)f * 3 * R2) * 4 + lane4;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int pl = 0; pl < 3; ++pl) {
        int x0, y0;
        float wx1, wy1;
        tap_coords(ua[pl], R, x0, wx1);
        tap_coords(va[pl], R, y0, wy1);
        // zero padding without branches: an out-of-range tap (index -1 or R) gets weight 0 and a clamped address
        const float wx0 = x0 >= 0 ? 1.0f - wx1 : 0.0f, wy0 = y0 >= 0 ? 1.0f - wy1 : 0.0f;
        wx1 = x0 + 1 < R ? wx1 : 0.0f;
        wy1 = y0 + 1 < R ? wy1 : 0.0f;
        const int xa = max(x0, 0), xb = min(x0 + 1, R - 1), ya = max(y0, 0) * R, yb = min(y0 + 1, R - 1) * R;
        const float4 *plane = base + pl * R2 * 4;
        const float4 t00 = __ldg(plane + (ya + xa) * 4), t01 = __ldg(plane + (ya + xb) * 4);
        const float4 t10 = __ldg(plane + (yb + xa) * 4), t11 = __ldg(plane + (yb + xb) * 4);
        const float w00 = wx0 * wy0, w01 = wx1 * wy0, w10 = wx0 * wy1, w11 = wx1 * wy1;
        // same accumulation order as before: (y0,x0), (y0,x1), (y1,x0), (y1,x1)
        float4 pacc;
        pacc.x = fmaf(t11.x, w11, fmaf(t10.x, w10, fmaf(t01.x, w01, t00.x * w00)));
        pacc.y = fmaf(t11.y, w11, fmaf(t10.y, w10, fmaf(t01.y, w01, t00.y * w00)));
        pacc.z = fmaf(t11.z, w11, fmaf(t10.z, w10, fmaf(t01.z, w01, t00.z * w00)));
        pacc.w = fmaf(t11.w, w11, fmaf(t10.w, w10, fmaf(t01.w, w01, t00.w * w00)));
        acc.x += pacc.x; acc.y += pacc.y; acc.z += pacc.z; acc.w += pacc.w;
    }
    // xyz columns of the heads + bias (x = [points, feats], renderer.py:165)
    float v[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float4 wb = s_wb[lane4 * 4 + k];
        v[k] = v[k] + (fmaf(wb.x, px, fmaf(wb.y, py, wb.z * pz)) + wb.w);
    }
    if (lane4 == 0) {                      // rot = normalize(rotation), eps 1e-12 (renderer.py:333)
        const float n = fmaxf(sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3]), 1e-12f);
        const float inv = 1.0f / n;        // one division; x * (1/n) is within 1 ulp of x / n
        reinterpret_cast<float4 *>(p.rot)[gid] = make_float4(v[0] * inv, v[1] * inv, v[2] * inv, v[3] * inv);
    } else if (lane4 == 1) {               // xyz = points + offset + transl (renderer.py:338); opacity raw
        const float *t = p.transl + 3 * f;
        p.xyz[3 * gid] = px + v[0] + t[0];
        p.xyz[3 * gid + 1] = py + v[1] + t[1];
        p.xyz[3 * gid + 2] = pz + v[2] + t[2];
        p.opacity[gid] = v[3];
    } else if (lane4 == 2) {               // scale raw
        p.scale[3 * gid] = v[0]; p.scale[3 * gid + 1] = v[1]; p.scale[3 * gid + 2] = v[2];
    } else {                               // color = sigmoid(shs) (renderer.py:335); 1 + e^-v in [1, inf): fast divide
        p.color[3 * gid] = __fdividef(1.0f, 1.0f + expf(-v[0]));
        p.color[3 * gid + 1] = __fdividef(1.0f, 1.0f + expf(-v[1]));
        p.color[3 * gid + 2] = __fdividef(1.0f, 1.0f + expf(-v[2]));
    }
}

}  // namespace

extern "C" int amav_triplane_project(int F, int C, int R, const float *tokens, int64_t token_frame_stride,
                                     const float *weight, float *proj, void *stream)
{
    AMAV_CHECK_ARG(F >= 0 && C > 0 && R > 0, "bad dims");
    AMAV_CHECK_ARG((R * R) % PJ_TEX == 0, "R*R must be a multiple of 128");
    AMAV_CHECK_ARG(C % (PJ_KGROUPS * 8) == 0, "C must be a multiple of 64");
    AMAV_CHECK_ARG(token_frame_stride >= (int64_t)C * 3 * R * R, "frame stride too small");
    if (F == 0) return AMAV_OK;
    AMAV_CHECK_ARG(tokens && weight && proj, "null buffer");
    AMAV_CHECK_ARG(((uintptr_t)tokens & 15) == 0 && (token_frame_stride & 3) == 0, "tokens must be 16-byte aligned");
    const size_t smem = ((size_t)C * 16 + (size_t)PJ_KGROUPS * 14 * PJ_TEX) * sizeof(float);
    AMAV_CHECK_ARG(smem <= 113 * 1024, "C too large for the shared-memory weight tile");
    amav_ensure_dyn_smem((const void *)project_kernel, (int)smem);
    const int64_t items = (int64_t)F * (R * R / PJ_TEX);                 // per plane
    AMAV_CHECK_ARG(items < ((int64_t)1 << 31), "too many tiles");
    const int per_plane = (2 * AMAV_SM_COUNT) / 3;                        // 98 x 3 planes = 294 of the 296 resident CTAs
    project_kernel<<<dim3((unsigned)(items < per_plane ? items : per_plane), 3), PJ_THREADS, smem, (cudaStream_t)stream>>>(
        F, C, R * R, tokens, token_frame_stride, weight, 3 + 3 * C, proj);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}

extern "C" int amav_sample_decode(int F, int N, int C, int R, float radius, const float *proj, const float *points,
                                  const float *transl, const float *weight, const float *bias, float *xyz,
                                  float *scale, float *rot, float *opacity, float *color, const int32_t *order, void *stream_)
{
    AMAV_CHECK_ARG(F >= 0 && N >= 0 && C > 0 && R > 0 && radius > 0.0f, "bad dims");
    if ((int64_t)F * N == 0) return AMAV_OK;
    AMAV_CHECK_ARG(proj && points && transl && weight && bias, "null input");
    AMAV_CHECK_ARG(xyz && scale && rot && opacity && color, "null output");
    cudaStream_t stream = (cudaStream_t)stream_;
    SampleParams p;
    p.F = F; p.N = N; p.R = R; p.radius = radius;
    p.proj = reinterpret_cast<const float4 *>(proj); p.points = points; p.transl = transl;
    p.weight = weight; p.bias = bias; p.ldw = 3 + 3 * C;
    p.xyz = xyz; p.scale = scale; p.rot = rot; p.opacity = opacity; p.color = color; p.order = order;
    AMAV_CHECK_ARG(F <= 65535 && (int64_t)3 * R * R * 4 < ((int64_t)1 << 31), "too many frames / planes too large for one call");
    sample_decode_kernel<<<dim3(amav_div_up((int64_t)N * 4, 256), F), 256, 0, stream>>>(p);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
