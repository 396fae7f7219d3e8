// SURVEY.md 8f-1 (first "next" row after the render path): the image losses every training step applies to the
// rendered frames -- l1_loss and ssim of src/utils/loss_utils.py:18,44-83, combined at
// src/models/lightning_model_wrapper.py:138-145 as  l1 + 0.1 * (1 - ssim)  -- and their gradient w.r.t. the
// rendered image, i.e. the dL/dimage the splat backward starts from.
//
// The reference runs 5 depth-wise 11x11 convolutions (mu1, mu2, E[x^2], E[y^2], E[xy]) plus ~15 elementwise kernels
// forward and autograd's mirror of all of them backward, each a full pass over [F,3,H,W] after an HWC->CHW permute.
// Here: two kernels, both reading the HWC frames the blend kernel wrote.
//   image_loss_fwd_kernel  16x16-pixel tile + 5-pixel halo in smem, separable 11-tap Gaussian (zero padding, as
//                          F.conv2d(padding=5)), the SSIM map, per-frame sums of |x - y| and of the map, and the three
//                          partial-derivative maps dS/dmu1, dS/dE[x^2], dS/dE[xy] (only when a gradient is wanted);
//   image_loss_bwd_kernel  the same separable filter over those three maps (the window is symmetric), combined with
//                          x, y and the L1 sign into dL/dx in HWC.
#include <math.h>

#include "common.cuh"

namespace {

constexpr int LT = 16;                 // tile edge
constexpr int LR = 5;                  // window radius (window_size 11)
constexpr int LH = LT + 2 * LR;        // 26: tile + halo
constexpr float SSIM_C1 = 0.01f * 0.01f, SSIM_C2 = 0.03f * 0.03f;

struct Window { float w[2 * LR + 1]; };

__device__ __forceinline__ float block_sum_256(float v, float *s_part)
{
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.0f;
    if (threadIdx.x == 0)
        for (int w = 0; w < 8; ++w) t += s_part[w];
    __syncthreads();
    return t;                          // valid on thread 0
}

// Separable 11-tap filter with register sliding windows: one work item produces LG consecutive outputs along the
// filter direction from LG + 10 inputs held in registers, so an output costs (LG + 10) / LG shared-memory loads per
// filtered quantity instead of 11 (the kernels were bound by the LDS pipe, not by FMAs).  All three channels of a tile
// go through one pass: 26 x 26 x 3 halo of both images -> horizontal pass (26 rows x 16 columns x 3 channels, five
// moments) -> vertical pass -> SSIM.
constexpr int LG = 4;                  // outputs per work item along the filter direction
constexpr int LCH = 3;                 // channels

__global__ void __launch_bounds__(256) image_loss_fwd_kernel(int H, int W, const float *__restrict__ img,
                                                              const float *__restrict__ gt, const Window win,
                                                              float *__restrict__ sums, float *__restrict__ maps)
{
    __shared__ float s_x[LCH][LH][LH + 1], s_y[LCH][LH][LH + 1];
    __shared__ float s_h[LCH][5][LH][LT + 4];          // row stride 20: the vertical pass's two row groups per warp hit disjoint banks
    __shared__ float s_part[8];
    const int f = blockIdx.z;
    const int tx0 = blockIdx.x * LT, ty0 = blockIdx.y * LT;
    const int64_t hw = (int64_t)H * W;
    const float *xf = img + (int64_t)f * hw * 3, *yf = gt + (int64_t)f * hw * 3;
    // halo of both images, all channels (HWC: a pixel's three floats are contiguous)
    for (int i = threadIdx.x; i < LH * LH * LCH; i += 256) {
        const int pix = i / LCH, c = i - pix * LCH;
        const int r = pix / LH, q = pix - r * LH;
        const int gy = ty0 + r - LR, gx = tx0 + q - LR;
        float a = 0.0f, b = 0.0f;
        if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
            a = xf[((int64_t)gy * W + gx) * 3 + c];
            b = yf[((int64_t)gy * W + gx) * 3 + c];
        }
        s_x[c][r][q] = a; s_y[c][r][q] = b;
    }
    __syncthreads();
    // horizontal pass: item = (channel, row, group of LG columns)
    for (int i = threadIdx.x; i < LCH * LH * (LT / LG); i += 256) {
        const int g = i % (LT / LG), r = (i / (LT / LG)) % LH, c = i / ((LT / LG) * LH);
        float a[LG + 2 * LR], b[LG + 2 * LR];
#pragma unroll
        for (int k = 0; k < LG + 2 * LR; ++k) { a[k] = s_x[c][r][g * LG + k]; b[k] = s_y[c][r][g * LG + k]; }
        float m[5][LG];
#pragma unroll
        for (int o = 0; o < LG; ++o) { m[0][o] = m[1][o] = m[2][o] = m[3][o] = m[4][o] = 0.0f; }
#pragma unroll
        for (int k = 0; k < LG + 2 * LR; ++k) {
            const float aa = a[k] * a[k], bb = b[k] * b[k], ab = a[k] * b[k];
#pragma unroll
            for (int o = 0; o < LG; ++o) {
                const int t = k - o;                           // tap index of input k for output o
                if (t >= 0 && t <= 2 * LR) {
                    const float w = win.w[t];
                    m[0][o] = fmaf(w, a[k], m[0][o]); m[1][o] = fmaf(w, b[k], m[1][o]);
                    m[2][o] = fmaf(w, aa, m[2][o]); m[3][o] = fmaf(w, bb, m[3][o]); m[4][o] = fmaf(w, ab, m[4][o]);
                }
            }
        }
#pragma unroll
        for (int q = 0; q < 5; ++q)
#pragma unroll
            for (int o = 0; o < LG; ++o) s_h[c][q][r][g * LG + o] = m[q][o];
    }
    __syncthreads();
    // vertical pass + SSIM: item = (channel, group of LG rows, column); consecutive threads = consecutive columns
    float l1 = 0.0f, ss = 0.0f;
    for (int i = threadIdx.x; i < LCH * (LT / LG) * LT; i += 256) {
        const int lx = i % LT, g = (i / LT) % (LT / LG), c = i / (LT * (LT / LG));
        float v[5][LG];
#pragma unroll
        for (int o = 0; o < LG; ++o) { v[0][o] = v[1][o] = v[2][o] = v[3][o] = v[4][o] = 0.0f; }
#pragma unroll
        for (int k = 0; k < LG + 2 * LR; ++k) {
            float h[5];
#pragma unroll
            for (int q = 0; q < 5; ++q) h[q] = s_h[c][q][g * LG + k][lx];
#pragma unroll
            for (int o = 0; o < LG; ++o) {
                const int t = k - o;
                if (t >= 0 && t <= 2 * LR) {
                    const float w = win.w[t];
#pragma unroll
                    for (int q = 0; q < 5; ++q) v[q][o] = fmaf(w, h[q], v[q][o]);
                }
            }
        }
        const int px = tx0 + lx;
#pragma unroll
        for (int o = 0; o < LG; ++o) {
            const int ly = g * LG + o, py = ty0 + ly;
            if (px < W && py < H) {
                const float mu1 = v[0][o], mu2 = v[1][o], e11 = v[2][o], e22 = v[3][o], e12 = v[4][o];
                const float x = s_x[c][ly + LR][lx + LR], y = s_y[c][ly + LR][lx + LR];
                l1 += fabsf(x - y);
                const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
                const float sig1 = e11 - mu1_sq, sig2 = e22 - mu2_sq, sig12 = e12 - mu12;
                const float A1 = 2.0f * mu12 + SSIM_C1, A2 = 2.0f * sig12 + SSIM_C2;
                const float B1 = mu1_sq + mu2_sq + SSIM_C1, B2 = sig1 + sig2 + SSIM_C2;
                const float inv = 1.0f / (B1 * B2);
                const float S = A1 * A2 * inv;
                ss += S;
                if (maps) {
                    // S as a function of (mu1, E[x^2], E[xy]) with sigma1^2 = E[x^2] - mu1^2, sigma12 = E[xy] - mu1 mu2
                    const float dS_de11 = -S / B2;
                    const float dS_de12 = 2.0f * A1 * inv;
                    const float dS_dmu1 = 2.0f * mu2 * (A2 - A1) * inv - S * 2.0f * mu1 * (B2 - B1) * inv;
                    float *mp = maps + ((int64_t)f * 3 + c) * 3 * hw + (int64_t)py * W + px;
                    mp[0] = dS_dmu1; mp[hw] = dS_de11; mp[2 * hw] = dS_de12;
                }
            }
        }
    }
    const float t1 = block_sum_256(l1, s_part);
    const float t2 = block_sum_256(ss, s_part);
    if (threadIdx.x == 0) { atomicAdd(sums + 2 * f, t1); atomicAdd(sums + 2 * f + 1, t2); }
}

// (One channel at a time and one tap per shared-memory load here: with the 41 KB the three-channel sliding-window
// variant needs, this kernel fell from 8 to 4 CTAs per SM and got slower, 0.34 -> 0.41 ms.)
// dL/dx = g_l1[f] * sign(x - y) + g_ssim[f] * (conv(dS/dmu1) + 2 x conv(dS/dE[x^2]) + y conv(dS/dE[xy]))
__global__ void __launch_bounds__(256) image_loss_bwd_kernel(int H, int W, const float *__restrict__ img,
                                                              const float *__restrict__ gt, const Window win,
                                                              const float *__restrict__ maps,
                                                              const float *__restrict__ g_l1,
                                                              const float *__restrict__ g_ssim, float *__restrict__ grad)
{
    __shared__ float s_m[3][LH][LH + 1];
    __shared__ float s_h[3][LH][LT + 1];
    const int f = blockIdx.z;
    const int tx0 = blockIdx.x * LT, ty0 = blockIdx.y * LT;
    const int lx = threadIdx.x & (LT - 1), ly = threadIdx.x >> 4;
    const int px = tx0 + lx, py = ty0 + ly;
    const bool inside = px < W && py < H;
    const int64_t hw = (int64_t)H * W;
    const float gl = g_l1[f], gs = g_ssim[f];
    for (int c = 0; c < 3; ++c) {
        __syncthreads();
        const float *m = maps + ((int64_t)f * 3 + c) * 3 * hw;
        for (int i = threadIdx.x; i < LH * LH; i += 256) {
            const int r = i / LH, q = i - r * LH;
            const int gy = ty0 + r - LR, gx = tx0 + q - LR;
            const bool ok = gx >= 0 && gx < W && gy >= 0 && gy < H;
            const int64_t o = (int64_t)gy * W + gx;
            s_m[0][r][q] = ok ? m[o] : 0.0f;
            s_m[1][r][q] = ok ? m[hw + o] : 0.0f;
            s_m[2][r][q] = ok ? m[2 * hw + o] : 0.0f;
        }
        __syncthreads();
        for (int i = threadIdx.x; i < LH * LT; i += 256) {
            const int r = i >> 4, q = i & (LT - 1);
            float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
            for (int k = 0; k <= 2 * LR; ++k) {
                const float w = win.w[k];
                a0 = fmaf(w, s_m[0][r][q + k], a0); a1 = fmaf(w, s_m[1][r][q + k], a1); a2 = fmaf(w, s_m[2][r][q + k], a2);
            }
            s_h[0][r][q] = a0; s_h[1][r][q] = a1; s_h[2][r][q] = a2;
        }
        __syncthreads();
        float c0 = 0.f, c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int k = 0; k <= 2 * LR; ++k) {
            const float w = win.w[k];
            c0 = fmaf(w, s_h[0][ly + k][lx], c0); c1 = fmaf(w, s_h[1][ly + k][lx], c1); c2 = fmaf(w, s_h[2][ly + k][lx], c2);
        }
        if (inside) {
            const int64_t o = ((int64_t)f * hw + (int64_t)py * W + px) * 3 + c;
            const float x = img[o], y = gt[o];
            const float d = x - y;
            const float sgn = d > 0.0f ? 1.0f : (d < 0.0f ? -1.0f : 0.0f);
            grad[o] = gl * sgn + gs * (c0 + 2.0f * x * c1 + y * c2);
        }
    }
}

Window make_window()
{
    // loss_utils.py:24-26: gauss = Tensor([exp(-(x - 5)^2 / (2 * 1.5^2))]); gauss / gauss.sum()  (float32)
    Window w;
    float sum = 0.0f;
    for (int k = 0; k <= 2 * LR; ++k) {
        w.w[k] = (float)exp(-(double)((k - LR) * (k - LR)) / (2.0 * 1.5 * 1.5));
        sum += w.w[k];
    }
    for (int k = 0; k <= 2 * LR; ++k) w.w[k] = w.w[k] / sum;
    return w;
}

}  // namespace

extern "C" int amav_image_loss_forward(int F, int H, int W, const float *img, const float *gt, float *sums, float *maps,
                                       void *stream)
{
    AMAV_CHECK_ARG(F >= 0 && F <= 65535 && H > 0 && W > 0, "bad dims");
    if (F == 0) return AMAV_OK;
    AMAV_CHECK_ARG(img && gt && sums, "null buffer");
    dim3 grid(amav_div_up(W, LT), amav_div_up(H, LT), F);
    image_loss_fwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(H, W, img, gt, make_window(), sums, maps);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}

extern "C" int amav_image_loss_backward(int F, int H, int W, const float *img, const float *gt, const float *maps,
                                        const float *g_l1, const float *g_ssim, float *grad, void *stream)
{
    AMAV_CHECK_ARG(F >= 0 && F <= 65535 && H > 0 && W > 0, "bad dims");
    if (F == 0) return AMAV_OK;
    AMAV_CHECK_ARG(img && gt && maps && g_l1 && g_ssim && grad, "null buffer");
    dim3 grid(amav_div_up(W, LT), amav_div_up(H, LT), F);
    image_loss_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(H, W, img, gt, make_window(), maps, g_l1, g_ssim, grad);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
