// Stage 4c backward: gradient of the blend w.r.t. the screen-space Gaussian records (upstream renderCUDA of
// backward.cu; reached in the reference through autograd of the GaussianRasterizer
// call at src/models/renderer.py:555-566).
//
// Compiled with -fmad=false and sharing amav_exp / the power expression with the forward kernel, so every
// (pixel, Gaussian) pair takes the same skip decisions here as it did in the forward pass.
//
// Upstream issues 9 global float atomicAdds per contributing (pixel, Gaussian) pair.  Here:
//   * persistent CTAs pull tiles from a device queue (same imbalance as the forward);
//   * a tile starts at the deepest list position any of its pixels reached (max n_contrib), not at the list end;
//   * each warp owns an 8x4 pixel rectangle and culls every 256-entry batch against it (same exact bound as the
//     forward), walks its list back to front, reduces the 9 per-pair terms over its 32 pixels with a 14-shuffle
//     split butterfly, and adds them to a per-batch accumulator in shared memory;
//   * after a batch the CTA flushes one record per (tile, Gaussian) with two float4 + one scalar global
//     reductions: ~1 global atomic record per (tile, Gaussian) instead of 9 per (pixel, Gaussian).
#include "blend_common.cuh"

namespace {

using namespace amav_blend;

struct BlendBwdParams {
    int F, N, W, H, gx, gy, hwc;
    const float4 *geom;
    const uint32_t *tile_start;
    const uint32_t *order;            // deep-tile list of amav_splat_scan, or NULL
    const int32_t *point_list;
    const float *final_T;
    const int32_t *n_contrib;
    const float *dL_dpix;
    float *dsum;                     // [F,N,12]: mean2D.xy (x 0.5W, 0.5H), conic A B C, opacity, rgb, 3 pad
    unsigned int *work_counter;
    uint32_t capacity;
};

constexpr int ACC_LD = 9;            // row stride of the smem accumulator (odd: conflict-free flush)

__global__ void __launch_bounds__(BLEND_THREADS, 3) blend_bwd_kernel(const BlendBwdParams p, const float *__restrict__ bg)
{
    // double-buffered batch state: thread t stages record t of a batch, zeroes its accumulator row and later flushes
    // it, so consecutive batches need ONE barrier between them (after staging); the flush of batch k overlaps the
    // other warps' walk over batch k+1
    __shared__ float4 s_recb[2][BLEND_THREADS * 3];          // 48-byte records, as in the forward kernel
    __shared__ float s_accb[2][BLEND_THREADS * ACC_LD];
    __shared__ int s_gidb[2][BLEND_THREADS];
    __shared__ unsigned char s_list[BLEND_THREADS / 32][BLEND_THREADS];
    __shared__ unsigned int s_next;
    __shared__ int s_maxc;
    const int tiles = p.gx * p.gy;
    const unsigned int total_tiles = (unsigned int)((int64_t)p.F * tiles);
    const float bg0 = bg[0], bg1 = bg[1], bg2 = bg[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sx = (warp & 1) * SUB_W, sy = (warp >> 1) * SUB_H;
    const int lx = sx + (lane & (SUB_W - 1)), ly = sy + (lane >> 3);
    const float alpha_min = 1.0f / 255.0f;
    const float ddelx_dx = 0.5f * (float)p.W, ddely_dy = 0.5f * (float)p.H;
    unsigned char *list = s_list[warp];
    const bool upper16 = lane & 16, upper8 = lane & 8, upper4 = lane & 4, upper2 = lane & 2, upper1 = lane & 1;
    const int64_t hw = (int64_t)p.H * p.W;
    __shared__ float s_expc[4];
    const ExpConsts ec = exp_consts(s_expc);

    for (;;) {
        if (threadIdx.x == 0) { s_next = atomicAdd(p.work_counter, 1u); s_maxc = 0; }
        __syncthreads();
        unsigned int t;
        if (!queue_tile(s_next, p.order, p.tile_start, total_tiles, t)) break;
        if (t == 0xffffffffu) { __syncthreads(); continue; }       // deep tile, already handed out
        const int f = (int)(t / tiles);
        const int tile = (int)(t - (unsigned int)f * tiles);
        const int ty = tile / p.gx, tx = tile - ty * p.gx;
        const int pxi = tx * 16 + lx, pyi = ty * 16 + ly;
        const bool inside = pxi < p.W && pyi < p.H;
        const float pxf = (float)pxi, pyf = (float)pyi;
        const float rx0 = (float)(tx * 16 + sx), ry0 = (float)(ty * 16 + sy);
        const uint32_t r0 = min(p.tile_start[t], p.capacity), r1 = min(p.tile_start[t + 1], p.capacity);
        const float4 *geom = p.geom + (int64_t)f * p.N * 3;
        float *dsum = p.dsum + (int64_t)f * p.N * 12;

        float T_final = 1.0f, g0 = 0.0f, g1 = 0.0f, g2 = 0.0f;
        int last_contributor = 0;
        if (inside) {
            const int64_t pix = (int64_t)pyi * p.W + pxi;
            T_final = p.final_T[(int64_t)f * hw + pix];
            last_contributor = p.n_contrib[(int64_t)f * hw + pix];
            if (p.hwc) {
                const float *g = p.dL_dpix + ((int64_t)f * hw + pix) * 3;
                g0 = g[0]; g1 = g[1]; g2 = g[2];
            } else {
                const float *g = p.dL_dpix + (int64_t)f * 3 * hw + pix;
                g0 = g[0]; g1 = g[hw]; g2 = g[2 * hw];
            }
        }
        int wmax = last_contributor;
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, d));
        if (lane == 0 && wmax > 0) atomicMax(&s_maxc, wmax);
        __syncthreads();
        const int nproc = min(s_maxc, (int)(r1 - r0));
        const float bg_dot = bg0 * g0 + bg1 * g1 + bg2 * g2;
        float T = T_final, accum0 = 0.0f, accum1 = 0.0f, accum2 = 0.0f, last_alpha = 0.0f;
        float lc0 = 0.0f, lc1 = 0.0f, lc2 = 0.0f;

        // flush of one staged batch: one record per (tile, Gaussian)
        auto flush = [&](int buf, int count) {
            if ((int)threadIdx.x < count) {
                const float *sa = s_accb[buf] + threadIdx.x * ACC_LD;
                const float4 q0 = make_float4(sa[0], sa[1], sa[2], sa[3]);
                const float4 q1 = make_float4(sa[4], sa[5], sa[6], sa[7]);
                const float q2 = sa[8];
                const bool nz = q0.x != 0.0f || q0.y != 0.0f || q0.z != 0.0f || q0.w != 0.0f || q1.x != 0.0f ||
                                q1.y != 0.0f || q1.z != 0.0f || q1.w != 0.0f || q2 != 0.0f;
                if (nz) {
                    float *dst = dsum + (int64_t)s_gidb[buf][threadIdx.x] * 12;
                    atomicAdd(reinterpret_cast<float4 *>(dst), q0);
                    atomicAdd(reinterpret_cast<float4 *>(dst) + 1, q1);
                    atomicAdd(dst + 8, q2);
                }
            }
        };
        int prev_nb = 0, it = 0;
        for (int b = (nproc - 1) >> 8; b >= 0 && nproc > 0; --b, ++it) {
            const int before = b * BLEND_THREADS;
            const int nb = min(BLEND_THREADS, nproc - before);
            const int buf = it & 1;
            float4 *s_rec = s_recb[buf];
            float *s_acc = s_accb[buf];
            // buffer `buf` was flushed by this same thread two batches ago: no barrier needed before re-staging it
            if ((int)threadIdx.x < nb) {
                const int g = p.point_list[r0 + before + threadIdx.x];
                const float4 *src = geom + 3 * (int64_t)g;
                s_rec[3 * threadIdx.x] = src[0];
                s_rec[3 * threadIdx.x + 1] = src[1];
                s_rec[3 * threadIdx.x + 2] = src[2];
                s_gidb[buf][threadIdx.x] = g;
            }
#pragma unroll
            for (int k = 0; k < ACC_LD; ++k) s_acc[threadIdx.x * ACC_LD + k] = 0.0f;
            __syncthreads();                                   // batch staged; every warp has finished the previous batch
            flush(buf ^ 1, prev_nb);
            prev_nb = nb;
            // this warp's list for the batch (list order), entries beyond the warp's deepest contributor dropped
            int cnt = 0;
            if (wmax > before) {
                const int lim = min(nb, wmax - before);
                for (int j0 = 0; j0 < lim; j0 += 32) {
                    const int j = j0 + lane;
                    const bool keep = j < lim && warp_rect_keep(s_rec[3 * j], s_rec[3 * j + 1], rx0, ry0);
                    const unsigned int m = __ballot_sync(0xffffffffu, keep);
                    if (keep) list[cnt + __popc(m & ((1u << lane) - 1u))] = (unsigned char)j;
                    cnt += __popc(m);
                }
            }
            __syncwarp();
            // Back-to-front walk in GROUPS OF FOUR list entries: the four T / accum recurrences run one after the other
            // (they are the sequential part), then the 4 x 8 per-pair terms are summed over the warp's 32 pixels by ONE
            // 31-shuffle split butterfly (lane l ends with term l & 7 of entry l >> 3) and the four ninth terms by 6 more:
            // 9.25 shuffles per entry instead of 19, one shared-memory atomic per lane, and the reductions of a group
            // overlap the recurrences of the next.
            for (int k = cnt - 1; k >= 0; k -= 4) {
                float vals[32], nin[4];
                int js[4];
                // phase A (independent per entry: the four exp chains overlap): alpha, G and the accept decision
                float4 ra[4], rb[4];
                float2 rc[4];
                float ddx[4], ddy[4], Gs[4], als[4];
                bool uses[4];
                bool any_use = false;
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    js[e] = k - e >= 0 ? (int)list[k - e] : -1;
                    const int j = max(js[e], 0);
                    const float4 *rec = s_rec + 3 * j;
                    ra[e] = rec[0];
                    rb[e] = rec[1];
                    const float4 c = rec[2];
                    rc[e] = make_float2(c.x, c.y);
                    ddx[e] = ra[e].x - pxf;
                    ddy[e] = ra[e].y - pyf;
                    const float power = fmaf(fmaf(ra[e].w, ddy[e], ra[e].z * ddx[e]), ddx[e], (rb[e].x * ddy[e]) * ddy[e]);
                    const bool inr = js[e] >= 0 && (before + j + 1 <= last_contributor) && power <= 0.0f && !(power < rb[e].z);
                    // same exp as the forward on every accepted pair (power in [pmin, 0]); the clamp only keeps the
                    // rejected lanes' garbage finite
                    const float G = amav_exp(fminf(fmaxf(power, -80.0f), 0.0f), ec);
                    const float alpha = fminf(0.99f, rb[e].y * G);
                    uses[e] = inr && !(alpha < alpha_min);
                    Gs[e] = uses[e] ? G : 0.0f;
                    als[e] = uses[e] ? alpha : 0.0f;
                    any_use |= uses[e];
                }
                if (!__any_sync(0xffffffffu, any_use)) continue;
                // phase B (the sequential part, branch-free): T and the colour behind each entry, then the 9 terms.  A
                // rejected pair has alpha = G = 0: T, accum, last colour stay, all its terms are 0.
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const bool use = uses[e];
                    const float alpha = als[e], G = Gs[e], dx = ddx[e], dy = ddy[e];
                    // 1 / (1 - alpha), 1 - alpha in [0.01, 1]: MUFU reciprocal (1 ulp) instead of upstream's
                    // two IEEE divisions (10 instructions each); gradients are compared at 1e-3
                    float inv_om;
                    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv_om) : "f"(1.0f - alpha));
                    T = T * inv_om;
                    const float w = alpha * T;
                    const float om = 1.0f - last_alpha;
                    const float n0 = fmaf(last_alpha, lc0, om * accum0), n1 = fmaf(last_alpha, lc1, om * accum1),
                                n2 = fmaf(last_alpha, lc2, om * accum2);
                    accum0 = use ? n0 : accum0; accum1 = use ? n1 : accum1; accum2 = use ? n2 : accum2;
                    lc0 = use ? rb[e].w : lc0; lc1 = use ? rc[e].x : lc1; lc2 = use ? rc[e].y : lc2;
                    last_alpha = use ? alpha : last_alpha;
                    float dL_dalpha = (lc0 - accum0) * g0;
                    dL_dalpha = fmaf(lc1 - accum1, g1, dL_dalpha);
                    dL_dalpha = fmaf(lc2 - accum2, g2, dL_dalpha);
                    dL_dalpha = dL_dalpha * T;
                    dL_dalpha = fmaf(-T_final * inv_om, bg_dot, dL_dalpha);
                    const float dL_dG = rb[e].y * dL_dalpha;
                    const float gdx = G * dx, gdy = G * dy;
                    // conic = (-2 Ah, -Bn, -2 Ch): dG/ddelx = -gdx*conic.x - gdy*conic.y
                    const float dG_ddelx = fmaf(2.0f * ra[e].z, gdx, ra[e].w * gdy);
                    const float dG_ddely = fmaf(2.0f * rb[e].x, gdy, ra[e].w * gdx);
                    const float hq = -0.5f * dL_dG;
                    vals[8 * e + 0] = dL_dG * dG_ddelx * ddelx_dx;
                    vals[8 * e + 1] = dL_dG * dG_ddely * ddely_dy;
                    vals[8 * e + 2] = hq * gdx * dx;
                    vals[8 * e + 3] = hq * gdx * dy;
                    vals[8 * e + 4] = hq * gdy * dy;
                    vals[8 * e + 5] = G * dL_dalpha;
                    vals[8 * e + 6] = w * g0;
                    vals[8 * e + 7] = w * g1;
                    nin[e] = w * g2;
                }
                // 32 values over 32 lanes: 16 + 8 + 4 + 2 + 1 shuffles; lane l ends with the warp total of vals[l]
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    const float send = upper16 ? vals[q] : vals[q + 16];
                    const float keep = upper16 ? vals[q + 16] : vals[q];
                    vals[q] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const float send = upper8 ? vals[q] : vals[q + 8];
                    const float keep = upper8 ? vals[q + 8] : vals[q];
                    vals[q] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float send = upper4 ? vals[q] : vals[q + 4];
                    const float keep = upper4 ? vals[q + 4] : vals[q];
                    vals[q] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
                }
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const float send = upper2 ? vals[q] : vals[q + 2];
                    const float keep = upper2 ? vals[q + 2] : vals[q];
                    vals[q] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
                }
                {
                    const float send = upper1 ? vals[0] : vals[1];
                    const float keep = upper1 ? vals[1] : vals[0];
                    vals[0] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
                }
                // the four ninth terms: lane l ends with the total of nin[l >> 3]
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const float send = upper16 ? nin[q] : nin[q + 2];
                    const float keep = upper16 ? nin[q + 2] : nin[q];
                    nin[q] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                }
                {
                    const float send = upper8 ? nin[0] : nin[1];
                    const float keep = upper8 ? nin[1] : nin[0];
                    nin[0] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
                }
                nin[0] += __shfl_xor_sync(0xffffffffu, nin[0], 4);
                nin[0] += __shfl_xor_sync(0xffffffffu, nin[0], 2);
                nin[0] += __shfl_xor_sync(0xffffffffu, nin[0], 1);
                const int e_mine = lane >> 3;
                const int j_mine = e_mine == 0 ? js[0] : e_mine == 1 ? js[1] : e_mine == 2 ? js[2] : js[3];
                if (j_mine >= 0) {
                    if (vals[0] != 0.0f) atomicAdd(&s_acc[j_mine * ACC_LD + (lane & 7)], vals[0]);
                    if ((lane & 7) == 0 && nin[0] != 0.0f) atomicAdd(&s_acc[j_mine * ACC_LD + 8], nin[0]);
                }
            }
        }
        __syncthreads();                                       // every warp is done with the last batch
        flush((it - 1) & 1, prev_nb);
        __syncthreads();
    }
}

}  // namespace

extern "C" int amav_splat_blend_backward(const AmavSplatDims *d, const float *geom, const uint32_t *tile_start,
                                         const int32_t *point_list, int64_t capacity, const float *bg, int hwc,
                                         const float *final_T, const int32_t *n_contrib, const float *dL_dcolor,
                                         float *dsum, uint32_t *work_counter, const uint32_t *tile_order, void *stream)
{
    AMAV_CHECK_ARG(d && d->frames >= 0 && d->width > 0 && d->height > 0, "bad dims");
    if (d->frames == 0) return AMAV_OK;
    AMAV_CHECK_ARG(tile_start && bg && final_T && n_contrib && dL_dcolor && dsum && work_counter, "null buffer");
    AMAV_CHECK_ARG(((uintptr_t)dsum & 15) == 0, "dsum must be 16-byte aligned");
    AMAV_CHECK_ARG(capacity >= 0, "bad capacity");
    BlendBwdParams p;
    p.F = d->frames; p.N = d->num_gaussians; p.W = d->width; p.H = d->height;
    p.gx = (d->width + 15) / 16; p.gy = (d->height + 15) / 16; p.hwc = hwc;
    p.geom = reinterpret_cast<const float4 *>(geom); p.tile_start = tile_start; p.point_list = point_list;
    p.final_T = final_T; p.n_contrib = n_contrib; p.dL_dpix = dL_dcolor; p.dsum = dsum;
    p.work_counter = work_counter; p.order = d->frames <= AMAV_DEEP_FIRST_MAX_FRAMES ? tile_order : nullptr;
    p.capacity = capacity > 0xffffffffll ? 0xffffffffu : (uint32_t)capacity;
    const int64_t total_tiles = (int64_t)d->frames * p.gx * p.gy;
    AMAV_CHECK_ARG(total_tiles < ((int64_t)1 << 31), "too many tiles");
    const int grid = (int)(total_tiles < AMAV_SM_COUNT * 4 ? total_tiles : AMAV_SM_COUNT * 4);
    blend_bwd_kernel<<<grid, BLEND_THREADS, 0, (cudaStream_t)stream>>>(p, bg);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
