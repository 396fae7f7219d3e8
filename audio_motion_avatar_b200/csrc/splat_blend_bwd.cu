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
//   * each warp owns an 8x8 pixel block (two pixels per thread, packed FP32) and culls every 256-entry batch against
//     it (same exact bound as the forward), walks its list back to front in groups of four entries, reduces the
//     4 x 9 per-pair terms over its 64 pixels with one 31-shuffle split butterfly (+ 6 for the ninth terms), and adds
//     them to a per-batch accumulator in shared memory;
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
constexpr int BW_THREADS = 128;      // 4 warps x (8 x 8 pixels): two pixels per thread, as in the forward kernel
#ifndef AMAV_BW_BATCH
#define AMAV_BW_BATCH 256
#endif
#ifndef AMAV_BW_CTAS
#define AMAV_BW_CTAS 4
#endif
constexpr int BW_BATCH = AMAV_BW_BATCH;   // staged list entries per batch (128 < BW_BATCH <= 256: two records per thread)
constexpr int BW_CTAS = AMAV_BW_CTAS;     // resident CTAs per SM the kernel is compiled for
static_assert(BW_BATCH > 128 && BW_BATCH <= 256 && BW_BATCH % 4 == 0, "batch size");
constexpr int BW_RW = 8, BW_RH = 8;

// Round-2 shape: every thread owns TWO pixels of one column (rows ly, ly + 4), so the per-pair arithmetic runs on packed
// FP32 instructions (FFMA2 / FMUL2 / FADD2) with the record fields as broadcast operands, and the warp reduction of the 9
// terms covers 64 pixels per butterfly.  Rejected pairs are walked as alpha = G = 0 pairs: upstream's recurrence
//     accum = last_alpha * last_colour + (1 - last_alpha) * accum;  last_colour = c;  last_alpha = alpha
// folds the pending pair in one step earlier and then multiplies by exactly 1, so accum, T and every later term come out
// bit-identical to skipping the pair -- and the walk needs no per-pixel select on its state (last_colour becomes the
// previous entry's colour for the whole warp: a broadcast operand).
__global__ void __launch_bounds__(BW_THREADS, BW_CTAS) blend_bwd_kernel(const BlendBwdParams p, const float *__restrict__ bg)
{
    // double-buffered batch state: thread t stages records t and t + 128 of a batch, zeroes their accumulator rows and
    // later flushes them, so consecutive batches need ONE barrier between them (after staging); the flush of batch k
    // overlaps the other warps' walk over batch k+1
    __shared__ float4 s_recb[2][BW_BATCH * 3];               // 48-byte records, as in the forward kernel
    __shared__ float s_accb[2][BW_BATCH * ACC_LD];
    __shared__ int s_gidb[2][BW_BATCH];
    __shared__ unsigned char s_list[BW_THREADS / 32][BW_BATCH];
    __shared__ unsigned int s_next;
    __shared__ int s_maxc;
    const int tiles = p.gx * p.gy;
    const unsigned int total_tiles = (unsigned int)((int64_t)p.F * tiles);
    const float bg0 = bg[0], bg1 = bg[1], bg2 = bg[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sx = (warp & 1) * BW_RW, sy = (warp >> 1) * BW_RH;
    const int lx = sx + (lane & 7), ly = sy + (lane >> 3);
    const float alpha_min = 1.0f / 255.0f;
    const float ddelx_dx = 0.5f * (float)p.W, ddely_dy = 0.5f * (float)p.H;
    unsigned char *list = s_list[warp];
    const bool upper16 = lane & 16, upper8 = lane & 8, upper4 = lane & 4, upper2 = lane & 2, upper1 = lane & 1;
    const int64_t hw = (int64_t)p.H * p.W;

    for (;;) {
        if (threadIdx.x == 0) { s_next = atomicAdd(p.work_counter, 1u); s_maxc = 0; }
        __syncthreads();
        unsigned int t;
        if (!queue_tile(s_next, p.order, p.tile_start, total_tiles, t)) break;
        if (t == 0xffffffffu) { __syncthreads(); continue; }       // deep tile, already handed out
        const int f = (int)(t / tiles);
        const int tile = (int)(t - (unsigned int)f * tiles);
        const int ty = tile / p.gx, tx = tile - ty * p.gx;
        const int pxi = tx * 16 + lx, py0 = ty * 16 + ly, py1 = py0 + 4;
        const bool in0 = pxi < p.W && py0 < p.H, in1 = pxi < p.W && py1 < p.H;
        const float pxf = (float)pxi;
        const float2 npy = make_float2(-(float)py0, -(float)py1);
        const float rx0 = (float)(tx * 16 + sx), ry0 = (float)(ty * 16 + sy);
        const uint32_t r0 = min(p.tile_start[t], p.capacity), r1 = min(p.tile_start[t + 1], p.capacity);
        const float4 *geom = p.geom + (int64_t)f * p.N * 3;
        float *dsum = p.dsum + (int64_t)f * p.N * 12;

        float2 T_final = make_float2(1.0f, 1.0f), g0 = make_float2(0.0f, 0.0f), g1 = g0, g2 = g0;
        int lastc0 = 0, lastc1 = 0;
        auto fetch = [&](bool inside, int pyi, float &Tf, float &a0, float &a1, float &a2, int &lc) {
            if (!inside) return;
            const int64_t pix = (int64_t)pyi * p.W + pxi;
            Tf = p.final_T[(int64_t)f * hw + pix];
            lc = p.n_contrib[(int64_t)f * hw + pix];
            if (p.hwc) {
                const float *g = p.dL_dpix + ((int64_t)f * hw + pix) * 3;
                a0 = g[0]; a1 = g[1]; a2 = g[2];
            } else {
                const float *g = p.dL_dpix + (int64_t)f * 3 * hw + pix;
                a0 = g[0]; a1 = g[hw]; a2 = g[2 * hw];
            }
        };
        fetch(in0, py0, T_final.x, g0.x, g1.x, g2.x, lastc0);
        fetch(in1, py1, T_final.y, g0.y, g1.y, g2.y, lastc1);
        int wmax = max(lastc0, lastc1);
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, d));
        if (lane == 0 && wmax > 0) atomicMax(&s_maxc, wmax);
        __syncthreads();
        const int nproc = min(s_maxc, (int)(r1 - r0));
        // -(bg . dL/dpix) T_final: the background term's factor
        const float2 nbgT = make_float2(-(bg0 * g0.x + bg1 * g1.x + bg2 * g2.x) * T_final.x, -(bg0 * g0.y + bg1 * g1.y + bg2 * g2.y) * T_final.y);
        float2 T = T_final, accum0 = make_float2(0.0f, 0.0f), accum1 = accum0, accum2 = accum0, last_alpha = accum0;
        float lc0 = 0.0f, lc1 = 0.0f, lc2 = 0.0f;                  // colour of the previous walked entry (warp-uniform)

        // flush of one staged batch: one record per (tile, Gaussian)
        auto flush = [&](int buf, int count) {
#pragma unroll
            for (int r = 0; r < (BW_BATCH + BW_THREADS - 1) / BW_THREADS; ++r) {
                const int row = threadIdx.x + r * BW_THREADS;
                if (row < count) {
                    const float *sa = s_accb[buf] + row * ACC_LD;
                    const float4 q0 = make_float4(sa[0], sa[1], sa[2], sa[3]);
                    const float4 q1 = make_float4(sa[4], sa[5], sa[6], sa[7]);
                    const float q2 = sa[8];
                    const bool nz = q0.x != 0.0f || q0.y != 0.0f || q0.z != 0.0f || q0.w != 0.0f || q1.x != 0.0f ||
                                    q1.y != 0.0f || q1.z != 0.0f || q1.w != 0.0f || q2 != 0.0f;
                    if (nz) {
                        float *dst = dsum + (int64_t)s_gidb[buf][row] * 12;
                        atomicAdd(reinterpret_cast<float4 *>(dst), q0);
                        atomicAdd(reinterpret_cast<float4 *>(dst) + 1, q1);
                        atomicAdd(dst + 8, q2);
                    }
                }
            }
        };
        int prev_nb = 0, it = 0;
        for (int b = (nproc - 1) / BW_BATCH; b >= 0 && nproc > 0; --b, ++it) {
            const int before = b * BW_BATCH;
            const int nb = min(BW_BATCH, nproc - before);
            const int buf = it & 1;
            float4 *s_rec = s_recb[buf];
            float *s_acc = s_accb[buf];
            // buffer `buf` was flushed by this same thread two batches ago: no barrier needed before re-staging it
            {
                const int j0 = threadIdx.x, j1 = threadIdx.x + BW_THREADS;
                const int ga = j0 < nb ? p.point_list[r0 + before + j0] : -1, gb = j1 < nb ? p.point_list[r0 + before + j1] : -1;
                float4 q0, q1, q2, q3, q4, q5;
                if (ga >= 0) { const float4 *src = geom + 3 * (int64_t)ga; q0 = src[0]; q1 = src[1]; q2 = src[2]; }
                if (gb >= 0) { const float4 *src = geom + 3 * (int64_t)gb; q3 = src[0]; q4 = src[1]; q5 = src[2]; }
                if (ga >= 0) { s_rec[3 * j0] = q0; s_rec[3 * j0 + 1] = q1; s_rec[3 * j0 + 2] = q2; s_gidb[buf][j0] = ga; }
                if (gb >= 0) { s_rec[3 * j1] = q3; s_rec[3 * j1 + 1] = q4; s_rec[3 * j1 + 2] = q5; s_gidb[buf][j1] = gb; }
#pragma unroll
                for (int k = 0; k < ACC_LD; ++k) {
                    s_acc[j0 * ACC_LD + k] = 0.0f;
                    if (j1 < BW_BATCH) s_acc[j1 * ACC_LD + k] = 0.0f;
                }
            }
            __syncthreads();                                   // batch staged; every warp has finished the previous batch
            flush(buf ^ 1, prev_nb);
            prev_nb = nb;
            // this warp's list for the batch (list order), entries beyond the warp's deepest contributor dropped
            int cnt = 0;
            if (wmax > before) {
                const int lim = min(nb, wmax - before);
                for (int j0 = 0; j0 < lim; j0 += 32) {
                    const int j = j0 + lane;
                    const bool keep = j < lim && warp_rect_keep<BW_RW, BW_RH>(s_rec[3 * j], s_rec[3 * j + 1], rx0, ry0);
                    const unsigned int m = __ballot_sync(0xffffffffu, keep);
                    if (keep) list[cnt + __popc(m & ((1u << lane) - 1u))] = (unsigned char)j;
                    cnt += __popc(m);
                }
            }
            __syncwarp();
            // Back-to-front walk in GROUPS OF FOUR list entries; per entry everything is packed over the thread's two pixels.
            // Then the 4 x 8 per-pair terms are summed over the warp's 64 pixels by ONE 31-shuffle split butterfly (lane l
            // ends with term l & 7 of entry l >> 3) and the four ninth terms by 6 more, one shared-memory atomic per lane.
            for (int k = cnt - 1; k >= 0; k -= 4) {
                float vals[32], nin[4];
                int js[4];
                bool any_use = false;
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    js[e] = k - e >= 0 ? (int)list[k - e] : -1;
                    const int j = max(js[e], 0);
                    const float4 *rec = s_rec + 3 * j;
                    const float4 a = rec[0];
                    const float4 bq = rec[1];
                    const float4 c = rec[2];
                    const float dx = a.x - pxf;
                    const float2 dy = __fadd2_rn(make_float2(a.y, a.y), npy);
                    const float adx = a.z * dx;
                    const float2 uu = __ffma2_rn(make_float2(a.w, a.w), dy, make_float2(adx, adx));
                    const float2 vv = __fmul2_rn(__fmul2_rn(make_float2(bq.x, bq.x), dy), dy);
                    const float2 power = __ffma2_rn(uu, make_float2(dx, dx), vv);       // the forward's operation sequence
                    const float2 Gr = amav_exp_f32x2(power);                            // and the forward's exp
                    const float ar0 = fminf(0.99f, bq.y * Gr.x), ar1 = fminf(0.99f, bq.y * Gr.y);
                    const int pos = before + j + 1;
                    const bool use0 = js[e] >= 0 && pos <= lastc0 && power.x <= 0.0f && !(power.x < bq.z) && !(ar0 < alpha_min);
                    const bool use1 = js[e] >= 0 && pos <= lastc1 && power.y <= 0.0f && !(power.y < bq.z) && !(ar1 < alpha_min);
                    any_use |= use0 | use1;
                    // a rejected pair is walked with alpha = G = 0 (see the kernel's header): no select on the state below
                    const float2 alpha = make_float2(use0 ? ar0 : 0.0f, use1 ? ar1 : 0.0f);
                    const float2 G = make_float2(use0 ? Gr.x : 0.0f, use1 ? Gr.y : 0.0f);
                    const float2 om = __ffma2_rn(alpha, make_float2(-1.0f, -1.0f), make_float2(1.0f, 1.0f));
                    // 1 / (1 - alpha), 1 - alpha in [0.01, 1]: MUFU reciprocal (1 ulp; exactly 1 for alpha = 0) instead of
                    // upstream's two IEEE divisions; gradients are compared at 1e-3
                    float2 inv_om;
                    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv_om.x) : "f"(om.x));
                    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv_om.y) : "f"(om.y));
                    T = __fmul2_rn(T, inv_om);
                    const float2 w = __fmul2_rn(alpha, T);
                    const float2 oml = __ffma2_rn(last_alpha, make_float2(-1.0f, -1.0f), make_float2(1.0f, 1.0f));
                    accum0 = __ffma2_rn(last_alpha, make_float2(lc0, lc0), __fmul2_rn(oml, accum0));
                    accum1 = __ffma2_rn(last_alpha, make_float2(lc1, lc1), __fmul2_rn(oml, accum1));
                    accum2 = __ffma2_rn(last_alpha, make_float2(lc2, lc2), __fmul2_rn(oml, accum2));
                    lc0 = bq.w; lc1 = c.x; lc2 = c.y;
                    last_alpha = alpha;
                    float2 dL = __fmul2_rn(__ffma2_rn(accum0, make_float2(-1.0f, -1.0f), make_float2(lc0, lc0)), g0);
                    dL = __ffma2_rn(__ffma2_rn(accum1, make_float2(-1.0f, -1.0f), make_float2(lc1, lc1)), g1, dL);
                    dL = __ffma2_rn(__ffma2_rn(accum2, make_float2(-1.0f, -1.0f), make_float2(lc2, lc2)), g2, dL);
                    dL = __fmul2_rn(dL, T);
                    dL = __ffma2_rn(nbgT, inv_om, dL);                                  // dL/dalpha of this pair
                    const float2 dL_dG = __fmul2_rn(make_float2(bq.y, bq.y), dL);
                    const float2 gdx = __fmul2_rn(G, make_float2(dx, dx)), gdy = __fmul2_rn(G, dy);
                    // conic = (-2 Ah, -Bn, -2 Ch): dG/ddelx = -gdx*conic.x - gdy*conic.y
                    const float a2 = 2.0f * a.z, c2 = 2.0f * bq.x;
                    const float2 dGx = __ffma2_rn(make_float2(a2, a2), gdx, __fmul2_rn(make_float2(a.w, a.w), gdy));
                    const float2 dGy = __ffma2_rn(make_float2(c2, c2), gdy, __fmul2_rn(make_float2(a.w, a.w), gdx));
                    const float2 hq = __fmul2_rn(dL_dG, make_float2(-0.5f, -0.5f));
                    const float2 t0 = __fmul2_rn(__fmul2_rn(dL_dG, dGx), make_float2(ddelx_dx, ddelx_dx));
                    const float2 t1 = __fmul2_rn(__fmul2_rn(dL_dG, dGy), make_float2(ddely_dy, ddely_dy));
                    const float2 hgx = __fmul2_rn(hq, gdx);
                    const float2 t2 = __fmul2_rn(hgx, make_float2(dx, dx)), t3 = __fmul2_rn(hgx, dy);
                    const float2 t4 = __fmul2_rn(__fmul2_rn(hq, gdy), dy);
                    const float2 t5 = __fmul2_rn(G, dL);
                    const float2 t6 = __fmul2_rn(w, g0), t7 = __fmul2_rn(w, g1), t8 = __fmul2_rn(w, g2);
                    vals[8 * e + 0] = t0.x + t0.y; vals[8 * e + 1] = t1.x + t1.y; vals[8 * e + 2] = t2.x + t2.y;
                    vals[8 * e + 3] = t3.x + t3.y; vals[8 * e + 4] = t4.x + t4.y; vals[8 * e + 5] = t5.x + t5.y;
                    vals[8 * e + 6] = t6.x + t6.y; vals[8 * e + 7] = t7.x + t7.y; nin[e] = t8.x + t8.y;
                }
                if (!__any_sync(0xffffffffu, any_use)) continue;
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
    const int grid = (int)(total_tiles < AMAV_SM_COUNT * BW_CTAS ? total_tiles : AMAV_SM_COUNT * BW_CTAS);
    blend_bwd_kernel<<<grid, BW_THREADS, 0, (cudaStream_t)stream>>>(p, bg);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
