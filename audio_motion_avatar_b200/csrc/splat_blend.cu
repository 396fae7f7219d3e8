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
#include <stdlib.h>

#include "blend_common.cuh"

namespace {

using namespace amav_blend;

struct BlendParams {
    int F, N, W, H, gx, gy, hwc, clamp01, vec_ok;
    const float4 *geom;
    const uint32_t *tile_start;
    const uint32_t *order;            // deep-tile list of amav_splat_scan, or NULL
    const int32_t *point_list;
    float bg0, bg1, bg2;
    float *out_color, *final_T, *out_invdepth;
    int32_t *n_contrib;
    unsigned int *work_counter;
    uint32_t capacity;
};

template <bool DEPTH, bool ROWS>
__global__ void __launch_bounds__(BLEND_THREADS, 5) blend_kernel(const BlendParams p, const float *__restrict__ bg)
{
    // one 48-byte record per staged entry: (x, y, -conic.x/2, -conic.y)(-conic.z/2, opacity, pmin, r)(g, b, depth, radius);
    // a single array so the list walk needs one address per entry (immediate offsets for the rest)
    __shared__ float4 s_rec[BLEND_THREADS * 3];
    // per warp: shared addresses (record base + 48 j) of the batch entries that touch its rectangle
    __shared__ __align__(4) unsigned short s_list[BLEND_THREADS / 32][BLEND_THREADS];
    __shared__ unsigned int s_next;
    __shared__ __align__(16) float s_out[ROWS ? 16 * 96 : 4];       // ROWS: HWC colour of a pair of tiles
    const int tiles = p.gx * p.gy;
    const unsigned int total_tiles = (unsigned int)((int64_t)p.F * tiles);
    const float bg0 = bg[0], bg1 = bg[1], bg2 = bg[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sx = (warp & 1) * SUB_W, sy = (warp >> 1) * SUB_H;
    const int lx = sx + (lane & (SUB_W - 1)), ly = sy + (lane >> 3);
    const float alpha_min = 1.0f / 255.0f;
    unsigned short *list = s_list[warp];
    // ROWS: the warp's list as a shared address pinned in a register (the output of a volatile asm is not
    // rematerialised; with the row-store epilogue ptxas otherwise rebuilds the pointer from tid inside the walk)
    uint32_t list_base = 0;
    if (ROWS) asm volatile("mov.u32 %0, %1;" : "=r"(list_base) : "r"((uint32_t)__cvta_generic_to_shared(s_list[warp])));
    // the per-warp lists hold the records' shared-memory addresses (16 bits: static shared memory of a CTA that is
    // not in a cluster starts at 0), so the list walk does no address arithmetic at all
    const uint32_t rec_base = (uint32_t)__cvta_generic_to_shared(s_rec);
    if (rec_base + 48u * BLEND_THREADS > 0x10000u) __trap();      // the lists hold 16-bit shared addresses
    __shared__ float s_expc[4];
    const ExpConsts ec = exp_consts(s_expc);

    for (;;) {
        // tiles differ in depth by orders of magnitude (the body covers ~20% of the image): CTAs pull the next
        // tile from a device counter instead of striding, so no SM idles behind a column of deep tiles
        if (threadIdx.x == 0) s_next = atomicAdd(p.work_counter, 1u);
        __syncthreads();
        // ROWS: the work unit is a PAIR of horizontally adjacent tiles (W is a multiple of 32), so that the colour leaves
        // as rows of 384 contiguous bytes = three full 128-byte lines; plain frame order (these are long launches)
        unsigned int t0;
        if (ROWS) {
            if (s_next >= total_tiles / 2) break;
            t0 = 2u * s_next;
        } else {
            if (!queue_tile(s_next, p.order, p.tile_start, total_tiles, t0)) break;
            if (t0 == 0xffffffffu) { __syncthreads(); continue; }  // deep tile, already handed out
        }
        const int f = (int)(t0 / tiles);
        const int tile0 = (int)(t0 - (unsigned int)f * tiles);
        const int ty = tile0 / p.gx;
#pragma unroll 1
        for (int half = 0; half < (ROWS ? 2 : 1); ++half) {
        const unsigned int t = t0 + half;
        const int tx = tile0 - ty * p.gx + half;
        const int pxi = tx * 16 + lx, pyi = ty * 16 + ly;
        const bool inside = pxi < p.W && pyi < p.H;
        float pxf, pyf;                             // pinned: ptxas otherwise re-converts them inside the list walk
        asm volatile("cvt.rn.f32.s32 %0, %1;" : "=f"(pxf) : "r"(pxi));
        asm volatile("cvt.rn.f32.s32 %0, %1;" : "=f"(pyf) : "r"(pyi));
        const float rx0 = (float)(tx * 16 + sx), ry0 = (float)(ty * 16 + sy);
        // ranges are clamped to the caller's entry capacity: after a bucket overflow the frame is incomplete (the
        // overflow flag tells the host) but nothing is read out of bounds
        const uint32_t r0 = min(p.tile_start[t], p.capacity), r1 = min(p.tile_start[t + 1], p.capacity);
        const float4 *geom = p.geom + (int64_t)f * p.N * 3;

        // A saturated pixel (upstream: done) is marked by T = 0 with its transmittance parked in Tsat: every later
        // pair then fails the Tn >= 1e-4 test by itself, so the list walk carries no flag (T > 0 while live: it only
        // ever takes values Tn >= 1e-4).  Pixels outside the image start saturated.
        float T = inside ? 1.0f : 0.0f, Tsat = 0.0f, C0 = 0.0f, C1 = 0.0f, C2 = 0.0f, Dv = 0.0f;
        int last = 0, before = 0;                 // `before`: list entries in earlier batches (upstream's contributor count)
        int todo = (int)(r1 - r0);
        for (uint32_t base = r0; base < r1; base += BLEND_THREADS, todo -= BLEND_THREADS, before += BLEND_THREADS) {
            if (__syncthreads_count(T == 0.0f) == BLEND_THREADS) break;
            if (base + threadIdx.x < r1) {
                const int g = p.point_list[base + threadIdx.x];
                const float4 *src = geom + 3 * (int64_t)g;
                s_rec[3 * threadIdx.x] = src[0];
                s_rec[3 * threadIdx.x + 1] = src[1];
                s_rec[3 * threadIdx.x + 2] = src[2];
            }
            __syncthreads();
            const int nb = todo < BLEND_THREADS ? todo : BLEND_THREADS;
            if (__all_sync(0xffffffffu, T == 0.0f)) continue;
            // 1. this warp's list: batch entries whose footprint can reach its 8x4 rectangle, in list order
            int cnt = 0;
            for (int j0 = 0; j0 < nb; j0 += 32) {
                const int j = j0 + lane;
                const bool keep = j < nb && warp_rect_keep(s_rec[3 * j], s_rec[3 * j + 1], rx0, ry0);
                const unsigned int m = __ballot_sync(0xffffffffu, keep);
                if (keep) list[cnt + __popc(m & ((1u << lane) - 1u))] = (unsigned short)(rec_base + 48u * (uint32_t)j);
                cnt += __popc(m);
            }
            __syncwarp();
            // 2. exact per-pixel blend over the warp's list (same arithmetic and tests as the oracle).  Two entries
            // per iteration: alpha of both is computed branch-free (it does not depend on T, and 96% of the listed
            // pairs need it on some lane anyway), so the two exp chains overlap; only the short T recurrence is serial.
            // A deep tile is one CTA's sequential walk, and small launches last as long as their deepest tile.
            int last_off = -1;
            auto pair_alpha = [&](uint32_t rec, float4 &b, float &alpha) -> bool {
                const float4 a = lds128<0>(rec);
                b = lds128<16>(rec);
                const float dx = a.x - pxf, dy = a.y - pyf;
                const float power = fmaf(fmaf(a.w, dy, a.z * dx), dx, (b.x * dy) * dy);
                alpha = fminf(0.99f, b.y * amav_exp(power, ec));
                // b.z = pmin: alpha surely < 1/255 below it (and exp is only meaningful above it)
                return !(power > 0.0f || power < b.z) && !(alpha < alpha_min);
            };
            auto accumulate = [&](uint32_t rec, uint32_t off, const float4 b, float alpha) {
                const float Tn = T * (1.0f - alpha);
                if (Tn < 0.0001f) { Tsat = fmaxf(Tsat, T); T = 0.0f; return; }
                const float w = alpha * T;
                const float2 c = lds64<32>(rec);
                C0 = fmaf(b.w, w, C0);
                C1 = fmaf(c.x, w, C1);
                C2 = fmaf(c.y, w, C2);
                if (DEPTH) Dv = fmaf(1.0f / lds64<40>(rec).x, w, Dv);
                T = Tn;
                last_off = (int)off;
            };
            int k = 0;
            for (; k + 1 < cnt; k += 2) {
                uint32_t off0, off1;
                if (ROWS) {                                             // two 16-bit list entries in one load
                    uint32_t both;
                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(both) : "r"(list_base + 2u * (uint32_t)k));
                    off0 = both & 0xffffu; off1 = both >> 16;
                } else {
                    off0 = list[k]; off1 = list[k + 1];
                }
                const uint32_t rec0 = off0, rec1 = off1;
                const float4 a0 = lds128<0>(rec0), a1 = lds128<0>(rec1);
                const float4 b0 = lds128<16>(rec0), b1 = lds128<16>(rec1);
                const float dx0 = a0.x - pxf, dx1 = a1.x - pxf, dy0 = a0.y - pyf, dy1 = a1.y - pyf;
                const float power0 = fmaf(fmaf(a0.w, dy0, a0.z * dx0), dx0, (b0.x * dy0) * dy0);
                const float power1 = fmaf(fmaf(a1.w, dy1, a1.z * dx1), dx1, (b1.x * dy1) * dy1);
                float e0, e1;
                amav_exp_x2(power0, power1, ec, e0, e1);
                const float alpha0 = fminf(0.99f, b0.y * e0), alpha1 = fminf(0.99f, b1.y * e1);
                const bool ok0 = !(power0 > 0.0f || power0 < b0.z) && !(alpha0 < alpha_min);
                const bool ok1 = !(power1 > 0.0f || power1 < b1.z) && !(alpha1 < alpha_min);
                if (ok0) accumulate(rec0, off0, b0, alpha0);
                if (ok1) accumulate(rec1, off1, b1, alpha1);
            }
            if (k < cnt) {
                const uint32_t off0 = list[k];
                float4 b0;
                float alpha0;
                if (pair_alpha(off0, b0, alpha0)) accumulate(off0, off0, b0, alpha0);
            }
            if (last_off >= 0) last = before + (last_off - (int)rec_base) / 48 + 1;
        }
        // colour: HWC tiles that lie inside the image horizontally go through shared memory and leave as 16 rows of
        // 192 contiguous bytes (12 float4 stores per row) instead of 256 x 3 scalar stores; what matters is the
        // multi-GPU gather, where `out_color` is another GPU's memory and NVLink moves full segments, not 12-byte pixels
        const int64_t hw = (int64_t)p.H * p.W;
        float o0 = 0.0f, o1 = 0.0f, o2 = 0.0f;
        if (inside) {
            if (T == 0.0f) T = Tsat;
            o0 = fmaf(T, bg0, C0); o1 = fmaf(T, bg1, C1); o2 = fmaf(T, bg2, C2);
            if (p.clamp01) {
                o0 = fminf(fmaxf(o0, 0.0f), 1.0f);
                o1 = fminf(fmaxf(o1, 0.0f), 1.0f);
                o2 = fminf(fmaxf(o2, 0.0f), 1.0f);
            }
            const int64_t pix = (int64_t)pyi * p.W + pxi;
            if (!ROWS) {
                if (p.hwc) {
                    float *o = p.out_color + ((int64_t)f * hw + pix) * 3;
                    o[0] = o0; o[1] = o1; o[2] = o2;
                } else {
                    float *o = p.out_color + (int64_t)f * 3 * hw + pix;
                    o[0] = o0; o[hw] = o1; o[2 * hw] = o2;
                }
            }
            if (p.final_T) p.final_T[(int64_t)f * hw + pix] = T;
            if (p.n_contrib) p.n_contrib[(int64_t)f * hw + pix] = last;
            if (DEPTH) p.out_invdepth[(int64_t)f * hw + pix] = Dv;
        }
        if (ROWS) {                                  // stage this tile's colour: row ly, 96 floats per row of the pair
            float *so = s_out + ly * 96 + half * 48 + lx * 3;
            so[0] = o0; so[1] = o1; so[2] = o2;
        }
        __syncthreads();
        }                                            // half
        if (ROWS) {
            // HWC colour of the tile pair: 16 rows x 384 contiguous bytes (24 float4 per row).  What matters is the
            // multi-GPU gather, where `out_color` is another GPU's memory: NVLink moves full lines, not 12-byte pixels
            const int64_t hw = (int64_t)p.H * p.W;
            const int tx0 = tile0 - ty * p.gx;
            for (int i = threadIdx.x; i < 16 * 24; i += BLEND_THREADS) {
                const int row = i / 24, q = i - row * 24;
                if (ty * 16 + row < p.H)
                    reinterpret_cast<float4 *>(p.out_color + ((int64_t)f * hw + (int64_t)(ty * 16 + row) * p.W + tx0 * 16) * 3)[q] =
                        reinterpret_cast<const float4 *>(s_out + row * 96)[q];
            }
            __syncthreads();                         // s_out is free for the next pair
        }
    }
}

}  // namespace

extern "C" int amav_splat_blend(const AmavSplatDims *d, const float *geom, const uint32_t *tile_start,
                                const int32_t *point_list, int64_t capacity, const float *bg, int hwc, int clamp01,
                                float *out_color, float *final_T, int32_t *n_contrib, float *out_invdepth,
                                uint32_t *work_counter, const uint32_t *tile_order, void *stream)
{
    AMAV_CHECK_ARG(d && d->frames >= 0 && d->width > 0 && d->height > 0, "bad dims");
    if (d->frames == 0) return AMAV_OK;
    AMAV_CHECK_ARG(tile_start && bg && out_color && work_counter, "null buffer");
    AMAV_CHECK_ARG((int64_t)d->frames * ((d->width + 15) / 16) * ((d->height + 15) / 16) < ((int64_t)1 << 31), "too many tiles");
    BlendParams p;
    p.F = d->frames; p.N = d->num_gaussians; p.W = d->width; p.H = d->height;
    p.gx = (d->width + 15) / 16; p.gy = (d->height + 15) / 16; p.hwc = hwc; p.clamp01 = clamp01;
    // Row stores (the colour of a pair of tiles leaves as 16 x 384 contiguous bytes) when the output buffer lives on
    // ANOTHER GPU (shard.PeerFrameGather: NVLink moves lines, not 12-byte pixels).  For local memory the plain
    // per-pixel stores are kept: the extra code costs the list walk ~4% through register allocation.
    int rows = 0;
    if (hwc && d->width % 32 == 0 && ((uintptr_t)out_color & 15) == 0 && !final_T && !n_contrib) {    // inference, whole tile pairs
        cudaPointerAttributes attr;
        int dev = -1;
        if (cudaGetDevice(&dev) == cudaSuccess && cudaPointerGetAttributes(&attr, out_color) == cudaSuccess)
            rows = attr.type == cudaMemoryTypeDevice && attr.device != dev;
        else
            cudaGetLastError();
        const char *force = getenv("AMAV_ROW_STORES");
        if (force) rows = force[0] == '1';
    }
    p.vec_ok = rows;
    p.geom = reinterpret_cast<const float4 *>(geom); p.tile_start = tile_start; p.point_list = point_list;
    p.out_color = out_color; p.final_T = final_T; p.n_contrib = n_contrib; p.out_invdepth = out_invdepth;
    p.work_counter = work_counter; p.order = d->frames <= AMAV_DEEP_FIRST_MAX_FRAMES ? tile_order : nullptr;
    AMAV_CHECK_ARG(capacity >= 0, "bad capacity");
    p.capacity = capacity > 0xffffffffll ? 0xffffffffu : (uint32_t)capacity; p.bg0 = p.bg1 = p.bg2 = 0.f;
    const int64_t total_tiles = (int64_t)d->frames * p.gx * p.gy;
    const int grid = (int)(total_tiles < AMAV_SM_COUNT * 8 ? total_tiles : AMAV_SM_COUNT * 8);
    if (out_invdepth) blend_kernel<true, false><<<grid, BLEND_THREADS, 0, (cudaStream_t)stream>>>(p, bg);
    else if (rows) blend_kernel<false, true><<<grid, BLEND_THREADS, 0, (cudaStream_t)stream>>>(p, bg);
    else blend_kernel<false, false><<<grid, BLEND_THREADS, 0, (cudaStream_t)stream>>>(p, bg);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
