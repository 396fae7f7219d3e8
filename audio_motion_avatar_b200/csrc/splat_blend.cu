// Stage 4c: front-to-back alpha blending of every tile of every frame (upstream renderCUDA,
// SURVEY.md Appendix A.3; reference call site src/models/renderer.py:555-568).
//
// Compiled with -fmad=false; the per-(pixel, Gaussian) arithmetic is the operation sequence of
// oracle/raster_oracle.c (amav_oracle_blend), including the polynomial exp, so colours, final_T and
// n_contrib are bit-identical to the oracle.
//
// Round-2 shape (the round-1 kernel spent 83 % of its issue slots at an FMA-pipe load of 40 %):
//   * one CTA = one 16x16 tile = 4 warps; a warp owns an 8x8 pixel block and every THREAD TWO PIXELS of one column
//     (rows ly and ly + 4), so the per-pair arithmetic runs on sm_100's packed FP32 instructions (FFMA2 / FMUL2 /
//     FADD2: one issue slot for both pixels, each half IEEE round-to-nearest -> still bit-exact); the record fields
//     are the broadcast (.F32) operands of those instructions, the pixel-dependent values the packed ones;
//   * the T recurrence is branch-free: a saturated pixel is marked by T <= 0 (its transmittance parked as -T), the
//     weight of a rejected pair is selected to 0, so the list walk has no divergent branch and no flag;
//   * persistent CTAs pull tiles from an atomic work counter, stage the tile's sorted records through shared memory
//     in batches of 256 (double-buffered with cp.async in the local-output instantiations: batch b + 1 arrives while
//     batch b is walked), each warp culls the batch against its block (exact bound of the quadratic form) into a list
//     of 16-bit shared addresses and walks the list four entries per iteration.
#include <stdlib.h>

#include "blend_common.cuh"

// row-store epilogue of the peer-gather instantiation: 1 = one TMA bulk store per 384-byte row, 0 = float4 stores
#ifndef AMAV_ROW_TILES
#define AMAV_ROW_TILES 2
#endif
#ifndef AMAV_ROWS_BULK_STORE
#define AMAV_ROWS_BULK_STORE 1
#endif
// 1: the staged batches are double-buffered: the records of batch b + 1 are fetched with cp.async while the warps walk
// batch b (one CTA barrier per batch, the gather latency hidden behind the walk); 0: one buffer, two barriers per batch
#ifndef AMAV_BLEND_DB
#define AMAV_BLEND_DB 1
#endif

namespace {

using namespace amav_blend;

__device__ __forceinline__ void cp_async16(uint32_t saddr, const void *gptr)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" :: "r"(saddr), "l"(gptr) : "memory");
}

constexpr int FWD_THREADS = 128;         // 4 warps x (8 x 8 pixels), 2 pixels per thread
constexpr int FWD_BATCH = 256;           // staged list entries per batch
constexpr int FWD_RW = 8, FWD_RH = 8;    // pixel block of one warp

struct BlendParams {
    int F, N, W, H, gx, gy, hwc, clamp01;
    const float4 *geom;
    const uint32_t *tile_start;
    const uint32_t *order;            // deep-tile list of amav_splat_scan, or NULL
    const int32_t *point_list;
    float *out_color, *final_T, *out_invdepth;
    int32_t *n_contrib;
    unsigned int *work_counter;
    uint32_t capacity;
    int deep_first;                   // hand the deep tiles out first (short launches)
    int ne_queue;                     // the queue is the scan's non-empty-tile list
};

// AUX: final_T / n_contrib / inverse depth are wanted (training, parity tests); the inference instantiations do not
// track the last contributor at all.
// ROWS: 0 = per-pixel stores; 2 or 4 = the work unit is a run of ROWS horizontally adjacent tiles whose colour leaves as rows of
// ROWS x 192 contiguous bytes (peer-gather instantiations)
template <bool AUX, int ROWS>
__global__ void __launch_bounds__(FWD_THREADS, 8) blend_kernel(const BlendParams p, const float *__restrict__ bg)
{
    // one 48-byte record per staged entry: (x, y, -conic.x/2, -conic.y)(-conic.z/2, opacity, pmin, r)(g, b, depth, radius);
    // a single array so the list walk needs one address per entry (immediate offsets for the rest)
    constexpr bool DB = AMAV_BLEND_DB && ROWS == 0;      // the row-store instantiations keep one buffer (their s_out needs the room)
    __shared__ float4 s_rec[FWD_BATCH * 3 * (DB ? 2 : 1)];
    // per warp: shared addresses (record base + 48 j) of the batch entries that touch its block
    __shared__ __align__(4) unsigned short s_list[FWD_THREADS / 32][FWD_BATCH];
    __shared__ unsigned int s_next;
    __shared__ __align__(16) float s_out[ROWS ? 16 * 48 * ROWS : 4];       // ROWS: HWC colour of the run of tiles
    const int tiles = p.gx * p.gy;
    const unsigned int total_tiles = (unsigned int)((int64_t)p.F * tiles);
    const float bg0 = bg[0], bg1 = bg[1], bg2 = bg[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sx = (warp & 1) * FWD_RW, sy = (warp >> 1) * FWD_RH;
    const int lx = sx + (lane & 7), ly = sy + (lane >> 3);          // second pixel of the thread: row ly + 4
    const float alpha_min = 1.0f / 255.0f;
    unsigned short *list = s_list[warp];
    // the per-warp lists hold the records' shared-memory addresses (16 bits: static shared memory of a CTA that is
    // not in a cluster starts at 0), so the list walk does no address arithmetic at all
    const uint32_t rec_base = (uint32_t)__cvta_generic_to_shared(s_rec);
    const uint32_t list_base = (uint32_t)__cvta_generic_to_shared(list);
    if (rec_base + 48u * FWD_BATCH * (DB ? 2 : 1) > 0x10000u) __trap();   // the lists hold 16-bit shared addresses
    unsigned int bb = 0;                      // DB: batches whose copy has been issued so far (CTA-uniform): batch x lives in buffer x & 1

    // NEQ (local output, tile lists from the scan): the queue holds the NON-EMPTY tiles only (four in five are empty at
    // 1024^2: a pop, two barriers and a 3 KB store each); the empty ones get their background here, without the queue
    const bool neq = !ROWS && p.ne_queue;
    const uint32_t *ne = p.order + 2 + total_tiles;              // [0] = count, list from [1] (scan_apply)
    unsigned int nA = 0, nB = 0, n_ne = 0;
    if (neq) {
        n_ne = ne[0];
        if (p.deep_first) { nA = p.order[0]; nB = p.order[1]; }
        const int64_t hw0 = (int64_t)p.H * p.W;
        for (unsigned int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
            if (p.tile_start[t + 1] != p.tile_start[t]) continue;
            const int f = (int)(t / tiles);
            const int tile = (int)(t - (unsigned int)f * tiles);
            const int ty = tile / p.gx, tx = tile - ty * p.gx;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int pxi = tx * 16 + (threadIdx.x & 15), pyi = ty * 16 + 8 * k + (threadIdx.x >> 4);
                if (pxi < p.W && pyi < p.H) {
                    float o0 = bg0, o1 = bg1, o2 = bg2;
                    if (p.clamp01) {
                        o0 = fminf(fmaxf(o0, 0.0f), 1.0f); o1 = fminf(fmaxf(o1, 0.0f), 1.0f); o2 = fminf(fmaxf(o2, 0.0f), 1.0f);
                    }
                    const int64_t pix = (int64_t)pyi * p.W + pxi;
                    if (p.hwc) {
                        float *o = p.out_color + ((int64_t)f * hw0 + pix) * 3;
                        o[0] = o0; o[1] = o1; o[2] = o2;
                    } else {
                        float *o = p.out_color + (int64_t)f * 3 * hw0 + pix;
                        o[0] = o0; o[hw0] = o1; o[2 * hw0] = o2;
                    }
                    if (AUX) {
                        if (p.final_T) p.final_T[(int64_t)f * hw0 + pix] = 1.0f;
                        if (p.n_contrib) p.n_contrib[(int64_t)f * hw0 + pix] = 0;
                        if (p.out_invdepth) p.out_invdepth[(int64_t)f * hw0 + pix] = 0.0f;
                    }
                }
            }
        }
    }
    for (;;) {
        // tiles differ in depth by orders of magnitude (the body covers ~20% of the image): CTAs pull the next
        // tile from a device counter instead of striding, so no SM idles behind a column of deep tiles
        if (threadIdx.x == 0) s_next = atomicAdd(p.work_counter, 1u);
        __syncthreads();
        // ROWS: the work unit is a PAIR of horizontally adjacent tiles (W is a multiple of 32), so that the colour leaves
        // as rows of 384 contiguous bytes = three full 128-byte lines; plain frame order (these are long launches)
        unsigned int t0;
        if (ROWS) {
            constexpr unsigned int RT = ROWS ? ROWS : 1;
            if (s_next >= total_tiles / RT) break;
            t0 = RT * s_next;
        } else if (neq) {
            const unsigned int q = s_next;
            if (q >= nA + nB + n_ne) break;
            if (q < nA) t0 = p.order[2 + q];
            else if (q < nA + nB) t0 = p.order[2 + (total_tiles - 1 - (q - nA))];
            else {
                t0 = ne[1 + (q - nA - nB)];
                if (p.deep_first && p.tile_start[t0 + 1] - p.tile_start[t0] >= (uint32_t)AMAV_DEEP1) { __syncthreads(); continue; }
            }
        } else {
            if (!queue_tile(s_next, p.deep_first ? p.order : nullptr, p.tile_start, total_tiles, t0)) break;
            if (t0 == 0xffffffffu) { __syncthreads(); continue; }  // deep tile, already handed out
        }
        const int f = (int)(t0 / tiles);
        const int tile0 = (int)(t0 - (unsigned int)f * tiles);
        const int ty = tile0 / p.gx;
#pragma unroll 1
        for (int half = 0; half < (ROWS ? ROWS : 1); ++half) {
        const unsigned int t = t0 + half;
        const int tx = tile0 - ty * p.gx + half;
        const int pxi = tx * 16 + lx, py0 = ty * 16 + ly, py1 = py0 + 4;
        const bool in0 = pxi < p.W && py0 < p.H, in1 = pxi < p.W && py1 < p.H;
        const float pxf = (float)pxi;
        const float2 npy = make_float2(-(float)py0, -(float)py1);   // a.y - py == a.y + (-py) exactly
        const float rx0 = (float)(tx * 16 + sx), ry0 = (float)(ty * 16 + sy);
        // ranges are clamped to the caller's entry capacity: after a bucket overflow the frame is incomplete (the
        // overflow flag tells the host) but nothing is read out of bounds
        const uint32_t r0 = min(p.tile_start[t], p.capacity), r1 = min(p.tile_start[t + 1], p.capacity);
        const float4 *geom = p.geom + (int64_t)f * p.N * 3;

        // A saturated pixel (upstream: done) is marked by T <= 0, its transmittance parked as -T: every later pair
        // then fails the Tn >= 1e-4 test by itself (T > 0 while live: it only ever takes values Tn >= 1e-4).
        // Pixels outside the image start saturated.
        float2 T = make_float2(in0 ? 1.0f : -1.0f, in1 ? 1.0f : -1.0f);
        float2 C0 = make_float2(0.0f, 0.0f), C1 = C0, C2 = C0, Dv = C0;
        int last0 = 0, last1 = 0;                  // 1-based list position of the pixel's last contributor (n_contrib)
        int todo = (int)(r1 - r0), before = 0;     // `before`: list entries in earlier batches
        // DB: buffer x & 1 may be overwritten once every warp has passed the barrier at the top of batch x - 1 (they are
        // done walking batch x - 2); the copy of batch x is issued right after that barrier, also across tile boundaries.
        int gn0 = -1, gn1 = -1;                    // Gaussians of the next batch to copy (this thread's two records)
        auto load_idx = [&](uint32_t bs, int left) {
            const int nbn = left < FWD_BATCH ? left : FWD_BATCH;
            gn0 = (int)threadIdx.x < nbn ? p.point_list[bs + threadIdx.x] : -1;
            gn1 = (int)threadIdx.x + FWD_THREADS < nbn ? p.point_list[bs + threadIdx.x + FWD_THREADS] : -1;
        };
        auto issue = [&]() {
            const uint32_t dst = rec_base + (bb & 1u) * (48u * FWD_BATCH) + 48u * threadIdx.x;
            if (gn0 >= 0) {
                const float4 *src = geom + 3 * (int64_t)gn0;
                cp_async16(dst, src); cp_async16(dst + 16u, src + 1); cp_async16(dst + 32u, src + 2);
            }
            if (gn1 >= 0) {
                const float4 *src = geom + 3 * (int64_t)gn1;
                cp_async16(dst + 48u * FWD_THREADS, src); cp_async16(dst + 48u * FWD_THREADS + 16u, src + 1);
                cp_async16(dst + 48u * FWD_THREADS + 32u, src + 2);
            }
            asm volatile("cp.async.commit_group;\n" ::: "memory");
            ++bb;
        };
        if (DB && r0 < r1) {
            load_idx(r0, todo);
            issue();
            if (r0 + FWD_BATCH < r1) load_idx(r0 + FWD_BATCH, todo - FWD_BATCH);
        }
        for (uint32_t base = r0; base < r1; base += FWD_BATCH, todo -= FWD_BATCH, before += FWD_BATCH) {
            if (DB) asm volatile("cp.async.wait_group 0;\n" ::: "memory");
            if (__syncthreads_count(T.x <= 0.0f && T.y <= 0.0f) == FWD_THREADS) break;
            const uint32_t cur_par = DB ? (bb - 1u) & 1u : 0u;
            const uint32_t cur_base = rec_base + cur_par * (48u * FWD_BATCH);
            const float4 *rec_cur = s_rec + cur_par * (FWD_BATCH * 3);
            const int nb = todo < FWD_BATCH ? todo : FWD_BATCH;
            if (DB) {
                if (base + FWD_BATCH < r1) {
                    issue();
                    if (base + 2 * FWD_BATCH < r1) load_idx(base + 2 * FWD_BATCH, todo - 2 * FWD_BATCH);
                }
            } else {                                             // stage: two records per thread, loads issued together
                const int j0 = threadIdx.x, j1 = threadIdx.x + FWD_THREADS;
                const int g0 = j0 < nb ? p.point_list[base + j0] : -1, g1 = j1 < nb ? p.point_list[base + j1] : -1;
                float4 q0, q1, q2, q3, q4, q5;
                if (g0 >= 0) { const float4 *src = geom + 3 * (int64_t)g0; q0 = src[0]; q1 = src[1]; q2 = src[2]; }
                if (g1 >= 0) { const float4 *src = geom + 3 * (int64_t)g1; q3 = src[0]; q4 = src[1]; q5 = src[2]; }
                if (g0 >= 0) { s_rec[3 * j0] = q0; s_rec[3 * j0 + 1] = q1; s_rec[3 * j0 + 2] = q2; }
                if (g1 >= 0) { s_rec[3 * j1] = q3; s_rec[3 * j1 + 1] = q4; s_rec[3 * j1 + 2] = q5; }
                __syncthreads();
            }
            if (__all_sync(0xffffffffu, T.x <= 0.0f && T.y <= 0.0f)) continue;
            // 1. this warp's list: batch entries whose footprint can reach its 8x8 block, in list order
            int cnt = 0;
            for (int j0 = 0; j0 < nb; j0 += 32) {
                const int j = j0 + lane;
                const bool keep = j < nb && warp_rect_keep<FWD_RW, FWD_RH>(rec_cur[3 * j], rec_cur[3 * j + 1], rx0, ry0);
                const unsigned int m = __ballot_sync(0xffffffffu, keep);
                if (keep) list[cnt + __popc(m & ((1u << lane) - 1u))] = (unsigned short)(cur_base + 48u * (uint32_t)j);
                cnt += __popc(m);
            }
            __syncwarp();
            // 2. exact per-pixel blend over the warp's list: the oracle's operation sequence, two pixels per instruction.
            int off0_last = -1, off1_last = -1;
            auto blend_entry = [&](const uint32_t rec) {
                const float4 a = lds128<0>(rec);
                const float4 b = lds128<16>(rec);
                const float dx = a.x - pxf;
                const float2 dy = __fadd2_rn(make_float2(a.y, a.y), npy);
                const float adx = a.z * dx;
                // power = fmaf(fmaf(Bn, dy, Ah*dx), dx, (Ch*dy)*dy)
                const float2 u = __ffma2_rn(make_float2(a.w, a.w), dy, make_float2(adx, adx));
                const float2 v = __fmul2_rn(__fmul2_rn(make_float2(b.x, b.x), dy), dy);
                const float2 power = __ffma2_rn(u, make_float2(dx, dx), v);
                const float2 e = amav_exp_f32x2(power);
                float2 alpha = __fmul2_rn(make_float2(b.y, b.y), e);
                alpha.x = fminf(0.99f, alpha.x);
                alpha.y = fminf(0.99f, alpha.y);
                // b.z = pmin: alpha surely < 1/255 below it (and exp is only meaningful above it)
                const bool ok0 = !(power.x > 0.0f || power.x < b.z) && !(alpha.x < alpha_min);
                const bool ok1 = !(power.y > 0.0f || power.y < b.z) && !(alpha.y < alpha_min);
                const float2 oma = __ffma2_rn(alpha, make_float2(-1.0f, -1.0f), make_float2(1.0f, 1.0f));   // 1 - alpha
                const float2 Tn = __fmul2_rn(T, oma);
                float2 w = __fmul2_rn(alpha, T);
                const bool live0 = ok0 && !(Tn.x < 0.0001f), live1 = ok1 && !(Tn.y < 0.0001f);
                w.x = live0 ? w.x : 0.0f;
                w.y = live1 ? w.y : 0.0f;
                const float2 c = lds64<32>(rec);
                C0 = __ffma2_rn(make_float2(b.w, b.w), w, C0);
                C1 = __ffma2_rn(make_float2(c.x, c.x), w, C1);
                C2 = __ffma2_rn(make_float2(c.y, c.y), w, C2);
                if (AUX && p.out_invdepth) {
                    const float inv = 1.0f / lds64<40>(rec).x;
                    Dv = __ffma2_rn(make_float2(inv, inv), w, Dv);
                }
                // an accepted pair either advances T or saturates the pixel (upstream's done = true): one guarded select
                if (ok0) T.x = live0 ? Tn.x : -fabsf(T.x);
                if (ok1) T.y = live1 ? Tn.y : -fabsf(T.y);
                if (AUX) {
                    if (live0) off0_last = (int)rec;
                    if (live1) off1_last = (int)rec;
                }
            };
            int k = 0;
#pragma unroll 1
            for (; k + 3 < cnt; k += 4) {
                const uint32_t la = list_base + 2u * (uint32_t)k;
                blend_entry(lds_u16<0>(la));
                blend_entry(lds_u16<2>(la));
                blend_entry(lds_u16<4>(la));
                blend_entry(lds_u16<6>(la));
            }
#pragma unroll 1
            for (; k < cnt; ++k) blend_entry(lds_u16<0>(list_base + 2u * (uint32_t)k));
            if (AUX) {
                if (off0_last >= 0) last0 = before + (off0_last - (int)cur_base) / 48 + 1;
                if (off1_last >= 0) last1 = before + (off1_last - (int)cur_base) / 48 + 1;
            }
        }
        // colour epilogue.  ROWS: HWC tile pairs go through shared memory and leave as 16 rows of 384 contiguous bytes
        const int64_t hw = (int64_t)p.H * p.W;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const bool inside = h ? in1 : in0;
            const int pyi = h ? py1 : py0;
            const float Tf = fabsf(h ? T.y : T.x);
            float o0 = 0.0f, o1 = 0.0f, o2 = 0.0f;
            if (inside) {
                o0 = fmaf(Tf, bg0, h ? C0.y : C0.x); o1 = fmaf(Tf, bg1, h ? C1.y : C1.x); o2 = fmaf(Tf, bg2, h ? C2.y : C2.x);
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
                if (AUX) {
                    if (p.final_T) p.final_T[(int64_t)f * hw + pix] = Tf;
                    if (p.n_contrib) p.n_contrib[(int64_t)f * hw + pix] = h ? last1 : last0;
                    if (p.out_invdepth) p.out_invdepth[(int64_t)f * hw + pix] = h ? Dv.y : Dv.x;
                }
            }
            if (ROWS) {                              // stage this tile's colour: 96 floats per row of the pair
                float *so = s_out + (ly + 4 * h) * (48 * ROWS) + half * 48 + lx * 3;
                so[0] = o0; so[1] = o1; so[2] = o2;
            }
        }
        __syncthreads();
        }                                            // half
        if (ROWS) {
            // HWC colour of the tile pair: 16 rows x 384 contiguous bytes (24 float4 per row).  What matters is the
            // multi-GPU gather, where `out_color` is another GPU's memory: NVLink moves full lines, not 12-byte pixels
            const int64_t hw = (int64_t)p.H * p.W;
            const int tx0 = tile0 - ty * p.gx;
#if AMAV_ROWS_BULK_STORE
            // one TMA bulk store (cp.async.bulk.global.shared::cta) per 384-byte row: the copy engine of the SM emits the
            // peer writes as whole bursts instead of 24 float4 stores per row from the LSU
            asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");      // s_out was written by ordinary stores
            __syncthreads();
            if (threadIdx.x < 16 && ty * 16 + (int)threadIdx.x < p.H) {
                float *dst = p.out_color + ((int64_t)f * hw + (int64_t)(ty * 16 + threadIdx.x) * p.W + tx0 * 16) * 3;
                const uint32_t src = (uint32_t)__cvta_generic_to_shared(s_out + threadIdx.x * (48 * ROWS));
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" :: "l"(dst), "r"(src), "n"(192 * ROWS) : "memory");
                asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");   // the source rows have been read
            }
#else
            for (int i = threadIdx.x; i < 16 * 12 * ROWS; i += FWD_THREADS) {
                const int row = i / (12 * ROWS), q = i - row * (12 * ROWS);
                if (ty * 16 + row < p.H)
                    reinterpret_cast<float4 *>(p.out_color + ((int64_t)f * hw + (int64_t)(ty * 16 + row) * p.W + tx0 * 16) * 3)[q] =
                        reinterpret_cast<const float4 *>(s_out + row * (48 * ROWS))[q];
            }
#endif
            __syncthreads();                         // s_out is free for the next pair
        }
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Local output (inference on one GPU, training, parity tests): WARP-INDEPENDENT workers.  The CTA kernel above keeps its
// four warps in lock step (two barriers per 256-entry batch; ncu: the barrier is its largest stall, issue slots 72 % used).
// Here a work unit is one 8x8 QUADRANT of a non-empty tile and every warp is its own worker: it pulls units from the
// device queue, streams the tile's records through a private two-stage cp.async ring (64 entries per stage, the point_list
// indices one stage further ahead in registers), culls each stage against its quadrant and walks the survivors -- no
// __syncthreads anywhere, a saturated quadrant stops reading at once, and the four quadrants of a deep tile run on four
// warps (L2 serves the record re-reads: 4 x 52 B per entry).  Empty tiles (four in five) never enter the queue: every
// warp first writes the background of its share of them (phase 0).
// ------------------------------------------------------------------------------------------------------------------
// Two measured alternatives to the default CTA-per-tile kernel, kept as compile-time switches (B200, 120 frames of 100k
// Gaussians at 1024^2; default 8.61 ms):  AMAV_BLEND_NE_QUEUE = 1: non-empty-tile queue + up-front background pass in the
// CTA kernel, 8.75 ms;  AMAV_BLEND_WARP_WORKERS = 1: the warp-independent workers below, 9.13 ms (13.8 ms when the queue is
// not in frame order: the four quadrant reads of a record must meet in L2).
#ifndef AMAV_BLEND_WARP_WORKERS
#define AMAV_BLEND_WARP_WORKERS 0
#endif
#ifndef AMAV_BLEND_NE_QUEUE
#define AMAV_BLEND_NE_QUEUE 0
#endif
constexpr int WK_THREADS = 128;
constexpr int WK_WARPS = WK_THREADS / 32;
constexpr int WK_CHUNK = 64;             // list entries per stage

template <bool AUX>
__global__ void __launch_bounds__(WK_THREADS, 8) blend_warp_kernel(const BlendParams p, const float *__restrict__ bg)
{
    __shared__ float4 s_rec[WK_WARPS][2][WK_CHUNK * 3];
    __shared__ __align__(4) unsigned short s_list[WK_WARPS][WK_CHUNK];
    const int tiles = p.gx * p.gy;
    const unsigned int total_tiles = (unsigned int)((int64_t)p.F * tiles);
    const float bg0 = bg[0], bg1 = bg[1], bg2 = bg[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float alpha_min = 1.0f / 255.0f;
    const uint32_t rec_base = (uint32_t)__cvta_generic_to_shared(&s_rec[warp][0][0]);     // stage 1 at + 48 * WK_CHUNK
    const uint32_t list_base = (uint32_t)__cvta_generic_to_shared(&s_list[warp][0]);
    if (rec_base + 2u * 48u * WK_CHUNK > 0x10000u) __trap();      // the lists hold 16-bit shared addresses
    const int64_t hw = (int64_t)p.H * p.W;
    const bool want_depth = AUX && p.out_invdepth != nullptr;

    auto store_pixel = [&](int f, int pxi, int pyi, float Tf, float c0, float c1, float c2, int last, float dv) {
        float o0 = fmaf(Tf, bg0, c0), o1 = fmaf(Tf, bg1, c1), o2 = fmaf(Tf, bg2, c2);
        if (p.clamp01) {
            o0 = fminf(fmaxf(o0, 0.0f), 1.0f);
            o1 = fminf(fmaxf(o1, 0.0f), 1.0f);
            o2 = fminf(fmaxf(o2, 0.0f), 1.0f);
        }
        const int64_t pix = (int64_t)pyi * p.W + pxi;
        if (p.hwc) {
            float *o = p.out_color + ((int64_t)f * hw + pix) * 3;
            o[0] = o0; o[1] = o1; o[2] = o2;
        } else {
            float *o = p.out_color + (int64_t)f * 3 * hw + pix;
            o[0] = o0; o[hw] = o1; o[2 * hw] = o2;
        }
        if (AUX) {
            if (p.final_T) p.final_T[(int64_t)f * hw + pix] = Tf;
            if (p.n_contrib) p.n_contrib[(int64_t)f * hw + pix] = last;
            if (p.out_invdepth) p.out_invdepth[(int64_t)f * hw + pix] = dv;
        }
    };

    // ---- phase 0: background of the empty tiles (no queue, no atomics): warp w of the grid takes tiles w, w + W, ... ----
    {
        const unsigned int gw = blockIdx.x * WK_WARPS + warp, nw = gridDim.x * WK_WARPS;
        for (unsigned int t = gw; t < total_tiles; t += nw) {
            if (p.tile_start[t + 1] != p.tile_start[t]) continue;
            const int f = (int)(t / tiles);
            const int tile = (int)(t - (unsigned int)f * tiles);
            const int ty = tile / p.gx, tx = tile - ty * p.gx;
#pragma unroll
            for (int k = 0; k < 8; ++k) {                        // 16 x 16 pixels, 16 lanes per row
                const int pxi = tx * 16 + (lane & 15), pyi = ty * 16 + 2 * k + (lane >> 4);
                if (pxi < p.W && pyi < p.H) store_pixel(f, pxi, pyi, 1.0f, 0.0f, 0.0f, 0.0f, 0, 0.0f);
            }
        }
    }

    // ---- phase 1: (non-empty tile, quadrant) units from the device queue ----
    // queue positions: the deep tiles first (order[0..1] lists, launches of <= AMAV_DEEP_FIRST_MAX_FRAMES frames), then the
    // non-empty list (deep tiles skipped there); a position is one tile = four consecutive units
    const uint32_t *ne = p.order + 2 + total_tiles;              // [0] = count, list from [1]
    const unsigned int n_ne = ne[0];
    const unsigned int nA = p.deep_first ? p.order[0] : 0u, nB = p.deep_first ? p.order[1] : 0u;
    const unsigned int n_units = 4u * (nA + nB + n_ne);
    unsigned int u_next = 0;
    if (lane == 0) u_next = atomicAdd(p.work_counter, 1u);
    for (;;) {
        const unsigned int u = __shfl_sync(0xffffffffu, u_next, 0);
        if (u >= n_units) break;
        if (lane == 0) u_next = atomicAdd(p.work_counter, 1u);  // the next pop is in flight while this unit is blended
        const unsigned int q = u >> 2;
        const int quad = (int)(u & 3u);
        unsigned int t;
        if (q < nA) t = p.order[2 + q];
        else if (q < nA + nB) t = p.order[2 + (total_tiles - 1 - (q - nA))];
        else {
            t = ne[1 + (q - nA - nB)];
            if (p.deep_first && p.tile_start[t + 1] - p.tile_start[t] >= (uint32_t)AMAV_DEEP1) continue;   // handed out above
        }
        const int f = (int)(t / tiles);
        const int tile = (int)(t - (unsigned int)f * tiles);
        const int ty = tile / p.gx, tx = tile - ty * p.gx;
        const int sx = (quad & 1) * FWD_RW, sy = (quad >> 1) * FWD_RH;
        const int pxi = tx * 16 + sx + (lane & 7), py0 = ty * 16 + sy + (lane >> 3), py1 = py0 + 4;
        const bool in0 = pxi < p.W && py0 < p.H, in1 = pxi < p.W && py1 < p.H;
        const float pxf = (float)pxi;
        const float2 npy = make_float2(-(float)py0, -(float)py1);
        const float rx0 = (float)(tx * 16 + sx), ry0 = (float)(ty * 16 + sy);
        const uint32_t r0 = min(p.tile_start[t], p.capacity), r1 = min(p.tile_start[t + 1], p.capacity);
        const float4 *geom = p.geom + (int64_t)f * p.N * 3;
        const int32_t *plist = p.point_list + r0;
        const int n = (int)(r1 - r0);
        const int nchunks = (n + WK_CHUNK - 1) / WK_CHUNK;

        float2 T = make_float2(in0 ? 1.0f : -1.0f, in1 ? 1.0f : -1.0f);
        float2 C0 = make_float2(0.0f, 0.0f), C1 = C0, C2 = C0, Dv = C0;
        int last0 = 0, last1 = 0;

        auto load_idx = [&](int c, int &a, int &b) {
            const int j = c * WK_CHUNK + lane;
            a = j < n ? __ldg(plist + j) : -1;
            b = j + 32 < n ? __ldg(plist + j + 32) : -1;
        };
        auto issue = [&](int c, int a, int b) {                 // records of chunk c -> stage c & 1 (one commit group)
            const uint32_t st = rec_base + (uint32_t)(c & 1) * (48u * WK_CHUNK) + 48u * (uint32_t)lane;
            if (a >= 0) {
                const float4 *src = geom + 3 * (int64_t)a;
                cp_async16(st, src); cp_async16(st + 16, src + 1); cp_async16(st + 32, src + 2);
            }
            if (b >= 0) {
                const float4 *src = geom + 3 * (int64_t)b;
                cp_async16(st + 48u * 32u, src); cp_async16(st + 48u * 32u + 16, src + 1); cp_async16(st + 48u * 32u + 32, src + 2);
            }
            asm volatile("cp.async.commit_group;\n" ::: "memory");
        };
        int ia, ib, ja = -1, jb = -1;
        load_idx(0, ia, ib);
        issue(0, ia, ib);
        if (nchunks > 1) load_idx(1, ja, jb);
        for (int c = 0; c < nchunks; ++c) {
            int ka = -1, kb = -1;
            if (c + 2 < nchunks) load_idx(c + 2, ka, kb);       // indices two stages ahead (registers)
            if (c + 1 < nchunks) issue(c + 1, ja, jb);          // records one stage ahead (cp.async)
            else asm volatile("cp.async.commit_group;\n" ::: "memory");
            asm volatile("cp.async.wait_group 1;\n" ::: "memory");
            __syncwarp();
            const uint32_t stage = rec_base + (uint32_t)(c & 1) * (48u * WK_CHUNK);
            const int nb = min(WK_CHUNK, n - c * WK_CHUNK);
            // 1. cull the stage against this quadrant (exact bound of the quadratic form), survivors in list order
            int cnt = 0;
#pragma unroll
            for (int j0 = 0; j0 < WK_CHUNK; j0 += 32) {
                const int j = j0 + lane;
                bool keep = false;
                if (j < nb) keep = warp_rect_keep<FWD_RW, FWD_RH>(lds128<0>(stage + 48u * (uint32_t)j), lds128<16>(stage + 48u * (uint32_t)j), rx0, ry0);
                const unsigned int m = __ballot_sync(0xffffffffu, keep);
                if (keep) s_list[warp][cnt + __popc(m & ((1u << lane) - 1u))] = (unsigned short)(stage + 48u * (uint32_t)j);
                cnt += __popc(m);
            }
            __syncwarp();
            // 2. exact per-pixel blend over the list: the oracle's operation sequence, two pixels per packed instruction
            int off0_last = -1, off1_last = -1;
            auto blend_entry = [&](const uint32_t rec) {
                const float4 a = lds128<0>(rec);
                const float4 b = lds128<16>(rec);
                const float dx = a.x - pxf;
                const float2 dy = __fadd2_rn(make_float2(a.y, a.y), npy);
                const float adx = a.z * dx;
                const float2 uu = __ffma2_rn(make_float2(a.w, a.w), dy, make_float2(adx, adx));
                const float2 vv = __fmul2_rn(__fmul2_rn(make_float2(b.x, b.x), dy), dy);
                const float2 power = __ffma2_rn(uu, make_float2(dx, dx), vv);
                const float2 e = amav_exp_f32x2(power);
                float2 alpha = __fmul2_rn(make_float2(b.y, b.y), e);
                alpha.x = fminf(0.99f, alpha.x);
                alpha.y = fminf(0.99f, alpha.y);
                const bool ok0 = !(power.x > 0.0f || power.x < b.z) && !(alpha.x < alpha_min);
                const bool ok1 = !(power.y > 0.0f || power.y < b.z) && !(alpha.y < alpha_min);
                const float2 oma = __ffma2_rn(alpha, make_float2(-1.0f, -1.0f), make_float2(1.0f, 1.0f));
                const float2 Tn = __fmul2_rn(T, oma);
                float2 w = __fmul2_rn(alpha, T);
                const bool live0 = ok0 && !(Tn.x < 0.0001f), live1 = ok1 && !(Tn.y < 0.0001f);
                w.x = live0 ? w.x : 0.0f;
                w.y = live1 ? w.y : 0.0f;
                const float2 cc = lds64<32>(rec);
                C0 = __ffma2_rn(make_float2(b.w, b.w), w, C0);
                C1 = __ffma2_rn(make_float2(cc.x, cc.x), w, C1);
                C2 = __ffma2_rn(make_float2(cc.y, cc.y), w, C2);
                if (want_depth) {
                    const float inv = 1.0f / lds64<40>(rec).x;
                    Dv = __ffma2_rn(make_float2(inv, inv), w, Dv);
                }
                if (ok0) T.x = live0 ? Tn.x : -fabsf(T.x);
                if (ok1) T.y = live1 ? Tn.y : -fabsf(T.y);
                if (AUX) {
                    if (live0) off0_last = (int)rec;
                    if (live1) off1_last = (int)rec;
                }
            };
            int k = 0;
#pragma unroll 1
            for (; k + 3 < cnt; k += 4) {
                const uint32_t la = list_base + 2u * (uint32_t)k;
                blend_entry(lds_u16<0>(la));
                blend_entry(lds_u16<2>(la));
                blend_entry(lds_u16<4>(la));
                blend_entry(lds_u16<6>(la));
            }
#pragma unroll 1
            for (; k < cnt; ++k) blend_entry(lds_u16<0>(list_base + 2u * (uint32_t)k));
            if (AUX) {
                if (off0_last >= 0) last0 = c * WK_CHUNK + (off0_last - (int)stage) / 48 + 1;
                if (off1_last >= 0) last1 = c * WK_CHUNK + (off1_last - (int)stage) / 48 + 1;
            }
            __syncwarp();                                        // every lane is done with this stage before it is refilled
            ja = ka; jb = kb;
            if (__all_sync(0xffffffffu, T.x <= 0.0f && T.y <= 0.0f)) break;       // quadrant saturated: stop reading
        }
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");   // nothing of this unit may land in the next unit's stages
        __syncwarp();
        if (in0) store_pixel(f, pxi, py0, fabsf(T.x), C0.x, C1.x, C2.x, last0, Dv.x);
        if (in1) store_pixel(f, pxi, py1, fabsf(T.y), C0.y, C1.y, C2.y, last1, Dv.y);
    }
}

// One-off per process: the opt-in attribute calls and device queries stay out of the per-call path.
int peer_pointer(const void *ptr)
{
    // last answer cached: a render loop hands the same output buffer in again and again
    static thread_local const void *last_ptr = nullptr;
    static thread_local int last_dev = -1, last_ans = 0;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return 0; }
    if (ptr == last_ptr && dev == last_dev) return last_ans;
    cudaPointerAttributes attr;
    int ans = 0;
    if (cudaPointerGetAttributes(&attr, ptr) == cudaSuccess) ans = attr.type == cudaMemoryTypeDevice && attr.device != dev;
    else cudaGetLastError();
    last_ptr = ptr; last_dev = dev; last_ans = ans;
    return ans;
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
    AMAV_CHECK_ARG(hwc >= AMAV_LAYOUT_CHW && hwc <= AMAV_LAYOUT_HWC_PIXELS, "bad layout");
    AMAV_CHECK_ARG((int64_t)d->frames * ((d->width + 15) / 16) * ((d->height + 15) / 16) < ((int64_t)1 << 31), "too many tiles");
    BlendParams p;
    p.F = d->frames; p.N = d->num_gaussians; p.W = d->width; p.H = d->height;
    p.gx = (d->width + 15) / 16; p.gy = (d->height + 15) / 16; p.hwc = hwc != AMAV_LAYOUT_CHW; p.clamp01 = clamp01;
    // Row stores (the colour of a pair of tiles leaves as 16 x 384 contiguous bytes) when the output buffer lives on
    // ANOTHER GPU (shard.PeerFrameGather: NVLink moves lines, not 12-byte pixels).  For local memory the plain
    // per-pixel stores are kept.  AMAV_LAYOUT_HWC asks the driver where the buffer lives (answer cached per pointer);
    // AMAV_LAYOUT_HWC_ROWS / _PIXELS let a caller who knows skip the query.
    int rows = 0;
    const bool rows_ok = d->width % 32 == 0 && ((uintptr_t)out_color & 15) == 0 && !final_T && !n_contrib && !out_invdepth;
    if (hwc == AMAV_LAYOUT_HWC_ROWS) {
        AMAV_CHECK_ARG(rows_ok, "row stores need width % 32 == 0, a 16-byte aligned buffer and no auxiliary outputs");
        rows = 1;
    } else if (hwc == AMAV_LAYOUT_HWC && rows_ok) {
        rows = peer_pointer(out_color);
    }
    p.geom = reinterpret_cast<const float4 *>(geom); p.tile_start = tile_start; p.point_list = point_list;
    p.out_color = out_color; p.final_T = final_T; p.n_contrib = n_contrib; p.out_invdepth = out_invdepth;
    p.work_counter = work_counter;
    p.deep_first = d->frames <= AMAV_DEEP_FIRST_MAX_FRAMES;
    AMAV_CHECK_ARG(capacity >= 0, "bad capacity");
    p.capacity = capacity > 0xffffffffll ? 0xffffffffu : (uint32_t)capacity;
    const int64_t total_tiles = (int64_t)d->frames * p.gx * p.gy;
    p.order = tile_order;
    p.ne_queue = AMAV_BLEND_NE_QUEUE && tile_order != nullptr && !rows;
    if (!tile_order) p.deep_first = 0;
    const bool aux = out_invdepth || final_T || n_contrib;
    if (rows || !tile_order || !AMAV_BLEND_WARP_WORKERS) {
        // CTA-per-tile kernel (the default): non-empty-tile queue for local output, tile pairs with row stores for a peer buffer
        const int grid = (int)(total_tiles < AMAV_SM_COUNT * 8 ? total_tiles : AMAV_SM_COUNT * 8);
        if (aux) blend_kernel<true, 0><<<grid, FWD_THREADS, 0, (cudaStream_t)stream>>>(p, bg);
        // (runs of 4 tiles = 768-byte rows, AMAV_ROW_TILES=4: measured on 8 GPUs, same 520 GB/s into rank 0 and a coarser queue:
        // 3.39 ms instead of 3.10 ms for the 15-frame shards)
        else if (rows && AMAV_ROW_TILES == 4 && d->width % 64 == 0) blend_kernel<false, 4><<<grid, FWD_THREADS, 0, (cudaStream_t)stream>>>(p, bg);
        else if (rows) blend_kernel<false, 2><<<grid, FWD_THREADS, 0, (cudaStream_t)stream>>>(p, bg);
        else blend_kernel<false, 0><<<grid, FWD_THREADS, 0, (cudaStream_t)stream>>>(p, bg);
    } else {
        // warp-independent workers over the same list (compile-time experiment)
        const int64_t want = (total_tiles + WK_WARPS - 1) / WK_WARPS;
        const int grid = (int)(want < AMAV_SM_COUNT * 8 ? want : AMAV_SM_COUNT * 8);
        if (aux) blend_warp_kernel<true><<<grid, WK_THREADS, 0, (cudaStream_t)stream>>>(p, bg);
        else blend_warp_kernel<false><<<grid, WK_THREADS, 0, (cudaStream_t)stream>>>(p, bg);
    }
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
