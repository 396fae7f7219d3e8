// Stage 2a on the 5th-generation tensor cores: the decoder heads applied to every triplane texel
// (reference: the five nn.Linear at src/models/renderer.py:165-171, fused with the query by linearity -- see triplane.cu).
//
//   D[128 texels x 16 slots] += A[128 x 8 channels] . B[8 x 16]        tcgen05.mma  kind::tf32, cta_group::1
//
// * accumulator in TMEM (32 columns allocated, 16 used), read back with tcgen05.ld for the epilogue;
// * fp32 accuracy out of TF32 operands: every fp32 value x is split into hi = x with the low 13 mantissa bits cleared
//   (exactly a TF32 number) and lo = x - hi (exact in fp32, <= 13 significant bits, so its own TF32 truncation loses
//   <= 2^-21 |x|); the product is hi.hi + lo.hi + hi.lo accumulated in fp32 -- three MMAs per K step.  Single-pass TF32
//   would cost 1.4e-3 on the decoded attributes, above north_star's 1e-3;
// * operands are staged in shared memory in the canonical no-swizzle K-major layout (8 x 16-byte core matrices): the
//   token tensor is channel-major ([C][3 R^2]: the MMA's K runs along the slow axis), so the staging threads read
//   coalesced rows of 128 texels and write transposed 16-byte chunks of 4 channels -- the transpose is the reason the
//   tile is not a plain TMA box copy (a TMA box of an MN-major fp32 tile lands in the MN-major canonical layout; that
//   variant needs no thread staging for `hi` but still needs the threads for `lo`);
// * two stages of 32 channels, one mbarrier per stage armed by tcgen05.commit: the staging of K block k+1 overlaps the
//   MMAs of block k; 72 KB of shared memory -> 3 CTAs per SM.
// The kernel stays HBM-bound by design (25 MB of tokens per frame against 0.2 GFLOP); what the tensor core buys is the
// issue slots: 3.5 instructions per token element instead of 14 FFMAs.
#include "common.cuh"

#include <mutex>

namespace {

__constant__ int c_slot_row_tc[16] = {3, 4, 5, 6, 0, 1, 2, 10, 7, 8, 9, -1, 11, 12, 13, -1};

constexpr int TC_THREADS = 256;
constexpr int TC_TEX = 128;                 // UMMA M
constexpr int TC_N = 16;                    // UMMA N (14 decoder outputs + 2 zero slots)
constexpr int TC_KB = 32;                   // channels per stage
constexpr int TC_A_BYTES = TC_TEX * TC_KB * 4;      // 16 KB
constexpr int TC_B_BYTES = TC_N * TC_KB * 4;        // 2 KB
constexpr int TC_STAGE_BYTES = 2 * TC_A_BYTES + 2 * TC_B_BYTES;     // A hi, A lo, B hi, B lo = 36 KB
constexpr int TC_SMEM = 2 * TC_STAGE_BYTES;                         // 72 KB
constexpr uint32_t TC_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TC_N >> 3) << 17) | ((uint32_t)(TC_TEX >> 4) << 24);
// c_format F32 | a_format TF32 | b_format TF32 | K-major A and B | N >> 3 | M >> 4   (cute::UMMA::InstrDescriptor)

__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes)
{
    // cute::UMMA::SmemDescriptor: start address >> 4 [0,14), leading byte offset >> 4 [16,30), stride byte offset >> 4
    // [32,46), version = 1 [46,48), layout type SWIZZLE_NONE = 0 [61,64)
    return (uint64_t)((addr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46);
}

__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
                 :: "r"(tmem_d), "l"(da), "l"(db), "r"(TC_IDESC), "r"(accumulate) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                 "@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}\n" :: "r"(bar), "r"(parity) : "memory");
}

__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

// Decoder weights of plane p, K block kb, split into hi / lo and laid out as the B operand wants them:
// wprep[((p * nkb + kb) * 2 + part) * 512 + (kc * 16 + n) * 4 + j] = part(W[slot_row[n]][3 + p C + kb 32 + kc 4 + j])
__global__ void weight_prep_kernel(int C, const float *__restrict__ weight, int ldw, float *__restrict__ wprep)
{
    const int nkb = C / TC_KB;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 3 * nkb * 512) return;
    const int j = i & 3, n = (i >> 2) & 15, kc = (i >> 6) & 7, kb = (i >> 9) % nkb, p = (i >> 9) / nkb;
    const int row = c_slot_row_tc[n];
    const float w = row < 0 ? 0.0f : weight[(int64_t)row * ldw + 3 + p * C + kb * TC_KB + kc * 4 + j];
    const float hi = tf32_hi(w);
    float *dst = wprep + (int64_t)(p * nkb + kb) * 1024 + (i & 511);
    dst[0] = hi;
    dst[512] = w - hi;
}

__global__ void __launch_bounds__(TC_THREADS, 3)
project_tc_kernel(int C, int R2, const float *__restrict__ tokens, int64_t frame_stride,
                  const float *__restrict__ wprep, float *__restrict__ proj)
{
    extern __shared__ __align__(128) uint8_t s_raw[];
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int tiles_per_plane = R2 / TC_TEX;
    const int tile = blockIdx.x % tiles_per_plane;
    const int p = (blockIdx.x / tiles_per_plane) % 3;
    const int f = blockIdx.x / (3 * tiles_per_plane);
    const int nkb = C / TC_KB;
    const uint32_t s_base = (uint32_t)__cvta_generic_to_shared(s_raw);
    const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(&s_bar[0]);

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;\n"
                     :: "r"((uint32_t)__cvta_generic_to_shared(&s_tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    if (tid == 32) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" :: "r"(bar0) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" :: "r"(bar0 + 8) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
    const uint32_t tmem = s_tmem;

    // staging role of this thread: texel m, channels [16 h, 16 h + 16) of every K block
    const int m = tid & (TC_TEX - 1), h = tid >> 7;
    const float *src = tokens + (int64_t)f * frame_stride + (int64_t)p * R2 + tile * TC_TEX + m + (int64_t)(16 * h) * (3 * R2);
    const float4 *wsrc = reinterpret_cast<const float4 *>(wprep + (int64_t)p * nkb * 1024) + tid;   // 256 float4 = hi + lo of a K block

    // two K blocks of loads are kept in flight per thread (register double buffer): the loads of block kb + 2 are issued
    // as soon as block kb has been written to shared memory
    float va[16], vb[16];
    float4 wa, wb;
    auto load = [&](int kb, float (&v)[16], float4 &wv) {
#pragma unroll
        for (int u = 0; u < 16; ++u) v[u] = __ldcs(src + (int64_t)(kb * TC_KB + u) * (3 * R2));
        wv = __ldg(wsrc + kb * 256);
    };
    auto step = [&](int kb, float (&v)[16], float4 &wv) {
        const int s = kb & 1;
        uint8_t *stage = s_raw + s * TC_STAGE_BYTES;
        if (kb >= 2) mbar_wait(bar0 + 8 * s, (uint32_t)((kb >> 1) - 1) & 1u);     // the MMAs of block kb - 2 have read this stage
        float4 *a_hi = reinterpret_cast<float4 *>(stage), *a_lo = reinterpret_cast<float4 *>(stage + TC_A_BYTES);
#pragma unroll
        for (int q = 0; q < 4; ++q) {                            // chunk kc = 4 h + q: 4 channels of texel m
            const float h0 = tf32_hi(v[4 * q]), h1 = tf32_hi(v[4 * q + 1]), h2 = tf32_hi(v[4 * q + 2]), h3 = tf32_hi(v[4 * q + 3]);
            const int idx = (4 * h + q) * TC_TEX + m;
            a_hi[idx] = make_float4(h0, h1, h2, h3);
            a_lo[idx] = make_float4(v[4 * q] - h0, v[4 * q + 1] - h1, v[4 * q + 2] - h2, v[4 * q + 3] - h3);
        }
        reinterpret_cast<float4 *>(stage + 2 * TC_A_BYTES)[tid] = wv;            // B hi (tid < 128) and B lo (tid >= 128)
        if (kb + 2 < nkb) load(kb + 2, v, wv);
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
            const uint32_t a_hi_s = s_base + s * TC_STAGE_BYTES, a_lo_s = a_hi_s + TC_A_BYTES;
            const uint32_t b_hi_s = a_hi_s + 2 * TC_A_BYTES, b_lo_s = b_hi_s + TC_B_BYTES;
#pragma unroll
            for (int j = 0; j < TC_KB / 8; ++j) {                // K = 8 per MMA: chunks 2j, 2j+1
                const uint64_t dah = smem_desc(a_hi_s + 2 * j * (TC_TEX * 16), TC_TEX * 16, 128);
                const uint64_t dal = smem_desc(a_lo_s + 2 * j * (TC_TEX * 16), TC_TEX * 16, 128);
                const uint64_t dbh = smem_desc(b_hi_s + 2 * j * (TC_N * 16), TC_N * 16, 128);
                const uint64_t dbl = smem_desc(b_lo_s + 2 * j * (TC_N * 16), TC_N * 16, 128);
                mma_tf32(tmem, dal, dbh, (kb | j) != 0);         // small terms first
                mma_tf32(tmem, dah, dbl, 1u);
                mma_tf32(tmem, dah, dbh, 1u);
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n"
                         :: "r"(bar0 + 8 * s) : "memory");
        }
    };
    load(0, va, wa);
    if (nkb > 1) load(1, vb, wb);
    for (int kb = 0; kb < nkb; kb += 2) {
        step(kb, va, wa);
        if (kb + 1 < nkb) step(kb + 1, vb, wb);
    }
    // the last commit covers every MMA issued before it
    mbar_wait(bar0 + 8 * ((nkb - 1) & 1), (uint32_t)((nkb - 1) >> 1) & 1u);
    asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
    if (warp < 4) {                                              // warp w reads TMEM lanes 32 w .. 32 w + 31 = its texels
        uint32_t r[16];
        const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                       "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                     : "r"(taddr) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
        float4 *out = reinterpret_cast<float4 *>(proj + (((int64_t)f * 3 + p) * R2 + tile * TC_TEX + tid) * 16);
#pragma unroll
        for (int q = 0; q < 4; ++q)
            out[q] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]), __uint_as_float(r[4 * q + 2]),
                                 __uint_as_float(r[4 * q + 3]));
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;\n" :: "r"(tmem) : "memory");
}

}  // namespace

extern "C" int64_t amav_triplane_project_tc_scratch_elems(int C) { return (int64_t)3 * (C / TC_KB) * 1024; }

extern "C" int amav_triplane_project_tc(int F, int C, int R, const float *tokens, int64_t token_frame_stride,
                                        const float *weight, float *wprep_scratch, float *proj, void *stream_)
{
    AMAV_CHECK_ARG(F >= 0 && C > 0 && R > 0, "bad dims");
    AMAV_CHECK_ARG((R * R) % TC_TEX == 0, "R*R must be a multiple of 128");
    AMAV_CHECK_ARG(C % TC_KB == 0, "C must be a multiple of 32");
    AMAV_CHECK_ARG(token_frame_stride >= (int64_t)C * 3 * R * R, "frame stride too small");
    if (F == 0) return AMAV_OK;
    AMAV_CHECK_ARG(tokens && weight && proj && wprep_scratch, "null buffer");
    AMAV_CHECK_ARG(((uintptr_t)wprep_scratch & 15) == 0 && ((uintptr_t)proj & 15) == 0, "buffers must be 16-byte aligned");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int n = 3 * (C / TC_KB) * 512;
    weight_prep_kernel<<<amav_div_up(n, 256), 256, 0, stream>>>(C, weight, 3 + 3 * C, wprep_scratch);
    AMAV_CHECK_LAUNCH();
    amav_ensure_dyn_smem((const void *)project_tc_kernel, TC_SMEM);
    const int64_t grid = (int64_t)F * 3 * (R * R / TC_TEX);
    AMAV_CHECK_ARG(grid < ((int64_t)1 << 31), "too many tiles");
    project_tc_kernel<<<(unsigned)grid, TC_THREADS, TC_SMEM, stream>>>(C, R * R, tokens, token_frame_stride, wprep_scratch, proj);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// Allocation-free variant (no counterpart in the reference): the split
// weights live in a file-scope device array instead of caller-owned scratch, for callers that cannot allocate (CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  192 KB per GPU that loads the library, held
// until exit.  Every call shares the array: weight_prep_kernel rewrites all of the 3 (C/32) 1024 floats the projection
// reads (nothing has to start at zero), and two calls that overlap on the GPU (different streams) race on it.
constexpr int TC_STATIC_MAX_C = 512;
constexpr int TC_STATIC_ELEMS = 3 * (TC_STATIC_MAX_C / TC_KB) * 1024;
static __device__ __align__(256) float g_wprep_static[TC_STATIC_ELEMS];

namespace {
constexpr int TC_MAX_DEVICES = 64;
std::mutex g_static_mu;
float *g_static_ptr[TC_MAX_DEVICES] = {};

// device address of g_wprep_static on the CURRENT device; the symbol's name in host code is only a handle, so the
// address is looked up once per device (with that device current) and cached
cudaError_t static_scratch(float **out)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= TC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[dev]) {
        float *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_wprep_static);
        if (e != cudaSuccess) return e;
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return cudaSuccess;
}
}  // namespace

extern "C" int amav_triplane_project_tc_static(int F, int C, int R, const float *tokens, int64_t token_frame_stride,
                                               const float *weight, float *proj, void *stream_)
{
    AMAV_CHECK_ARG(F >= 0 && C > 0 && R > 0, "bad dims");
    AMAV_CHECK_ARG(C <= TC_STATIC_MAX_C, "C exceeds the static scratch (512 channels max): use amav_triplane_project_tc with caller-owned scratch");
    AMAV_CHECK_ARG((R * R) % TC_TEX == 0, "R*R must be a multiple of 128");
    AMAV_CHECK_ARG(C % TC_KB == 0, "C must be a multiple of 32");
    AMAV_CHECK_ARG(token_frame_stride >= (int64_t)C * 3 * R * R, "frame stride too small");
    if (F == 0) return AMAV_OK;
    AMAV_CHECK_ARG(tokens && weight && proj, "null buffer");
    AMAV_CHECK_ARG(((uintptr_t)proj & 15) == 0, "buffers must be 16-byte aligned");
    const int64_t grid = (int64_t)F * 3 * (R * R / TC_TEX);
    AMAV_CHECK_ARG(grid < ((int64_t)1 << 31), "too many tiles");
    float *wprep = nullptr;
    cudaError_t e = static_scratch(&wprep);
    if (e != cudaSuccess || !wprep) {
        amav_set_error("%s: cudaGetSymbolAddress(static scratch) failed: %s", __func__, cudaGetErrorString(e));
        return AMAV_ECUDA;
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    const int n = 3 * (C / TC_KB) * 512;
    weight_prep_kernel<<<amav_div_up(n, 256), 256, 0, stream>>>(C, weight, 3 + 3 * C, wprep);
    AMAV_CHECK_LAUNCH();
    e = amav_ensure_dyn_smem((const void *)project_tc_kernel, TC_SMEM);
    if (e != cudaSuccess) {
        amav_set_error("%s: cudaFuncSetAttribute failed: %s", __func__, cudaGetErrorString(e));
        return AMAV_ECUDA;
    }
    project_tc_kernel<<<(unsigned)grid, TC_THREADS, TC_SMEM, stream>>>(C, R * R, tokens, token_frame_stride, wprep, proj);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
