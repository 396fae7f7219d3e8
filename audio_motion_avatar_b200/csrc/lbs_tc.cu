// Stage 3, the one GEMM-shaped piece of the SMPL-X forward on the 5th-generation tensor cores:
//
//   v_posed[f][c] = v_template[c] + sum_s shape[f][s] shapedirs[c][s] + sum_k pose_feature[f][k] posedirs[k][c]
//
// (smplx.lbs.lbs: blend_shapes + the pose-corrective offsets; reference call site src/models/renderer.py:261-274.)  All three
// terms are ONE product D = A' . B' with the K axis extended: A' = [pose_feature | shape | 1] ([F, 486 + S + 1]),
// B' = [posedirs ; shapedirs^T ; v_template] ([486 + S + 1, 3V]).  At 120 frames that is 3.9 GFLOP against 61 MB of posedirs:
// compute-bound on the FP32 FMA pipe (the round-1 kernel took 0.25 ms), a few microseconds of tensor-core time.
//
//   tcgen05.mma.cta_group::1.kind::tf32, M = 128 frames x N = 64 columns x K = 8 per instruction, accumulator in TMEM (64
//   columns), 3xTF32 (every operand split into a TF32 hi part and an exact fp32 remainder lo: hi.hi + lo.hi + hi.lo), so the
//   result keeps fp32-level accuracy (template coordinates are ~1 m: single-pass TF32 would cost 2e-4 m);
//   operands staged by the threads into the canonical no-swizzle K-major core-matrix layout (A: pose features are K-major in
//   memory already; B: posedirs rows are read coalesced along the columns and written as transposed 16-byte chunks);
//   two 32-deep K stages, one mbarrier per stage armed by tcgen05.commit; the 128 x 64 tile leaves through shared memory as
//   coalesced row segments.
#include "common.cuh"

namespace {

constexpr int LT_THREADS = 256;
constexpr int LT_M = 128, LT_N = 64, LT_KB = 32;
constexpr int LT_A_BYTES = LT_M * LT_KB * 4;            // 16 KB
constexpr int LT_B_BYTES = LT_N * LT_KB * 4;            // 8 KB
constexpr int LT_STAGE_BYTES = 2 * LT_A_BYTES + 2 * LT_B_BYTES;   // 48 KB
constexpr int LT_SMEM = 2 * LT_STAGE_BYTES;             // 96 KB (the output tile reuses stage 0)
constexpr uint32_t LT_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(LT_N >> 3) << 17) | ((uint32_t)(LT_M >> 4) << 24);

__device__ __forceinline__ uint64_t lt_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes)
{
    return (uint64_t)((addr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46);
}

__device__ __forceinline__ void lt_mma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
                 :: "r"(tmem_d), "l"(da), "l"(db), "r"(LT_IDESC), "r"(accumulate) : "memory");
}

__device__ __forceinline__ void lt_wait(uint32_t bar, uint32_t parity)
{
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                 "@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}\n" :: "r"(bar), "r"(parity) : "memory");
}

__device__ __forceinline__ float lt_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

// grid = (column tiles of 64, frame tiles of 128)
__global__ void __launch_bounds__(LT_THREADS, 2)
pose_blend_tc_kernel(int F, int V, int S, int K, const float *__restrict__ pose_feature, const float *__restrict__ shape_comp,
                     const float *__restrict__ v_template, const float *__restrict__ shapedirs,
                     const float *__restrict__ posedirs, float *__restrict__ v_posed)
{
    extern __shared__ __align__(128) uint8_t s_raw[];
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int C3 = 3 * V;
    const int col0 = blockIdx.x * LT_N, f0 = blockIdx.y * LT_M;
    const int KX = K + S + 1;                                  // extended K: pose features, shape coefficients, the constant 1
    const int nkb = (KX + LT_KB - 1) / LT_KB;
    const uint32_t s_base = (uint32_t)__cvta_generic_to_shared(s_raw);
    const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(&s_bar[0]);

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;\n"
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

    // staging roles.  A: frame row m, 16 consecutive k of the block.  B: column n, 8 consecutive k of the block.
    const int am = tid & (LT_M - 1), ah = tid >> 7;
    const int bn = tid & (LT_N - 1), bq = tid >> 6;
    const int fa = f0 + am, cb = col0 + bn;
    auto a_elem = [&](int k) -> float {
        if (fa >= F || k >= KX) return 0.0f;
        if (k < K) return __ldg(pose_feature + (int64_t)fa * K + k);
        if (k < K + S) return __ldg(shape_comp + (int64_t)fa * S + (k - K));
        return 1.0f;
    };
    auto b_elem = [&](int k) -> float {
        if (cb >= C3 || k >= KX) return 0.0f;
        if (k < K) return __ldg(posedirs + (int64_t)k * C3 + cb);
        if (k < K + S) return __ldg(shapedirs + (int64_t)cb * S + (k - K));
        return __ldg(v_template + cb);
    };
    float va[16], vb[8];
    auto load = [&](int kb) {
#pragma unroll
        for (int u = 0; u < 16; ++u) va[u] = a_elem(kb * LT_KB + 16 * ah + u);
#pragma unroll
        for (int u = 0; u < 8; ++u) vb[u] = b_elem(kb * LT_KB + 8 * bq + u);
    };
    load(0);
    for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb & 1;
        uint8_t *stage = s_raw + s * LT_STAGE_BYTES;
        if (kb >= 2) lt_wait(bar0 + 8 * s, (uint32_t)((kb >> 1) - 1) & 1u);       // the MMAs of block kb - 2 have read this stage
        float4 *a_hi = reinterpret_cast<float4 *>(stage), *a_lo = reinterpret_cast<float4 *>(stage + LT_A_BYTES);
        float4 *b_hi = reinterpret_cast<float4 *>(stage + 2 * LT_A_BYTES), *b_lo = reinterpret_cast<float4 *>(stage + 2 * LT_A_BYTES + LT_B_BYTES);
#pragma unroll
        for (int q = 0; q < 4; ++q) {                            // A chunk kc = 4 ah + q: address kc * (128 * 16) + m * 16
            const float h0 = lt_hi(va[4 * q]), h1 = lt_hi(va[4 * q + 1]), h2 = lt_hi(va[4 * q + 2]), h3 = lt_hi(va[4 * q + 3]);
            const int idx = (4 * ah + q) * LT_M + am;
            a_hi[idx] = make_float4(h0, h1, h2, h3);
            a_lo[idx] = make_float4(va[4 * q] - h0, va[4 * q + 1] - h1, va[4 * q + 2] - h2, va[4 * q + 3] - h3);
        }
#pragma unroll
        for (int q = 0; q < 2; ++q) {                            // B chunk kc = 2 bq + q: address kc * (64 * 16) + n * 16
            const float h0 = lt_hi(vb[4 * q]), h1 = lt_hi(vb[4 * q + 1]), h2 = lt_hi(vb[4 * q + 2]), h3 = lt_hi(vb[4 * q + 3]);
            const int idx = (2 * bq + q) * LT_N + bn;
            b_hi[idx] = make_float4(h0, h1, h2, h3);
            b_lo[idx] = make_float4(vb[4 * q] - h0, vb[4 * q + 1] - h1, vb[4 * q + 2] - h2, vb[4 * q + 3] - h3);
        }
        if (kb + 1 < nkb) load(kb + 1);                          // in flight during the barrier and the MMA issue
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
            const uint32_t a_hi_s = s_base + s * LT_STAGE_BYTES, a_lo_s = a_hi_s + LT_A_BYTES;
            const uint32_t b_hi_s = a_hi_s + 2 * LT_A_BYTES, b_lo_s = b_hi_s + LT_B_BYTES;
#pragma unroll
            for (int j = 0; j < LT_KB / 8; ++j) {
                const uint64_t dah = lt_desc(a_hi_s + 2 * j * (LT_M * 16), LT_M * 16, 128);
                const uint64_t dal = lt_desc(a_lo_s + 2 * j * (LT_M * 16), LT_M * 16, 128);
                const uint64_t dbh = lt_desc(b_hi_s + 2 * j * (LT_N * 16), LT_N * 16, 128);
                const uint64_t dbl = lt_desc(b_lo_s + 2 * j * (LT_N * 16), LT_N * 16, 128);
                lt_mma(tmem, dal, dbh, (kb | j) != 0);
                lt_mma(tmem, dah, dbl, 1u);
                lt_mma(tmem, dah, dbh, 1u);
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n"
                         :: "r"(bar0 + 8 * s) : "memory");
        }
    }
    lt_wait(bar0 + 8 * ((nkb - 1) & 1), (uint32_t)((nkb - 1) >> 1) & 1u);        // the last commit covers every MMA
    if (nkb >= 2) lt_wait(bar0 + 8 * (nkb & 1), (uint32_t)((nkb - 2) >> 1) & 1u);
    asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
    // epilogue: TMEM -> registers (warp w owns TMEM lanes 32 w ..) -> shared [128][65] -> coalesced row segments
    float *s_out = reinterpret_cast<float *>(s_raw);
    if (warp < 4) {
#pragma unroll
        for (int part = 0; part < 4; ++part) {
            uint32_t r[16];
            const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(part * 16);
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                           "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                         : "r"(taddr) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
#pragma unroll
            for (int q = 0; q < 16; ++q) s_out[tid * (LT_N + 1) + part * 16 + q] = __uint_as_float(r[q]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    for (int i = tid; i < LT_M * LT_N; i += LT_THREADS) {
        const int m = i >> 6, n = i & (LT_N - 1);
        if (f0 + m < F && col0 + n < C3) v_posed[(int64_t)(f0 + m) * C3 + col0 + n] = s_out[m * (LT_N + 1) + n];
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;\n" :: "r"(tmem) : "memory");
}

}  // namespace

// Internal to the library (declared in common.cuh): the tensor-core form of pose_blend_kernel.
int amav_pose_blend_tc(int F, int V, int S, int K, const float *pose_feature, const float *shape_comp, const float *v_template,
                       const float *shapedirs, const float *posedirs, float *v_posed, cudaStream_t stream)
{
    amav_ensure_dyn_smem((const void *)pose_blend_tc_kernel, LT_SMEM);
    dim3 grid(amav_div_up((int64_t)3 * V, LT_N), amav_div_up(F, LT_M));
    pose_blend_tc_kernel<<<grid, LT_THREADS, LT_SMEM, stream>>>(F, V, S, K, pose_feature, shape_comp, v_template, shapedirs,
                                                                posedirs, v_posed);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
