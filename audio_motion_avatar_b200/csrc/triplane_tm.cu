// Producer-side layout of the triplane.  The upstream transformer computes its output token-major,
// hidden_states [B, 3*R*R, C] = proj_out(...), and only then pays `.permute(0, 2, 1).contiguous()` to hand the
// renderer the channel-major [B, C, 3*R*R] tensor (src/models/transformers.py:1070-1073): 25 MB read + 25 MB written per
// frame for a transposition the projection does not need.  project_tm_kernel consumes the token-major tensor as it
// is: same output (projected planes [F,3,R*R,16], triplane.cu) from tokens_tm [F, 3*R*R, C].
//
//   CTA = 64 texels of one plane x all C channels, 256 threads.  A texel's C channels are contiguous (2 KB), so a
//   [64 texels x 64 channels] chunk is loaded as 256-byte row segments (float4 per thread, streaming loads, the next
//   chunk prefetched into registers during the FMAs) and parked in shared memory with a row stride of 65 floats; the
//   FMAs then run with lane = texel (conflict-free column reads) and the weights broadcast from smem [C][16], the
//   work decomposition of project_kernel: warp w owns 8 of the chunk's 64 channels, 2 texels x 14 accumulators per lane,
//   cross-warp reduction through shared memory at the end.  HBM-bound: the 25.2 MB/frame of tokens are read once.
#include "common.cuh"

namespace {

__constant__ int c_slot_row_tm[16] = {3, 4, 5, 6, 0, 1, 2, 10, 7, 8, 9, -1, 11, 12, 13, -1};

constexpr int TM_THREADS = 256;
constexpr int TM_TEX = 64;             // texels per CTA
constexpr int TM_KC = 64;              // channels per staged chunk
constexpr int TM_LD = TM_KC + 1;       // row stride of the staged chunk (floats)
constexpr int TM_WARPS = TM_THREADS / 32;
constexpr int TM_CPW = TM_KC / TM_WARPS;   // channels per warp per chunk

__device__ __forceinline__ float4 ldcs4(const float *p) { return __ldcs(reinterpret_cast<const float4 *>(p)); }

__global__ void __launch_bounds__(TM_THREADS, 3)
project_tm_kernel(int C, int R2, const float *__restrict__ tokens, int64_t frame_stride,
                  const float *__restrict__ weight, int ldw, float *__restrict__ proj)
{
    extern __shared__ float4 s_dyn_tm[];
    float *s_w = reinterpret_cast<float *>(s_dyn_tm);                // [C][16]
    float *s_t = s_w + (size_t)C * 16;                               // [TM_TEX][TM_LD], later red[8][14][TM_TEX]
    const int tiles_per_plane = R2 / TM_TEX;
    const int tile = blockIdx.x % tiles_per_plane;
    const int p = (blockIdx.x / tiles_per_plane) % 3;
    const int f = blockIdx.x / (3 * tiles_per_plane);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // this thread's share of a chunk: 4 float4 = texel rows (threadIdx.x >> 4) + 16 q, channels 4 * (threadIdx.x & 15)
    const int lrow = threadIdx.x >> 4, lcol = (threadIdx.x & 15) * 4;
    const float *src = tokens + (int64_t)f * frame_stride + ((int64_t)p * R2 + (int64_t)tile * TM_TEX + lrow) * C + lcol;
    float4 v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) v[q] = ldcs4(src + (int64_t)(16 * q) * C);

    for (int i = threadIdx.x; i < C * 16; i += TM_THREADS) {
        const int c = i >> 4, s = i & 15;
        const int row = c_slot_row_tm[s];
        s_w[i] = row < 0 ? 0.0f : weight[(int64_t)row * ldw + 3 + p * C + c];
    }

    float acc[2][14];
#pragma unroll
    for (int t = 0; t < 2; ++t)
#pragma unroll
        for (int o = 0; o < 14; ++o) acc[t][o] = 0.0f;

    for (int c0 = 0; c0 < C; c0 += TM_KC) {
        __syncthreads();                                             // previous chunk consumed (first: nothing)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float *d = s_t + (lrow + 16 * q) * TM_LD + lcol;
            d[0] = v[q].x; d[1] = v[q].y; d[2] = v[q].z; d[3] = v[q].w;
        }
        __syncthreads();                                             // chunk (and, first time, the weights) visible
        if (c0 + TM_KC < C) {
#pragma unroll
            for (int q = 0; q < 4; ++q) v[q] = ldcs4(src + (int64_t)(16 * q) * C + c0 + TM_KC);
        }
#pragma unroll
        for (int u = 0; u < TM_CPW; ++u) {
            const int cc = warp * TM_CPW + u;
            const float4 *wp = reinterpret_cast<const float4 *>(s_w + (size_t)(c0 + cc) * 16);
            const float4 w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3];
            const float w[14] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x, w2.y, w2.z, w3.x, w3.y, w3.z};
            const float t0 = s_t[lane * TM_LD + cc], t1 = s_t[(lane + 32) * TM_LD + cc];
#pragma unroll
            for (int o = 0; o < 14; ++o) {
                acc[0][o] = fmaf(t0, w[o], acc[0][o]);
                acc[1][o] = fmaf(t1, w[o], acc[1][o]);
            }
        }
    }
    __syncthreads();
    // cross-warp (channel slice) reduction through shared memory: red[warp][o][texel]
    float *s_red = s_t;
#pragma unroll
    for (int o = 0; o < 14; ++o) {
        s_red[((size_t)warp * 14 + o) * TM_TEX + lane] = acc[0][o];
        s_red[((size_t)warp * 14 + o) * TM_TEX + lane + 32] = acc[1][o];
    }
    __syncthreads();
    // thread -> (texel, quarter of the 16 slots): one float4 store per thread, a texel's 64 bytes from 4 threads
    const int tex = threadIdx.x & (TM_TEX - 1), quarter = threadIdx.x >> 6;
    // compact output index of each slot of this quarter (slots 11 and 15 are zero padding)
    float r[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int slot = quarter * 4 + k;
        const int o = slot < 11 ? slot : slot - 1;                   // slots 0..10 -> 0..10, 12..14 -> 11..13
        if (slot != 11 && slot != 15) {
            float s = 0.0f;
            for (int g = 0; g < TM_WARPS; ++g) s += s_red[((size_t)g * 14 + o) * TM_TEX + tex];
            r[k] = s;
        }
    }
    float *out = proj + (((int64_t)f * 3 + p) * R2 + (int64_t)tile * TM_TEX + tex) * 16;
    reinterpret_cast<float4 *>(out)[quarter] = make_float4(r[0], r[1], r[2], r[3]);
}

}  // namespace

extern "C" int amav_triplane_project_tm(int F, int C, int R, const float *tokens_tm, int64_t token_frame_stride,
                                        const float *weight, float *proj, void *stream)
{
    AMAV_CHECK_ARG(F >= 0 && C > 0 && R > 0, "bad dims");
    AMAV_CHECK_ARG((R * R) % TM_TEX == 0, "R*R must be a multiple of 64");
    AMAV_CHECK_ARG(C % TM_KC == 0, "C must be a multiple of 64");
    AMAV_CHECK_ARG(token_frame_stride >= (int64_t)C * 3 * R * R, "frame stride too small");
    if (F == 0) return AMAV_OK;
    AMAV_CHECK_ARG(tokens_tm && weight && proj, "null buffer");
    AMAV_CHECK_ARG(((uintptr_t)tokens_tm & 15) == 0 && (token_frame_stride & 3) == 0, "tokens must be 16-byte aligned");
    const size_t tile_floats = (size_t)TM_TEX * TM_LD, red_floats = (size_t)TM_WARPS * 14 * TM_TEX;
    const size_t smem = ((size_t)C * 16 + (tile_floats > red_floats ? tile_floats : red_floats)) * sizeof(float);
    AMAV_CHECK_ARG(smem <= 200 * 1024, "C too large for the shared-memory weight tile");
    cudaFuncSetAttribute(project_tm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int64_t grid = (int64_t)F * 3 * (R * R / TM_TEX);
    AMAV_CHECK_ARG(grid < ((int64_t)1 << 31), "too many tiles");
    project_tm_kernel<<<(unsigned)grid, TM_THREADS, smem, (cudaStream_t)stream>>>(C, R * R, tokens_tm, token_frame_stride,
                                                                                  weight, 3 + 3 * C, proj);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
