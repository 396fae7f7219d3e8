// Producer-side layout of the triplane.  The upstream transformer computes its output token-major,
// hidden_states [B, 3*R*R, C] = proj_out(...), and only then pays `.permute(0, 2, 1).contiguous()` to hand the
// renderer the channel-major [B, C, 3*R*R] tensor (src/models/transformers.py:1070-1073): 25 MB read + 25 MB written per
// frame for a transposition the projection does not need.  project_tm_kernel consumes the token-major tensor as it
// is: same output (projected planes [F,3,R*R,16], triplane.cu) from tokens_tm [F, 3*R*R, C].
//
//   CTA = 128 texels of one plane x all C channels, 256 threads.  A texel's C channels are contiguous (2 KB), so a
//   [128 texels x 32 channels] chunk is 128-byte row segments: cp.async (16 bytes per request, L1 bypassed) drops them
//   into a four-slot ring of shared-memory tiles (row stride 36 floats), three chunks in flight while one is in the FMAs
//   -- no staging registers.  The FMAs run with lane = texel (4 texels per lane, LDS.128 of 4 channels at a time:
//   4*lane + c covers all 32 banks per quarter-warp) and the weights broadcast from smem [C][16], the work decomposition
//   of project_kernel: warp w owns 4 of the chunk's 32 channels, 4 texels x 14 accumulators per lane, cross-warp
//   reduction through shared memory at the end.  HBM-bound: the 25.2 MB/frame of tokens are read once.
#include "common.cuh"

namespace {

__constant__ int c_slot_row_tm[16] = {3, 4, 5, 6, 0, 1, 2, 10, 7, 8, 9, -1, 11, 12, 13, -1};

constexpr int TM_THREADS = 256;
constexpr int TM_TEX = 128;            // texels per CTA
constexpr int TM_KC = 32;              // channels per staged chunk
constexpr int TM_STAGES = 4;           // staged chunks in shared memory (TM_STAGES - 1 in flight during the FMAs)
constexpr int TM_LD = TM_KC + 4;       // row stride of the staged chunk (floats; rows stay 16-byte aligned)
constexpr int TM_WARPS = TM_THREADS / 32;
constexpr int TM_CPW = TM_KC / TM_WARPS;   // channels per warp per chunk (4 = one float4)
constexpr int TM_RPT = TM_KC / 4;          // 16-byte requests per texel row of a chunk
constexpr int TM_RSTEP = TM_THREADS / TM_RPT;   // texel rows covered by one request per thread
constexpr int TM_REQ = TM_TEX / TM_RSTEP;       // requests per thread and chunk
constexpr int TM_TPL = TM_TEX / 32;        // texels per lane
constexpr int TM_TILE = TM_TEX * TM_LD;    // floats per staged chunk

__device__ __forceinline__ float4 ldcs4(const float *p) { return __ldcs(reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ void cp_async16(float *smem_dst, const float *gmem_src)
{
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__global__ void __launch_bounds__(TM_THREADS, 2)
project_tm_kernel(int C, int R2, const float *__restrict__ tokens, int64_t frame_stride,
                  const float *__restrict__ weight, int ldw, float *__restrict__ proj)
{
    extern __shared__ float4 s_dyn_tm[];
    float *s_w = reinterpret_cast<float *>(s_dyn_tm);                // [C][16]
    float *s_t = s_w + (size_t)C * 16;                               // TM_STAGES x [TM_TEX][TM_LD], later red[8][14][TM_TEX]
    const int tiles_per_plane = R2 / TM_TEX;
    const int tile = blockIdx.x % tiles_per_plane;
    const int p = (blockIdx.x / tiles_per_plane) % 3;
    const int f = blockIdx.x / (3 * tiles_per_plane);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // this thread's share of a chunk: TM_REQ requests of 16 bytes = texel rows lrow + TM_RSTEP q, channels lcol..lcol+3
    const int lrow = threadIdx.x / TM_RPT, lcol = (threadIdx.x % TM_RPT) * 4;
    const float *src = tokens + (int64_t)f * frame_stride + ((int64_t)p * R2 + (int64_t)tile * TM_TEX + lrow) * C + lcol;
    float *dst = s_t + lrow * TM_LD + lcol;
    const int nchunks = C / TM_KC;
#pragma unroll
    for (int st = 0; st < TM_STAGES - 1; ++st) {
        if (st < nchunks) {
#pragma unroll
            for (int q = 0; q < TM_REQ; ++q)
                cp_async16(dst + st * TM_TILE + TM_RSTEP * q * TM_LD, src + (int64_t)(TM_RSTEP * q) * C + st * TM_KC);
        }
        cp_async_commit();                                           // one group per chunk slot, empty or not
    }

    for (int i = threadIdx.x; i < C * 16; i += TM_THREADS) {
        const int c = i >> 4, s = i & 15;
        const int row = c_slot_row_tm[s];
        s_w[i] = row < 0 ? 0.0f : weight[(int64_t)row * ldw + 3 + p * C + c];
    }

    float acc[TM_TPL][14];
#pragma unroll
    for (int t = 0; t < TM_TPL; ++t)
#pragma unroll
        for (int o = 0; o < 14; ++o) acc[t][o] = 0.0f;

    int buf = 0;
    for (int k = 0; k < nchunks; ++k) {
        cp_async_wait<TM_STAGES - 2>();  // chunk k has landed (this thread's requests) ...
        __syncthreads();                 // ... everybody's; first time also the weights; buffer of chunk k-1 is free
        {
            const int kn = k + TM_STAGES - 1;
            if (kn < nchunks) {
                float *dn = dst + (kn % TM_STAGES) * TM_TILE;
#pragma unroll
                for (int q = 0; q < TM_REQ; ++q)
                    cp_async16(dn + TM_RSTEP * q * TM_LD, src + (int64_t)(TM_RSTEP * q) * C + kn * TM_KC);
            }
            cp_async_commit();
        }
        const float *tb = s_t + buf * TM_TILE + warp * TM_CPW;
        const int c0 = k * TM_KC;
#pragma unroll
        for (int h = 0; h < TM_CPW / 4; ++h) {
            float4 tv[TM_TPL];
#pragma unroll
            for (int t = 0; t < TM_TPL; ++t)
                tv[t] = *reinterpret_cast<const float4 *>(tb + (lane + 32 * t) * TM_LD + 4 * h);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float4 *wp = reinterpret_cast<const float4 *>(s_w + (size_t)(c0 + warp * TM_CPW + 4 * h + u) * 16);
                const float4 w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3];
                const float w[14] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x, w2.y, w2.z, w3.x, w3.y, w3.z};
#pragma unroll
                for (int t = 0; t < TM_TPL; ++t) {
                    const float x = u == 0 ? tv[t].x : u == 1 ? tv[t].y : u == 2 ? tv[t].z : tv[t].w;
#pragma unroll
                    for (int o = 0; o < 14; ++o) acc[t][o] = fmaf(x, w[o], acc[t][o]);
                }
            }
        }
        buf = buf + 1 == TM_STAGES ? 0 : buf + 1;
    }
    cp_async_wait<0>();
    __syncthreads();
    // cross-warp (channel slice) reduction through shared memory: red[warp][o][texel]
    float *s_red = s_t;
#pragma unroll
    for (int o = 0; o < 14; ++o)
#pragma unroll
        for (int t = 0; t < TM_TPL; ++t) s_red[((size_t)warp * 14 + o) * TM_TEX + lane + 32 * t] = acc[t][o];
    __syncthreads();
    // thread -> (texel, half of the 16 slots): two float4 stores per thread (slots 11 and 15 are zero padding)
    const int tex = threadIdx.x & (TM_TEX - 1), half = threadIdx.x >> 7;
    float r[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int slot = half * 8 + k;
        const int o = slot < 11 ? slot : slot - 1;                   // slots 0..10 -> 0..10, 12..14 -> 11..13
        float s = 0.0f;
        if (slot != 11 && slot != 15)
            for (int g = 0; g < TM_WARPS; ++g) s += s_red[((size_t)g * 14 + o) * TM_TEX + tex];
        r[k] = s;
    }
    float *out = proj + (((int64_t)f * 3 + p) * R2 + (int64_t)tile * TM_TEX + tex) * 16;
    reinterpret_cast<float4 *>(out)[half * 2] = make_float4(r[0], r[1], r[2], r[3]);
    reinterpret_cast<float4 *>(out)[half * 2 + 1] = make_float4(r[4], r[5], r[6], r[7]);
}


// ---------------------------------------------------------------------------------------------
// Backward of the token-major projection (autograd of the five Linear heads on the feature columns, in the producer's
// layout): d_tokens_tm[f][p*R2+t][c] = sum_slot W[row(slot)][3+p*C+c] * d_proj[f][p][t][slot] and
// d_weight[row(slot)][3+p*C+c] += sum_{f,t} d_proj[f][p][t][slot] * tokens_tm[f][p*R2+t][c].
// Both: CTA = TB_TEX texels of one plane, thread = (texel half, float4 channel column), the d_proj tile broadcast from
// shared memory, 4 channels x 14 slots of weights (tokens kernel) or accumulators (weight kernel) in registers; a
// texel's C channels leave / arrive as one contiguous row.
constexpr int TB_THREADS = 256;
constexpr int TB_TEX = 128;
constexpr int TB_COLS = 128;           // float4 channel columns handled per pass (C = 512: one pass)

__global__ void __launch_bounds__(TB_THREADS)
project_tm_bwd_tokens_kernel(int C, int R2, const float *__restrict__ d_proj, const float *__restrict__ weight, int ldw,
                             float *__restrict__ d_tokens, int64_t frame_stride, int accumulate)
{
    __shared__ float4 s_dp[TB_TEX * 4];                              // [texel][16]
    const int tiles_per_plane = R2 / TB_TEX;
    const int tile = blockIdx.x % tiles_per_plane;
    const int p = (blockIdx.x / tiles_per_plane) % 3;
    const int f = blockIdx.x / (3 * tiles_per_plane);
    const float4 *dp = reinterpret_cast<const float4 *>(d_proj) + (((int64_t)f * 3 + p) * R2 + (int64_t)tile * TB_TEX) * 4;
    for (int i = threadIdx.x; i < TB_TEX * 4; i += TB_THREADS) s_dp[i] = dp[i];
    __syncthreads();
    const int half = threadIdx.x / TB_COLS, col0 = threadIdx.x % TB_COLS;
    float *dst = d_tokens + (int64_t)f * frame_stride + ((int64_t)p * R2 + (int64_t)tile * TB_TEX) * C;
    for (int col = col0; col < C / 4; col += TB_COLS) {
        float w[4][14];                                              // [channel of the float4][compact slot]
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int s = 0; s < 16; ++s) {
                const int row = c_slot_row_tm[s];
                if (row >= 0) w[j][s < 11 ? s : s - 1] = weight[(int64_t)row * ldw + 3 + p * C + 4 * col + j];
            }
        for (int t = half * (TB_TEX / 2); t < (half + 1) * (TB_TEX / 2); ++t) {
            const float4 d0 = s_dp[4 * t], d1 = s_dp[4 * t + 1], d2 = s_dp[4 * t + 2], d3 = s_dp[4 * t + 3];
            const float d[14] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w, d2.x, d2.y, d2.z, d3.x, d3.y, d3.z};
            float o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float a = 0.0f;
#pragma unroll
                for (int s = 0; s < 14; ++s) a = fmaf(w[j][s], d[s], a);
                o[j] = a;
            }
            float4 *q = reinterpret_cast<float4 *>(dst + (int64_t)t * C) + col;
            if (accumulate) {
                const float4 old = *q;
                o[0] += old.x; o[1] += old.y; o[2] += old.z; o[3] += old.w;
            }
            *q = make_float4(o[0], o[1], o[2], o[3]);
        }
    }
}

__global__ void __launch_bounds__(TB_THREADS)
project_tm_bwd_weight_kernel(int C, int R2, const float *__restrict__ d_proj, const float *__restrict__ tokens,
                             int64_t frame_stride, float *__restrict__ d_weight, int ldw)
{
    __shared__ float4 s_dp[TB_TEX * 4];                              // [texel][16]
    const int tiles_per_plane = R2 / TB_TEX;
    const int tile = blockIdx.x % tiles_per_plane;
    const int p = (blockIdx.x / tiles_per_plane) % 3;
    const int f = blockIdx.x / (3 * tiles_per_plane);
    const float4 *dp = reinterpret_cast<const float4 *>(d_proj) + (((int64_t)f * 3 + p) * R2 + (int64_t)tile * TB_TEX) * 4;
    for (int i = threadIdx.x; i < TB_TEX * 4; i += TB_THREADS) s_dp[i] = dp[i];
    __syncthreads();
    const int half = threadIdx.x / TB_COLS, col0 = threadIdx.x % TB_COLS;
    const float *src = tokens + (int64_t)f * frame_stride + ((int64_t)p * R2 + (int64_t)tile * TB_TEX) * C;
    for (int col = col0; col < C / 4; col += TB_COLS) {
        float acc[4][14];
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int s = 0; s < 14; ++s) acc[j][s] = 0.0f;
        for (int t = half * (TB_TEX / 2); t < (half + 1) * (TB_TEX / 2); ++t) {
            const float4 x = ldcs4(src + (int64_t)t * C + 4 * col);
            const float4 d0 = s_dp[4 * t], d1 = s_dp[4 * t + 1], d2 = s_dp[4 * t + 2], d3 = s_dp[4 * t + 3];
            const float d[14] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w, d2.x, d2.y, d2.z, d3.x, d3.y, d3.z};
            const float xv[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int s = 0; s < 14; ++s) acc[j][s] = fmaf(xv[j], d[s], acc[j][s]);
        }
        // one reduction per (CTA half, channel, slot): 2 x 56 per column into the [14, 3+3C] gradient
#pragma unroll
        for (int s = 0; s < 16; ++s) {
            const int row = c_slot_row_tm[s];
            if (row < 0) continue;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                atomicAdd(d_weight + (int64_t)row * ldw + 3 + p * C + 4 * col + j, acc[j][s < 11 ? s : s - 1]);
        }
    }
}

}  // namespace

extern "C" int amav_triplane_project_tm(int F, int C, int R, const float *tokens_tm, int64_t token_frame_stride,
                                        const float *weight, float *proj, void *stream)
{
    AMAV_CHECK_ARG(F >= 0 && C > 0 && R > 0, "bad dims");
    AMAV_CHECK_ARG((R * R) % TM_TEX == 0, "R*R must be a multiple of 128");
    AMAV_CHECK_ARG(C % 64 == 0, "C must be a multiple of 64");
    AMAV_CHECK_ARG(token_frame_stride >= (int64_t)C * 3 * R * R, "frame stride too small");
    if (F == 0) return AMAV_OK;
    AMAV_CHECK_ARG(tokens_tm && weight && proj, "null buffer");
    AMAV_CHECK_ARG(((uintptr_t)tokens_tm & 15) == 0 && (token_frame_stride & 3) == 0, "tokens must be 16-byte aligned");
    const size_t tile_floats = (size_t)TM_STAGES * TM_TILE, red_floats = (size_t)TM_WARPS * 14 * TM_TEX;
    const size_t smem = ((size_t)C * 16 + (tile_floats > red_floats ? tile_floats : red_floats)) * sizeof(float);
    AMAV_CHECK_ARG(smem <= 200 * 1024, "C too large for the shared-memory weight tile");
    amav_ensure_dyn_smem((const void *)project_tm_kernel, (int)smem);
    const int64_t grid = (int64_t)F * 3 * (R * R / TM_TEX);
    AMAV_CHECK_ARG(grid < ((int64_t)1 << 31), "too many tiles");
    project_tm_kernel<<<(unsigned)grid, TM_THREADS, smem, (cudaStream_t)stream>>>(C, R * R, tokens_tm, token_frame_stride,
                                                                                  weight, 3 + 3 * C, proj);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}

extern "C" int amav_triplane_project_backward_tm(int F, int C, int R, const float *d_proj, const float *tokens_tm,
                                                 int64_t token_frame_stride, const float *weight, float *d_tokens_tm,
                                                 int64_t d_token_frame_stride, int accumulate_tokens, float *d_weight,
                                                 void *stream_)
{
    AMAV_CHECK_ARG(F >= 0 && C > 0 && R > 0, "bad dims");
    AMAV_CHECK_ARG((R * R) % TB_TEX == 0, "R*R must be a multiple of 128");
    AMAV_CHECK_ARG(C % 64 == 0, "C must be a multiple of 64");
    if (F == 0) return AMAV_OK;
    AMAV_CHECK_ARG(d_proj && weight, "null buffer");
    AMAV_CHECK_ARG(((uintptr_t)d_proj & 15) == 0, "d_proj must be 16-byte aligned");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int64_t grid = (int64_t)F * 3 * (R * R / TB_TEX);
    AMAV_CHECK_ARG(grid < ((int64_t)1 << 31), "too many tiles");
    if (d_tokens_tm) {
        AMAV_CHECK_ARG(d_token_frame_stride >= (int64_t)C * 3 * R * R && (d_token_frame_stride & 3) == 0 &&
                       ((uintptr_t)d_tokens_tm & 15) == 0, "bad d_tokens stride / alignment");
        project_tm_bwd_tokens_kernel<<<(unsigned)grid, TB_THREADS, 0, stream>>>(C, R * R, d_proj, weight, 3 + 3 * C, d_tokens_tm,
                                                                               d_token_frame_stride, accumulate_tokens);
        AMAV_CHECK_LAUNCH();
    }
    if (d_weight) {
        AMAV_CHECK_ARG(tokens_tm && token_frame_stride >= (int64_t)C * 3 * R * R && (token_frame_stride & 3) == 0 &&
                       ((uintptr_t)tokens_tm & 15) == 0, "bad tokens stride / alignment");
        project_tm_bwd_weight_kernel<<<(unsigned)grid, TB_THREADS, 0, stream>>>(C, R * R, d_proj, tokens_tm, token_frame_stride,
                                                                               d_weight, 3 + 3 * C);
        AMAV_CHECK_LAUNCH();
    }
    return AMAV_OK;
}
