// baseline/naive_raster.cu -- COMPARATOR, NOT PRODUCT CODE.
//
// A plain CUDA rasteriser with the STRUCTURE of upstream diff-gaussian-rasterization's forward pass (the package the
// reference calls at src/models/renderer.py:516-566; its source is not available offline, SURVEY.md Appendix A is the
// spec): per FRAME  preprocess -> cub inclusive scan of tiles_touched -> device-to-host read of the entry count ->
// key/value emission -> cub 64-bit radix sort on (tile << 32 | depth bits) -> tile ranges -> one 16x16 block per tile
// blending with expf, 256-entry shared-memory batches and a block-wide done count.
// bench.py times it next to the product path so that "x times the reference's render throughput" has a GPU-side
// denominator.  Nothing under audio_motion_avatar_b200/ links or loads this file.
#include <cub/cub.cuh>
#include <cuda_runtime.h>
#include <stdint.h>

namespace {

struct Scratch {
    float *depth = nullptr, *xy = nullptr, *conic = nullptr;
    int *radii = nullptr, *rect = nullptr;
    uint32_t *touched = nullptr, *offsets = nullptr;
    uint64_t *keys = nullptr, *keys2 = nullptr;
    uint32_t *vals = nullptr, *vals2 = nullptr;
    int2 *ranges = nullptr;
    void *cub_tmp = nullptr;
    size_t cub_bytes = 0;
    int64_t capN = 0, capD = 0, capT = 0;
};
Scratch g;

__global__ void nr_preprocess(int N, const float *__restrict__ means, const float *__restrict__ scales,
                              const float *__restrict__ rots, const float *__restrict__ opac, const float *__restrict__ view,
                              const float *__restrict__ proj, int W, int H, float tanx, float tany, float *depth, int *radii,
                              float *xy, float *conic, int *rect, uint32_t *touched)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    radii[i] = 0; touched[i] = 0;
    const float px = means[3 * i], py = means[3 * i + 1], pz = means[3 * i + 2];
    const float vx = view[0] * px + view[4] * py + view[8] * pz + view[12];
    const float vy = view[1] * px + view[5] * py + view[9] * pz + view[13];
    const float vz = view[2] * px + view[6] * py + view[10] * pz + view[14];
    if (!(vz > 0.2f)) return;
    const float hx = proj[0] * px + proj[4] * py + proj[8] * pz + proj[12];
    const float hy = proj[1] * px + proj[5] * py + proj[9] * pz + proj[13];
    const float hw = proj[3] * px + proj[7] * py + proj[11] * pz + proj[15];
    const float pw = 1.0f / (hw + 0.0000001f);
    const float sx = scales[3 * i], sy = scales[3 * i + 1], sz = scales[3 * i + 2];
    const float r = rots[4 * i], x = rots[4 * i + 1], y = rots[4 * i + 2], z = rots[4 * i + 3];
    const float R[9] = {1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                        2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                        2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)};
    float M[9];
    for (int a = 0; a < 3; ++a) { M[3 * a] = R[3 * a] * sx; M[3 * a + 1] = R[3 * a + 1] * sy; M[3 * a + 2] = R[3 * a + 2] * sz; }
    float S[9];
    for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) S[3 * a + b] = M[3 * a] * M[3 * b] + M[3 * a + 1] * M[3 * b + 1] + M[3 * a + 2] * M[3 * b + 2];
    const float fx = W / (2.0f * tanx), fy = H / (2.0f * tany);
    const float tx = fminf(1.3f * tanx, fmaxf(-1.3f * tanx, vx / vz)) * vz, ty = fminf(1.3f * tany, fmaxf(-1.3f * tany, vy / vz)) * vz;
    const float J00 = fx / vz, J02 = -(fx * tx) / (vz * vz), J11 = fy / vz, J12 = -(fy * ty) / (vz * vz);
    float A[6];
    for (int c = 0; c < 3; ++c) { A[c] = J00 * view[4 * c] + J02 * view[4 * c + 2]; A[3 + c] = J11 * view[4 * c + 1] + J12 * view[4 * c + 2]; }
    float u[3], v[3];
    for (int c = 0; c < 3; ++c) { u[c] = A[0] * S[c] + A[1] * S[3 + c] + A[2] * S[6 + c]; v[c] = A[3] * S[c] + A[4] * S[3 + c] + A[5] * S[6 + c]; }
    float ca = u[0] * A[0] + u[1] * A[1] + u[2] * A[2] + 0.3f;
    const float cb = u[0] * A[3] + u[1] * A[4] + u[2] * A[5];
    float cc = v[0] * A[3] + v[1] * A[4] + v[2] * A[5] + 0.3f;
    const float det = ca * cc - cb * cb;
    if (det == 0.0f) return;
    const float inv = 1.0f / det;
    const float mid = 0.5f * (ca + cc);
    const float lam = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
    const float radf = ceilf(3.0f * sqrtf(lam));
    const float ix = ((hx * pw + 1.0f) * W - 1.0f) * 0.5f, iy = ((hy * pw + 1.0f) * H - 1.0f) * 0.5f;
    if (!(fabsf(ix) < 1e9f) || !(fabsf(iy) < 1e9f) || !(radf < 1e9f)) return;
    const int rad = (int)radf, gx = (W + 15) / 16, gy = (H + 15) / 16;
    const int x0 = min(gx, max(0, (int)((ix - rad) / 16.0f))), x1 = min(gx, max(0, (int)((ix + rad + 15) / 16.0f)));
    const int y0 = min(gy, max(0, (int)((iy - rad) / 16.0f))), y1 = min(gy, max(0, (int)((iy + rad + 15) / 16.0f)));
    if ((x1 - x0) * (y1 - y0) == 0) return;
    depth[i] = vz; radii[i] = rad; xy[2 * i] = ix; xy[2 * i + 1] = iy;
    conic[4 * i] = cc * inv; conic[4 * i + 1] = -cb * inv; conic[4 * i + 2] = ca * inv; conic[4 * i + 3] = opac[i];
    rect[4 * i] = x0; rect[4 * i + 1] = y0; rect[4 * i + 2] = x1; rect[4 * i + 3] = y1;
    touched[i] = (uint32_t)((x1 - x0) * (y1 - y0));
}

__global__ void nr_emit(int N, int gx, const int *__restrict__ radii, const int *__restrict__ rect, const float *__restrict__ depth,
                        const uint32_t *__restrict__ offsets, uint64_t *keys, uint32_t *vals)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N || radii[i] <= 0) return;
    uint32_t off = i == 0 ? 0u : offsets[i - 1];
    const uint64_t d = (uint64_t)__float_as_uint(depth[i]);
    for (int y = rect[4 * i + 1]; y < rect[4 * i + 3]; ++y)
        for (int x = rect[4 * i]; x < rect[4 * i + 2]; ++x) {
            keys[off] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | d;
            vals[off] = (uint32_t)i;
            ++off;
        }
}

__global__ void nr_ranges(int64_t D, const uint64_t *__restrict__ keys, int2 *ranges)
{
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= D) return;
    const uint32_t t = (uint32_t)(keys[k] >> 32);
    if (k == 0) ranges[t].x = 0;
    else {
        const uint32_t p = (uint32_t)(keys[k - 1] >> 32);
        if (p != t) { ranges[p].y = (int)k; ranges[t].x = (int)k; }
    }
    if (k == D - 1) ranges[t].y = (int)D;
}

__global__ void __launch_bounds__(256) nr_render(int W, int H, const int2 *__restrict__ ranges, const uint32_t *__restrict__ list,
                                                  const float *__restrict__ xy, const float *__restrict__ conic,
                                                  const float *__restrict__ rgb, const float *__restrict__ bg, float *out)
{
    __shared__ int s_id[256];
    __shared__ float2 s_xy[256];
    __shared__ float4 s_co[256];
    const int gx = (W + 15) / 16;
    const int tile = blockIdx.y * gx + blockIdx.x;
    const int px = blockIdx.x * 16 + (threadIdx.x & 15), py = blockIdx.y * 16 + (threadIdx.x >> 4);
    const bool inside = px < W && py < H;
    const int2 rg = ranges[tile];
    bool done = !inside;
    float T = 1.0f, C[3] = {0.f, 0.f, 0.f};
    int todo = rg.y - rg.x;
    for (int base = rg.x; base < rg.y; base += 256, todo -= 256) {
        if (__syncthreads_count(done) == 256) break;
        if (base + (int)threadIdx.x < rg.y) {
            const int id = (int)list[base + threadIdx.x];
            s_id[threadIdx.x] = id;
            s_xy[threadIdx.x] = make_float2(xy[2 * id], xy[2 * id + 1]);
            s_co[threadIdx.x] = make_float4(conic[4 * id], conic[4 * id + 1], conic[4 * id + 2], conic[4 * id + 3]);
        }
        __syncthreads();
        for (int j = 0; !done && j < min(256, todo); ++j) {
            const float dx = s_xy[j].x - (float)px, dy = s_xy[j].y - (float)py;
            const float4 co = s_co[j];
            const float power = -0.5f * (co.x * dx * dx + co.z * dy * dy) - co.y * dx * dy;
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, co.w * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            const float Tn = T * (1.0f - alpha);
            if (Tn < 0.0001f) { done = true; continue; }
            const int id = s_id[j];
            for (int ch = 0; ch < 3; ++ch) C[ch] += rgb[3 * id + ch] * alpha * T;
            T = Tn;
        }
    }
    if (inside)
        for (int ch = 0; ch < 3; ++ch) out[(size_t)ch * H * W + (size_t)py * W + px] = fminf(fmaxf(C[ch] + T * bg[ch], 0.0f), 1.0f);
}

template <typename T> void grow(T *&p, int64_t n) { if (p) cudaFree(p); cudaMalloc(&p, sizeof(T) * (size_t)n); }

}  // namespace

// One frame, activated inputs (scales, opacities as the rasteriser takes them).  Returns the entry count D, or -1.
extern "C" int64_t naive_raster_forward(int N, const float *means, const float *scales, const float *rots, const float *opac,
                                        const float *colors, const float *view, const float *proj, float tanx, float tany,
                                        const float *bg, int W, int H, float *out_chw, void *stream_)
{
    cudaStream_t st = (cudaStream_t)stream_;
    const int gx = (W + 15) / 16, gy = (H + 15) / 16, tiles = gx * gy;
    if (N > g.capN) {
        g.capN = N;
        grow(g.depth, N); grow(g.xy, 2 * (int64_t)N); grow(g.conic, 4 * (int64_t)N); grow(g.radii, N); grow(g.rect, 4 * (int64_t)N);
        grow(g.touched, N); grow(g.offsets, N);
    }
    if (tiles > g.capT) { g.capT = tiles; grow(g.ranges, tiles); }
    nr_preprocess<<<(N + 255) / 256, 256, 0, st>>>(N, means, scales, rots, opac, view, proj, W, H, tanx, tany, g.depth, g.radii, g.xy,
                                                   g.conic, g.rect, g.touched);
    size_t need = 0;
    cub::DeviceScan::InclusiveSum(nullptr, need, g.touched, g.offsets, N, st);
    if (need > g.cub_bytes) { if (g.cub_tmp) cudaFree(g.cub_tmp); cudaMalloc(&g.cub_tmp, need); g.cub_bytes = need; }
    cub::DeviceScan::InclusiveSum(g.cub_tmp, g.cub_bytes, g.touched, g.offsets, N, st);
    uint32_t D = 0;
    cudaMemcpyAsync(&D, g.offsets + (N - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
    cudaStreamSynchronize(st);                                    // upstream reads the count back too
    if (D > g.capD) { g.capD = (int64_t)(D * 1.25) + 1024; grow(g.keys, g.capD); grow(g.keys2, g.capD); grow(g.vals, g.capD); grow(g.vals2, g.capD); }
    cudaMemsetAsync(g.ranges, 0, sizeof(int2) * tiles, st);
    if (D > 0) {
        nr_emit<<<(N + 255) / 256, 256, 0, st>>>(N, gx, g.radii, g.rect, g.depth, g.offsets, g.keys, g.vals);
        int bits = 32; for (int t = tiles; t > 0; t >>= 1) ++bits;
        cub::DeviceRadixSort::SortPairs(nullptr, need, g.keys, g.keys2, g.vals, g.vals2, (int)D, 0, bits, st);
        if (need > g.cub_bytes) { if (g.cub_tmp) cudaFree(g.cub_tmp); cudaMalloc(&g.cub_tmp, need); g.cub_bytes = need; }
        cub::DeviceRadixSort::SortPairs(g.cub_tmp, g.cub_bytes, g.keys, g.keys2, g.vals, g.vals2, (int)D, 0, bits, st);
        nr_ranges<<<(D + 255) / 256, 256, 0, st>>>((int64_t)D, g.keys2, g.ranges);
    }
    nr_render<<<dim3(gx, gy), 256, 0, st>>>(W, H, g.ranges, g.vals2, g.xy, g.conic, colors, bg, out_chw);
    return cudaGetLastError() == cudaSuccess ? (int64_t)D : -1;
}
