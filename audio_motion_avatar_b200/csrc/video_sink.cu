// Video sink: the demo's per-frame host loop `(frame * 255).astype(np.uint8)` + cv2.cvtColor(RGB2BGR)
// + the even/odd interleave of two sequences (src/main2.py:226-239, :332-363) as one HBM-bound pass on the device:
// 12 bytes read and 3 written per pixel, so a quarter of the bytes cross PCIe / NVLink afterwards.
//   thread = 4 consecutive pixels of one frame: three float4 streaming loads (48 contiguous bytes, the warp covers
//   1536 contiguous bytes), three 32-bit stores (12 contiguous bytes).  Frames whose H*W is not a multiple of 4 take
//   the scalar instantiation (the frame bases are then not 16-byte aligned).
#include "common.cuh"

namespace {

// numpy: float32 * 255 (one fp32 rounding), astype(uint8) truncates toward zero.  The renderer's frames are already
// clamped to [0,1] (renderer.py:568); other inputs saturate at 0 / 255 here, where numpy would wrap.
__device__ __forceinline__ uint32_t to_u8(float v)
{
    const float s = __fmul_rn(v, 255.0f);
    return (uint32_t)min(max(__float2int_rz(s), 0), 255);
}

__device__ __forceinline__ float4 ld_stream(const float4 *p)
{
    float4 v;
    asm volatile("ld.global.cs.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

template <bool VEC>
__global__ void __launch_bounds__(256) frames_to_u8_kernel(int64_t HW, const float *__restrict__ frames,
                                                           int64_t in_frame_stride, uint8_t *__restrict__ out,
                                                           int64_t out_frame_stride, int out_first, int out_step,
                                                           int bgr)
{
    const int f = blockIdx.y;
    const float *src = frames + (int64_t)f * in_frame_stride;
    uint8_t *dst = out + ((int64_t)out_first + (int64_t)f * out_step) * out_frame_stride;
    const int64_t p0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (p0 >= HW) return;
    if (VEC) {
        const float4 *s4 = reinterpret_cast<const float4 *>(src + p0 * 3);
        const float4 a = ld_stream(s4), b = ld_stream(s4 + 1), c = ld_stream(s4 + 2);
        // pixels: (a.x a.y a.z) (a.w b.x b.y) (b.z b.w c.x) (c.y c.z c.w)
        uint32_t q[12] = {to_u8(a.x), to_u8(a.y), to_u8(a.z), to_u8(a.w), to_u8(b.x), to_u8(b.y),
                          to_u8(b.z), to_u8(b.w), to_u8(c.x), to_u8(c.y), to_u8(c.z), to_u8(c.w)};
        if (bgr) {
#pragma unroll
            for (int k = 0; k < 12; k += 3) { const uint32_t t = q[k]; q[k] = q[k + 2]; q[k + 2] = t; }
        }
        uint32_t *d4 = reinterpret_cast<uint32_t *>(dst + p0 * 3);
#pragma unroll
        for (int k = 0; k < 3; ++k)
            __stcs(d4 + k, q[4 * k] | (q[4 * k + 1] << 8) | (q[4 * k + 2] << 16) | (q[4 * k + 3] << 24));
    } else {
        const int64_t p1 = min(p0 + 4, HW);
        for (int64_t p = p0; p < p1; ++p) {
            const uint32_t r = to_u8(src[p * 3]), g = to_u8(src[p * 3 + 1]), b = to_u8(src[p * 3 + 2]);
            dst[p * 3] = (uint8_t)(bgr ? b : r);
            dst[p * 3 + 1] = (uint8_t)g;
            dst[p * 3 + 2] = (uint8_t)(bgr ? r : b);
        }
    }
}

}  // namespace

extern "C" int amav_frames_to_u8(int F, int H, int W, const float *frames, int64_t in_frame_stride, uint8_t *out,
                                 int64_t out_frame_stride, int out_first, int out_step, int bgr, void *stream)
{
    AMAV_CHECK_ARG(F >= 0 && F <= 65535 && H >= 0 && W >= 0, "bad dims");
    const int64_t HW = (int64_t)H * W;
    if (F == 0 || HW == 0) return AMAV_OK;
    AMAV_CHECK_ARG(frames && out, "null buffer");
    AMAV_CHECK_ARG(in_frame_stride >= HW * 3 && out_frame_stride >= HW * 3, "frame stride smaller than a frame");
    AMAV_CHECK_ARG(out_first >= 0 && out_step >= 1, "bad output frame placement");
    const bool vec = (HW % 4 == 0) && (in_frame_stride % 4 == 0) && (out_frame_stride % 4 == 0) &&
                     (((uintptr_t)frames & 15) == 0) && (((uintptr_t)out & 3) == 0);
    dim3 grid(amav_div_up(amav_div_up(HW, 4), 256), F);
    if (vec)
        frames_to_u8_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(HW, frames, in_frame_stride, out,
                                                                           out_frame_stride, out_first, out_step, bgr);
    else
        frames_to_u8_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(HW, frames, in_frame_stride, out,
                                                                            out_frame_stride, out_first, out_step, bgr);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
