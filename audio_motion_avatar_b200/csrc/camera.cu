// Camera set-up of render_one for all frames of a call in one launch (reference src/models/renderer.py:486-510,
// src/utils/graphic_utils.py:67-78 getWorld2View2_torch, :103-136 getProjectionMatrix_torch, :144-145 focal2fov_torch).
//
// The reference builds Rt = [R | t; 0 0 0 1] from the extrinsic, inverts it twice in fp32 (C2W and back), transposes,
// forms K_ndc from Python floats (double precision), takes math.tan of the fp32 half-angle in double, and multiplies
// view x proj with a batched fp32 matmul.  Here: inv(inv(Rt)) = Rt (the reference's double inversion only adds ~1e-7
// of rounding noise to it), K_ndc and tan in double, the 4x4 product in fp32 FMAs.  Differences from the reference's
// matrices are ~1e-7 relative (tests: <= 1e-5 against the reference's own golden vectors).
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(128) camera_kernel(int F, const float *__restrict__ K, const float *__restrict__ E,
                                                     int width, int height, double znear, double zfar,
                                                     float *__restrict__ view, float *__restrict__ full,
                                                     float *__restrict__ tanfov)
{
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= F) return;
    const float *k = K + 9 * (int64_t)f, *e = E + 16 * (int64_t)f;
    // viewmatrix = Rt^T: rows 0..2 of Rt are rows 0..2 of the extrinsic, row 3 = (0,0,0,1)
    float v[16];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) v[c * 4 + r] = e[r * 4 + c];
    v[3] = 0.0f; v[7] = 0.0f; v[11] = 0.0f; v[15] = 1.0f;
    const float fx = k[0], fy = k[4], cx = k[2], cy = k[5];
    const float fovx = 2.0f * atanf((float)width / (2.0f * fx));
    const float fovy = 2.0f * atanf((float)height / (2.0f * fy));
    tanfov[2 * f] = (float)tan((double)(fovx * 0.5f));
    tanfov[2 * f + 1] = (float)tan((double)(fovy * 0.5f));
    // projmatrix = K_ndc^T
    float p[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) p[i] = 0.0f;
    p[0] = (float)(2.0 * (double)fx / width);
    p[2 * 4 + 0] = (float)((2.0 * (double)cx - width) / width);
    p[1 * 4 + 1] = (float)(2.0 * (double)fy / height);
    p[2 * 4 + 1] = (float)((2.0 * (double)cy - height) / height);
    p[2 * 4 + 2] = (float)(zfar / (zfar - znear));
    p[3 * 4 + 2] = (float)(-zfar * znear / (zfar - znear));
    p[2 * 4 + 3] = 1.0f;
    float *vo = view + 16 * (int64_t)f, *fo = full + 16 * (int64_t)f;
#pragma unroll
    for (int i = 0; i < 16; ++i) vo[i] = v[i];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float s = v[r * 4] * p[c];
#pragma unroll
            for (int q = 1; q < 4; ++q) s = fmaf(v[r * 4 + q], p[q * 4 + c], s);
            fo[r * 4 + c] = s;
        }
}

}  // namespace

extern "C" int amav_camera_setup(int F, const float *K, const float *E, int width, int height, double znear, double zfar,
                                 float *view, float *full, float *tanfov, void *stream)
{
    AMAV_CHECK_ARG(F >= 0 && width > 0 && height > 0 && zfar > znear && znear > 0.0, "bad dims");
    if (F == 0) return AMAV_OK;
    AMAV_CHECK_ARG(K && E && view && full && tanfov, "null buffer");
    camera_kernel<<<amav_div_up(F, 128), 128, 0, (cudaStream_t)stream>>>(F, K, E, width, height, znear, zfar,
                                                                        view, full, tanfov);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
