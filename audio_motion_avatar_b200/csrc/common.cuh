// Shared helpers for libamav_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "amav.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libamav_b200 is written for sm_100a (B200) only"
#endif

#define AMAV_SM_COUNT 148
// tiles with at least this many entries are listed by amav_splat_scan and handed out first by the blend kernels
#define AMAV_DEEP0 4096
#define AMAV_DEEP1 2048
// ... in launches of at most this many frames: in long launches the tail is a small share, and pulling the deep
// tiles of every frame forward costs more in L2 misses on the records (measured at 120 frames: 11.8 -> 12.5 ms)
#define AMAV_DEEP_FIRST_MAX_FRAMES 32

void amav_set_error(const char *fmt, ...);
void amav_count_launches(int n);
// opt-in to `bytes` of dynamic shared memory for `func` on the current device, once (capi.cu)
cudaError_t amav_ensure_dyn_smem(const void *func, int bytes);

#define AMAV_CHECK_ARG(cond, msg)                                   \
    do {                                                            \
        if (!(cond)) {                                              \
            amav_set_error("%s: %s", __func__, msg);                \
            return AMAV_EINVAL;                                     \
        }                                                           \
    } while (0)

#define AMAV_CHECK_LAUNCH()                                                         \
    do {                                                                            \
        cudaError_t e__ = cudaGetLastError();                                       \
        if (e__ != cudaSuccess) {                                                   \
            amav_set_error("%s: CUDA error: %s", __func__, cudaGetErrorString(e__)); \
            return AMAV_ECUDA;                                                      \
        }                                                                           \
        amav_count_launches(1);                                                     \
    } while (0)

// tensor-core form of the pose/shape blend of the SMPL-X forward (lbs_tc.cu); 0 on success
int amav_pose_blend_tc(int F, int V, int S, int K, const float *pose_feature, const float *shape_comp, const float *v_template,
                       const float *shapedirs, const float *posedirs, float *v_posed, cudaStream_t stream);

static inline int amav_div_up(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }
