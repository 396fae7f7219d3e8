// Stage 3: SMPL-X forward (linear blend skinning of the Gaussian anchor sites) for all frames of a call.
//
// Replaces smplx==0.1.28 SMPLX.forward -> lbs.lbs as called at src/models/renderer.py:261-274 (~60 small
// PyTorch kernels and a 54-step Python loop per call) and the densify gather of renderer.py:276-288
// (pytorch3d SubdivideMeshes + randperm subset) by four launches per call:
//   lbs_pose_kernel    one CTA per frame: Rodrigues, pose feature, joints, kinematic chain -> A[f][j] (3x4)
//   pose_blend_kernel  v_posed = v_template + shapedirs.shape + pose_feature.posedirs for 16 frames per CTA
//                      (the [F,486]x[486,3V] product; posedirs is read from HBM once per 16 frames)
//   skin_kernel        verts = (sum_j W[v,j] A[f][j]) . [v_posed; 1]
//   densify_kernel     sites = sum_k wgt[i][k] * verts[par[i][k]]
// plus a once-per-model precompute (joint regressor folded through the shape basis, transposed weights).
// `transl` is NOT applied here (renderer.py:248-249,271); covariances are never skinned.
#include "common.cuh"

// 1: the pose/shape blend runs on the tensor cores (lbs_tc.cu); 0: the FP32-FMA pose_blend_kernel below
#ifndef AMAV_POSE_BLEND_TC
#define AMAV_POSE_BLEND_TC 1
#endif

namespace {

constexpr int MAXJ = 64;

// ---------------------------------------------------------------------------------------------
// once per model: J_template[j][k] = sum_v Jreg[j][v] v_template[v][k];
//                 J_dirs[j][k][s]  = sum_v Jreg[j][v] shapedirs[v][k][s];  weights_T[j][v] = lbs_weights[v][j]
__global__ void __launch_bounds__(256) lbs_precompute_kernel(int V, int J, int S, const float *__restrict__ v_template,
                                                             const float *__restrict__ shapedirs,
                                                             const float *__restrict__ Jreg,
                                                             const float *__restrict__ lbs_weights,
                                                             float *__restrict__ J_template, float *__restrict__ J_dirs,
                                                             float *__restrict__ weights_T)
{
    const int j = blockIdx.x;
    __shared__ float s_part[8];
    const int nout = 3 + 3 * S;
    for (int o = 0; o < nout; ++o) {
        float acc = 0.0f;
        for (int v = threadIdx.x; v < V; v += blockDim.x) {
            const float r = Jreg[(int64_t)j * V + v];
            const float x = o < 3 ? v_template[3 * v + o] : shapedirs[(int64_t)v * 3 * S + (o - 3)];
            acc = fmaf(r, x, acc);
        }
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
        if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            float t = 0.0f;
            for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_part[w];
            if (o < 3) J_template[3 * j + o] = t;
            else J_dirs[(int64_t)j * 3 * S + (o - 3)] = t;
        }
        __syncthreads();
    }
    for (int v = threadIdx.x; v < V; v += blockDim.x) weights_T[(int64_t)j * V + v] = lbs_weights[(int64_t)v * J + j];
}

// ---------------------------------------------------------------------------------------------
// one CTA (64 threads) per frame.  SIXD = false: full_pose [F,J,3] axis-angles through smplx's Rodrigues;
// SIXD = true: full_pose [F,J,6] in the 6D rotation representation the SMPL-X parameter decoder's heads emit, turned
// into R by Gram-Schmidt (pytorch3d.transforms.rotation_6d_to_matrix: rows b1 = a1/|a1|, b2 = (a2 - (b1.a2) b1)/|.|,
// b3 = b1 x b2) -- the reference goes 6D -> matrix -> matrix_to_axis_angle -> batch_rodrigues -> the same matrix
// (src/models/smplx_decoder.py:106-127 then renderer.py:104-118, :261-274).
template <bool SIXD>
__global__ void __launch_bounds__(64) lbs_pose_kernel(int J, int S, const float *__restrict__ full_pose,
                                                      const float *__restrict__ shape_comp,
                                                      const float *__restrict__ J_template,
                                                      const float *__restrict__ J_dirs,
                                                      const int32_t *__restrict__ parents,
                                                      float *__restrict__ pose_feature, float *__restrict__ joints,
                                                      float *__restrict__ A)
{
    const int f = blockIdx.x;
    __shared__ float s_R[MAXJ][9];
    __shared__ float s_J[MAXJ][3];
    __shared__ float s_G[MAXJ][12];
    __shared__ int s_par[MAXJ];
    const int t = threadIdx.x;
    if (t < J) {
        float R[9];
        if (SIXD) {
            const float *d6 = full_pose + ((int64_t)f * J + t) * 6;
            const float a1x = d6[0], a1y = d6[1], a1z = d6[2], a2x = d6[3], a2y = d6[4], a2z = d6[5];
            // F.normalize: v / max(||v||, 1e-12)
            const float n1 = fmaxf(sqrtf(a1x * a1x + a1y * a1y + a1z * a1z), 1e-12f);
            const float b1x = a1x / n1, b1y = a1y / n1, b1z = a1z / n1;
            const float dt = b1x * a2x + b1y * a2y + b1z * a2z;
            const float ux = a2x - dt * b1x, uy = a2y - dt * b1y, uz = a2z - dt * b1z;
            const float n2 = fmaxf(sqrtf(ux * ux + uy * uy + uz * uz), 1e-12f);
            const float b2x = ux / n2, b2y = uy / n2, b2z = uz / n2;
            R[0] = b1x; R[1] = b1y; R[2] = b1z;
            R[3] = b2x; R[4] = b2y; R[5] = b2z;
            R[6] = b1y * b2z - b1z * b2y; R[7] = b1z * b2x - b1x * b2z; R[8] = b1x * b2y - b1y * b2x;
        } else {
        // smplx.lbs.batch_rodrigues: angle = ||r + 1e-8||, dir = r / angle, R = I + sin K + (1-cos) K K
        const float rx = full_pose[((int64_t)f * J + t) * 3], ry = full_pose[((int64_t)f * J + t) * 3 + 1],
                    rz = full_pose[((int64_t)f * J + t) * 3 + 2];
        const float ex = rx + 1e-8f, ey = ry + 1e-8f, ez = rz + 1e-8f;
        const float angle = sqrtf(ex * ex + ey * ey + ez * ez);
        const float x = rx / angle, y = ry / angle, z = rz / angle;
        float sn, cs;
        sincosf(angle, &sn, &cs);
        const float oc = 1.0f - cs;
        // K = [[0,-z,y],[z,0,-x],[-y,x,0]];  K K = [[-(y2+z2), xy, xz],[xy, -(x2+z2), yz],[xz, yz, -(x2+y2)]]
        R[0] = 1.0f + oc * (-(y * y + z * z)); R[1] = sn * (-z) + oc * (x * y);   R[2] = sn * y + oc * (x * z);
        R[3] = sn * z + oc * (x * y);          R[4] = 1.0f + oc * (-(x * x + z * z)); R[5] = sn * (-x) + oc * (y * z);
        R[6] = sn * (-y) + oc * (x * z);       R[7] = sn * x + oc * (y * z);      R[8] = 1.0f + oc * (-(x * x + y * y));
        }
#pragma unroll
        for (int k = 0; k < 9; ++k) s_R[t][k] = R[k];
        if (t > 0) {
#pragma unroll
            for (int k = 0; k < 9; ++k)
                pose_feature[(int64_t)f * (J - 1) * 9 + (t - 1) * 9 + k] = R[k] - ((k == 0 || k == 4 || k == 8) ? 1.0f : 0.0f);
        }
        // joints = J_regressor . v_shaped  (regressor folded through the shape basis)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float a = J_template[3 * t + k];
            for (int s = 0; s < S; ++s) a = fmaf(J_dirs[((int64_t)t * 3 + k) * S + s], shape_comp[(int64_t)f * S + s], a);
            s_J[t][k] = a;
            joints[((int64_t)f * J + t) * 3 + k] = a;
        }
        s_par[t] = parents[t];
    }
    __syncthreads();
    // kinematic chain (smplx.lbs.batch_rigid_transform): G_j = G_parent . [R_j | J_j - J_parent]; 12 lanes, one
    // element each; parents precede children.
    if (t < 12) {
        const int r = t >> 2, c = t & 3;
        for (int j = 0; j < J; ++j) {
            const int par = s_par[j];
            float loc[3];   // column c of the local transform [R_j | rel_j]
            if (c < 3) { loc[0] = s_R[j][c]; loc[1] = s_R[j][3 + c]; loc[2] = s_R[j][6 + c]; }
            else {
                loc[0] = s_J[j][0] - (par >= 0 ? s_J[par][0] : 0.0f);
                loc[1] = s_J[j][1] - (par >= 0 ? s_J[par][1] : 0.0f);
                loc[2] = s_J[j][2] - (par >= 0 ? s_J[par][2] : 0.0f);
            }
            float g;
            if (par < 0) g = loc[r];
            else {
                g = s_G[par][r * 4] * loc[0] + s_G[par][r * 4 + 1] * loc[1] + s_G[par][r * 4 + 2] * loc[2];
                if (c == 3) g += s_G[par][r * 4 + 3];
            }
            __syncwarp(0xfffu);
            s_G[j][t] = g;
            __syncwarp(0xfffu);
        }
    }
    __syncthreads();
    // A_j = G_j - [0 | G_j(3x3) . J_j]
    if (t < J) {
        float *out = A + ((int64_t)f * J + t) * 12;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const float g0 = s_G[t][r * 4], g1 = s_G[t][r * 4 + 1], g2 = s_G[t][r * 4 + 2], g3 = s_G[t][r * 4 + 3];
            out[r * 4] = g0; out[r * 4 + 1] = g1; out[r * 4 + 2] = g2;
            out[r * 4 + 3] = g3 - (g0 * s_J[t][0] + g1 * s_J[t][1] + g2 * s_J[t][2]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
#if !AMAV_POSE_BLEND_TC
// v_posed[f][v] = v_template[v] + shapedirs[v].shape_comp[f] + pose_feature[f].posedirs[:, 3v:3v+3]
// Thread = vertex (3 columns), CTA = 128 vertices x PB_F frames; pose features of the frame group in
// shared memory as [k][PB_F] so one LDS.128 feeds four frames.
constexpr int PB_F = 16;
constexpr int PB_THREADS = 128;       // (64-thread CTAs reach more SMs for a 4-frame call but were slower: 0.10 -> 0.14 ms)

__global__ void __launch_bounds__(PB_THREADS) pose_blend_kernel(int F, int V, int S, int K,
                                                                 const float *__restrict__ pose_feature,
                                                                 const float *__restrict__ shape_comp,
                                                                 const float *__restrict__ v_template,
                                                                 const float *__restrict__ shapedirs,
                                                                 const float *__restrict__ posedirs,
                                                                 float *__restrict__ v_posed)
{
    extern __shared__ float4 s_dyn4[];
    float *s_pf = reinterpret_cast<float *>(s_dyn4);      // [K][PB_F]
    float *s_sc = s_pf + (size_t)K * PB_F;                // [S][PB_F]
    const int f0 = blockIdx.y * PB_F;
    for (int i = threadIdx.x; i < K * PB_F; i += PB_THREADS) {
        const int k = i / PB_F, fb = i - k * PB_F;
        s_pf[i] = (f0 + fb < F) ? pose_feature[(int64_t)(f0 + fb) * K + k] : 0.0f;
    }
    for (int i = threadIdx.x; i < S * PB_F; i += PB_THREADS) {
        const int k = i / PB_F, fb = i - k * PB_F;
        s_sc[i] = (f0 + fb < F) ? shape_comp[(int64_t)(f0 + fb) * S + k] : 0.0f;
    }
    __syncthreads();
    const int v = blockIdx.x * PB_THREADS + threadIdx.x;
    if (v >= V) return;
    float acc[PB_F][3];
    {   // v_shaped first (blend_shapes), then the pose offsets are added on top
        const float t0 = v_template[3 * v], t1 = v_template[3 * v + 1], t2 = v_template[3 * v + 2];
#pragma unroll
        for (int fb = 0; fb < PB_F; ++fb) { acc[fb][0] = 0.0f; acc[fb][1] = 0.0f; acc[fb][2] = 0.0f; }
        for (int s = 0; s < S; ++s) {
            const float d0 = shapedirs[((int64_t)v * 3) * S + s], d1 = shapedirs[((int64_t)v * 3 + 1) * S + s],
                        d2 = shapedirs[((int64_t)v * 3 + 2) * S + s];
#pragma unroll
            for (int fb = 0; fb < PB_F; ++fb) {
                const float c = s_sc[s * PB_F + fb];
                acc[fb][0] = fmaf(c, d0, acc[fb][0]); acc[fb][1] = fmaf(c, d1, acc[fb][1]); acc[fb][2] = fmaf(c, d2, acc[fb][2]);
            }
        }
#pragma unroll
        for (int fb = 0; fb < PB_F; ++fb) { acc[fb][0] += t0; acc[fb][1] += t1; acc[fb][2] += t2; }
    }
    float off[PB_F][3];
#pragma unroll
    for (int fb = 0; fb < PB_F; ++fb) { off[fb][0] = 0.0f; off[fb][1] = 0.0f; off[fb][2] = 0.0f; }
    const float *pd = posedirs + (int64_t)3 * v;
    const int64_t ld = (int64_t)3 * V;
#pragma unroll 2
    for (int k = 0; k < K; ++k) {
        const float p0 = __ldg(pd + k * ld), p1 = __ldg(pd + k * ld + 1), p2 = __ldg(pd + k * ld + 2);
        const float4 *pf4 = reinterpret_cast<const float4 *>(s_pf + (size_t)k * PB_F);
#pragma unroll
        for (int q = 0; q < PB_F / 4; ++q) {
            const float4 c = pf4[q];
            off[4 * q][0] = fmaf(c.x, p0, off[4 * q][0]); off[4 * q][1] = fmaf(c.x, p1, off[4 * q][1]); off[4 * q][2] = fmaf(c.x, p2, off[4 * q][2]);
            off[4 * q + 1][0] = fmaf(c.y, p0, off[4 * q + 1][0]); off[4 * q + 1][1] = fmaf(c.y, p1, off[4 * q + 1][1]); off[4 * q + 1][2] = fmaf(c.y, p2, off[4 * q + 1][2]);
            off[4 * q + 2][0] = fmaf(c.z, p0, off[4 * q + 2][0]); off[4 * q + 2][1] = fmaf(c.z, p1, off[4 * q + 2][1]); off[4 * q + 2][2] = fmaf(c.z, p2, off[4 * q + 2][2]);
            off[4 * q + 3][0] = fmaf(c.w, p0, off[4 * q + 3][0]); off[4 * q + 3][1] = fmaf(c.w, p1, off[4 * q + 3][1]); off[4 * q + 3][2] = fmaf(c.w, p2, off[4 * q + 3][2]);
        }
    }
#pragma unroll
    for (int fb = 0; fb < PB_F; ++fb) {
        if (f0 + fb < F) {
            float *o = v_posed + ((int64_t)(f0 + fb) * V + v) * 3;
            o[0] = acc[fb][0] + off[fb][0]; o[1] = acc[fb][1] + off[fb][1]; o[2] = acc[fb][2] + off[fb][2];
        }
    }
}
#endif  // !AMAV_POSE_BLEND_TC

// verts[f][v] = (sum_j W[v][j] A[f][j]) . [v_posed; 1]; zero weights are skipped (adding w*A = 0 is exact, so
// the dense sum in joint order is unchanged).
__global__ void __launch_bounds__(256) skin_kernel(int V, int J, const float *__restrict__ weights_T,
                                                    const float *__restrict__ A, const float *__restrict__ v_posed,
                                                    float *__restrict__ verts)
{
    __shared__ float4 s_A[MAXJ * 3];
    const int f = blockIdx.y;
    for (int i = threadIdx.x; i < J * 3; i += blockDim.x) s_A[i] = reinterpret_cast<const float4 *>(A)[(int64_t)f * J * 3 + i];
    __syncthreads();
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    float4 T0 = make_float4(0.f, 0.f, 0.f, 0.f), T1 = T0, T2 = T0;
    for (int j = 0; j < J; ++j) {
        const float w = __ldg(weights_T + (int64_t)j * V + v);
        if (w != 0.0f) {
            const float4 a0 = s_A[3 * j], a1 = s_A[3 * j + 1], a2 = s_A[3 * j + 2];
            T0.x = fmaf(w, a0.x, T0.x); T0.y = fmaf(w, a0.y, T0.y); T0.z = fmaf(w, a0.z, T0.z); T0.w = fmaf(w, a0.w, T0.w);
            T1.x = fmaf(w, a1.x, T1.x); T1.y = fmaf(w, a1.y, T1.y); T1.z = fmaf(w, a1.z, T1.z); T1.w = fmaf(w, a1.w, T1.w);
            T2.x = fmaf(w, a2.x, T2.x); T2.y = fmaf(w, a2.y, T2.y); T2.z = fmaf(w, a2.z, T2.z); T2.w = fmaf(w, a2.w, T2.w);
        }
    }
    const float *vp = v_posed + ((int64_t)f * V + v) * 3;
    const float x = vp[0], y = vp[1], z = vp[2];
    float *o = verts + ((int64_t)f * V + v) * 3;
    o[0] = T0.x * x + T0.y * y + T0.z * z + T0.w;
    o[1] = T1.x * x + T1.y * y + T1.z * z + T1.w;
    o[2] = T2.x * x + T2.y * y + T2.z * z + T2.w;
}

// sites[f][i] = sum_k wgt[i][k] * verts[f][par[i][k]].  A thread keeps its site's table row in registers for DEN_FB
// frames (P == 4, the level-2 table: one int4 + one float4 load); the sums run in the same order as one frame at a time.
constexpr int DEN_FB = 8;
#ifndef AMAV_DEN_UNROLL
#define AMAV_DEN_UNROLL 1
#endif
constexpr int DEN_UNROLL = AMAV_DEN_UNROLL;      // frames of a thread's loop whose 12 gathers are issued together

__global__ void __launch_bounds__(256) densify_kernel(int F, int V, int N, int P, const float *__restrict__ verts,
                                                       const int32_t *__restrict__ par, const float *__restrict__ wgt,
                                                       float *__restrict__ sites, int vec4,
                                                       const int32_t *__restrict__ order)
{
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    const int f0 = blockIdx.y * DEN_FB, f1 = min(F, f0 + DEN_FB);
    if (slot >= N) return;
    const int i = order ? __ldg(order + slot) : slot;        // locality order: neighbouring lanes share parent vertices
    if (vec4) {
        const int4 q = __ldg(reinterpret_cast<const int4 *>(par) + i);
        const float4 w = __ldg(reinterpret_cast<const float4 *>(wgt) + i);
        const int qq[4] = {q.x, q.y, q.z, q.w};
        const float ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll DEN_UNROLL
        for (int f = f0; f < f1; ++f) {
            const float *vb = verts + (int64_t)f * V * 3;
            float x = 0.0f, y = 0.0f, z = 0.0f;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (ww[k] != 0.0f) {
                    x = fmaf(ww[k], vb[3 * qq[k]], x); y = fmaf(ww[k], vb[3 * qq[k] + 1], y); z = fmaf(ww[k], vb[3 * qq[k] + 2], z);
                }
            float *o = sites + ((int64_t)f * N + i) * 3;
            o[0] = x; o[1] = y; o[2] = z;
        }
        return;
    }
    for (int f = f0; f < f1; ++f) {
        const float *vb = verts + (int64_t)f * V * 3;
        float x = 0.0f, y = 0.0f, z = 0.0f;
        for (int k = 0; k < P; ++k) {
            const float w = __ldg(wgt + (int64_t)i * P + k);
            if (w != 0.0f) {
                const int q = __ldg(par + (int64_t)i * P + k);
                x = fmaf(w, vb[3 * q], x); y = fmaf(w, vb[3 * q + 1], y); z = fmaf(w, vb[3 * q + 2], z);
            }
        }
        float *o = sites + ((int64_t)f * N + i) * 3;
        o[0] = x; o[1] = y; o[2] = z;
    }
}

}  // namespace

extern "C" int amav_lbs_precompute(int V, int J, int S, const float *v_template, const float *shapedirs,
                                   const float *J_regressor, const float *lbs_weights, float *J_template,
                                   float *J_dirs, float *weights_T, void *stream)
{
    AMAV_CHECK_ARG(V > 0 && J > 0 && J <= MAXJ && S >= 0, "bad dims (J <= 64)");
    AMAV_CHECK_ARG(v_template && shapedirs && J_regressor && lbs_weights && J_template && J_dirs && weights_T, "null buffer");
    lbs_precompute_kernel<<<J, 256, 0, (cudaStream_t)stream>>>(V, J, S, v_template, shapedirs, J_regressor, lbs_weights,
                                                               J_template, J_dirs, weights_T);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}

static int lbs_forward_impl(bool sixd, int F, int V, int J, int S, const float *full_pose, const float *shape_comp,
                                const float *v_template, const float *shapedirs, const float *posedirs,
                                const int32_t *parents, const float *J_template, const float *J_dirs,
                                const float *weights_T, float *pose_feature, float *joints, float *A, float *v_posed,
                                float *verts, void *stream_)
{
    AMAV_CHECK_ARG(F >= 0 && V > 0 && J > 1 && J <= MAXJ && S >= 0, "bad dims (1 < J <= 64)");
    if (F == 0) return AMAV_OK;
    AMAV_CHECK_ARG(full_pose && shape_comp && v_template && shapedirs && posedirs && parents, "null input");
    AMAV_CHECK_ARG(J_template && J_dirs && weights_T, "null precomputed table (amav_lbs_precompute)");
    AMAV_CHECK_ARG(pose_feature && joints && A && v_posed && verts, "null work/output buffer");
    AMAV_CHECK_ARG(((uintptr_t)A & 15) == 0, "A must be 16-byte aligned");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int K = (J - 1) * 9;
    if (sixd)
        lbs_pose_kernel<true><<<F, 64, 0, stream>>>(J, S, full_pose, shape_comp, J_template, J_dirs, parents, pose_feature, joints, A);
    else
        lbs_pose_kernel<false><<<F, 64, 0, stream>>>(J, S, full_pose, shape_comp, J_template, J_dirs, parents, pose_feature, joints, A);
#if AMAV_POSE_BLEND_TC
    // v_posed on the tensor cores (lbs_tc.cu: tcgen05.mma kind::tf32, 3xTF32, TMEM accumulator)
    AMAV_CHECK_ARG(F <= 65535 * 128, "F too large for one call");
    if (amav_pose_blend_tc(F, V, S, K, pose_feature, shape_comp, v_template, shapedirs, posedirs, v_posed, stream) != 0) {
        amav_set_error("%s: pose_blend_tc launch failed", __func__);
        return AMAV_ECUDA;
    }
#else
    const size_t smem = (size_t)(K + S) * PB_F * sizeof(float);
    AMAV_CHECK_ARG(smem <= 200 * 1024, "pose basis too large");
    amav_ensure_dyn_smem((const void *)pose_blend_kernel, (int)smem);
    dim3 gb(amav_div_up(V, PB_THREADS), amav_div_up(F, PB_F));
    pose_blend_kernel<<<gb, PB_THREADS, smem, stream>>>(F, V, S, K, pose_feature, shape_comp, v_template, shapedirs,
                                                        posedirs, v_posed);
#endif
    AMAV_CHECK_ARG(F <= 65535, "F too large for one call");
    dim3 gs(amav_div_up(V, 256), F);
    skin_kernel<<<gs, 256, 0, stream>>>(V, J, weights_T, A, v_posed, verts);
    amav_count_launches(2);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}

extern "C" int amav_lbs_forward(int F, int V, int J, int S, const float *full_pose, const float *shape_comp,
                                const float *v_template, const float *shapedirs, const float *posedirs,
                                const int32_t *parents, const float *J_template, const float *J_dirs,
                                const float *weights_T, float *pose_feature, float *joints, float *A, float *v_posed,
                                float *verts, void *stream)
{
    return lbs_forward_impl(false, F, V, J, S, full_pose, shape_comp, v_template, shapedirs, posedirs, parents, J_template, J_dirs,
                            weights_T, pose_feature, joints, A, v_posed, verts, stream);
}

extern "C" int amav_lbs_forward_6d(int F, int V, int J, int S, const float *pose_6d, const float *shape_comp,
                                const float *v_template, const float *shapedirs, const float *posedirs,
                                const int32_t *parents, const float *J_template, const float *J_dirs,
                                const float *weights_T, float *pose_feature, float *joints, float *A, float *v_posed,
                                float *verts, void *stream)
{
    return lbs_forward_impl(true, F, V, J, S, pose_6d, shape_comp, v_template, shapedirs, posedirs, parents, J_template, J_dirs,
                            weights_T, pose_feature, joints, A, v_posed, verts, stream);
}

extern "C" int amav_densify_gather(int F, int V, int N, int P, const float *verts, const int32_t *par,
                                   const float *wgt, float *sites, const int32_t *order, void *stream)
{
    AMAV_CHECK_ARG(F >= 0 && F <= 65535 && V > 0 && N >= 0 && P >= 1, "bad dims");
    if ((int64_t)F * N == 0) return AMAV_OK;
    AMAV_CHECK_ARG(verts && par && wgt && sites, "null buffer");
    dim3 g(amav_div_up(N, 256), amav_div_up(F, DEN_FB));
    const int vec4 = P == 4 && ((uintptr_t)par & 15) == 0 && ((uintptr_t)wgt & 15) == 0;
    densify_kernel<<<g, 256, 0, (cudaStream_t)stream>>>(F, V, N, P, verts, par, wgt, sites, vec4, order);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}
