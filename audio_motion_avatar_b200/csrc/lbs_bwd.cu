// Stage 3 backward: gradient of the SMPL-X forward (smplx==0.1.28 lbs.lbs, reached in the reference through autograd
// from src/models/renderer.py:261-274 when the SMPL-X parameters are predicted, i.e. stage-2 training) and of the
// densify gather (renderer.py:276-288), for all frames of a call:
//   densify_bwd_kernel     d_verts[par] += wgt * d_sites
//   skin_bwd_kernel        d_v_posed = T_v^T d_verts;  d_A[f][j] += sum_v W[v][j] d_verts[v] (x) [v_posed[v]; 1]
//                          (per-warp joint skipping + split-butterfly reduction, one flush per CTA)
//   pose_blend_bwd_kernel  d_pose_feature = posedirs . d_v_posed,  d_shape = shapedirs . d_v_posed
//                          (posedirs streamed once per 16 frames, as in the forward)
//   chain_bwd_kernel       one CTA per frame: relative transforms -> kinematic chain -> Rodrigues, joints -> shape
#include "common.cuh"

namespace {

constexpr int MAXJ = 64;

__global__ void __launch_bounds__(256) densify_bwd_kernel(int V, int N, int P, const float *__restrict__ d_sites,
                                                           const int32_t *__restrict__ par, const float *__restrict__ wgt,
                                                           float *__restrict__ d_verts)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int f = blockIdx.y;
    if (i >= N) return;
    const float *g = d_sites + ((int64_t)f * N + i) * 3;
    const float gx = g[0], gy = g[1], gz = g[2];
    if (gx == 0.0f && gy == 0.0f && gz == 0.0f) return;
    float *vb = d_verts + (int64_t)f * V * 3;
    for (int k = 0; k < P; ++k) {
        const float w = __ldg(wgt + (int64_t)i * P + k);
        if (w != 0.0f) {
            const int q = __ldg(par + (int64_t)i * P + k);
            atomicAdd(vb + 3 * q, w * gx); atomicAdd(vb + 3 * q + 1, w * gy); atomicAdd(vb + 3 * q + 2, w * gz);
        }
    }
}

// 16 values over 32 lanes: afterwards lane L holds the warp sum of slot ((L>>4)&1)*8+((L>>3)&1)*4+((L>>2)&1)*2+((L>>1)&1)
__device__ __forceinline__ float butterfly16(float (&a)[16], int lane)
{
    const bool up16 = lane & 16, up8 = lane & 8, up4 = lane & 4, up2 = lane & 2;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const float send = up16 ? a[q] : a[q + 8];
        const float keep = up16 ? a[q + 8] : a[q];
        a[q] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float send = up8 ? a[q] : a[q + 4];
        const float keep = up8 ? a[q + 4] : a[q];
        a[q] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        const float send = up4 ? a[q] : a[q + 2];
        const float keep = up4 ? a[q + 2] : a[q];
        a[q] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    const float send = up2 ? a[0] : a[1];
    const float keep = up2 ? a[1] : a[0];
    float r = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    r += __shfl_xor_sync(0xffffffffu, r, 1);
    return r;
}

__device__ __forceinline__ int butterfly16_slot(int lane)
{
    return ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
}

__global__ void __launch_bounds__(256) skin_bwd_kernel(int V, int J, const float *__restrict__ weights_T,
                                                        const float *__restrict__ A, const float *__restrict__ v_posed,
                                                        const float *__restrict__ d_verts, float *__restrict__ d_v_posed,
                                                        float *__restrict__ d_A)
{
    __shared__ float4 s_A[MAXJ * 3];
    __shared__ float s_dA[MAXJ * 12];
    const int f = blockIdx.y;
    for (int i = threadIdx.x; i < J * 3; i += blockDim.x) s_A[i] = reinterpret_cast<const float4 *>(A)[(int64_t)f * J * 3 + i];
    for (int i = threadIdx.x; i < J * 12; i += blockDim.x) s_dA[i] = 0.0f;
    __syncthreads();
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool live = v < V;
    float gx = 0.f, gy = 0.f, gz = 0.f, x = 0.f, y = 0.f, z = 0.f;
    if (live) {
        const float *g = d_verts + ((int64_t)f * V + v) * 3;
        gx = g[0]; gy = g[1]; gz = g[2];
        const float *vp = v_posed + ((int64_t)f * V + v) * 3;
        x = vp[0]; y = vp[1]; z = vp[2];
    }
    const int slot = butterfly16_slot(lane);
    float dx = 0.f, dy = 0.f, dz = 0.f;
    for (int j = 0; j < J; ++j) {
        const float w = live ? __ldg(weights_T + (int64_t)j * V + v) : 0.0f;
        if (!__any_sync(0xffffffffu, w != 0.0f)) continue;
        const float4 a0 = s_A[3 * j], a1 = s_A[3 * j + 1], a2 = s_A[3 * j + 2];
        // d_v_posed += w * A3^T g
        dx = fmaf(w, a0.x * gx + a1.x * gy + a2.x * gz, dx);
        dy = fmaf(w, a0.y * gx + a1.y * gy + a2.y * gz, dy);
        dz = fmaf(w, a0.z * gx + a1.z * gy + a2.z * gz, dz);
        const float wx = w * gx, wy = w * gy, wz = w * gz;
        float a[16] = {wx * x, wx * y, wx * z, wx, wy * x, wy * y, wy * z, wy, wz * x, wz * y, wz * z, wz, 0.f, 0.f, 0.f, 0.f};
        const float r = butterfly16(a, lane);
        if ((lane & 1) == 0 && slot < 12) atomicAdd(&s_dA[j * 12 + slot], r);
    }
    if (live) {
        float *o = d_v_posed + ((int64_t)f * V + v) * 3;
        o[0] = dx; o[1] = dy; o[2] = dz;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < J * 12; i += blockDim.x)
        if (s_dA[i] != 0.0f) atomicAdd(d_A + (int64_t)f * J * 12 + i, s_dA[i]);
}

// d_pose_feature[f][k] += sum_i posedirs[k][i] d_v_posed[f][i];  d_shape[f][s] += sum_{v,c} shapedirs[v][c][s] d_v_posed[f][v][c]
constexpr int PBB_F = 16;
constexpr int PBB_THREADS = 128;

__global__ void __launch_bounds__(PBB_THREADS) pose_blend_bwd_kernel(int F, int V, int S, int K,
                                                                      const float *__restrict__ d_v_posed,
                                                                      const float *__restrict__ shapedirs,
                                                                      const float *__restrict__ posedirs,
                                                                      float *__restrict__ d_pose_feature,
                                                                      float *__restrict__ d_shape)
{
    extern __shared__ float s_out[];                 // [K + S][PBB_F]
    const int f0 = blockIdx.y * PBB_F;
    for (int i = threadIdx.x; i < (K + S) * PBB_F; i += PBB_THREADS) s_out[i] = 0.0f;
    __syncthreads();
    const int v = blockIdx.x * PBB_THREADS + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool live = v < V;
    float d[PBB_F][3];
#pragma unroll
    for (int fb = 0; fb < PBB_F; ++fb) {
        const bool ok = live && f0 + fb < F;
        const float *g = d_v_posed + ((int64_t)(f0 + fb) * V + (live ? v : 0)) * 3;
        d[fb][0] = ok ? g[0] : 0.0f; d[fb][1] = ok ? g[1] : 0.0f; d[fb][2] = ok ? g[2] : 0.0f;
    }
    const int slot = butterfly16_slot(lane);
    const float *pd = posedirs + (int64_t)3 * (live ? v : 0);
    const int64_t ld = (int64_t)3 * V;
    // blockIdx.z splits the K pose-feature rows (few frames = few CTAs otherwise: 82 for a 4-subject training step)
    const int kchunk = (K + (int)gridDim.z - 1) / (int)gridDim.z;
    const int k_begin = (int)blockIdx.z * kchunk, k_end = min(K, k_begin + kchunk);
    for (int k = k_begin; k < k_end; ++k) {
        const float p0 = live ? __ldg(pd + k * ld) : 0.0f, p1 = live ? __ldg(pd + k * ld + 1) : 0.0f,
                    p2 = live ? __ldg(pd + k * ld + 2) : 0.0f;
        float a[16];
#pragma unroll
        for (int fb = 0; fb < PBB_F; ++fb) a[fb] = fmaf(p0, d[fb][0], fmaf(p1, d[fb][1], p2 * d[fb][2]));
        const float r = butterfly16(a, lane);
        if ((lane & 1) == 0) atomicAdd(&s_out[k * PBB_F + slot], r);
    }
    for (int s = 0; s < (blockIdx.z == 0 ? S : 0); ++s) {
        const float *sd = shapedirs + (int64_t)(live ? v : 0) * 3 * S + s;
        const float p0 = live ? __ldg(sd) : 0.0f, p1 = live ? __ldg(sd + S) : 0.0f, p2 = live ? __ldg(sd + 2 * S) : 0.0f;
        float a[16];
#pragma unroll
        for (int fb = 0; fb < PBB_F; ++fb) a[fb] = fmaf(p0, d[fb][0], fmaf(p1, d[fb][1], p2 * d[fb][2]));
        const float r = butterfly16(a, lane);
        if ((lane & 1) == 0) atomicAdd(&s_out[(K + s) * PBB_F + slot], r);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < (K + S) * PBB_F; i += PBB_THREADS) {
        const int k = i / PBB_F, fb = i - k * PBB_F;
        const float val = s_out[i];
        if (f0 + fb < F && val != 0.0f) {
            if (k < K) atomicAdd(d_pose_feature + (int64_t)(f0 + fb) * K + k, val);
            else atomicAdd(d_shape + (int64_t)(f0 + fb) * S + (k - K), val);
        }
    }
}

// One CTA (64 threads) per frame.
// SIXD: full_pose is [F,J,6] (the 6D heads, amav_lbs_forward_6d) and d_full_pose [F,J,6]: Gram-Schmidt backward in
// place of the Rodrigues backward.
template <bool SIXD>
__global__ void __launch_bounds__(64) chain_bwd_kernel(int J, int S, const float *__restrict__ full_pose,
                                                       const float *__restrict__ shape_comp,
                                                       const float *__restrict__ J_template,
                                                       const float *__restrict__ J_dirs,
                                                       const int32_t *__restrict__ parents,
                                                       const float *__restrict__ d_A,
                                                       const float *__restrict__ d_pose_feature,
                                                       float *__restrict__ d_full_pose, float *__restrict__ d_shape)
{
    const int f = blockIdx.x;
    __shared__ float s_R[MAXJ][9], s_J[MAXJ][3], s_G[MAXJ][12], s_dG[MAXJ][12], s_dR[MAXJ][9], s_dJ[MAXJ][3];
    __shared__ float s_rod[MAXJ][8];      // x, y, z (= r / angle), sin, cos, angle, -, -
    __shared__ int s_par[MAXJ];
    const int t = threadIdx.x;
    if (t < J) {
        if (SIXD) {
            const float *d6 = full_pose + ((int64_t)f * J + t) * 6;
            const float a1x = d6[0], a1y = d6[1], a1z = d6[2], a2x = d6[3], a2y = d6[4], a2z = d6[5];
            const float n1 = fmaxf(sqrtf(a1x * a1x + a1y * a1y + a1z * a1z), 1e-12f);
            const float b1x = a1x / n1, b1y = a1y / n1, b1z = a1z / n1;
            const float dt = b1x * a2x + b1y * a2y + b1z * a2z;
            const float ux = a2x - dt * b1x, uy = a2y - dt * b1y, uz = a2z - dt * b1z;
            const float n2 = fmaxf(sqrtf(ux * ux + uy * uy + uz * uz), 1e-12f);
            const float b2x = ux / n2, b2y = uy / n2, b2z = uz / n2;
            s_R[t][0] = b1x; s_R[t][1] = b1y; s_R[t][2] = b1z;
            s_R[t][3] = b2x; s_R[t][4] = b2y; s_R[t][5] = b2z;
            s_R[t][6] = b1y * b2z - b1z * b2y; s_R[t][7] = b1z * b2x - b1x * b2z; s_R[t][8] = b1x * b2y - b1y * b2x;
            s_rod[t][0] = n1; s_rod[t][1] = n2; s_rod[t][2] = dt;
        } else {
        const float rx = full_pose[((int64_t)f * J + t) * 3], ry = full_pose[((int64_t)f * J + t) * 3 + 1],
                    rz = full_pose[((int64_t)f * J + t) * 3 + 2];
        const float ex = rx + 1e-8f, ey = ry + 1e-8f, ez = rz + 1e-8f;
        const float angle = sqrtf(ex * ex + ey * ey + ez * ez);
        const float x = rx / angle, y = ry / angle, z = rz / angle;
        float sn, cs;
        sincosf(angle, &sn, &cs);
        const float oc = 1.0f - cs;
        s_R[t][0] = 1.0f + oc * (-(y * y + z * z)); s_R[t][1] = sn * (-z) + oc * (x * y); s_R[t][2] = sn * y + oc * (x * z);
        s_R[t][3] = sn * z + oc * (x * y); s_R[t][4] = 1.0f + oc * (-(x * x + z * z)); s_R[t][5] = sn * (-x) + oc * (y * z);
        s_R[t][6] = sn * (-y) + oc * (x * z); s_R[t][7] = sn * x + oc * (y * z); s_R[t][8] = 1.0f + oc * (-(x * x + y * y));
        s_rod[t][0] = x; s_rod[t][1] = y; s_rod[t][2] = z; s_rod[t][3] = sn; s_rod[t][4] = cs; s_rod[t][5] = angle;
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float a = J_template[3 * t + k];
            for (int s = 0; s < S; ++s) a = fmaf(J_dirs[((int64_t)t * 3 + k) * S + s], shape_comp[(int64_t)f * S + s], a);
            s_J[t][k] = a;
            s_dJ[t][k] = 0.0f;
        }
        s_par[t] = parents[t];
    }
    __syncthreads();
    if (t == 0) {
        // forward chain (values needed by the backward), then the backward chain, serially: 2 x 55 small steps
        for (int j = 0; j < J; ++j) {
            const int par = s_par[j];
            if (par < 0) {
                for (int r = 0; r < 3; ++r) {
                    for (int c = 0; c < 3; ++c) s_G[j][r * 4 + c] = s_R[j][r * 3 + c];
                    s_G[j][r * 4 + 3] = s_J[j][r];
                }
            } else {
                const float rel[3] = {s_J[j][0] - s_J[par][0], s_J[j][1] - s_J[par][1], s_J[j][2] - s_J[par][2]};
                for (int r = 0; r < 3; ++r) {
                    const float g0 = s_G[par][r * 4], g1 = s_G[par][r * 4 + 1], g2 = s_G[par][r * 4 + 2];
                    for (int c = 0; c < 3; ++c) s_G[j][r * 4 + c] = g0 * s_R[j][c] + g1 * s_R[j][3 + c] + g2 * s_R[j][6 + c];
                    s_G[j][r * 4 + 3] = g0 * rel[0] + g1 * rel[1] + g2 * rel[2] + s_G[par][r * 4 + 3];
                }
            }
        }
        // A_j = [G3 | Gt - G3 J_j]
        const float *dA = d_A + (int64_t)f * J * 12;
        for (int j = 0; j < J; ++j) {
            for (int r = 0; r < 3; ++r) {
                const float dAt = dA[j * 12 + r * 4 + 3];
                for (int c = 0; c < 3; ++c) {
                    s_dG[j][r * 4 + c] = dA[j * 12 + r * 4 + c] - dAt * s_J[j][c];
                    s_dJ[j][c] -= s_G[j][r * 4 + c] * dAt;
                }
                s_dG[j][r * 4 + 3] = dAt;
            }
        }
        for (int j = J - 1; j >= 0; --j) {
            const int par = s_par[j];
            if (par < 0) {
                for (int r = 0; r < 3; ++r) {
                    for (int c = 0; c < 3; ++c) s_dR[j][r * 3 + c] = s_dG[j][r * 4 + c];
                    s_dJ[j][r] += s_dG[j][r * 4 + 3];
                }
                continue;
            }
            const float rel[3] = {s_J[j][0] - s_J[par][0], s_J[j][1] - s_J[par][1], s_J[j][2] - s_J[par][2]};
            // dR_j = G3_p^T dG3_j;  drel = G3_p^T dGt_j
            for (int a = 0; a < 3; ++a) {
                for (int c = 0; c < 3; ++c)
                    s_dR[j][a * 3 + c] = s_G[par][a] * s_dG[j][c] + s_G[par][4 + a] * s_dG[j][4 + c] + s_G[par][8 + a] * s_dG[j][8 + c];
                const float dr = s_G[par][a] * s_dG[j][3] + s_G[par][4 + a] * s_dG[j][7] + s_G[par][8 + a] * s_dG[j][11];
                s_dJ[j][a] += dr;
                s_dJ[par][a] -= dr;
            }
            // dG3_p += dG3_j R_j^T + dGt_j (x) rel;  dGt_p += dGt_j
            for (int r = 0; r < 3; ++r) {
                for (int a = 0; a < 3; ++a)
                    s_dG[par][r * 4 + a] += s_dG[j][r * 4] * s_R[j][a * 3] + s_dG[j][r * 4 + 1] * s_R[j][a * 3 + 1] +
                                            s_dG[j][r * 4 + 2] * s_R[j][a * 3 + 2] + s_dG[j][r * 4 + 3] * rel[a];
                s_dG[par][r * 4 + 3] += s_dG[j][r * 4 + 3];
            }
        }
    }
    __syncthreads();
    if (t < J) {
        // pose feature = R_j - I (j >= 1) feeds the pose blend shapes
        float dR[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) dR[k] = s_dR[t][k] + (t > 0 ? d_pose_feature[(int64_t)f * (J - 1) * 9 + (t - 1) * 9 + k] : 0.0f);
        if (SIXD) {
            // R rows: b1 = a1/|a1|, b2 = u/|u| with u = a2 - (b1.a2) b1, b3 = b1 x b2
            const float n1 = s_rod[t][0], n2 = s_rod[t][1], dt = s_rod[t][2];
            const float b1x = s_R[t][0], b1y = s_R[t][1], b1z = s_R[t][2], b2x = s_R[t][3], b2y = s_R[t][4], b2z = s_R[t][5];
            const float *d6 = full_pose + ((int64_t)f * J + t) * 6;
            const float a2x = d6[3], a2y = d6[4], a2z = d6[5];
            // b3 = b1 x b2: d/db1 = b2 x g3, d/db2 = g3 x b1
            float gb1x = dR[0] + (b2y * dR[8] - b2z * dR[7]), gb1y = dR[1] + (b2z * dR[6] - b2x * dR[8]),
                  gb1z = dR[2] + (b2x * dR[7] - b2y * dR[6]);
            const float gb2x = dR[3] + (dR[7] * b1z - dR[8] * b1y), gb2y = dR[4] + (dR[8] * b1x - dR[6] * b1z),
                        gb2z = dR[5] + (dR[6] * b1y - dR[7] * b1x);
            const float p2 = b2x * gb2x + b2y * gb2y + b2z * gb2z;
            const float gux = (gb2x - b2x * p2) / n2, guy = (gb2y - b2y * p2) / n2, guz = (gb2z - b2z * p2) / n2;
            const float q = b1x * gux + b1y * guy + b1z * guz;
            gb1x -= q * a2x + dt * gux; gb1y -= q * a2y + dt * guy; gb1z -= q * a2z + dt * guz;
            const float p1 = b1x * gb1x + b1y * gb1y + b1z * gb1z;
            float *o = d_full_pose + ((int64_t)f * J + t) * 6;
            o[0] = (gb1x - b1x * p1) / n1; o[1] = (gb1y - b1y * p1) / n1; o[2] = (gb1z - b1z * p1) / n1;
            o[3] = gux - b1x * q; o[4] = guy - b1y * q; o[5] = guz - b1z * q;
        } else {
        const float x = s_rod[t][0], y = s_rod[t][1], z = s_rod[t][2], sn = s_rod[t][3], cs = s_rod[t][4], angle = s_rod[t][5];
        const float oc = 1.0f - cs;
        // R = I + sn K + oc M, K = [[0,-z,y],[z,0,-x],[-y,x,0]], M = K K
        float dnx = sn * (dR[7] - dR[5]), dny = sn * (dR[2] - dR[6]), dnz = sn * (dR[3] - dR[1]);
        dnx += oc * (dR[1] * y + dR[2] * z + dR[3] * y + dR[6] * z - 2.0f * x * (dR[4] + dR[8]));
        dny += oc * (dR[1] * x + dR[3] * x + dR[5] * z + dR[7] * z - 2.0f * y * (dR[0] + dR[8]));
        dnz += oc * (dR[2] * x + dR[6] * x + dR[5] * y + dR[7] * y - 2.0f * z * (dR[0] + dR[4]));
        const float dRK = -z * dR[1] + y * dR[2] + z * dR[3] - x * dR[5] - y * dR[6] + x * dR[7];
        const float dRM = -(y * y + z * z) * dR[0] + x * y * (dR[1] + dR[3]) + x * z * (dR[2] + dR[6]) - (x * x + z * z) * dR[4] +
                          y * z * (dR[5] + dR[7]) - (x * x + y * y) * dR[8];
        const float dth = cs * dRK + sn * dRM;
        // n = r / angle, angle = |r + 1e-8|
        const float rx = x * angle, ry = y * angle, rz = z * angle;
        const float dth_t = dth - (dnx * rx + dny * ry + dnz * rz) / (angle * angle);
        float *o = d_full_pose + ((int64_t)f * J + t) * 3;
        o[0] = dnx / angle + dth_t * (rx + 1e-8f) / angle;
        o[1] = dny / angle + dth_t * (ry + 1e-8f) / angle;
        o[2] = dnz / angle + dth_t * (rz + 1e-8f) / angle;
        }
    }
    // joints = J_template + J_dirs . shape
    for (int s = t; s < S; s += blockDim.x) {
        float a = 0.0f;
        for (int j = 0; j < J; ++j)
#pragma unroll
            for (int k = 0; k < 3; ++k) a = fmaf(J_dirs[((int64_t)j * 3 + k) * S + s], s_dJ[j][k], a);
        atomicAdd(d_shape + (int64_t)f * S + s, a);
    }
}

}  // namespace

extern "C" int amav_densify_gather_backward(int F, int V, int N, int P, const float *d_sites, const int32_t *par,
                                            const float *wgt, float *d_verts, void *stream)
{
    AMAV_CHECK_ARG(F >= 0 && F <= 65535 && V > 0 && N >= 0 && P >= 1, "bad dims");
    if ((int64_t)F * N == 0) return AMAV_OK;
    AMAV_CHECK_ARG(d_sites && par && wgt && d_verts, "null buffer");
    dim3 g(amav_div_up(N, 256), F);
    densify_bwd_kernel<<<g, 256, 0, (cudaStream_t)stream>>>(V, N, P, d_sites, par, wgt, d_verts);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}

static int lbs_backward_impl(bool sixd, int F, int V, int J, int S, const float *full_pose, const float *shape_comp,
                                 const float *shapedirs, const float *posedirs, const int32_t *parents,
                                 const float *J_template, const float *J_dirs, const float *weights_T, const float *A,
                                 const float *v_posed, const float *d_verts, float *d_v_posed, float *d_A,
                                 float *d_pose_feature, float *d_full_pose, float *d_shape, void *stream_)
{
    AMAV_CHECK_ARG(F >= 0 && F <= 65535 && V > 0 && J > 1 && J <= MAXJ && S >= 0, "bad dims (1 < J <= 64)");
    if (F == 0) return AMAV_OK;
    AMAV_CHECK_ARG(full_pose && shape_comp && shapedirs && posedirs && parents && J_template && J_dirs && weights_T, "null input");
    AMAV_CHECK_ARG(A && v_posed && d_verts, "null forward state / gradient");
    AMAV_CHECK_ARG(d_v_posed && d_A && d_pose_feature && d_full_pose && d_shape, "null work/output buffer");
    AMAV_CHECK_ARG(((uintptr_t)A & 15) == 0, "A must be 16-byte aligned");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int K = (J - 1) * 9;
    dim3 gs(amav_div_up(V, 256), F);
    skin_bwd_kernel<<<gs, 256, 0, stream>>>(V, J, weights_T, A, v_posed, d_verts, d_v_posed, d_A);
    const size_t smem = (size_t)(K + S) * PBB_F * sizeof(float);
    AMAV_CHECK_ARG(smem <= 200 * 1024, "pose basis too large");
    amav_ensure_dyn_smem((const void *)pose_blend_bwd_kernel, (int)smem);
    dim3 gb(amav_div_up(V, PBB_THREADS), amav_div_up(F, PBB_F));
    const int ksplit = (4 * AMAV_SM_COUNT + (int)(gb.x * gb.y) - 1) / (int)(gb.x * gb.y);
    gb.z = ksplit < 1 ? 1 : (ksplit > 8 ? 8 : ksplit);
    pose_blend_bwd_kernel<<<gb, PBB_THREADS, smem, stream>>>(F, V, S, K, d_v_posed, shapedirs, posedirs, d_pose_feature, d_shape);
    if (sixd)
        chain_bwd_kernel<true><<<F, 64, 0, stream>>>(J, S, full_pose, shape_comp, J_template, J_dirs, parents, d_A,
                                                     d_pose_feature, d_full_pose, d_shape);
    else
        chain_bwd_kernel<false><<<F, 64, 0, stream>>>(J, S, full_pose, shape_comp, J_template, J_dirs, parents, d_A,
                                                      d_pose_feature, d_full_pose, d_shape);
    amav_count_launches(2);
    AMAV_CHECK_LAUNCH();
    return AMAV_OK;
}

extern "C" int amav_lbs_backward(int F, int V, int J, int S, const float *full_pose, const float *shape_comp,
                                 const float *shapedirs, const float *posedirs, const int32_t *parents,
                                 const float *J_template, const float *J_dirs, const float *weights_T, const float *A,
                                 const float *v_posed, const float *d_verts, float *d_v_posed, float *d_A,
                                 float *d_pose_feature, float *d_full_pose, float *d_shape, void *stream)
{
    return lbs_backward_impl(false, F, V, J, S, full_pose, shape_comp, shapedirs, posedirs, parents, J_template, J_dirs, weights_T, A,
                             v_posed, d_verts, d_v_posed, d_A, d_pose_feature, d_full_pose, d_shape, stream);
}

extern "C" int amav_lbs_backward_6d(int F, int V, int J, int S, const float *pose_6d, const float *shape_comp,
                                 const float *shapedirs, const float *posedirs, const int32_t *parents,
                                 const float *J_template, const float *J_dirs, const float *weights_T, const float *A,
                                 const float *v_posed, const float *d_verts, float *d_v_posed, float *d_A,
                                 float *d_pose_feature, float *d_pose_6d, float *d_shape, void *stream)
{
    return lbs_backward_impl(true, F, V, J, S, pose_6d, shape_comp, shapedirs, posedirs, parents, J_template, J_dirs, weights_T, A,
                             v_posed, d_verts, d_v_posed, d_A, d_pose_feature, d_pose_6d, d_shape, stream);
}
