/*
 * amav.h -- C ABI of libamav_b200.so: the B200 (sm_100a) render path of
 * liubingqi7/audio-motion-avatar's src/models/renderer.py.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name starts with h_;
 *   - DEVICE CONTRACT: kernels are launched on the calling thread's CURRENT CUDA device (cudaSetDevice) and on
 *     `stream`, which must belong to that device; every buffer must be memory of that device, with the one documented
 *     exception of amav_splat_blend's out_color, which may be a peer mapping of another GPU's memory.  The Python
 *     mirror makes the tensors' device current around every call (audio_motion_avatar_b200/_lib.py: call);
 *   - the caller allocates every input, output and workspace buffer; the library never
 *     allocates, frees or keeps state between calls (except the last-error string);
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*); no call synchronises
 *     the device;
 *   - return value: 0 on success, negative AMAV_E* on failure; amav_last_error() describes it;
 *   - F = frames in the call, N = Gaussians per frame, tiles = ceil(W/16)*ceil(H/16);
 *   - matrices are the 16 floats the reference hands to GaussianRasterizationSettings
 *     (renderer.py:523-524): already transposed, element (row r, col c) at m[c*4+r].
 *
 * Each entry point cites the reference interface it replaces (file:line under the reference repo).
 */
#ifndef AMAV_H
#define AMAV_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AMAV_OK 0
#define AMAV_EINVAL (-1)   /* bad argument                           */
#define AMAV_ECUDA (-2)    /* CUDA launch / runtime error            */
#define AMAV_ECAP (-3)     /* caller-provided capacity too small     */

/* colour layout argument (`hwc`) of amav_splat_blend / amav_splat_blend_backward */
#define AMAV_LAYOUT_CHW 0         /* [F,3,H,W]: what GaussianRasterizer returns per frame (renderer.py:557) */
#define AMAV_LAYOUT_HWC 1         /* [F,H,W,3]: what render_batch returns (renderer.py:477-479); the library asks the driver
                                     whether out_color is another GPU's memory and then stores whole rows (cached per pointer) */
#define AMAV_LAYOUT_HWC_ROWS 2    /* HWC, row stores forced (the caller knows the buffer is a peer mapping); forward only */
#define AMAV_LAYOUT_HWC_PIXELS 3  /* HWC, per-pixel stores forced; forward only */

#define AMAV_TILE 16       /* tile edge in pixels (upstream BLOCK_X/BLOCK_Y) */
#define AMAV_DEC_OUT 14    /* decoded attributes per Gaussian: xyz3 rot4 scale3 opacity1 shs3 */
#define AMAV_DEC_PAD 16    /* padded row of the projected triplane */

const char *amav_last_error(void);
int amav_version(void);
/* number of CUDA kernels this library has launched since it was loaded (all threads) */
int64_t amav_launch_count(void);

/* ---------------------------------------------------------------------------------------------
 * Stage 1: triplane projection.  Replaces the channel dimension of F.grid_sample + 5x nn.Linear
 * (renderer.py:305-312 and :165-171) by projecting the triplane through the decoder heads first
 * (the decode is linear, so decode(bilinear(T)) == bilinear(decode(T))).
 *   tokens  [F, C, 3*R*R] fp32, channel-major (renderer.py:85-91: plane p of frame f is
 *           tokens[f][:, p*R*R:(p+1)*R*R]); frame stride = token_frame_stride floats
 *   weight  [14, 3+3C] fp32: rows = xyz(3) rot(4) scale(3) opacity(1) shs(3) (renderer.py:51-55)
 *   proj    [F, 3, R*R, 16] fp32 out: proj[f][p][texel][slot] = sum_c weight[row(slot)][3+p*C+c]*tokens[f][c][p*R*R+texel]
 *           slots: 0..3 rotation (rows 3..6), 4..6 xyz (rows 0..2), 7 opacity (row 10), 8..10 scaling (rows 7..9),
 *           12..14 shs (rows 11..13), 11 and 15 zero.  C % 64 == 0, (R*R) % 128 == 0.
 */
int amav_triplane_project(int F, int C, int R, const float *tokens, int64_t token_frame_stride,
                          const float *weight, float *proj, void *stream);

/* amav_triplane_project on the 5th-generation tensor cores (csrc/triplane_tc.cu): same inputs and output as
 * amav_triplane_project; tcgen05.mma kind::tf32 with the fp32 operands split into TF32 hi + lo parts (3 MMAs per K step,
 * fp32-level accuracy), accumulator in TMEM.  Replaces the five nn.Linear of src/models/renderer.py:165-171 (cuBLAS SGEMM
 * there).  wprep_scratch: amav_triplane_project_tc_scratch_elems(C) floats of caller-owned scratch (the split weights). */
int64_t amav_triplane_project_tc_scratch_elems(int C);
int amav_triplane_project_tc(int F, int C, int R, const float *tokens, int64_t token_frame_stride,
                             const float *weight, float *wprep_scratch, float *proj, void *stream);

/* Allocation-free variant of amav_triplane_project_tc.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  Same inputs, output and arithmetic as amav_triplane_project_tc; the split-weight
 * scratch is a file-scope device array of the library (192 KB, resident on every GPU that loads the library until the
 * process exits) instead of a caller buffer, so C <= AMAV_TC_STATIC_MAX_C (larger C: AMAV_EINVAL with a message naming
 * the limit).  The call makes no cudaMalloc / cudaFree / allocating library call.  Every call on a device shares the
 * array: each call rewrites all of the part it reads (nothing relies on earlier contents), and calls that OVERLAP on the
 * GPU (different streams, or two host threads on one device without ordering) race on it -- serialise them. */
#define AMAV_TC_STATIC_MAX_C 512
int amav_triplane_project_tc_static(int F, int C, int R, const float *tokens, int64_t token_frame_stride,
                                    const float *weight, float *proj, void *stream);

/* Same projection from the producer's own layout: tokens_tm [F, 3*R*R, C] token-major, i.e. the upstream
 * transformer's hidden_states before its `.permute(0, 2, 1).contiguous()` (src/models/transformers.py:1070-1073; the
 * residual it adds afterwards is added in the same layout).  Texel (p, t) of frame f is the C contiguous floats at
 * tokens_tm[f][p*R*R + t]; frame stride in floats.  Output identical in layout to amav_triplane_project and equal up to
 * fp32 summation order.  C % 64 == 0, (R*R) % 128 == 0.
 */
int amav_triplane_project_tm(int F, int C, int R, const float *tokens_tm, int64_t token_frame_stride,
                             const float *weight, float *proj, void *stream);

/* Stage 2: bilinear query of the projected triplane + xyz/bias terms + construct_gaussians
 * (renderer.py:292-317, :165-171, :319-346).
 *   points [F,N,3]  posed, pre-transl sites;  transl [F,3];  weight [14,3+3C]; bias [14]
 *   outputs (the `gaussians` dict of renderer.py:337-344, each [F,N,k]):
 *   xyz[3] = points+offset+transl, scale[3] raw, rot[4] normalised (eps 1e-12), opacity[1] raw,
 *   color[3] = sigmoid(shs).
 *   order [N] int32 or NULL: a permutation of the site indices; thread group i processes site order[i] (results are
 *   written at the site's own index, so the output does not depend on it).  A locality order fixed at set-up (Morton
 *   order of the rest-pose sites) lets neighbouring lanes share texels; the reference's randperm order shares none.
 */
int amav_sample_decode(int F, int N, int C, int R, float radius, const float *proj, const float *points,
                       const float *transl, const float *weight, const float *bias,
                       float *xyz, float *scale, float *rot, float *opacity, float *color, const int32_t *order, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Stage 3: SMPL-X forward (smplx==0.1.28 SMPLX.forward -> lbs.lbs; call site renderer.py:261-274)
 * for F frames in three launches, and the densify gather (renderer.py:276-288).
 *
 * amav_lbs_precompute: once per body model (renderer.py:206-225 builds the model once too).
 *   v_template [V,3]; shapedirs [V,3,S]; J_regressor [J,V]; lbs_weights [V,J]
 *   out: J_template [J,3] = J_regressor.v_template; J_dirs [J,3,S] = J_regressor.shapedirs
 *        (so joints = J_template + J_dirs.shape = J_regressor.v_shaped); weights_T [J,V].
 * amav_lbs_forward:
 *   full_pose [F,J*3] axis-angle incl. pose_mean, joint order of smplx.SMPLX.forward;
 *   shape_comp [F,S] = [betas, expression]; posedirs [(J-1)*9, 3V]; parents [J] int32 (parents[0] = -1)
 *   work: pose_feature [F,(J-1)*9], joints [F,J,3], A [F,J,12] (3x4 relative transforms, 16-byte aligned),
 *         v_posed [F,V,3]
 *   verts [F,V,3] out (no transl: renderer.py:248-249,271)
 */
int amav_lbs_precompute(int V, int J, int S, const float *v_template, const float *shapedirs,
                        const float *J_regressor, const float *lbs_weights, float *J_template,
                        float *J_dirs, float *weights_T, void *stream);
int amav_lbs_forward(int F, int V, int J, int S, const float *full_pose, const float *shape_comp,
                     const float *v_template, const float *shapedirs, const float *posedirs,
                     const int32_t *parents, const float *J_template, const float *J_dirs,
                     const float *weights_T, float *pose_feature, float *joints, float *A, float *v_posed,
                     float *verts, void *stream);

/* Same forward with the pose given as pose_6d [F,J,6], the 6D rotation representation the SMPL-X parameter decoder's
 * heads emit (src/models/smplx_decoder.py:106-127), in the joint order of full_pose.  R_j comes from
 * Gram-Schmidt (pytorch3d.transforms.rotation_6d_to_matrix: the 6 numbers are the first two ROWS of R) instead of the
 * reference's 6D -> matrix -> matrix_to_axis_angle -> batch_rodrigues round trip, which reproduces the same matrix up to
 * fp32 rounding.  Only valid for a body model whose pose_mean is zero (flat_hand_mean); the host layer checks.
 * Backward: amav_lbs_backward_6d. */
int amav_lbs_forward_6d(int F, int V, int J, int S, const float *pose_6d, const float *shape_comp,
                        const float *v_template, const float *shapedirs, const float *posedirs,
                        const int32_t *parents, const float *J_template, const float *J_dirs,
                        const float *weights_T, float *pose_feature, float *joints, float *A, float *v_posed,
                        float *verts, void *stream);

/* sites[f][i] = sum_k wgt[i][k] * verts[f][par[i][k]], k < P (flattened SubdivideMeshes levels +
 * randperm subset, renderer.py:280-288).  P = 1 with wgt = 1 is the plain gather.  order: as in amav_sample_decode. */
int amav_densify_gather(int F, int V, int N, int P, const float *verts, const int32_t *par,
                        const float *wgt, float *sites, const int32_t *order, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Stage 4: tile-based Gaussian splatting (diff_gaussian_rasterization forward as called at
 * renderer.py:516-566, batched over frames; no per-frame host sync).
 */
typedef struct AmavSplatDims {
    int32_t frames;          /* F */
    int32_t num_gaussians;   /* N per Gaussian set */
    int32_t views_per_set;   /* frame f reads Gaussian set f / views_per_set (render_multi_view, renderer.py:431-445); 1 otherwise */
    int32_t width, height;
    int32_t activate;        /* 1: apply render_one's activations (renderer.py:532-533,547) to raw scale/opacity/colour */
    float scale_modifier;    /* renderer.py:522 (1.0) */
} AmavSplatDims;

/* 4a  per-Gaussian projection (upstream preprocessCUDA) + per-tile counting.
 *   view, proj [F,16]; tanfov [F,2] = (tanfovx, tanfovy)
 *   geom   [F,N,12] fp32: (x, y, -conic.x/2, -conic.y) (-conic.z/2, opacity, pmin, r) (g, b, depth, radius);
 *          pmin = log((1/255)/opacity) - 1e-4: blend's conservative pre-test before exp()
 *   rect   [F,N,4] uint16: tile rectangle (minx, miny, maxx, maxy); all zero when culled
 *   tiles_touched [F,N] int32
 *   tile_count [F*tiles] uint32: MUST be zeroed by the caller; incremented once per (Gaussian, tile)
 */
int amav_splat_preprocess(const AmavSplatDims *d, const float *means3d, const float *scales,
                          const float *rots, const float *opacities, const float *colors,
                          const float *view, const float *proj, const float *tanfov,
                          float *geom, uint16_t *rect, int32_t *tiles_touched, uint32_t *tile_count,
                          void *stream);

/* 4b  binning.  amav_splat_scan: exclusive scan of tile_count -> tile_start [F*tiles+1]
 *     (tile_start[last] = D, the number of (Gaussian, tile) entries; the only value the host may
 *     want to read back, once per call, to size `entries`).  amav_splat_bin: scatter of
 *     (depth_bits<<32 | gaussian) entries into per-tile buckets + per-tile sort.
 *     Together equivalent to upstream's InclusiveSum + duplicateWithKeys + SortPairs +
 *     identifyTileRanges: tile t's sorted run is point_list[tile_start[t] : tile_start[t+1]], ordered
 *     by (depth bits, Gaussian index) = a stable sort of (tile<<32 | depth) keys.
 *   scan_tmp   [amav_splat_scan_tmp_elems(F*tiles)] uint32 scratch.  From element amav_splat_scan_order_offset(F*tiles)
 *              on it holds the deep-tile list the blend kernels take as `tile_order`: [0] = number of tiles with
 *              >= 4096 entries (listed from the front), [1] = number with 2048..4095 (listed from the back), then
 *              F*tiles words of list.  Valid until the next amav_splat_scan on the same scratch.
 *   tile_cursor[F*tiles] uint32: MUST be zeroed by the caller (scratch: bucket cursors, then the large-bucket sort's queue head)
 *   entries    [capacity] uint64 scratch; point_list [capacity] int32 out (Gaussian index within the frame)
 *   overflow   [2] int32, MUST be zeroed by the caller: [0] is set to 1 (and binning truncated) when D > capacity;
 *              [1] is the sort kernel's tile queue head
 */
int64_t amav_splat_scan_tmp_elems(int64_t num_tiles_total);
int64_t amav_splat_scan_order_offset(int64_t num_tiles_total);
int amav_splat_scan(const AmavSplatDims *d, const uint32_t *tile_count, uint32_t *tile_start,
                    uint32_t *scan_tmp, void *stream);
int amav_splat_bin(const AmavSplatDims *d, const float *geom, const uint16_t *rect,
                   const uint32_t *tile_start, uint32_t *tile_cursor, uint64_t *entries,
                   int32_t *point_list, int64_t capacity, int32_t *overflow, void *stream);

/* 4c  blend (upstream renderCUDA) over all tiles of all frames.
 *   bg [3]; out_color [F,3,H,W] (hwc=AMAV_LAYOUT_CHW) or [F,H,W,3] (AMAV_LAYOUT_HWC*, the layout render_batch returns,
 *   renderer.py:477-479); clamp01=1 applies renderer.py:568.
 *   final_T [F,H,W], n_contrib [F,H,W] int32, out_invdepth [F,H,W]: each may be NULL (inference).
 *   capacity: entries in point_list (same value as given to amav_splat_bin); tile ranges are clamped to it.
 *   work_counter [1] uint32: MUST be zeroed by the caller (tiles are handed to CTAs dynamically).
 *   tile_order: the deep-tile list written by amav_splat_scan (see scan_tmp), or NULL for plain frame order.  A tile
 *   is one CTA's sequential walk, so the deepest tiles go first (in launches of up to 32 frames, where the tail
 *   matters); the result does not depend on the order.
 */
int amav_splat_blend(const AmavSplatDims *d, const float *geom, const uint32_t *tile_start,
                     const int32_t *point_list, int64_t capacity, const float *bg, int hwc, int clamp01,
                     float *out_color, float *final_T, int32_t *n_contrib, float *out_invdepth,
                     uint32_t *work_counter, const uint32_t *tile_order, void *stream);

/* Camera set-up of render_one for F frames (renderer.py:486-510; graphic_utils.py:67-78,103-136,144-145):
 *   K [F,3,3] intrinsics, E [F,4,4] world-to-camera extrinsics ->
 *   view [F,4,4] = getWorld2View2_torch(R, T)^T, full [F,4,4] = view x getProjectionMatrix_torch(...)^T (the two
 *   matrices GaussianRasterizationSettings takes, already transposed), tanfov [F,2] = tan(fov/2) of focal2fov_torch.
 *   The reference's double fp32 inversion of Rt is replaced by Rt itself (~1e-7 relative difference). */
int amav_camera_setup(int F, const float *K, const float *E, int width, int height, double znear, double zfar,
                      float *view, float *full, float *tanfov, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Stage 4 backward (diff_gaussian_rasterization's _RasterizeGaussians.backward, reached in the reference
 * through autograd from renderer.py:555-568).  Needs the forward's geom, tile_start,
 * point_list, final_T and n_contrib.
 *
 * amav_splat_blend_backward: dL/d(colour output) -> per-(frame, Gaussian) screen-space sums
 *   dL_dcolor [F,3,H,W] (hwc=0) or [F,H,W,3] (hwc=1): gradient w.r.t. the blend's colour output (the [0,1] clamp
 *             of renderer.py:568 is the identity for colours and backgrounds inside [0,1])
 *   dsum      [F,N,12] fp32, MUST be zeroed by the caller, 16-byte aligned:
 *             0,1 dL/dmean2D (x 0.5*W, 0.5*H as upstream), 2..4 dL/dconic (A, B, C; B holds half the true
 *             derivative, upstream's convention), 5 dL/dopacity, 6..8 dL/drgb, 9..11 unused
 *   work_counter [1] uint32: MUST be zeroed by the caller; tile_order as in amav_splat_blend (or NULL).
 * amav_splat_preprocess_backward: sums -> gradients of the rasteriser inputs, through render_one's activations when
 *   d->activate (renderer.py:532-533,547), summed over the views_per_set views that share a Gaussian set.
 *   d_means3d/d_scales/d_colors [G,N,3], d_rots [G,N,4] (16-byte aligned), d_opacities [G,N], G = F/views_per_set.
 */
int amav_splat_blend_backward(const AmavSplatDims *d, const float *geom, const uint32_t *tile_start,
                              const int32_t *point_list, int64_t capacity, const float *bg, int hwc,
                              const float *final_T, const int32_t *n_contrib, const float *dL_dcolor,
                              float *dsum, uint32_t *work_counter, const uint32_t *tile_order, void *stream);
int amav_splat_preprocess_backward(const AmavSplatDims *d, const float *means3d, const float *scales,
                                   const float *rots, const float *opacities, const float *colors,
                                   const float *view, const float *proj, const float *tanfov,
                                   const float *geom, const float *dsum, float *d_means3d, float *d_scales,
                                   float *d_rots, float *d_opacities, float *d_colors, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Stages 1+2 backward (autograd of renderer.py:292-317 grid_sample, :165-171 the five Linear heads and
 * :319-346 construct_gaussians in the reference).
 *
 * amav_sample_decode_backward: gradients w.r.t. the `gaussians` dict entries (g_*: [F,N,k], each may be NULL = 0)
 *   -> d_proj   [F,3,R*R,16], MUST be zeroed by the caller: gradient of the projected planes (float4 reductions)
 *      d_points [F,N,3] (may be NULL): direct term + xyz columns of the heads + bilinear position derivative
 *                (zero outside the clamp of renderer.py:298)
 *      d_transl [F,3] (may be NULL), MUST be zeroed by the caller
 *      d_wb     [14,4], MUST be zeroed by the caller: dL/dW[:, 0:3] in columns 0..2 and dL/dbias in column 3
 * amav_triplane_project_backward:
 *      d_tokens [F,C,3*R*R] (frame stride d_token_frame_stride; may be NULL) = W_feat^T . d_proj, overwritten, or
 *                added to when accumulate_tokens != 0
 *      d_weight [14,3+3C] (may be NULL; then tokens may be NULL): feature columns 3.. are ADDED to (zero it first)
 */
int amav_sample_decode_backward(int F, int N, int C, int R, float radius, const float *proj, const float *points,
                                const float *weight, const float *bias, const float *g_xyz, const float *g_scale,
                                const float *g_rot, const float *g_opacity, const float *g_color, float *d_proj,
                                float *d_points, float *d_transl, float *d_wb, void *stream);
int amav_triplane_project_backward(int F, int C, int R, const float *d_proj, const float *tokens,
                                   int64_t token_frame_stride, const float *weight, float *d_tokens,
                                   int64_t d_token_frame_stride, int accumulate_tokens, float *d_weight, void *stream);

/* amav_triplane_project_backward_tm: the same two products for the producer's token-major layout (amav_triplane_project_tm):
 * tokens_tm, d_tokens_tm [F, 3*R*R, C].  C % 64 == 0, (R*R) % 128 == 0. */
int amav_triplane_project_backward_tm(int F, int C, int R, const float *d_proj, const float *tokens_tm,
                                      int64_t token_frame_stride, const float *weight, float *d_tokens_tm,
                                      int64_t d_token_frame_stride, int accumulate_tokens, float *d_weight, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Stage 3 backward (autograd of smplx.lbs.lbs and of the densify gather in the reference, renderer.py:261-288;
 * exercised when the SMPL-X parameters are predicted, i.e. stage-2 training).
 *
 * amav_densify_gather_backward: d_verts [F,V,3] (MUST be zeroed by the caller) += wgt * d_sites [F,N,3].
 * amav_lbs_backward: d_verts [F,V,3] -> d_full_pose [F,J*3] (overwritten), d_shape [F,S] (MUST be zeroed).
 *   A, v_posed: the forward's work buffers (amav_lbs_forward); d_v_posed [F,V,3] scratch;
 *   d_A [F,J,12] and d_pose_feature [F,(J-1)*9] scratch, MUST be zeroed by the caller.
 */
/* amav_lbs_backward_6d: the same backward for the 6D pose entry (amav_lbs_forward_6d): pose_6d [F,J,6] in,
 * d_pose_6d [F,J,6] out (overwritten) through the Gram-Schmidt construction instead of Rodrigues -- the gradient the
 * reference gets by autograd through rotation_6d_to_matrix -> matrix_to_axis_angle -> batch_rodrigues
 * (smplx_decoder.py:106-127), i.e. what reaches the SMPL-X parameter decoder in stage-2 training. */
int amav_lbs_backward_6d(int F, int V, int J, int S, const float *pose_6d, const float *shape_comp,
                         const float *shapedirs, const float *posedirs, const int32_t *parents,
                         const float *J_template, const float *J_dirs, const float *weights_T, const float *A,
                         const float *v_posed, const float *d_verts, float *d_v_posed, float *d_A,
                         float *d_pose_feature, float *d_pose_6d, float *d_shape, void *stream);
int amav_densify_gather_backward(int F, int V, int N, int P, const float *d_sites, const int32_t *par,
                                 const float *wgt, float *d_verts, void *stream);
int amav_lbs_backward(int F, int V, int J, int S, const float *full_pose, const float *shape_comp,
                      const float *shapedirs, const float *posedirs, const int32_t *parents,
                      const float *J_template, const float *J_dirs, const float *weights_T, const float *A,
                      const float *v_posed, const float *d_verts, float *d_v_posed, float *d_A,
                      float *d_pose_feature, float *d_full_pose, float *d_shape, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Image losses of the training step (src/utils/loss_utils.py:18 l1_loss, :44-83 ssim; combined at
 * lightning_model_wrapper.py:138-145) on the HWC frames the blend writes, and their gradient w.r.t. the rendered image.
 *   img, gt [F,H,W,3];  11x11 Gaussian window (sigma 1.5), zero padding, C1 = 0.01^2, C2 = 0.03^2.
 * amav_image_loss_forward:  sums [F,2], MUST be zeroed by the caller: [f][0] += sum |img - gt|, [f][1] += sum of the SSIM
 *                           map over the frame's H*W*3 values; maps [F,3,3,H,W] (may be NULL when no gradient is wanted):
 *                           dS/dmu1, dS/dE[x^2], dS/dE[xy] per channel.
 * amav_image_loss_backward: grad [F,H,W,3] = g_l1[f] * sign(img - gt) + g_ssim[f] * d(sum SSIM map)/d(img)
 *                           (g_l1, g_ssim: DEVICE arrays [F]; e.g. 1/n and -0.1/n for l1 + 0.1 * (1 - ssim)).
 */
int amav_image_loss_forward(int F, int H, int W, const float *img, const float *gt, float *sums, float *maps,
                            void *stream);
int amav_image_loss_backward(int F, int H, int W, const float *img, const float *gt, const float *maps,
                             const float *g_l1, const float *g_ssim, float *grad, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Video sink: the demo's host loop `(frame * 255).astype(np.uint8)`, cv2.cvtColor(RGB2BGR) and the
 * even/odd interleave of two rendered sequences (src/main2.py:226-239, :332-363) as one pass on the device.
 *   frames [F,H,W,3] fp32 (what render_batch returns), frame stride in floats;
 *   out    uint8 [*,H,W,3], frame stride in bytes: input frame f lands in output frame out_first + f * out_step
 *          (out_step = 2 with out_first = 0 / 1 interleaves two sequences); bgr != 0 swaps channels 0 and 2.
 *   Conversion: fp32 product with 255, truncation toward zero (numpy astype), saturated to [0,255].
 */
int amav_frames_to_u8(int F, int H, int W, const float *frames, int64_t in_frame_stride, uint8_t *out,
                      int64_t out_frame_stride, int out_first, int out_step, int bgr, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* AMAV_H */
