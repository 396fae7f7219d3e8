"""Drop-in for the reference's ``src/models/renderer.py`` (liubingqi7/audio-motion-avatar): same names,
argument meaning and return shapes -- ``Renderer``, ``render_batch``, ``render_one``, ``render_multi_view``,
``SCALE_BIAS``, ``OPACITY_BIAS``, ``SUBDEVIDE_VERTS`` -- on top of libamav_b200.so.

Differences from the reference, all deliberate:
* all frames of a call go through ONE batched pass per stage (no Python loop over frames, renderer.py:475-477;
  no per-frame host sync, renderer.py:504-505 / graphic_utils.py:124-125);
* the triplane is projected through the linear heads before it is sampled (fact 5) instead of materialising
  [F,N,1536] features;
* the PTv3 point encoder/refiner (renderer.py:34-47,143-151) is bypassed: its last layer is zero-initialised so
  it is an exact no-op at random init, and spconv is not available; the conv triplane upsampler is not built;
* ``cfg.num_gaussians`` (optional) overrides SUBDEVIDE_VERTS so 100k/500k sites are possible;
* there is no CPU path: tensors are moved to ``cfg.device`` (a CUDA device) and errors are raised otherwise.
"""
from __future__ import annotations

import math
import warnings
from typing import Dict, Optional

import numpy as np
import torch
import torch.nn as nn

from . import _lib, stages, synthetic
from .splat import SplatWorkspace, splat_forward, splat_render

SUBDEVIDE_VERTS = {0: 10000, 1: 30000, 2: 30000}      # renderer.py:14-18 (sic)
SCALE_BIAS = 3.9                                        # renderer.py:428
OPACITY_BIAS = 0.0                                      # renderer.py:429
ZNEAR, ZFAR = 0.01, 100.0                               # renderer.py:492-493


def inverse_sigmoid(x: float) -> float:
    """src/utils/math_utils.py:7-11."""
    return math.log(x / (1.0 - x))


# ------------------------------------------------------------------------------------------------
# camera set-up of render_one (renderer.py:486-510; graphic_utils.py:67-78,103-136,144-145), batched
# ------------------------------------------------------------------------------------------------
def camera_matrices(K: torch.Tensor, E: torch.Tensor, height: int, width: int):
    """K [F,3,3], E [F,4,4] on the device -> (viewmatrix [F,4,4], projmatrix [F,4,4], tanfov [F,2]) exactly as
    render_one hands them to GaussianRasterizationSettings (already transposed), without a host sync.

    One launch of ``amav_camera_setup`` for all frames (csrc/camera.cu): Rt from the extrinsic (the reference's two
    fp32 inversions of it cancel up to ~1e-7 of rounding noise), K_ndc and tan of the half-angle in double as the
    reference's Python floats / math.tan, view x proj in fp32."""
    K = K.reshape(-1, 3, 3).to(torch.float32).contiguous()
    E = E.reshape(-1, 4, 4).to(torch.float32).contiguous()
    Fr = int(K.shape[0])
    if E.shape[0] != Fr:
        raise ValueError("K and E differ in frame count")
    dev = K.device
    view = torch.empty(Fr, 4, 4, device=dev)
    full = torch.empty(Fr, 4, 4, device=dev)
    tanfov = torch.empty(Fr, 2, device=dev)
    _lib.call("amav_camera_setup", Fr, _lib.ptr(K), _lib.ptr(E), int(width), int(height), ZNEAR, ZFAR,
              _lib.ptr(view), _lib.ptr(full), _lib.ptr(tanfov), _lib.stream_ptr())
    return view, full, tanfov


_BG_CACHE: Dict[tuple, torch.Tensor] = {}


def _background(bg_color, device) -> torch.Tensor:
    if bg_color is None:
        bg_color = [1, 1, 1]
    if isinstance(bg_color, torch.Tensor):
        return bg_color.to(device, torch.float32).reshape(3)
    key = (str(device), tuple(float(c) for c in bg_color))
    t = _BG_CACHE.get(key)                      # a Python-list background is uploaded once per device, not per call
    if t is None:
        t = _BG_CACHE[key] = torch.tensor(list(key[1]), dtype=torch.float32, device=device)
    return t


_WORKSPACES: Dict[tuple, SplatWorkspace] = {}


def _workspace(device) -> SplatWorkspace:
    """The inference workspace (bucket buffers, tile counters, overflow flag) of the CURRENT STREAM of ``device``: calls
    enqueued on one stream are ordered and can share it; two streams (or a CUDA-graph capture, which runs on its own
    stream) render concurrently and get their own buffers.  The C ABI itself is re-entrant."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    key = (idx, torch.cuda.current_stream(idx).cuda_stream)
    ws = _WORKSPACES.get(key)
    if ws is None:
        ws = _WORKSPACES[key] = SplatWorkspace()
    return ws


_DEFERRED: Optional[list] = None


class deferred_overflow_checks:
    """``with deferred_overflow_checks() as chk:`` -- inference calls inside the block do not read their bucket
    overflow flag back (one host sync per call, which also keeps the CPU from enqueueing the next call's kernels
    while this one renders); ``chk.overflowed()`` reads all of them at once afterwards.  If it returns True the
    frames of the block are incomplete: render them again outside the block (the sizing read then grows the buffers).
    Used where one logical render is cut into pieces (``shard.render_and_gather``)."""

    def __enter__(self):
        global _DEFERRED
        self._prev, _DEFERRED = _DEFERRED, []
        self._flags = _DEFERRED
        return self

    def __exit__(self, *exc):
        global _DEFERRED
        _DEFERRED = self._prev

    def overflowed(self) -> bool:
        if not self._flags:
            return False
        return bool(torch.stack([f.reshape(()) for f in self._flags]).any().item())


def _render_frames(xyz, rot, scale, opacity, color, K, E, args, bg_color, debug, views_per_set, hwc,
                   out=None, defer_check=False, cameras=None):
    """All frames of a call in one batched splat pass.  Gaussian tensors [G,N,k] (pre-activation scale and
    opacity as in the reference's ``gaussians`` dict), cameras [G*views_per_set,...].

    Host syncs: the first call on a device reads the entry count D back once to size the bucket buffer; later
    calls trust the cached capacity (grown by 25%) and only read the overflow flag AFTER everything is enqueued
    (``defer_check`` leaves even that to the caller, see ``Renderer.render_sequence``).  An overflow re-renders
    with the sizing read."""
    if not getattr(args, "rgb", True):
        # renderer.py:541-545 feeds 3-channel colours to eval_sh(sh_degree=3), which asserts on the coefficient
        # count (graphic_utils.py:722): the branch is dead in the reference too.
        raise ValueError("args.rgb=False (SH colours) is not supported: the reference's own SH branch asserts "
                         "for 3-channel colours; use rgb=True")
    dev = xyz.device
    if dev.type != "cuda":
        raise _lib.AmavError("rendering needs CUDA tensors (no CPU path)")
    height, width = int(args.image_size[0]), int(args.image_size[1])
    view, full, tanfov = cameras if cameras is not None else camera_matrices(K.to(dev), E.to(dev), height, width)
    bg = _background(bg_color, dev)
    activate = True
    if debug:                                        # renderer.py:535-537
        scale = torch.full_like(scale, 0.01)
        opacity = torch.full_like(opacity, 0.1)
        color = color.clamp(0.0, 1.0)
        activate = False
    if torch.is_grad_enabled() and any(t.requires_grad for t in (xyz, rot, scale, opacity, color)):
        # training: autograd node with its own buffers (the forward's lists and per-pixel state feed the backward)
        # the [0,1] clamp (renderer.py:568) stays a torch op here so that autograd masks dL/dimage exactly as the
        # reference's does wherever the blended colour leaves [0,1] (a caller-supplied bg_color, saturated pixels)
        return splat_render(xyz, scale, rot, opacity, color, view, full, tanfov, bg, height, width,
                            views_per_set=views_per_set, activate=activate, hwc=hwc, clamp01=False).clamp(0.0, 1.0)
    ws = _workspace(dev)

    def run(sync_free):
        return splat_forward(xyz, scale, rot, opacity, color, view, full, tanfov, bg, height, width,
                             views_per_set=views_per_set, activate=activate, hwc=hwc, clamp01=True,
                             workspace=ws, sync_free=sync_free, out=out)

    res = run(sync_free=ws.covers((int(xyz.shape[-2]), height, width, views_per_set), int(view.shape[0])))
    if res["D"] < 0 and not defer_check and _DEFERRED is not None:
        _DEFERRED.append(ws.overflow.clone())           # the flag buffer is reused (zeroed) by the next call
    elif res["D"] < 0 and not defer_check:
        if int(ws.overflow.item()) != 0:
            res = run(sync_free=False)
    return res["color"]


def render_batch(gaussians, K, E, args, bg_color=None, debug=False, out=None):
    """renderer.py:447-479.  gaussians: dict of [B*T,N,k]; K [B,T,3,3]; E [B,T,4,4] -> [B,T,H,W,3] in [0,1].
    ``out`` (inference only): a contiguous [B*T,H,W,3] fp32 device buffer the frames are written into -- it may be
    another GPU's memory mapped into this process (``shard.PeerFrameGather``)."""
    B, T = E.shape[0], E.shape[1]
    g = lambda k, d: gaussians[k].reshape(B * T, -1, d).float()
    img = _render_frames(g("xyz", 3), g("rot", 4), g("scale", 3), g("opacity", 1), g("color", 3),
                         K.reshape(-1, 3, 3), E.reshape(-1, 4, 4), args, bg_color, debug, 1, True, out=out)
    return img.reshape(B, T, int(args.image_size[0]), int(args.image_size[1]), 3)


def render_multi_view(gaussians, K, E, args, bg_color=None, debug=False):
    """renderer.py:431-445.  gaussians [B,N,k] rendered from T views each, without the reference's T-fold
    ``expand().reshape()`` copies: frame f reads Gaussian set f // T."""
    B, T = E.shape[0], E.shape[1]
    g = lambda k, d: gaussians[k].reshape(B, -1, d).float()
    img = _render_frames(g("xyz", 3), g("rot", 4), g("scale", 3), g("opacity", 1), g("color", 3),
                         K.reshape(-1, 3, 3), E.reshape(-1, 4, 4), args, bg_color, debug, T, True)
    return img.reshape(B, T, int(args.image_size[0]), int(args.image_size[1]), 3)


def render_one(xyzs, rots, scales, opacities, colors, K, E, args, bg_color=None, debug=False):
    """renderer.py:481-570: one frame, returns [3,H,W] clamped to [0,1]."""
    img = _render_frames(xyzs[None].float(), rots[None].float(), scales[None].float(), opacities[None].float(),
                         colors[None].float(), K[None], E[None], args, bg_color, debug, 1, False)
    return img[0]


# ------------------------------------------------------------------------------------------------
# body model loading (what smplx.SMPLX(model_path, ...) does at renderer.py:206-225)
# ------------------------------------------------------------------------------------------------
def load_smplx_model(path: str, num_betas: int = 10, num_expression_coeffs: int = 10,
                     flat_hand_mean: bool = False) -> Dict[str, torch.Tensor]:
    """Read ``SMPLX_NEUTRAL.npz`` (the file smplx==0.1.28 loads; ``path`` may be the model folder, a
    ``smplx`` sub-folder or the file) into the buffer dict used by ``stages.BodyModel``:
    shapedirs = [shape basis[:num_betas], expression basis[:num_expression_coeffs]] (smplx keeps the
    expression basis at offset 300), posedirs reshaped to [(J-1)*9, 3V], parents from kintree_table[0]
    with parents[0] = -1, pose_mean = zeros except the hand means unless ``flat_hand_mean``."""
    import os
    cands = [path, os.path.join(path, "SMPLX_NEUTRAL.npz"), os.path.join(path, "smplx", "SMPLX_NEUTRAL.npz")]
    file = next((c for c in cands if os.path.isfile(c)), None)
    if file is None:
        raise FileNotFoundError(f"no SMPL-X model file under {path!r} (looked for SMPLX_NEUTRAL.npz)")
    d = np.load(file, allow_pickle=True)
    V = d["v_template"].shape[0]
    sd = np.asarray(d["shapedirs"], dtype=np.float32)
    shape_basis = sd[:, :, :num_betas]
    expr_basis = sd[:, :, 300:300 + num_expression_coeffs]
    posedirs = np.asarray(d["posedirs"], dtype=np.float32).reshape(V * 3, -1).T
    parents = np.asarray(d["kintree_table"][0], dtype=np.int64).copy()
    parents[0] = -1
    J = parents.shape[0]
    pose_mean = np.zeros(J * 3, dtype=np.float32)
    if not flat_hand_mean:
        pose_mean[25 * 3:40 * 3] = np.asarray(d["hands_meanl"], dtype=np.float32).reshape(-1)
        pose_mean[40 * 3:55 * 3] = np.asarray(d["hands_meanr"], dtype=np.float32).reshape(-1)
    return {
        "v_template": torch.from_numpy(np.asarray(d["v_template"], dtype=np.float32)),
        "faces": torch.from_numpy(np.asarray(d["f"], dtype=np.int64)),
        "shapedirs": torch.from_numpy(np.ascontiguousarray(np.concatenate([shape_basis, expr_basis], axis=2))),
        "posedirs": torch.from_numpy(np.ascontiguousarray(posedirs)),
        "J_regressor": torch.from_numpy(np.asarray(d["J_regressor"], dtype=np.float32)),
        "parents": torch.from_numpy(parents),
        "lbs_weights": torch.from_numpy(np.asarray(d["weights"], dtype=np.float32)),
        "pose_mean": torch.from_numpy(pose_mean),
    }


class _Decoder(nn.Module):
    pass


class GraphedForward:
    """``Renderer.forward`` (inference) captured in a CUDA graph for fixed shapes.

    The tensors given to the constructor ARE the graph's static inputs (``copy_inputs=False``: zero-copy, the producer
    refills them in place before ``replay()``) or are cloned (``copy_inputs=True``); ``__call__`` copies any tensor that
    is not the static one into it first.  ``images`` / ``gaussians`` are the static outputs (overwritten by the next
    replay).  Bucket capacity is fixed from a sizing run; the overflow flag is read by ``overflowed()`` (one host sync)
    and ``recapture()`` re-sizes, exactly as the eager path's deferred checks."""

    def __init__(self, renderer, tokens, cams, params, out=None, copy_inputs: bool = False, warmup: int = 2):
        dev = renderer._device
        own = (lambda t: t.to(dev).clone()) if copy_inputs else (lambda t: t if t.device == dev else t.to(dev))
        self.renderer = renderer
        self.tokens = own(tokens)
        self.cams = {k: own(v) for k, v in cams.items()}
        self.params = {k: own(v) for k, v in params.items()}
        self.out = out
        self.warmup = warmup
        self.stream = torch.cuda.Stream(dev)
        self.graph = None
        self._capture()

    def _capture(self):
        r, dev = self.renderer, self.renderer._device
        if _lib._TIMER is not None:
            raise _lib.AmavError("StageTimer records timing events, which a stream capture does not allow")
        self.stream.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(self.stream), torch.no_grad():
            for _ in range(max(1, self.warmup)):         # sizes this stream's workspace (the one read of D) and warms up
                r(self.tokens, self.cams, None, self.params, out=self.out)
            self.stream.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            n0 = _lib.launch_count()
            with deferred_overflow_checks() as chk:
                with torch.cuda.graph(self.graph, stream=self.stream):
                    res = r(self.tokens, self.cams, None, self.params, out=self.out)
            self.launches = _lib.launch_count() - n0     # library kernels inside one replay
            self._chk = chk
        self.images, self.gaussians = res[0], res[1]
        torch.cuda.current_stream(dev).wait_stream(self.stream)

    def replay(self):
        self.graph.replay()
        return self.images, self.gaussians

    def overflowed(self) -> bool:
        return self._chk.overflowed()

    def recapture(self):
        """After an overflow: drop the cached capacity of the capture stream's workspace and capture again."""
        dev = self.renderer._device
        with torch.cuda.stream(self.stream):
            _workspace(dev).capacity = 0
        self._capture()

    def __call__(self, tokens=None, cams=None, params=None, check: bool = True):
        def put(dst, src):
            if src is not None and src.data_ptr() != dst.data_ptr():
                dst.copy_(src, non_blocking=True)
        put(self.tokens, tokens)
        for k, v in (cams or {}).items():
            put(self.cams[k], v)
        for k, v in (params or {}).items():
            put(self.params[k], v)
        self.replay()
        if check and self.overflowed():
            self.recapture()
            self.replay()
        return self.images, self.gaussians


class Renderer(nn.Module):
    """renderer.py:20-346.  ``Renderer(cfg, smpl_decoder=None, body_model=None)``; ``body_model`` (a buffer dict as
    from ``synthetic.make_body_model`` / ``load_smplx_model``) replaces ``cfg.smplx_model_path`` when no SMPL-X
    file is available."""

    def __init__(self, cfg=None, smpl_decoder=None, body_model: Optional[Dict[str, torch.Tensor]] = None):
        super().__init__()
        self.cfg = cfg
        dev = torch.device(getattr(cfg, "device", "cuda"))
        if dev.type != "cuda":
            raise _lib.AmavError("Renderer needs cfg.device to be a CUDA device (no CPU path)")
        _lib.lib()                                   # fail loudly now if the extension is missing
        self._device = dev
        if body_model is None:
            body_model = load_smplx_model(cfg.smplx_model_path, 10, cfg.num_expression_coeffs,
                                          bool(getattr(cfg, "flat_hand_mean", False)))
        self.smplx_model = stages.BodyModel(body_model, dev)
        steps = int(getattr(cfg, "subdivide_steps", 2))
        self.num_verts = int(getattr(cfg, "num_gaussians", 0) or SUBDEVIDE_VERTS[steps])
        self._dense_par = self._dense_wgt = None
        if getattr(cfg, "densify_smplx_verts", False):
            self.init_smplx_subdivider(subdivide_steps=steps)
        if getattr(cfg, "predict_smplx_params", False):
            self.smpl_decoder = smpl_decoder
        else:
            self.smpl_decoder = None                 # the reference leaves it unset and crashes at :104
        if getattr(cfg, "upsample_triplane", False):
            raise NotImplementedError("upsample_triplane is outside the render path built here (SURVEY.md a-13); "
                                      "hand the renderer a ready triplane and set upsample_triplane: false")
        if not getattr(cfg, "no_point_refiner", True):
            warnings.warn("point_encoder/point_refiner (PTv3) are bypassed: zero-initialised last layer makes them an "
                          "exact no-op at random init; trained refiner weights are ignored", stacklevel=2)
        C = int(cfg.triplane_feature_dim)
        dec = _Decoder()
        dec.xyz_layer = nn.Linear(C * 3 + 3, 3)
        dec.rotation_layer = nn.Linear(C * 3 + 3, 4)
        dec.scaling_layer = nn.Linear(C * 3 + 3, 3)
        dec.opacity_layer = nn.Linear(C * 3 + 3, 1)
        dec.shs_layer = nn.Linear(C * 3 + 3, 3)
        for layer in (dec.xyz_layer, dec.rotation_layer, dec.scaling_layer, dec.opacity_layer, dec.shs_layer):
            nn.init.constant_(layer.weight, 0)
            nn.init.constant_(layer.bias, 0)
        with torch.no_grad():                        # renderer.py:57-71
            dec.rotation_layer.bias[0] = 1.0
            dec.scaling_layer.bias.fill_(-1.0)
            dec.opacity_layer.bias.fill_(inverse_sigmoid(0.1))
        self.gaussian_decoder = dec.to(dev)
        self._fixed_idx = None
        # processing order of the per-site kernels (stages.locality_order); known whenever the site set is fixed.  Opt-in
        # (cfg.site_locality_order): measured on B200 at 100k sites x 120 frames it LOSES -- sample_decode 0.92 -> 1.10 ms,
        # densify 0.20 -> 0.22 ms -- the scattered 12..16-byte result stores and point loads cost more than the shared
        # texels save
        self._site_order = None
        if not getattr(cfg, "densify_smplx_verts", False):
            self._site_order = stages.locality_order(self.smplx_model.v_template, float(cfg.radius))

    # -- densify table (renderer.py:227-243) ------------------------------------------------------
    def init_smplx_subdivider(self, subdivide_steps=2):
        """The reference keeps ``max(1, subdivide_steps)`` SubdivideMeshes objects; here the levels are
        flattened once into a (parents, weights) table over the base vertices."""
        levels = max(1, int(subdivide_steps))
        faces = self.smplx_model.faces
        faces = faces.cpu().numpy() if isinstance(faces, torch.Tensor) else np.asarray(faces)
        par, wgt = synthetic.densify_table(self.smplx_model.V, faces, levels)
        self._dense_par = torch.from_numpy(par).to(self._device)
        self._dense_wgt = torch.from_numpy(wgt).to(self._device)

    def fix_site_subset(self, seed: Optional[int] = None) -> torch.Tensor:
        """Draw the densified-vertex subset once (the reference redraws ``torch.randperm`` on the global CPU
        generator every forward, renderer.py:287); used for sequences that must keep the same sites."""
        n = int(self._dense_par.shape[0])
        g = None if seed is None else torch.Generator().manual_seed(seed)
        self._fixed_idx = torch.randperm(n, generator=g)[:self.num_verts].to(self._device)
        # the subset's rows of the densify table, gathered once instead of on every forward
        self._fixed_par = self._dense_par.index_select(0, self._fixed_idx).contiguous()
        self._fixed_wgt = self._dense_wgt.index_select(0, self._fixed_idx).contiguous()
        rest = stages.densify_gather(self.smplx_model.v_template[None].contiguous(), self._fixed_par, self._fixed_wgt)[0]
        self._site_order = stages.locality_order(rest, float(self.cfg.radius))
        return self._fixed_idx

    def get_smpl_vertices(self, smpl_params):
        """renderer.py:245-290: posed vertices (no transl), optionally densified + subset -> [B*T, N, 3].

        A dict that carries ``pose_6d`` [..., J, 6] (stages.pose_6d_from_decoder_heads: the SMPL-X parameter decoder's
        6D heads, smplx_decoder.py:106-127) instead of the axis-angle keys goes through amav_lbs_forward_6d, without the
        reference's matrix -> axis-angle -> Rodrigues round trip (gradients by amav_lbs_backward_6d)."""
        if "pose_6d" in smpl_params:
            shape = torch.cat([smpl_params["betas"].reshape(-1, smpl_params["betas"].shape[-1]),
                               smpl_params["expression"].reshape(-1, smpl_params["expression"].shape[-1])], dim=1)
            verts = stages.lbs_vertices_6d(self.smplx_model, smpl_params["pose_6d"].to(self._device),
                                          shape.to(self._device, torch.float32))
        else:
            full, shape = self.smplx_model.full_pose(smpl_params)
            verts = stages.lbs_vertices(self.smplx_model, full, shape)
        if getattr(self.cfg, "densify_smplx_verts", False):
            order = None
            if self._fixed_idx is not None:
                par, wgt = self._fixed_par, self._fixed_wgt
                order = self._site_order if getattr(self.cfg, "site_locality_order", False) else None
            else:
                if torch.cuda.is_current_stream_capturing():
                    raise _lib.AmavError("a CUDA-graph capture cannot redraw torch.randperm on the CPU generator "
                                         "(renderer.py:287): call fix_site_subset() first")
                idx = torch.randperm(int(self._dense_par.shape[0]))[:self.num_verts].to(self._device)
                par = self._dense_par.index_select(0, idx).contiguous()
                wgt = self._dense_wgt.index_select(0, idx).contiguous()
            verts = stages.densify_sites(verts, par, wgt, order)
        return verts

    def _decoder_wb(self):
        return stages.decoder_matrix(self.gaussian_decoder)

    def sample_from_triplane(self, triplane_features, points):
        """renderer.py:292-317 ([F,3,C,Hp,Wp] or unbatched, points [F,N,3]) -> [F,N,3C] sampled features.
        Kept for API compatibility (the forward pass never materialises this tensor): it is computed with the
        projection kernel against an identity 'decoder', 6 channels at a time (slow; not the hot path)."""
        batched = points.ndim == 3
        if not batched:
            triplane_features, points = triplane_features[None], points[None]
        Fr, Np, C, Hp, Wp = triplane_features.shape
        tokens = triplane_features.permute(0, 2, 1, 3, 4).reshape(Fr, C, Np * Hp * Wp).contiguous().float().to(self._device)
        points = points.float().to(self._device)
        N = points.shape[1]
        out = torch.empty(Fr, N, Np * C, device=self._device)
        zeros3 = torch.zeros(Fr, 3, device=self._device)
        bias = torch.zeros(14, device=self._device)
        rows = (0, 1, 2, 7, 8, 9)                    # heads that come back un-activated: xyz offset, scaling
        for p in range(Np):
            for c0 in range(0, C, 6):
                n = min(6, C - c0)
                W = torch.zeros(14, 3 + 3 * C, device=self._device)
                for k in range(n):
                    W[rows[k], 3 + p * C + c0 + k] = 1.0
                proj = stages.triplane_project(tokens, W, C, Hp)
                g = stages.sample_decode(proj, points, zeros3, W, bias, C, Hp, float(self.cfg.radius))
                vals = torch.cat([g["xyz"] - points, g["scale"]], dim=-1)
                out[:, :, p * C + c0:p * C + c0 + n] = vals[:, :, :n]
        return out if batched else out.squeeze(0)

    def construct_gaussians(self, gaussian_params, points, smpl_params):
        """renderer.py:319-346 on already-decoded heads (API compatibility; forward() fuses this)."""
        rot = torch.nn.functional.normalize(gaussian_params["rotation"], dim=-1)
        color = torch.sigmoid(gaussian_params["shs"])
        return {"xyz": points + gaussian_params["xyz_offset"] + smpl_params["transl"].reshape(-1, 1, 3),
                "scale": gaussian_params["scaling"], "rot": rot, "opacity": gaussian_params["opacity"],
                "color": color, "shs": color}

    def decode_gaussians(self, triplane_features, points, transl):
        """Stages 1+2 for [F,C,3*R*R] tokens and [F,N,3] posed sites -> the ``gaussians`` dict.
        ``cfg.triplane_token_major`` (extension): tokens are [F,3*R*R,C], the layout the upstream
        transformer computes in before its permute+contiguous (transformers.py:1070-1073)."""
        C, R = int(self.cfg.triplane_feature_dim), int(self.cfg.triplane_resolution)
        W, b = self._decoder_wb()
        tm = bool(getattr(self.cfg, "triplane_token_major", False))
        if torch.is_grad_enabled() and any(t.requires_grad for t in (triplane_features, points, transl, W, b)):
            return stages.decode_gaussians_autograd(triplane_features, points, transl, W, b, C, R, float(self.cfg.radius),
                                                    token_major=tm, order=self._order_for(points))
        proj = stages.triplane_project(triplane_features, W, C, R, token_major=tm)
        return stages.sample_decode(proj, points, transl, W, b, C, R, float(self.cfg.radius), order=self._order_for(points))

    def _order_for(self, points):
        """The locality order applies to the renderer's own site set (densified + fixed subset, or the base vertices)."""
        o = self._site_order if getattr(self.cfg, "site_locality_order", False) else None
        if o is None or (getattr(self.cfg, "densify_smplx_verts", False) and self._fixed_idx is None):
            return None
        return o if int(points.shape[-2]) == int(o.shape[0]) else None

    def forward(self, triplane_features, cam_params, smpl_tokens=None, smpl_params_gt=None, out=None):
        """renderer.py:73-204.  triplane_features [B,T,C,3*R*R]; cam_params {'intrinsic' [B,T,3,3], 'extrinsic'
        [B,T,4,4]}; returns (rendered_images [B,T,H,W,3], gaussians[, pred_smpl_params]).  ``out`` (extension,
        inference only): [B*T,H,W,3] buffer to render into, see ``render_batch``."""
        dev = self._device
        if smpl_tokens is not None:
            B, T = smpl_tokens.shape[0], smpl_tokens.shape[1]
        elif smpl_params_gt is not None:
            B, T = triplane_features.shape[0], triplane_features.shape[1]
        else:
            raise ValueError("Renderer.forward needs smpl_tokens (with a smpl_decoder) or smpl_params_gt")
        pred_smpl_params = None
        if self.smpl_decoder is not None and smpl_tokens is not None:
            pred_smpl_params = self.smpl_decoder(smpl_tokens.reshape(B * T, *smpl_tokens.shape[2:]))
            for key in list(pred_smpl_params.keys()):           # renderer.py:111-118
                v = pred_smpl_params[key]
                if key in ("body_pose", "left_hand_pose", "right_hand_pose"):
                    pred_smpl_params[key] = v.reshape(B, T, *v.shape[1:])
                elif key in ("betas", "transl", "global_orient", "expression", "jaw_pose", "leye_pose", "reye_pose", "focal"):
                    pred_smpl_params[key] = v.reshape(B, T, -1)
        smpl_params = smpl_params_gt if smpl_params_gt is not None else pred_smpl_params
        if smpl_params is None:
            raise ValueError("no SMPL-X parameters: pass smpl_params_gt or build the Renderer with a smpl_decoder "
                             "and cfg.predict_smplx_params=True")
        C, R = int(self.cfg.triplane_feature_dim), int(self.cfg.triplane_resolution)
        if getattr(self.cfg, "triplane_token_major", False):
            tokens = triplane_features.reshape(B * T, 3 * R * R, C)
        else:
            tokens = triplane_features.reshape(B * T, C, 3 * R * R)
        if tokens.device != dev or tokens.dtype != torch.float32:
            tokens = tokens.to(dev, torch.float32)
        points = self.get_smpl_vertices(smpl_params)                                   # [F,N,3]
        transl = smpl_params["transl"].reshape(B * T, 3).to(dev, torch.float32)
        gaussians = self.decode_gaussians(tokens, points, transl)
        rendered = render_batch(gaussians, cam_params["intrinsic"], cam_params["extrinsic"], self.cfg, debug=False, out=out)
        if getattr(self.cfg, "predict_smplx_params", False):
            return rendered, gaussians, pred_smpl_params
        return rendered, gaussians

    def capture(self, triplane_features, cam_params, smpl_params, out=None, copy_inputs: bool = False):
        """CUDA-graph the inference forward for these shapes: returns a ``GraphedForward`` whose ``replay()`` re-runs the
        whole path (LBS -> densify -> project -> sample/decode -> camera -> preprocess -> scan -> bin -> blend) as ONE
        graph launch -- no per-kernel launch cost, no host work, no host sync.  The reference hands the renderer T = 6
        frames per call (configs/model/triplane_audio_net.yaml:4): at that size the eager call is launch-bound."""
        return GraphedForward(self, triplane_features, cam_params, smpl_params, out=out, copy_inputs=copy_inputs)

    # -- host-resident sequences ------------------------------------------------------------------
    def render_sequence(self, triplane_features, cam_params, smpl_params, chunk: int = 8, out=None,
                        out_dtype: torch.dtype = torch.float32, bgr: bool = False):
        """Render a [B,T] frame sequence whose triplane tokens live in (pinned) HOST memory and return the
        frames in pinned host memory: the stage-2 demo loop of the reference (main2.py:179-239) hands the
        renderer one clip after another and copies every frame back for the video sink.

        Frames are independent, so the sequence is cut into ``chunk``-frame pieces that flow through three
        streams: host->device copy of chunk k+1, render of chunk k, device->host copy of chunk k-1.
        LBS for the whole sequence runs once up front (its inputs are a few KB per frame).

        ``out_dtype=torch.uint8`` applies the demo's video sink on the device before the copy back
        (``(frame * 255).astype(np.uint8)`` and, with ``bgr``, cv2's RGB2BGR; main2.py:351-353): a quarter of
        the bytes leave the GPU (stages.frames_to_u8)."""
        if out_dtype not in (torch.float32, torch.uint8):
            raise ValueError("out_dtype must be torch.float32 or torch.uint8")
        as_u8 = out_dtype == torch.uint8
        dev = self._device
        B, T = triplane_features.shape[0], triplane_features.shape[1]
        Fr = B * T
        C, R = int(self.cfg.triplane_feature_dim), int(self.cfg.triplane_resolution)
        H, W = int(self.cfg.image_size[0]), int(self.cfg.image_size[1])
        tok = triplane_features.reshape(Fr, C, 3 * R * R)
        if tok.dtype != torch.float32:
            raise ValueError("triplane_features must be fp32")
        if out is None:
            out = torch.empty(B, T, H, W, 3, dtype=out_dtype, pin_memory=True)
        if out.dtype != out_dtype:
            raise ValueError(f"out must be {out_dtype}")
        out_f = out.reshape(Fr, H, W, 3)
        main = torch.cuda.current_stream(dev)
        params = {k: v.to(dev, non_blocking=True) for k, v in smpl_params.items()}
        K = cam_params["intrinsic"].reshape(Fr, 3, 3).to(dev, non_blocking=True)
        E = cam_params["extrinsic"].reshape(Fr, 4, 4).to(dev, non_blocking=True)
        points = self.get_smpl_vertices(params)
        transl = params["transl"].reshape(Fr, 3).float()
        view, full, tanfov = camera_matrices(K, E, H, W)           # whole sequence at once, no per-chunk glue
        bg = _background(None, dev)
        Wd, bd = self._decoder_wb()
        Wd, bd = Wd.detach().contiguous(), bd.detach().contiguous()
        st = self._seq_state(chunk, C, R, H, W)
        h2d, d2h = st["h2d"], st["d2h"]
        h2d.wait_stream(main)
        nchunks = (Fr + chunk - 1) // chunk
        pending = []
        for k in range(nchunks):
            f0, f1 = k * chunk, min(Fr, (k + 1) * chunk)
            n = f1 - f0
            slot = k % st["depth"]
            with torch.cuda.stream(h2d):
                if st["tok_free"][slot] is not None:
                    h2d.wait_event(st["tok_free"][slot])           # render of chunk k-depth has consumed the buffer
                st["tok"][slot][:n].copy_(tok[f0:f1], non_blocking=True)
                ready = torch.cuda.Event()
                ready.record(h2d)
            main.wait_event(ready)
            if st["img_free"][slot] is not None:
                main.wait_event(st["img_free"][slot])              # D2H of chunk k-depth has drained the image buffer
            proj = stages.triplane_project(st["tok"][slot][:n], Wd, C, R, out=st["proj"][:n])
            consumed = torch.cuda.Event()
            consumed.record(main)
            st["tok_free"][slot] = consumed
            g = stages.sample_decode(proj, points[f0:f1], transl[f0:f1], Wd, bd, C, R, float(self.cfg.radius),
                                     order=self._order_for(points))
            img = st["img"][slot][:n]
            _render_frames(g["xyz"], g["rot"], g["scale"], g["opacity"], g["color"], None, None, self.cfg,
                           bg, False, 1, True, out=img, defer_check=True,
                           cameras=(view[f0:f1], full[f0:f1], tanfov[f0:f1]))
            ws = _workspace(dev)
            pending.append((k, ws.overflow.clone() if ws.overflow is not None else None))
            if as_u8:
                if st.get("img8") is None:
                    st["img8"] = [torch.empty(chunk, H, W, 3, dtype=torch.uint8, device=dev) for _ in range(st["depth"])]
                img = stages.frames_to_u8(img, bgr=bgr, out=st["img8"][slot][:n])
            rendered = torch.cuda.Event()
            rendered.record(main)
            with torch.cuda.stream(d2h):
                d2h.wait_event(rendered)
                out_f[f0:f1].copy_(img, non_blocking=True)
                drained = torch.cuda.Event()
                drained.record(d2h)
            st["img_free"][slot] = drained
        main.wait_stream(d2h)
        main.synchronize()
        bad = [k for k, flag in pending if flag is not None and int(flag.item()) != 0]
        for k in bad:                                              # bucket overflow: redo the chunk with the sizing read
            f0, f1 = k * chunk, min(Fr, (k + 1) * chunk)
            tk = tok[f0:f1].to(dev)
            g = self.decode_gaussians(tk, points[f0:f1], transl[f0:f1])
            _workspace(dev).capacity = 0
            img = _render_frames(g["xyz"], g["rot"], g["scale"], g["opacity"], g["color"], K[f0:f1], E[f0:f1], self.cfg,
                                 None, False, 1, True)
            out_f[f0:f1].copy_(stages.frames_to_u8(img, bgr=bgr) if as_u8 else img)
        return out

    def _seq_state(self, chunk, C, R, H, W, depth: int = 3):
        key = (chunk, C, R, H, W, depth)
        st = getattr(self, "_seq", None)
        if st is None or st["key"] != key:
            dev = self._device
            st = {"key": key, "depth": depth, "h2d": torch.cuda.Stream(dev), "d2h": torch.cuda.Stream(dev),
                  "tok": [torch.empty(chunk, C, 3 * R * R, device=dev) for _ in range(depth)],
                  "img": [torch.empty(chunk, H, W, 3, device=dev) for _ in range(depth)],
                  "proj": torch.empty(chunk, 3, R * R, 16, device=dev),
                  "img8": None, "tok_free": [None] * depth, "img_free": [None] * depth}
            self._seq = st
        else:
            st["tok_free"] = [None] * st["depth"]
            st["img_free"] = [None] * st["depth"]
        return st
