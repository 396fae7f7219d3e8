"""Host wrappers of stages 1-3 over the C ABI (include/amav.h): SMPL-X forward + densify, triplane
projection, sample/decode.  Every function takes CUDA fp32 tensors and enqueues on the current stream;
there is no CPU path."""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr

POSE_KEYS = ("global_orient", "body_pose", "jaw_pose", "leye_pose", "reye_pose", "left_hand_pose", "right_hand_pose")


def _f32c(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to(torch.float32).contiguous()


class BodyModel:
    """Device-resident constants of the SMPL-X(-shaped) body model + the once-per-model precompute
    (amav_lbs_precompute).  Mirrors the buffers of ``smplx.SMPLX`` that the reference builds at
    renderer.py:206-225: v_template, shapedirs (betas+expression), posedirs, J_regressor, parents,
    lbs_weights, pose_mean, faces."""

    def __init__(self, model: Dict[str, torch.Tensor], device):
        dev = torch.device(device)
        if dev.type != "cuda":
            raise _lib.AmavError("BodyModel needs a CUDA device (no CPU path)")
        self.device = dev
        self.v_template = _f32c(model["v_template"]).to(dev)
        self.shapedirs = _f32c(model["shapedirs"]).to(dev)
        self.posedirs = _f32c(model["posedirs"]).to(dev)
        self.J_regressor = _f32c(model["J_regressor"]).to(dev)
        self.lbs_weights = _f32c(model["lbs_weights"]).to(dev)
        self.parents = model["parents"].to(torch.int32).contiguous().to(dev)
        self.pose_mean = _f32c(model["pose_mean"]).to(dev)
        self.faces = model.get("faces")
        self.V = int(self.v_template.shape[0])
        self.J = int(self.parents.shape[0])
        self.S = int(self.shapedirs.shape[2])
        if self.posedirs.shape != ((self.J - 1) * 9, 3 * self.V):
            raise ValueError(f"posedirs must be [(J-1)*9, 3V], got {tuple(self.posedirs.shape)}")
        self.J_template = torch.empty(self.J, 3, device=dev)
        self.J_dirs = torch.empty(self.J, 3, self.S, device=dev)
        self.weights_T = torch.empty(self.J, self.V, device=dev)
        call("amav_lbs_precompute", self.V, self.J, self.S, ptr(self.v_template), ptr(self.shapedirs),
             ptr(self.J_regressor), ptr(self.lbs_weights), ptr(self.J_template), ptr(self.J_dirs),
             ptr(self.weights_T), stream_ptr())

    def full_pose(self, params: Dict[str, torch.Tensor]) -> Tuple[torch.Tensor, torch.Tensor]:
        """[F,J*3] axis-angles in smplx.SMPLX.forward's joint order (+ pose_mean) and [F,S] shape
        coefficients [betas, expression] from the reference's parameter dict (renderer.py:250-258)."""
        Fr = params["global_orient"].numel() // 3
        parts = [params[k].reshape(Fr, -1).to(self.device, torch.float32) for k in POSE_KEYS]
        full = torch.cat(parts, dim=1) + self.pose_mean[None]
        shape = torch.cat([params["betas"].reshape(Fr, -1), params["expression"].reshape(Fr, -1)], dim=1)
        return full.contiguous(), shape.to(self.device, torch.float32).contiguous()


def lbs_forward(body: BodyModel, full_pose: torch.Tensor, shape_comp: torch.Tensor,
                return_aux: bool = False):
    """Posed vertices [F,V,3] without transl (renderer.py:261-274 -> smplx lbs)."""
    Fr = int(full_pose.shape[0])
    dev = body.device
    if full_pose.shape[1] != body.J * 3 or shape_comp.shape[1] != body.S:
        raise ValueError("full_pose must be [F,J*3] and shape_comp [F,S]")
    pf = torch.empty(Fr, (body.J - 1) * 9, device=dev)
    joints = torch.empty(Fr, body.J, 3, device=dev)
    A = torch.empty(Fr, body.J, 12, device=dev)
    v_posed = torch.empty(Fr, body.V, 3, device=dev)
    verts = torch.empty(Fr, body.V, 3, device=dev)
    call("amav_lbs_forward", Fr, body.V, body.J, body.S, ptr(_f32c(full_pose)), ptr(_f32c(shape_comp)),
         ptr(body.v_template), ptr(body.shapedirs), ptr(body.posedirs), ptr(body.parents), ptr(body.J_template),
         ptr(body.J_dirs), ptr(body.weights_T), ptr(pf), ptr(joints), ptr(A), ptr(v_posed), ptr(verts), stream_ptr())
    if return_aux:
        return verts, {"pose_feature": pf, "joints": joints, "A": A, "v_posed": v_posed}
    return verts


def pose_6d_from_decoder_heads(body_root_6d: torch.Tensor, body_6d: torch.Tensor, hand_6d: torch.Tensor,
                               jaw_6d: torch.Tensor, leye_6d: torch.Tensor, reye_6d: torch.Tensor) -> torch.Tensor:
    """[F,J,6] in smplx.SMPLX.forward's joint order (global, 21 body, jaw, leye, reye, 15 left hand, 15 right hand)
    from the six 6D heads of the reference's SMPLXDecoder (smplx_decoder.py:106-127: `pred_hand_pose_6d` holds the left
    hand first, then the right)."""
    Fr = body_root_6d.shape[0]
    hand = hand_6d.reshape(Fr, -1, 6)
    nh = hand.shape[1] // 2
    parts = [body_root_6d.reshape(Fr, 1, 6), body_6d.reshape(Fr, -1, 6), jaw_6d.reshape(Fr, 1, 6),
             leye_6d.reshape(Fr, 1, 6), reye_6d.reshape(Fr, 1, 6), hand[:, :nh], hand[:, nh:]]
    return torch.cat(parts, dim=1).contiguous()


def lbs_forward_6d(body: BodyModel, pose_6d: torch.Tensor, shape_comp: torch.Tensor, return_aux: bool = False):
    """Posed vertices [F,V,3] from 6D rotations [F,J,6] (amav_lbs_forward_6d).  The axis-angle pose_mean
    of the body model cannot be composed with a rotation given as a matrix, so it must be zero (flat_hand_mean=True,
    as the reference builds its model when it predicts the parameters).  Plain forward; ``lbs_vertices_6d`` is the
    autograd-aware call."""
    pose_6d = pose_6d.reshape(-1, body.J, 6)
    Fr = int(pose_6d.shape[0])
    if shape_comp.shape != (Fr, body.S):
        raise ValueError("shape_comp must be [F,S]")
    if not hasattr(body, "_pose_mean_zero"):
        body._pose_mean_zero = bool((body.pose_mean == 0).all().item())
    if not body._pose_mean_zero:
        raise ValueError("the 6D pose path needs a body model with pose_mean == 0 (flat_hand_mean=True)")
    dev = body.device
    pf = torch.empty(Fr, (body.J - 1) * 9, device=dev)
    joints = torch.empty(Fr, body.J, 3, device=dev)
    A = torch.empty(Fr, body.J, 12, device=dev)
    v_posed = torch.empty(Fr, body.V, 3, device=dev)
    verts = torch.empty(Fr, body.V, 3, device=dev)
    call("amav_lbs_forward_6d", Fr, body.V, body.J, body.S, ptr(_f32c(pose_6d)), ptr(_f32c(shape_comp)),
         ptr(body.v_template), ptr(body.shapedirs), ptr(body.posedirs), ptr(body.parents), ptr(body.J_template),
         ptr(body.J_dirs), ptr(body.weights_T), ptr(pf), ptr(joints), ptr(A), ptr(v_posed), ptr(verts), stream_ptr())
    if return_aux:
        return verts, {"pose_feature": pf, "joints": joints, "A": A, "v_posed": v_posed}
    return verts


class Lbs6dFunction(torch.autograd.Function):
    """(pose_6d [F,J,6], shape_comp [F,S]) -> posed vertices [F,V,3]; gradients by amav_lbs_backward_6d."""

    @staticmethod
    def forward(ctx, body, pose_6d, shape_comp):
        verts, aux = lbs_forward_6d(body, pose_6d, shape_comp, return_aux=True)
        ctx.body = body
        ctx.save_for_backward(pose_6d, shape_comp, aux["A"], aux["v_posed"])
        return verts

    @staticmethod
    def backward(ctx, d_verts):
        body = ctx.body
        pose_6d, shape_comp, A, v_posed = ctx.saved_tensors
        Fr, dev = int(A.shape[0]), body.device
        d_v_posed = torch.empty(Fr, body.V, 3, device=dev)
        d_A = torch.zeros(Fr, body.J, 12, device=dev)
        d_pf = torch.zeros(Fr, (body.J - 1) * 9, device=dev)
        d_pose = torch.empty(Fr, body.J, 6, device=dev)
        d_shape = torch.zeros(Fr, body.S, device=dev)
        call("amav_lbs_backward_6d", Fr, body.V, body.J, body.S, ptr(_f32c(pose_6d)), ptr(_f32c(shape_comp)),
             ptr(body.shapedirs), ptr(body.posedirs), ptr(body.parents), ptr(body.J_template), ptr(body.J_dirs),
             ptr(body.weights_T), ptr(A), ptr(v_posed), ptr(_f32c(d_verts)), ptr(d_v_posed), ptr(d_A), ptr(d_pf),
             ptr(d_pose), ptr(d_shape), stream_ptr())
        return None, d_pose.reshape(pose_6d.shape), d_shape


def lbs_vertices_6d(body: BodyModel, pose_6d: torch.Tensor, shape_comp: torch.Tensor) -> torch.Tensor:
    """Autograd-aware 6D LBS: the plain forward when nothing requires a gradient."""
    if torch.is_grad_enabled() and (pose_6d.requires_grad or shape_comp.requires_grad):
        return Lbs6dFunction.apply(body, pose_6d.reshape(-1, body.J, 6), shape_comp)
    return lbs_forward_6d(body, pose_6d, shape_comp)


def frames_to_u8(frames: torch.Tensor, bgr: bool = False, out: Optional[torch.Tensor] = None,
                 out_first: int = 0, out_step: int = 1) -> torch.Tensor:
    """Video sink (amav_frames_to_u8; main2.py:226-239, :332-363): fp32 frames [F,H,W,3] -> uint8 [F,H,W,3] as
    `(frame * 255).astype(np.uint8)`, optionally RGB->BGR.  With ``out`` [F_out,H,W,3] frame f is written to
    out[out_first + f * out_step] (out_step = 2 interleaves two sequences the way the demo's combined_seq does)."""
    if frames.dim() != 4 or frames.shape[-1] != 3 or frames.dtype != torch.float32:
        raise ValueError("frames must be fp32 [F,H,W,3]")
    Fr, H, W = int(frames.shape[0]), int(frames.shape[1]), int(frames.shape[2])
    if out is None:
        if out_first != 0 or out_step != 1:
            raise ValueError("out_first / out_step need an explicit out buffer")
        out = torch.empty(Fr, H, W, 3, dtype=torch.uint8, device=frames.device)
    if out.dtype != torch.uint8 or out.dim() != 4 or tuple(out.shape[1:]) != (H, W, 3):
        raise ValueError("out must be uint8 [F_out,H,W,3]")
    if Fr and out_first + (Fr - 1) * out_step >= out.shape[0]:
        raise ValueError("out has too few frames for out_first / out_step")
    call("amav_frames_to_u8", Fr, H, W, ptr(frames), H * W * 3, ptr(out), H * W * 3, out_first, out_step,
         int(bool(bgr)), stream_ptr())
    return out


def densify_gather(verts: torch.Tensor, par: torch.Tensor, wgt: torch.Tensor, order: Optional[torch.Tensor] = None) -> torch.Tensor:
    """sites[f,i] = sum_k wgt[i,k] * verts[f, par[i,k]] (flattened SubdivideMeshes levels + subset,
    renderer.py:276-288).  par int32 [N,P], wgt fp32 [N,P]."""
    Fr, V = int(verts.shape[0]), int(verts.shape[1])
    N, P = int(par.shape[0]), int(par.shape[1])
    sites = torch.empty(Fr, N, 3, device=verts.device)
    call("amav_densify_gather", Fr, V, N, P, ptr(_f32c(verts)), ptr(par), ptr(wgt), ptr(sites), ptr(order), stream_ptr())
    return sites


def decoder_matrix(decoder) -> Tuple[torch.Tensor, torch.Tensor]:
    """[14, 3+3C] weight and [14] bias in head order xyz, rotation, scaling, opacity, shs
    (renderer.py:51-55) from an ``nn.Module`` with ``*_layer`` Linears or a state-dict-like mapping."""
    names = ("xyz", "rotation", "scaling", "opacity", "shs")
    if isinstance(decoder, dict):
        w = [decoder[f"{n}_layer.weight"] for n in names]
        b = [decoder[f"{n}_layer.bias"] for n in names]
    else:
        w = [getattr(decoder, f"{n}_layer").weight for n in names]
        b = [getattr(decoder, f"{n}_layer").bias for n in names]
    return torch.cat(w, dim=0), torch.cat(b, dim=0)


PROJECT_IMPL = "ffma"      # "ffma": FP32 FMA kernel (csrc/triplane.cu); "tc": tcgen05 3xTF32 kernel (csrc/triplane_tc.cu)


def triplane_project(tokens: torch.Tensor, weight: torch.Tensor, C: int, R: int,
                     out: Optional[torch.Tensor] = None, token_major: bool = False, impl: Optional[str] = None) -> torch.Tensor:
    """tokens [F, C, 3*R*R] (channel-major, renderer.py:85-91; the frame stride may exceed C*3*R*R)
    -> projected planes [F, 3, R*R, 16].  ``token_major``: tokens [F, 3*R*R, C], the layout the upstream
    transformer computes in before its permute+contiguous (transformers.py:1070-1073).
    ``impl``: "ffma" or "tc" (tensor cores: tcgen05.mma kind::tf32, 3xTF32 split, accumulator in TMEM); default PROJECT_IMPL."""
    if token_major:
        if tokens.dim() != 3 or tokens.shape[1] != 3 * R * R or tokens.shape[2] != C:
            raise ValueError(f"token-major tokens must be [F,{3 * R * R},{C}], got {tuple(tokens.shape)}")
        if tokens.dtype != torch.float32 or not tokens.is_cuda:
            raise _lib.AmavError("tokens must be CUDA fp32")
        if tokens.stride(2) != 1 or tokens.stride(1) != C:
            tokens = tokens.contiguous()
        Fr = int(tokens.shape[0])
        fstride = int(tokens.stride(0)) if Fr > 1 else C * 3 * R * R
        if out is None:
            out = torch.empty(Fr, 3, R * R, 16, device=tokens.device)
        call("amav_triplane_project_tm", Fr, C, R, ctypes_ptr(tokens), fstride, ptr(_f32c(weight)), ptr(out), stream_ptr())
        return out
    if tokens.dim() != 3 or tokens.shape[1] != C or tokens.shape[2] != 3 * R * R:
        raise ValueError(f"tokens must be [F,{C},{3 * R * R}], got {tuple(tokens.shape)}")
    if tokens.dtype != torch.float32 or not tokens.is_cuda:
        raise _lib.AmavError("tokens must be CUDA fp32")
    if tokens.stride(2) != 1 or tokens.stride(1) != 3 * R * R:
        tokens = tokens.contiguous()
    Fr = int(tokens.shape[0])
    fstride = int(tokens.stride(0)) if Fr > 1 else C * 3 * R * R
    if out is None:
        out = torch.empty(Fr, 3, R * R, 16, device=tokens.device)
    if (impl or PROJECT_IMPL) == "tc":
        if C % 32 != 0:
            raise ValueError("the tensor-core projection needs C % 32 == 0")
        scratch = torch.empty(3 * (C // 32) * 1024, device=tokens.device)
        call("amav_triplane_project_tc", Fr, C, R, ctypes_ptr(tokens), fstride, ptr(_f32c(weight)), ptr(scratch), ptr(out), stream_ptr())
        return out
    call("amav_triplane_project", Fr, C, R, ctypes_ptr(tokens), fstride, ptr(_f32c(weight)), ptr(out), stream_ptr())
    return out


def ctypes_ptr(t: torch.Tensor):
    """Device pointer of a tensor whose inner dims are dense but whose outer stride may be padded."""
    import ctypes
    return ctypes.c_void_p(t.data_ptr())


def sample_decode(proj: torch.Tensor, points: torch.Tensor, transl: torch.Tensor, weight: torch.Tensor,
                  bias: torch.Tensor, C: int, R: int, radius: float, order: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
    """Bilinear query of the projected planes at ``points`` [F,N,3] + xyz/bias terms + construct_gaussians
    (renderer.py:292-346).  Returns the reference's ``gaussians`` dict (each [F,N,k]; shs is color)."""
    Fr, N = int(points.shape[0]), int(points.shape[1])
    dev = points.device
    xyz = torch.empty(Fr, N, 3, device=dev)
    scale = torch.empty(Fr, N, 3, device=dev)
    rot = torch.empty(Fr, N, 4, device=dev)
    opacity = torch.empty(Fr, N, 1, device=dev)
    color = torch.empty(Fr, N, 3, device=dev)
    call("amav_sample_decode", Fr, N, C, R, float(radius), ptr(proj), ptr(_f32c(points)),
         ptr(_f32c(transl.reshape(Fr, 3))), ptr(_f32c(weight)), ptr(_f32c(bias)), ptr(xyz), ptr(scale), ptr(rot),
         ptr(opacity), ptr(color), ptr(order), stream_ptr())
    return {"xyz": xyz, "scale": scale, "rot": rot, "opacity": opacity, "color": color, "shs": color}


# ------------------------------------------------------------------------------------------------
# backward of stages 1+2 and the autograd node the Renderer uses in training
# ------------------------------------------------------------------------------------------------
def sample_decode_backward(proj, points, weight, bias, grads: Dict[str, Optional[torch.Tensor]], C: int, R: int,
                           radius: float, want_points: bool = True):
    """grads: dL/d{xyz,scale,rot,opacity,color} ([F,N,k] or None).  Returns (d_proj [F,3,R*R,16], d_points [F,N,3] or
    None, d_transl [F,3], d_wb [14,4])."""
    Fr, N = int(points.shape[0]), int(points.shape[1])
    dev = points.device
    d_proj = torch.zeros(Fr, 3, R * R, 16, device=dev)
    d_points = torch.empty(Fr, N, 3, device=dev) if want_points else None
    d_transl = torch.zeros(Fr, 3, device=dev)
    d_wb = torch.zeros(14, 4, device=dev)
    g = lambda k: None if grads.get(k) is None else _f32c(grads[k])
    gx, gs, gr, go, gc = g("xyz"), g("scale"), g("rot"), g("opacity"), g("color")
    call("amav_sample_decode_backward", Fr, N, C, R, float(radius), ptr(proj), ptr(_f32c(points)), ptr(_f32c(weight)),
         ptr(_f32c(bias)), ptr(gx), ptr(gs), ptr(gr), ptr(go), ptr(gc), ptr(d_proj), ptr(d_points), ptr(d_transl),
         ptr(d_wb), stream_ptr())
    return d_proj, d_points, d_transl, d_wb


def triplane_project_backward(d_proj, tokens, weight, C: int, R: int, want_tokens: bool = True, want_weight: bool = True,
                              token_major: bool = False):
    """d_proj [F,3,R*R,16] -> (d_tokens [F,C,3*R*R] ([F,3*R*R,C] when ``token_major``) or None, d_weight [14,3+3C]
    feature columns or None)."""
    Fr = int(d_proj.shape[0])
    dev = d_proj.device
    shape = (Fr, 3 * R * R, C) if token_major else (Fr, C, 3 * R * R)
    d_tokens = torch.empty(shape, device=dev) if want_tokens else None
    d_weight = torch.zeros(14, 3 + 3 * C, device=dev) if want_weight else None
    tok_ptr, fstride = None, C * 3 * R * R
    if want_weight:
        if tokens.stride(2) != 1 or tokens.stride(1) != shape[2]:
            tokens = tokens.contiguous()
        tok_ptr = ctypes_ptr(tokens)
        fstride = int(tokens.stride(0)) if Fr > 1 else C * 3 * R * R
    call("amav_triplane_project_backward_tm" if token_major else "amav_triplane_project_backward", Fr, C, R, ptr(d_proj),
         tok_ptr, fstride, ptr(_f32c(weight)), ptr(d_tokens), C * 3 * R * R, 0, ptr(d_weight), stream_ptr())
    return d_tokens, d_weight


class DecodeFunction(torch.autograd.Function):
    """tokens [F,C,3R^2] ([F,3R^2,C] when ``token_major``), points [F,N,3], transl [F,3], weight [14,3+3C], bias [14] -> (xyz, scale, rot, opacity, color):
    project-then-sample forward, and its hand-written backward (no [F,N,1536] feature tensor in either direction)."""

    @staticmethod
    def forward(ctx, tokens, points, transl, weight, bias, C, R, radius, token_major=False, order=None):
        proj = triplane_project(tokens, weight.detach(), C, R, token_major=token_major)
        g = sample_decode(proj, points, transl, weight.detach(), bias.detach(), C, R, radius, order=order)
        ctx.save_for_backward(tokens, points, weight, bias, proj)
        ctx.dims = (C, R, radius)
        ctx.token_major = bool(token_major)
        return g["xyz"], g["scale"], g["rot"], g["opacity"], g["color"]

    @staticmethod
    def backward(ctx, g_xyz, g_scale, g_rot, g_opacity, g_color):
        tokens, points, weight, bias, proj = ctx.saved_tensors
        C, R, radius = ctx.dims
        need_tok, need_pts, need_tr, need_w, need_b = ctx.needs_input_grad[:5]
        d_proj, d_points, d_transl, d_wb = sample_decode_backward(
            proj, points, weight, bias, {"xyz": g_xyz, "scale": g_scale, "rot": g_rot, "opacity": g_opacity, "color": g_color},
            C, R, radius, want_points=need_pts)
        d_tokens = d_weight = None
        if need_tok or need_w:
            d_tokens, d_weight = triplane_project_backward(d_proj, tokens, weight, C, R, want_tokens=need_tok, want_weight=need_w,
                                                           token_major=ctx.token_major)
            if d_weight is not None:
                d_weight[:, :3] = d_wb[:, :3]
            if d_tokens is not None and d_tokens.shape != tokens.shape:
                d_tokens = d_tokens.reshape(tokens.shape)
        return (d_tokens, d_points, d_transl.reshape(-1, 3) if need_tr else None, d_weight,
                d_wb[:, 3].contiguous() if need_b else None, None, None, None, None, None)


def decode_gaussians_autograd(tokens, points, transl, weight, bias, C: int, R: int, radius: float,
                              token_major: bool = False, order: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
    xyz, scale, rot, opacity, color = DecodeFunction.apply(tokens, points, transl, weight, bias, int(C), int(R), float(radius),
                                                           bool(token_major), order)
    return {"xyz": xyz, "scale": scale, "rot": rot, "opacity": opacity, "color": color, "shs": color}


# ------------------------------------------------------------------------------------------------
# backward of stage 3 (LBS + densify) as autograd nodes
# ------------------------------------------------------------------------------------------------
class LbsFunction(torch.autograd.Function):
    """(full_pose [F,J*3], shape_comp [F,S]) -> posed vertices [F,V,3]; gradients by amav_lbs_backward."""

    @staticmethod
    def forward(ctx, body, full_pose, shape_comp):
        verts, aux = lbs_forward(body, full_pose, shape_comp, return_aux=True)
        ctx.body = body
        ctx.save_for_backward(full_pose, shape_comp, aux["A"], aux["v_posed"])
        return verts

    @staticmethod
    def backward(ctx, d_verts):
        body = ctx.body
        full_pose, shape_comp, A, v_posed = ctx.saved_tensors
        Fr, dev = int(full_pose.shape[0]), body.device
        d_v_posed = torch.empty(Fr, body.V, 3, device=dev)
        d_A = torch.zeros(Fr, body.J, 12, device=dev)
        d_pf = torch.zeros(Fr, (body.J - 1) * 9, device=dev)
        d_pose = torch.empty(Fr, body.J * 3, device=dev)
        d_shape = torch.zeros(Fr, body.S, device=dev)
        call("amav_lbs_backward", Fr, body.V, body.J, body.S, ptr(_f32c(full_pose)), ptr(_f32c(shape_comp)),
             ptr(body.shapedirs), ptr(body.posedirs), ptr(body.parents), ptr(body.J_template), ptr(body.J_dirs),
             ptr(body.weights_T), ptr(A), ptr(v_posed), ptr(_f32c(d_verts)), ptr(d_v_posed), ptr(d_A), ptr(d_pf),
             ptr(d_pose), ptr(d_shape), stream_ptr())
        return None, d_pose, d_shape


class DensifyFunction(torch.autograd.Function):
    """verts [F,V,3] -> sites [F,N,3] through the fixed (parents, weights) table."""

    @staticmethod
    def forward(ctx, verts, par, wgt):
        ctx.save_for_backward(par, wgt)
        ctx.V = int(verts.shape[1])
        return densify_gather(verts, par, wgt)

    @staticmethod
    def backward(ctx, d_sites):
        par, wgt = ctx.saved_tensors
        Fr, N, P = int(d_sites.shape[0]), int(par.shape[0]), int(par.shape[1])
        d_verts = torch.zeros(Fr, ctx.V, 3, device=d_sites.device)
        call("amav_densify_gather_backward", Fr, ctx.V, N, P, ptr(_f32c(d_sites)), ptr(par), ptr(wgt), ptr(d_verts), stream_ptr())
        return d_verts, None, None


def lbs_vertices(body: BodyModel, full_pose: torch.Tensor, shape_comp: torch.Tensor) -> torch.Tensor:
    """Autograd-aware LBS: the plain forward when nothing requires a gradient."""
    if torch.is_grad_enabled() and (full_pose.requires_grad or shape_comp.requires_grad):
        return LbsFunction.apply(body, full_pose, shape_comp)
    return lbs_forward(body, full_pose, shape_comp)


def densify_sites(verts: torch.Tensor, par: torch.Tensor, wgt: torch.Tensor, order: Optional[torch.Tensor] = None) -> torch.Tensor:
    if torch.is_grad_enabled() and verts.requires_grad:
        return DensifyFunction.apply(verts, par, wgt)
    return densify_gather(verts, par, wgt, order)


def locality_order(rest_sites: torch.Tensor, extent: float) -> torch.Tensor:
    """Processing order for the per-site kernels: Morton (Z-curve) order of the rest-pose site positions [N,3] inside
    [-extent, extent]^3, 10 bits per axis.  int32 [N] on the sites' device; a permutation, so results do not depend on it."""
    q = ((rest_sites / extent).clamp(-1, 1) * 0.5 + 0.5) * 1023.0
    q = q.to(torch.int64).clamp(0, 1023)

    def spread(v):                                   # 10 bits -> every third bit
        v = (v | (v << 16)) & 0x030000FF
        v = (v | (v << 8)) & 0x0300F00F
        v = (v | (v << 4)) & 0x030C30C3
        v = (v | (v << 2)) & 0x09249249
        return v

    code = spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)
    return torch.argsort(code, stable=True).to(torch.int32).contiguous()
