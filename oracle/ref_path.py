"""oracle/ref_path.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement (torch CPU fp32 tensors used as plain arrays) of the reference's render path,
src/models/renderer.py of liubingqi7/audio-motion-avatar.  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import this module.

Pinning status
--------------
* sample_from_triplane / decoder heads / construct_gaussians / camera set-up: PINNED against
  outputs of the reference's own functions, imported in the build container with stub modules
  (tests/golden/make_golden.py -> tests/golden/ref_*.npz; checked by tests/test_oracle_golden.py).
* LBS (smplx==0.1.28, requirements.txt:12), subdivision (pytorch3d.ops.SubdivideMeshes,
  requirements.txt:10) and the rasteriser (diff-gaussian-rasterization, README.md:117-121) live in
  third-party packages absent from /root/reference and not installable offline, and the reference
  ships no tests or fixtures for them: PARITY UNPINNED for those three; they restate the packages'
  published algorithms and are checked by invariants (tests/test_oracle_invariants.py).
"""
from __future__ import annotations

import ctypes
import math
import os
import subprocess
from typing import Dict, Optional, Tuple

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))

SCALE_BIAS = 3.9      # renderer.py:428
OPACITY_BIAS = 0.0    # renderer.py:429


# ----------------------------------------------------------------------------------------
# a-4  triplane query (renderer.py:292-317) -- explicit bilinear, no grid_sample
# ----------------------------------------------------------------------------------------
def sample_from_triplane(triplane: torch.Tensor, points: torch.Tensor, radius: float) -> torch.Tensor:
    """triplane [F,3,C,Hp,Wp], points [F,N,3] -> [F,N,3C] (plane-major features).

    pos = clamp(points/radius, -1, 1) (renderer.py:298); planes sample (x,y),(x,z),(y,z)
    (renderer.py:300-303) with the first coordinate indexing W; bilinear, align_corners=False,
    zero padding (renderer.py:305-310): ix = ((gx+1)*W-1)/2, out-of-range taps contribute 0.
    """
    Fr, Np, C, Hp, Wp = triplane.shape
    N = points.shape[1]
    pos = torch.clamp(points / radius, -1.0, 1.0)
    out = torch.empty(Fr, N, Np * C, dtype=triplane.dtype)
    pairs = ((0, 1), (0, 2), (1, 2))
    for p, (a, b) in enumerate(pairs):
        ix = ((pos[..., a] + 1.0) * Wp - 1.0) / 2.0
        iy = ((pos[..., b] + 1.0) * Hp - 1.0) / 2.0
        x0 = torch.floor(ix)
        y0 = torch.floor(iy)
        wx1 = ix - x0
        wy1 = iy - y0
        wx0 = 1.0 - wx1
        wy0 = 1.0 - wy1
        acc = torch.zeros(Fr, N, C, dtype=triplane.dtype)
        plane = triplane[:, p].reshape(Fr, C, Hp * Wp)
        for dy, wy in ((0, wy0), (1, wy1)):
            for dx, wx in ((0, wx0), (1, wx1)):
                xi = (x0 + dx).long()
                yi = (y0 + dy).long()
                ok = (xi >= 0) & (xi < Wp) & (yi >= 0) & (yi < Hp)
                lin = (yi.clamp(0, Hp - 1) * Wp + xi.clamp(0, Wp - 1))            # [F,N]
                tex = torch.gather(plane, 2, lin[:, None, :].expand(Fr, C, N))    # [F,C,N]
                w = (wx * wy) * ok.to(triplane.dtype)
                acc += tex.permute(0, 2, 1) * w[..., None]
        out[:, :, p * C:(p + 1) * C] = acc
    return out


# ----------------------------------------------------------------------------------------
# a-5 / a-6  decoder heads and construct_gaussians (renderer.py:165-171, 319-346)
# ----------------------------------------------------------------------------------------
HEADS = (("xyz", 3), ("rotation", 4), ("scaling", 3), ("opacity", 1), ("shs", 3))


def decode_heads(points: torch.Tensor, feats: torch.Tensor, weights: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """x = [points, feats] (renderer.py:165); five nn.Linear heads (renderer.py:167-171)."""
    x = torch.cat([points, feats], dim=-1)
    out = {}
    for name, _ in HEADS:
        out[name] = x @ weights[f"{name}_layer.weight"].T + weights[f"{name}_layer.bias"]
    return out


def construct_gaussians(dec: Dict[str, torch.Tensor], points: torch.Tensor, transl: torch.Tensor) -> Dict[str, torch.Tensor]:
    """renderer.py:319-346: rot = normalize(rotation) (eps 1e-12), color = sigmoid(shs),
    xyz = points + offset + transl; scale / opacity are passed on raw."""
    rot = dec["rotation"]
    rot = rot / rot.norm(dim=-1, keepdim=True).clamp_min(1e-12)
    color = torch.sigmoid(dec["shs"])
    return {
        "xyz": points + dec["xyz"] + transl.reshape(-1, 1, 3),
        "scale": dec["scaling"],
        "rot": rot,
        "opacity": dec["opacity"],
        "color": color,
        "shs": color,
    }


# ----------------------------------------------------------------------------------------
# a-3  SMPL-X forward = smplx.lbs.lbs (smplx==0.1.28; call site renderer.py:261-274)
# ----------------------------------------------------------------------------------------
def batch_rodrigues(rot_vecs: torch.Tensor) -> torch.Tensor:
    """smplx.lbs.batch_rodrigues: angle = ||r + 1e-8||, R = I + sin K + (1-cos) K K."""
    n = rot_vecs.shape[0]
    angle = torch.norm(rot_vecs + 1e-8, dim=1, keepdim=True)
    rot_dir = rot_vecs / angle
    cos = torch.cos(angle)[:, None]
    sin = torch.sin(angle)[:, None]
    rx, ry, rz = torch.split(rot_dir, 1, dim=1)
    zeros = torch.zeros(n, 1, dtype=rot_vecs.dtype)
    K = torch.cat([zeros, -rz, ry, rz, zeros, -rx, -ry, rx, zeros], dim=1).view(n, 3, 3)
    ident = torch.eye(3, dtype=rot_vecs.dtype)[None]
    return ident + sin * K + (1 - cos) * torch.bmm(K, K)


def full_pose_from_params(p: Dict[str, torch.Tensor], pose_mean: torch.Tensor) -> torch.Tensor:
    """smplx.SMPLX.forward joint order: global, body(21), jaw, leye, reye, lhand(15), rhand(15);
    then ``full_pose += pose_mean``.  Inputs are already flattened to [F, ...]."""
    Fr = p["global_orient"].shape[0]
    full = torch.cat([
        p["global_orient"].reshape(Fr, 1, 3), p["body_pose"].reshape(Fr, 21, 3),
        p["jaw_pose"].reshape(Fr, 1, 3), p["leye_pose"].reshape(Fr, 1, 3), p["reye_pose"].reshape(Fr, 1, 3),
        p["left_hand_pose"].reshape(Fr, 15, 3), p["right_hand_pose"].reshape(Fr, 15, 3)], dim=1)
    return full.reshape(Fr, 165) + pose_mean[None]


def lbs_vertices(params: Dict[str, torch.Tensor], model: Dict[str, torch.Tensor],
                 return_transforms: bool = False):
    """Posed vertices [F,V,3] WITHOUT transl (renderer.py:248-249,271: transl is commented out).

    v_shaped = v_template + shapedirs.[betas,expr];  J = J_regressor.v_shaped;
    v_posed = v_shaped + vec(R[1:]-I).posedirs;  chain G_j = G_parent.[R_j | J_j - J_parent];
    A_j = G_j - [0 | G_j.J_j];  T_v = sum_j W_vj A_j;  verts = (T_v.[v_posed;1])[:3].
    """
    Fr = params["global_orient"].shape[0]
    full_pose = full_pose_from_params(params, model["pose_mean"])
    shape_comp = torch.cat([params["betas"].reshape(Fr, -1), params["expression"].reshape(Fr, -1)], dim=-1)
    v_shaped = model["v_template"][None] + torch.einsum("bl,mkl->bmk", shape_comp, model["shapedirs"])
    J = torch.einsum("bik,ji->bjk", v_shaped, model["J_regressor"])
    rot = batch_rodrigues(full_pose.reshape(-1, 3)).view(Fr, -1, 3, 3)
    pose_feature = (rot[:, 1:] - torch.eye(3)[None, None]).reshape(Fr, -1)
    v_posed = v_shaped + torch.matmul(pose_feature, model["posedirs"]).view(Fr, -1, 3)
    parents = model["parents"].tolist()
    nj = len(parents)
    rel = J.clone()
    rel[:, 1:] = J[:, 1:] - J[:, parents[1:]]
    tm = torch.zeros(Fr, nj, 4, 4)
    tm[:, :, :3, :3] = rot
    tm[:, :, :3, 3] = rel
    tm[:, :, 3, 3] = 1.0
    chain = [tm[:, 0]]
    for j in range(1, nj):
        chain.append(torch.matmul(chain[parents[j]], tm[:, j]))
    G = torch.stack(chain, dim=1)
    Jh = torch.cat([J, torch.zeros(Fr, nj, 1)], dim=2)[..., None]          # [F,J,4,1]
    corr = torch.matmul(G, Jh)                                             # [F,J,4,1]
    A = G.clone()
    A[:, :, :, 3:4] = G[:, :, :, 3:4] - corr
    T = torch.matmul(model["lbs_weights"][None].expand(Fr, -1, -1), A.view(Fr, nj, 16)).view(Fr, -1, 4, 4)
    vh = torch.cat([v_posed, torch.ones(Fr, v_posed.shape[1], 1)], dim=2)[..., None]
    verts = torch.matmul(T, vh)[:, :, :3, 0]
    if return_transforms:
        return verts, A, J
    return verts


# ----------------------------------------------------------------------------------------
# 8f-3  SMPLXDecoder tail: 6D -> matrix -> axis-angle (src/models/smplx_decoder.py:106-127 calls
# pytorch3d.transforms.rotation_6d_to_matrix / matrix_to_axis_angle; pytorch3d is absent here, PARITY UNPINNED:
# restated from the package's published algorithm, checked by tests/test_oracle_invariants.py)
# ----------------------------------------------------------------------------------------
def rotation_6d_to_matrix(d6: torch.Tensor) -> torch.Tensor:
    """pytorch3d.transforms.rotation_6d_to_matrix: Gram-Schmidt of the two 3-vectors; b1, b2, b3 are the ROWS."""
    a1, a2 = d6[..., :3], d6[..., 3:]
    b1 = torch.nn.functional.normalize(a1, dim=-1)
    b2 = a2 - (b1 * a2).sum(-1, keepdim=True) * b1
    b2 = torch.nn.functional.normalize(b2, dim=-1)
    b3 = torch.cross(b1, b2, dim=-1)
    return torch.stack((b1, b2, b3), dim=-2)


def _sqrt_positive_part(x: torch.Tensor) -> torch.Tensor:
    ret = torch.zeros_like(x)
    pos = x > 0
    ret[pos] = torch.sqrt(x[pos])
    return ret


def matrix_to_quaternion(matrix: torch.Tensor) -> torch.Tensor:
    """pytorch3d.transforms.matrix_to_quaternion (real part first): four candidate quaternions, the one with the
    largest denominator is taken; result standardised to a non-negative real part."""
    batch = matrix.shape[:-2]
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = torch.unbind(matrix.reshape(batch + (9,)), dim=-1)
    q_abs = _sqrt_positive_part(torch.stack([1.0 + m00 + m11 + m22, 1.0 + m00 - m11 - m22,
                                             1.0 - m00 + m11 - m22, 1.0 - m00 - m11 + m22], dim=-1))
    quat_by_rijk = torch.stack([
        torch.stack([q_abs[..., 0] ** 2, m21 - m12, m02 - m20, m10 - m01], dim=-1),
        torch.stack([m21 - m12, q_abs[..., 1] ** 2, m10 + m01, m02 + m20], dim=-1),
        torch.stack([m02 - m20, m10 + m01, q_abs[..., 2] ** 2, m12 + m21], dim=-1),
        torch.stack([m10 - m01, m20 + m02, m21 + m12, q_abs[..., 3] ** 2], dim=-1)], dim=-2)
    flr = torch.tensor(0.1, dtype=q_abs.dtype)
    cand = quat_by_rijk / (2.0 * q_abs[..., None].max(flr))
    pick = torch.nn.functional.one_hot(q_abs.argmax(dim=-1), num_classes=4) > 0.5
    out = cand[pick, :].reshape(batch + (4,))
    return torch.where(out[..., 0:1] < 0, -out, out)


def quaternion_to_axis_angle(q: torch.Tensor) -> torch.Tensor:
    """pytorch3d.transforms.quaternion_to_axis_angle."""
    norms = torch.norm(q[..., 1:], p=2, dim=-1, keepdim=True)
    half = torch.atan2(norms, q[..., :1])
    angles = 2 * half
    small = angles.abs() < 1e-6
    soa = torch.empty_like(angles)
    soa[~small] = torch.sin(half[~small]) / angles[~small]
    soa[small] = 0.5 - (angles[small] * angles[small]) / 48
    return q[..., 1:] / soa


def matrix_to_axis_angle(matrix: torch.Tensor) -> torch.Tensor:
    return quaternion_to_axis_angle(matrix_to_quaternion(matrix))


def params_from_pose_6d(pose_6d: torch.Tensor, betas: torch.Tensor, expression: torch.Tensor) -> Dict[str, torch.Tensor]:
    """The reference's route for predicted poses: every joint's 6D head -> matrix -> axis-angle
    (smplx_decoder.py:106-143), in the dict layout get_smpl_vertices reads (renderer.py:250-258).
    pose_6d [F,55,6] in SMPLX.forward's joint order."""
    Fr = pose_6d.shape[0]
    aa = matrix_to_axis_angle(rotation_6d_to_matrix(pose_6d))              # [F,55,3]
    return {"global_orient": aa[:, 0], "body_pose": aa[:, 1:22], "jaw_pose": aa[:, 22], "leye_pose": aa[:, 23],
            "reye_pose": aa[:, 24], "left_hand_pose": aa[:, 25:40], "right_hand_pose": aa[:, 40:55],
            "betas": betas.reshape(Fr, -1), "expression": expression.reshape(Fr, -1)}


# ----------------------------------------------------------------------------------------
# 8f-4  video sink of the demo (src/main2.py:351-353): (frame * 255).astype(np.uint8), optional cv2 RGB2BGR
# ----------------------------------------------------------------------------------------
def frames_to_u8(frames: np.ndarray, bgr: bool = False) -> np.ndarray:
    u8 = (np.asarray(frames, dtype=np.float32) * 255).astype(np.uint8)
    return u8[..., ::-1].copy() if bgr else u8


def subdivide_vertices(verts: torch.Tensor, level_edges) -> torch.Tensor:
    """pytorch3d SubdivideMeshes on the vertices only: per level, append the mean of each edge's
    endpoints (``verts[:, edges].mean(dim=2)``) after the existing vertices (renderer.py:282-285)."""
    for edges in level_edges:
        e = torch.as_tensor(edges, dtype=torch.long)
        mid = verts[:, e].mean(dim=2)
        verts = torch.cat([verts, mid], dim=1)
    return verts


# ----------------------------------------------------------------------------------------
# a-8  camera set-up of render_one (renderer.py:486-510; graphic_utils.py:67-78,103-136,144-145)
# ----------------------------------------------------------------------------------------
def camera_setup(K: torch.Tensor, E: torch.Tensor, height: int, width: int):
    """Returns (viewmatrix [4,4], projmatrix [4,4], tanfovx, tanfovy, campos [3]) exactly as
    render_one hands them to GaussianRasterizationSettings (already transposed)."""
    znear, zfar = 0.01, 100.0
    R = E[:3, :3].reshape(3, 3).transpose(1, 0)
    T = E[:3, 3]
    fovy = 2 * torch.atan(height / (2 * K[1, 1]))
    fovx = 2 * torch.atan(width / (2 * K[0, 0]))
    tanfovx = math.tan(fovx * 0.5)
    tanfovy = math.tan(fovy * 0.5)
    Rt = torch.zeros(4, 4)
    Rt[:3, :3] = R.transpose(0, 1)
    Rt[:3, 3] = T
    Rt[3, 3] = 1.0
    C2W = torch.inverse(Rt)
    C2W[:3, 3] = (C2W[:3, 3] + torch.tensor([.0, .0, .0])) * 1.0
    view = torch.inverse(C2W).transpose(0, 1)
    fx, fy, px, py = K[0, 0].item(), K[1, 1].item(), K[0, 2].item(), K[1, 2].item()
    K_ndc = torch.tensor([
        [2 * fx / width, 0, (2 * px - width) / width, 0],
        [0, 2 * fy / height, (2 * py - height) / height, 0],
        [0, 0, zfar / (zfar - znear), -zfar * znear / (zfar - znear)],
        [0, 0, 1, 0]]).float()
    proj = K_ndc.transpose(0, 1)
    full = view.unsqueeze(0).bmm(proj.unsqueeze(0)).squeeze(0)
    campos = view.inverse()[3, :3]
    return view.contiguous(), full.contiguous(), tanfovx, tanfovy, campos


def activate_for_raster(g: Dict[str, torch.Tensor]) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """renderer.py:532-533,547: scales = min(exp(s-3.9), 0.1), opac = sigmoid(o-0), colour = clamp(c,0,1)."""
    scales = torch.min(torch.exp(g["scale"] - SCALE_BIAS), torch.tensor(0.1))
    opac = torch.sigmoid(g["opacity"] - OPACITY_BIAS)
    col = torch.clamp(g["color"], 0.0, 1.0)
    return scales, opac, col


# ----------------------------------------------------------------------------------------
# a-9  rasteriser: ctypes binding of oracle/raster_oracle.c
# ----------------------------------------------------------------------------------------
_LIBS: Dict[str, ctypes.CDLL] = {}
_VARIANT = "exact"          # which build raster_* calls go to: "exact" | "upstream" | "fast" (see oracle/Makefile)


def _host_tag() -> str:
    """The fast build is compiled -march=native, so its file name carries a hash of this host's CPU flags: a .so
    built in the CPU container is not run on a GPU box with another CPU (it is rebuilt there; gcc is in the image)."""
    import hashlib
    try:
        with open("/proc/cpuinfo") as fh:
            flags = next((ln for ln in fh if ln.startswith("flags")), "")
    except OSError:
        flags = ""
    return hashlib.sha1(flags.encode()).hexdigest()[:10]


def build_c_oracle(force: bool = False, variant: str = "exact") -> str:
    src = os.path.join(_HERE, "raster_oracle.c")
    if variant == "exact":
        so, target = os.path.join(_HERE, "libamav_oracle.so"), ["libamav_oracle.so"]
    elif variant == "upstream":
        so, target = os.path.join(_HERE, "libamav_oracle_upstream.so"), ["libamav_oracle_upstream.so"]
    elif variant == "fast":
        name = f"libamav_oracle_fast_{_host_tag()}.so"
        so, target = os.path.join(_HERE, name), ["fast", f"FAST_OUT={name}"]
    else:
        raise ValueError(variant)
    if force or not os.path.exists(so) or (os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(so)):
        subprocess.check_call(["make", "-s", "-C", _HERE] + target)
    return so


def _lib(variant: Optional[str] = None):
    variant = variant or _VARIANT
    L = _LIBS.get(variant)
    if L is None:
        L = ctypes.CDLL(build_c_oracle(variant=variant))
        L.amav_oracle_scan.restype = ctypes.c_int64
        L.amav_oracle_exp.restype = ctypes.c_float
        L.amav_oracle_exp.argtypes = [ctypes.c_float]
        _LIBS[variant] = L
    return L


class oracle_variant:
    """``with oracle_variant("upstream"):`` routes the raster_* calls below to another build of raster_oracle.c."""

    def __init__(self, variant: str):
        self.variant = variant

    def __enter__(self):
        global _VARIANT
        self._prev, _VARIANT = _VARIANT, self.variant
        _lib(self.variant)
        return self

    def __exit__(self, *exc):
        global _VARIANT
        _VARIANT = self._prev


def _p(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def _f32(t) -> np.ndarray:
    return np.ascontiguousarray(t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else t, dtype=np.float32)


def raster_preprocess(means3d, scales, rots, opacities, view, proj, W, H, tanfovx, tanfovy, scale_modifier=1.0):
    m, s, r, o = _f32(means3d), _f32(scales), _f32(rots), _f32(opacities).reshape(-1)
    v, p = _f32(view).reshape(16), _f32(proj).reshape(16)
    N = m.shape[0]
    out = dict(depth=np.zeros(N, np.float32), radii=np.zeros(N, np.int32), xy=np.zeros((N, 2), np.float32),
               conic_opacity=np.zeros((N, 4), np.float32), cov3d=np.zeros((N, 6), np.float32),
               rect=np.zeros((N, 4), np.int32), tiles_touched=np.zeros(N, np.int32))
    _lib().amav_oracle_preprocess(
        ctypes.c_int(N), _p(m), _p(s), _p(r), _p(o), _p(v), _p(p), ctypes.c_int(W), ctypes.c_int(H),
        ctypes.c_float(tanfovx), ctypes.c_float(tanfovy), ctypes.c_float(scale_modifier),
        _p(out["depth"]), _p(out["radii"]), _p(out["xy"]), _p(out["conic_opacity"]), _p(out["cov3d"]),
        _p(out["rect"]), _p(out["tiles_touched"]))
    return out


def raster_bin(pre: Dict[str, np.ndarray], W: int, H: int):
    """scan -> emit -> stable sort -> ranges.  Returns dict(keys_unsorted, keys, point_list, ranges, D)."""
    N = pre["depth"].shape[0]
    gx, gy = (W + 15) // 16, (H + 15) // 16
    offsets = np.zeros(max(N, 1), np.int64)
    D = int(_lib().amav_oracle_scan(ctypes.c_int(N), _p(pre["tiles_touched"]), _p(offsets)))
    keys = np.zeros(max(D, 1), np.uint64)
    vals = np.zeros(max(D, 1), np.int32)
    _lib().amav_oracle_emit_keys(ctypes.c_int(N), _p(pre["radii"]), _p(pre["rect"]), _p(pre["depth"]),
                                 _p(offsets), ctypes.c_int(W), _p(keys), _p(vals))
    unsorted = keys[:D].copy()
    _lib().amav_oracle_sort_pairs(ctypes.c_int64(D), _p(keys), _p(vals))
    ranges = np.zeros((gx * gy, 2), np.int32)
    _lib().amav_oracle_tile_ranges(ctypes.c_int64(D), _p(keys), ctypes.c_int(gx * gy), _p(ranges))
    return dict(keys_unsorted=unsorted, keys=keys[:D], point_list=vals[:D], ranges=ranges, D=D, offsets=offsets[:N])


def raster_blend(pre, binned, colors, bg, W, H, want_invdepth=True):
    rgb, bgc = _f32(colors), _f32(bg)
    out = dict(color=np.zeros((3, H, W), np.float32), final_T=np.zeros((H, W), np.float32),
               n_contrib=np.zeros((H, W), np.int32), invdepth=np.zeros((H, W), np.float32))
    pl = binned["point_list"] if binned["D"] > 0 else np.zeros(1, np.int32)
    _lib().amav_oracle_blend(ctypes.c_int(W), ctypes.c_int(H), _p(binned["ranges"]), _p(pl), _p(pre["xy"]),
                             _p(pre["conic_opacity"]), _p(rgb), _p(pre["depth"]), _p(bgc), _p(out["color"]),
                             _p(out["final_T"]), _p(out["n_contrib"]),
                             _p(out["invdepth"]) if want_invdepth else None)
    return out


def rasterize(means3d, scales, rots, opacities, colors, view, proj, tanfovx, tanfovy, bg, W, H):
    """Full forward of the rasteriser for one frame; returns (pre, binned, blended) dicts."""
    pre = raster_preprocess(means3d, scales, rots, opacities, view, proj, W, H, tanfovx, tanfovy)
    binned = raster_bin(pre, W, H)
    out = raster_blend(pre, binned, colors, bg, W, H)
    return pre, binned, out


def render_one(g_frame: Dict[str, torch.Tensor], K: torch.Tensor, E: torch.Tensor, height: int, width: int,
               bg=(1.0, 1.0, 1.0)) -> torch.Tensor:
    """renderer.py:481-570 for one frame (rgb=True branch); returns [3,H,W] clamped to [0,1]."""
    view, full, tfx, tfy, _ = camera_setup(K, E, height, width)
    scales, opac, col = activate_for_raster(g_frame)
    _, _, out = rasterize(g_frame["xyz"], scales, g_frame["rot"], opac, col, view, full, tfx, tfy, bg, width, height)
    return torch.from_numpy(out["color"]).clamp(0, 1)


# ----------------------------------------------------------------------------------------
# the reference's own op sequence for query + decode, for the timed CPU baseline
# ----------------------------------------------------------------------------------------
def sample_from_triplane_torch(triplane: torch.Tensor, points: torch.Tensor, radius: float) -> torch.Tensor:
    """renderer.py:292-317 with the reference's own PyTorch calls (F.grid_sample on [(F*3),C,Hp,Wp], then the
    '(B Np) Cp () N -> B N (Np Cp)' transpose), used where the CPU baseline must cost what the reference costs.
    Checked against ``sample_from_triplane`` (the explicit restatement) in tests/test_oracle_golden.py."""
    Fr, Np, Cc, Hp, Wp = triplane.shape
    pos = torch.clamp(points / radius, -1, 1)
    idx = torch.stack((pos[..., [0, 1]], pos[..., [0, 2]], pos[..., [1, 2]]), dim=-3)           # [F,3,N,2]
    out = torch.nn.functional.grid_sample(triplane.reshape(Fr * Np, Cc, Hp, Wp), idx.reshape(Fr * Np, 1, -1, 2),
                                          align_corners=False, mode="bilinear")                # [(F 3),C,1,N]
    return out.reshape(Fr, Np, Cc, -1).permute(0, 3, 1, 2).reshape(Fr, -1, Np * Cc)


def reference_frames(model, weights, tokens, params, K, E, height, width, channels, res, radius=1.4,
                     level_edges=None, site_idx=None, bg=(1.0, 1.0, 1.0), differentiable=False):
    """The reference's Renderer.forward (renderer.py:73-204) for F frames on the CPU: LBS -> (subdivide + subset)
    -> grid_sample twice (the reference samples at initial and refined points, renderer.py:136,158; the refiner
    offset is exactly zero at init) -> five Linear heads -> construct_gaussians -> per-frame render_one.
    tokens [F,C,3*R*R]; params flattened to [F,...].  Returns images [F,H,W,3]."""
    pts = lbs_vertices(params, model)
    if level_edges is not None:
        pts = subdivide_vertices(pts, level_edges)
    if site_idx is not None:
        pts = pts[:, site_idx]
    Fr = tokens.shape[0]
    tri = tokens.reshape(Fr, channels, 3, res, res).permute(0, 2, 1, 3, 4)
    sample_from_triplane_torch(tri, pts, radius)                                               # renderer.py:136
    feats = sample_from_triplane_torch(tri, pts, radius)                                       # renderer.py:158
    g = construct_gaussians(decode_heads(pts, feats, weights), pts, params["transl"])
    one = render_one_autograd if differentiable else render_one      # autograd through the C rasteriser's backward
    imgs = [one({k: v[f] for k, v in g.items()}, K[f], E[f], height, width, bg).permute(1, 2, 0) for f in range(Fr)]
    return torch.stack(imgs), g


# ----------------------------------------------------------------------------------------
# a-10  rasteriser backward: ctypes binding of the C restatement
# ----------------------------------------------------------------------------------------
def raster_backward(means3d, scales, rots, opacities, colors, view, proj, tanfovx, tanfovy, bg, W, H,
                    pre, binned, out, dL_dcolor_out, scale_modifier=1.0):
    """Gradients of a scalar loss w.r.t. the rasteriser's inputs (activated scales/opacities, colours_precomp)
    given dL/d(colour output) [3,H,W] and the forward's (pre, binned, out) dicts.  Returns a dict with
    means3d [N,3], scales [N,3], rots [N,4], opacities [N,1], colors [N,3] plus the screen-space sums
    mean2d [N,2], conic [N,3]."""
    m, s, r = _f32(means3d), _f32(scales), _f32(rots)
    v, p = _f32(view).reshape(16), _f32(proj).reshape(16)
    rgb, bgc, g = _f32(colors), _f32(bg), _f32(dL_dcolor_out)
    N = m.shape[0]
    d2 = np.zeros((N, 2), np.float32); dcon = np.zeros((N, 3), np.float32)
    dop = np.zeros(N, np.float32); dcol = np.zeros((N, 3), np.float32)
    pl = binned["point_list"] if binned["D"] > 0 else np.zeros(1, np.int32)
    _lib().amav_oracle_blend_backward(
        ctypes.c_int(W), ctypes.c_int(H), ctypes.c_int(N), _p(binned["ranges"]), _p(pl), _p(pre["xy"]),
        _p(pre["conic_opacity"]), _p(rgb), _p(bgc), _p(out["final_T"]), _p(out["n_contrib"]), _p(g),
        _p(d2), _p(dcon), _p(dop), _p(dcol))
    dm = np.zeros((N, 3), np.float32); ds = np.zeros((N, 3), np.float32); dr = np.zeros((N, 4), np.float32)
    _lib().amav_oracle_preprocess_backward(
        ctypes.c_int(N), _p(m), _p(s), _p(r), _p(v), _p(p), ctypes.c_int(W), ctypes.c_int(H),
        ctypes.c_float(tanfovx), ctypes.c_float(tanfovy), ctypes.c_float(scale_modifier), _p(pre["radii"]),
        _p(d2), _p(dcon), _p(dm), _p(ds), _p(dr))
    return dict(means3d=dm, scales=ds, rots=dr, opacities=dop.reshape(N, 1), colors=dcol, mean2d=d2, conic=dcon)


class _OracleRaster(torch.autograd.Function):
    """One frame of the C rasteriser as an autograd node (forward = amav_oracle_*, backward = Appendix B), so the
    oracle's full path (LBS -> query -> decode -> activations -> rasteriser) is differentiable on the CPU the way
    the reference's is through diff_gaussian_rasterization's autograd Function."""

    @staticmethod
    def forward(ctx, means3d, scales, rots, opacities, colors, view, proj, tanfovx, tanfovy, bg, W, H):
        pre, binned, out = rasterize(means3d, scales, rots, opacities, colors, view, proj, tanfovx, tanfovy, bg, W, H)
        ctx.save_for_backward(means3d, scales, rots, opacities, colors, view, proj)
        ctx.misc = (tanfovx, tanfovy, bg, W, H, pre, binned, out)
        return torch.from_numpy(out["color"].copy())

    @staticmethod
    def backward(ctx, grad):
        means3d, scales, rots, opacities, colors, view, proj = ctx.saved_tensors
        tanfovx, tanfovy, bg, W, H, pre, binned, out = ctx.misc
        g = raster_backward(means3d, scales, rots, opacities, colors, view, proj, tanfovx, tanfovy, bg, W, H,
                            pre, binned, out, grad.contiguous())
        t = lambda k: torch.from_numpy(g[k])
        return (t("means3d"), t("scales"), t("rots"), t("opacities"), t("colors"), None, None, None, None, None, None, None)


def render_one_autograd(g_frame: Dict[str, torch.Tensor], K, E, height, width, bg=(1.0, 1.0, 1.0)) -> torch.Tensor:
    """render_one (renderer.py:481-570) for one frame with gradients to the Gaussian attributes; [3,H,W] in [0,1]."""
    view, full, tfx, tfy, _ = camera_setup(K, E, height, width)
    scales, opac, col = activate_for_raster(g_frame)
    img = _OracleRaster.apply(g_frame["xyz"], scales, g_frame["rot"], opac, col, view, full, tfx, tfy, bg, width, height)
    return img.clamp(0, 1)
