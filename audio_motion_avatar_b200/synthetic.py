"""Deterministic synthetic world for the render path.

No SMPL-X model file, checkpoint or dataset is available offline, so the inputs of the
reference's render path (src/models/renderer.py:73-204 in the reference) are generated here:

* an SMPL-X-*shaped* body model (V=10,475 vertices, J=55 joints with the real SMPL-X
  kinematic ``parents``, dense ``lbs_weights [V,55]`` with 4 non-zeros per row,
  ``shapedirs [V,3,20]``, ``posedirs [486,3V]``, ``J_regressor [55,V]``) whose surface is a
  union of closed capsules (torso, head, arms, legs);
* edge-midpoint subdivision tables equivalent to ``pytorch3d.ops.SubdivideMeshes`` applied
  ``levels`` times (reference call sites renderer.py:236-243,280-285);
* random triplane tokens, decoder weights, smooth pose sequences and pinhole cameras.

Everything is generated on the CPU from ``torch.Generator().manual_seed(seed)`` / numpy
``default_rng(seed)`` and is a pure function of its arguments.
"""
from __future__ import annotations

import math
from typing import Dict, List, Tuple

import numpy as np
import torch

NUM_JOINTS = 55
NUM_VERTS = 10475
NUM_BETAS = 10

# smplx kinematic tree (smplx==0.1.28, SMPLX_NEUTRAL kintree_table[0]); -1 marks the root.
SMPLX_PARENTS = [-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 15, 15, 15,
                 20, 25, 26, 20, 28, 29, 20, 31, 32, 20, 34, 35, 20, 37, 38,
                 21, 40, 41, 21, 43, 44, 21, 46, 47, 21, 49, 50, 21, 52, 53]


def _rest_joints() -> np.ndarray:
    """Rest-pose joint positions of the capsule body, metres, x right / y up / z front."""
    J = np.zeros((NUM_JOINTS, 3), dtype=np.float64)
    body = {
        0: (0, -0.25, 0), 1: (0.09, -0.33, 0), 2: (-0.09, -0.33, 0), 3: (0, -0.12, 0),
        4: (0.10, -0.72, 0), 5: (-0.10, -0.72, 0), 6: (0, 0.02, 0), 7: (0.10, -1.10, 0),
        8: (-0.10, -1.10, 0), 9: (0, 0.10, 0), 10: (0.10, -1.15, 0.10), 11: (-0.10, -1.15, 0.10),
        12: (0, 0.30, 0), 13: (0.07, 0.22, 0), 14: (-0.07, 0.22, 0), 15: (0, 0.40, 0),
        16: (0.19, 0.25, 0), 17: (-0.19, 0.25, 0), 18: (0.45, 0.25, 0), 19: (-0.45, 0.25, 0),
        20: (0.70, 0.25, 0), 21: (-0.70, 0.25, 0), 22: (0, 0.38, 0.05), 23: (0.03, 0.45, 0.08),
        24: (-0.03, 0.45, 0.08),
    }
    for k, v in body.items():
        J[k] = v
    finger_z = [0.03, 0.01, -0.03, -0.01, 0.05]  # index, middle, pinky, ring, thumb
    knuckle_x = [0.78, 0.82, 0.86]
    for side, base, sign in ((0, 25, 1.0), (1, 40, -1.0)):
        for f in range(5):
            for k in range(3):
                J[base + 3 * f + k] = (sign * knuckle_x[k], 0.25, finger_z[f])
    return J


def _capsule(p0, p1, r0, r1, rings: int, segs: int, flat: float = 1.0) -> Tuple[np.ndarray, np.ndarray]:
    """Closed lat-long surface around the segment p0->p1: 2 poles + rings*segs vertices.

    Radius goes r0 -> r1 along the axis with rounded caps; ``flat`` squashes the second
    cross-section axis (an elliptic cross-section).  Returns (verts [2+rings*segs,3], faces).
    """
    p0 = np.asarray(p0, dtype=np.float64)
    p1 = np.asarray(p1, dtype=np.float64)
    axis = p1 - p0
    length = np.linalg.norm(axis)
    a = axis / length
    helper = np.array([0.0, 0.0, 1.0]) if abs(a[2]) < 0.9 else np.array([1.0, 0.0, 0.0])
    u = np.cross(a, helper)
    u /= np.linalg.norm(u)
    v = np.cross(a, u)
    verts = [p0 - a * r0, p1 + a * r1]
    for r in range(rings):
        t = (r + 0.5) / rings                       # (0,1) along the capsule incl. caps
        s = -r0 + t * (length + r0 + r1)            # signed position along the axis from p0
        rad_here = r0 + (r1 - r0) * min(max(s / length, 0.0), 1.0)
        if s < 0.0:
            rad = math.sqrt(max(r0 * r0 - s * s, 1e-8))
        elif s > length:
            rad = math.sqrt(max(r1 * r1 - (s - length) ** 2, 1e-8))
        else:
            rad = rad_here
        c = p0 + a * s
        for k in range(segs):
            th = 2.0 * math.pi * (k + 0.5 * (r & 1)) / segs
            verts.append(c + rad * (math.cos(th) * u + flat * math.sin(th) * v))
    faces = []
    ring = lambda r, k: 2 + r * segs + (k % segs)
    for k in range(segs):
        faces.append((0, ring(0, k + 1), ring(0, k)))
        faces.append((1, ring(rings - 1, k), ring(rings - 1, k + 1)))
    for r in range(rings - 1):
        for k in range(segs):
            faces.append((ring(r, k), ring(r, k + 1), ring(r + 1, k)))
            faces.append((ring(r, k + 1), ring(r + 1, k + 1), ring(r + 1, k)))
    return np.asarray(verts), np.asarray(faces, dtype=np.int64)


def make_body_template() -> Tuple[np.ndarray, np.ndarray]:
    """v_template [10475,3] float32 and faces [20926,3] int64 of the capsule body."""
    parts = [
        _capsule((0, -0.30, 0), (0, 0.28, 0), 0.17, 0.15, 60, 58, flat=0.65),     # torso
        _capsule((0, 0.40, 0.01), (0, 0.52, 0.01), 0.095, 0.095, 40, 38),          # head
        _capsule((0.17, 0.25, 0), (0.88, 0.25, 0), 0.05, 0.03, 45, 30),            # left arm
        _capsule((-0.17, 0.25, 0), (-0.88, 0.25, 0), 0.05, 0.03, 45, 30),          # right arm
        _capsule((0.10, -0.30, 0), (0.10, -1.12, 0), 0.085, 0.045, 46, 30),        # left leg
        _capsule((-0.10, -0.30, 0), (-0.10, -1.12, 0), 0.085, 0.045, 46, 30),      # right leg
    ]
    verts, faces, base = [], [], 0
    for v, f in parts:
        verts.append(v)
        faces.append(f + base)
        base += v.shape[0]
    verts = np.concatenate(verts)
    faces = np.concatenate(faces)
    # top up to exactly NUM_VERTS with 1->3 face splits (keeps every part closed, genus 0)
    k = 0
    while verts.shape[0] < NUM_VERTS:
        fi = 7 + 997 * k
        a, b, c = faces[fi]
        n = verts.shape[0]
        verts = np.concatenate([verts, verts[[a, b, c]].mean(0, keepdims=True)])
        faces[fi] = (a, b, n)
        faces = np.concatenate([faces, np.array([(b, c, n), (c, a, n)], dtype=np.int64)])
        k += 1
    assert verts.shape[0] == NUM_VERTS, verts.shape
    return verts.astype(np.float32), faces


def make_body_model(seed: int = 0, flat_hand_mean: bool = True, num_expression_coeffs: int = 10) -> Dict[str, torch.Tensor]:
    """All constant tensors of the SMPL-X-shaped model (names follow smplx.SMPLX buffers)."""
    rng = np.random.default_rng(seed)
    v_template, faces = make_body_template()
    J = _rest_joints()
    V = v_template.shape[0]
    d = np.linalg.norm(v_template[:, None, :].astype(np.float64) - J[None], axis=-1)    # [V,55]
    # skinning weights: 4 nearest joints, softmax(-d/tau)
    near = np.argsort(d, axis=1)[:, :4]
    w4 = np.exp(-(np.take_along_axis(d, near, 1) - d.min(1, keepdims=True)) / 0.05)
    w4 /= w4.sum(1, keepdims=True)
    lbs_weights = np.zeros((V, NUM_JOINTS), dtype=np.float32)
    np.put_along_axis(lbs_weights, near, w4.astype(np.float32), 1)
    # joint regressor: normalised Gaussian bumps around each joint
    jr = np.exp(-0.5 * (d.T / 0.05) ** 2) + 1e-12
    jr /= jr.sum(1, keepdims=True)
    n_shape = NUM_BETAS + num_expression_coeffs
    shapedirs = rng.normal(0.0, 0.01, size=(V, 3, n_shape)).astype(np.float32)
    posedirs = rng.normal(0.0, 0.001, size=((NUM_JOINTS - 1) * 9, V * 3)).astype(np.float32)
    pose_mean = np.zeros(NUM_JOINTS * 3, dtype=np.float32)
    if not flat_hand_mean:
        pose_mean[25 * 3:] = rng.normal(0.0, 0.1, size=30 * 3).astype(np.float32)
    return {
        "v_template": torch.from_numpy(v_template),
        "faces": torch.from_numpy(faces),
        "shapedirs": torch.from_numpy(shapedirs),
        "posedirs": torch.from_numpy(posedirs),
        "J_regressor": torch.from_numpy(jr.astype(np.float32)),
        "parents": torch.tensor(SMPLX_PARENTS, dtype=torch.int64),
        "lbs_weights": torch.from_numpy(lbs_weights),
        "pose_mean": torch.from_numpy(pose_mean),
    }


# --------------------------------------------------------------------------------------
# Edge-midpoint subdivision (what pytorch3d.ops.SubdivideMeshes computes)
# --------------------------------------------------------------------------------------
def subdivide_once(num_verts: int, faces: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """One SubdivideMeshes step on the topology only.

    Returns ``edges [E,2]`` (unique, ``v0 < v1``, ascending by ``v0 * V + v1``: the order of
    ``Meshes.edges_packed()``) and the 4x faces.  New vertex ``num_verts + e`` is the midpoint
    of edge ``e``; faces become ``(v0,e01,e20) (v1,e12,e01) (v2,e20,e12) (e01,e12,e20)``.
    Reference call sites: renderer.py:236-243 (template) and :280-285 (posed vertices).
    """
    f = faces.astype(np.int64)
    e_all = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]], axis=0)
    e_all.sort(axis=1)
    key = e_all[:, 0] * num_verts + e_all[:, 1]
    ukey, inv = np.unique(key, return_inverse=True)
    edges = np.stack([ukey // num_verts, ukey % num_verts], axis=1)
    nf = f.shape[0]
    e01 = inv[:nf] + num_verts
    e12 = inv[nf:2 * nf] + num_verts
    e20 = inv[2 * nf:] + num_verts
    new_faces = np.concatenate([
        np.stack([f[:, 0], e01, e20], 1),
        np.stack([f[:, 1], e12, e01], 1),
        np.stack([f[:, 2], e20, e12], 1),
        np.stack([e01, e12, e20], 1),
    ], axis=0)
    return edges, new_faces


def subdivision_levels(num_verts: int, faces: np.ndarray, levels: int) -> List[np.ndarray]:
    """Edge lists of ``levels`` successive subdivisions (level l indexes level l-1 vertices)."""
    out, V, f = [], num_verts, faces
    for _ in range(levels):
        edges, f = subdivide_once(V, f)
        out.append(edges)
        V += edges.shape[0]
    return out


def densify_table(num_verts: int, faces: np.ndarray, levels: int) -> Tuple[np.ndarray, np.ndarray]:
    """Flatten ``levels`` subdivisions into one linear map onto the base vertices.

    Returns ``parents [V_dense, 2**levels] int32`` and ``weights [V_dense, 2**levels] float32``
    with ``dense[i] = sum_k weights[i,k] * base[parents[i,k]]``; unused slots have weight 0.
    Weights are exact dyadic fractions, so this equals the level-by-level ``(a+b)/2`` up to
    fp32 rounding of the sum.
    """
    P = 2 ** levels
    par = np.zeros((num_verts, P), dtype=np.int32)
    wgt = np.zeros((num_verts, P), dtype=np.float32)
    par[:, 0] = np.arange(num_verts)
    wgt[:, 0] = 1.0
    for edges in subdivision_levels(num_verts, faces, levels):
        a, b = edges[:, 0], edges[:, 1]
        half = P // 2
        # every existing row uses at most P/2 slots before the last level; pack a|b halves
        pa, wa = _compact(par[a], wgt[a], half)
        pb, wb = _compact(par[b], wgt[b], half)
        npar = np.concatenate([pa, pb], axis=1)
        nwgt = np.concatenate([wa, wb], axis=1) * np.float32(0.5)
        par = np.concatenate([par, npar], axis=0)
        wgt = np.concatenate([wgt, nwgt], axis=0)
    return par, wgt


def _compact(par: np.ndarray, wgt: np.ndarray, width: int) -> Tuple[np.ndarray, np.ndarray]:
    """Move the non-zero slots of each row to the front and keep ``width`` columns."""
    order = np.argsort(-(wgt > 0).astype(np.int8), axis=1, kind="stable")
    par = np.take_along_axis(par, order, 1)
    wgt = np.take_along_axis(wgt, order, 1)
    assert not (wgt[:, width:] > 0).any(), "subdivision parent count exceeds table width"
    return par[:, :width], wgt[:, :width]


# --------------------------------------------------------------------------------------
# Per-sequence inputs
# --------------------------------------------------------------------------------------
def make_triplane_tokens(frames: int, channels: int = 512, res: int = 64, seed: int = 0,
                         drift: float = 0.05, pin_memory: bool = False) -> torch.Tensor:
    """Triplane tokens ``[1, F, C, 3*res*res]`` fp32, channel-major as the upstream transformer
    emits them (transformers.py:1071, triplane_audio_net.py:261).  Frame f = base + drift*noise_f."""
    g = torch.Generator().manual_seed(seed)
    base = torch.randn(channels, 3 * res * res, generator=g)
    out = torch.empty(1, frames, channels, 3 * res * res, pin_memory=pin_memory)
    for f in range(frames):
        out[0, f] = base + drift * torch.randn(channels, 3 * res * res, generator=g)
    return out


def make_decoder_weights(channels: int = 512, seed: int = 0) -> Dict[str, torch.Tensor]:
    """Seeded weights for the five linear heads (renderer.py:51-55), nn.Linear-style U(+-1/sqrt(K)),
    with the reference's constant biases (renderer.py:57-71)."""
    g = torch.Generator().manual_seed(seed + 1)
    K = 3 * channels + 3
    bound = 1.0 / math.sqrt(K)
    heads = {"xyz": 3, "rotation": 4, "scaling": 3, "opacity": 1, "shs": 3}
    bias = {
        "xyz": torch.zeros(3),
        "rotation": torch.tensor([1.0, 0.0, 0.0, 0.0]),
        "scaling": torch.full((3,), -1.0),
        "opacity": torch.tensor([math.log(0.1 / 0.9)]),
        "shs": torch.zeros(3),
    }
    out = {}
    for name, k in heads.items():
        w = (torch.rand(k, K, generator=g) * 2.0 - 1.0) * bound
        if name == "xyz":
            w = w * 0.01      # keep decoded offsets at the millimetre scale
        out[f"{name}_layer.weight"] = w
        out[f"{name}_layer.bias"] = bias[name].clone()
    return out


def make_pose_sequence(batch: int, frames: int, seed: int = 0, num_expression_coeffs: int = 10,
                       transl=(0.0, 0.35, 2.4)) -> Dict[str, torch.Tensor]:
    """SMPL-X parameter dict with the reference's shapes (renderer.py:250-258,338):
    axis-angles ~ N(0, 0.2^2) smoothed over time, betas fixed per subject."""
    g = torch.Generator().manual_seed(seed + 2)

    def smooth(*shape):
        x = torch.randn(batch, frames + 8, *shape, generator=g) * 0.2
        k = torch.ones(9) / 9.0
        flat = x.reshape(batch, frames + 8, -1).permute(0, 2, 1)
        flat = torch.nn.functional.conv1d(flat.reshape(-1, 1, frames + 8), k.view(1, 1, 9)).reshape(batch, -1, frames)
        return (flat.permute(0, 2, 1) * 3.0).reshape(batch, frames, *shape).contiguous()

    betas = torch.randn(batch, 1, NUM_BETAS, generator=g).expand(-1, frames, -1).contiguous()
    return {
        "global_orient": smooth(3),
        "body_pose": smooth(21, 3),
        "betas": betas,
        "left_hand_pose": smooth(15, 3),
        "right_hand_pose": smooth(15, 3),
        "jaw_pose": smooth(3),
        "leye_pose": smooth(3),
        "reye_pose": smooth(3),
        "expression": torch.randn(batch, frames, num_expression_coeffs, generator=g) * 0.5,
        "transl": torch.tensor(transl).view(1, 1, 3).expand(batch, frames, 3).contiguous(),
    }


def make_cameras(batch: int, frames: int, height: int, width: int, orbit: bool = False,
                 distance: float = 2.4) -> Dict[str, torch.Tensor]:
    """``{'intrinsic' [B,T,3,3], 'extrinsic' [B,T,4,4]}``.  Default: E = I as the TED loader
    (dataset_speech_vid.py:314-317), fx=fy=1.1*W.  ``orbit``: cameras on a circle of radius
    ``distance`` around the origin looking at it (training views)."""
    K = torch.zeros(batch, frames, 3, 3)
    K[..., 0, 0] = 1.1 * width
    K[..., 1, 1] = 1.1 * width
    K[..., 0, 2] = width / 2.0
    K[..., 1, 2] = height / 2.0
    K[..., 2, 2] = 1.0
    E = torch.eye(4).repeat(batch, frames, 1, 1)
    if orbit:
        for t in range(frames):
            ang = 2.0 * math.pi * t / frames
            c, s = math.cos(ang), math.sin(ang)
            R = torch.tensor([[c, 0.0, -s], [0.0, 1.0, 0.0], [s, 0.0, c]])   # world->camera rotation
            E[:, t, :3, :3] = R
            E[:, t, :3, 3] = torch.tensor([0.0, 0.0, distance])
    return {"intrinsic": K, "extrinsic": E}
