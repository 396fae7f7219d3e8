"""Invariant checks that pin the un-pinnable restatements (LBS, subdivision, rasteriser): the reference
ships no fixtures for them and their packages are absent offline."""
import math

import numpy as np
import pytest
import torch

from audio_motion_avatar_b200 import synthetic as S
from oracle import ref_path as O


@pytest.fixture(scope="module")
def model():
    return S.make_body_model(seed=0)


def _zero_params(Fr):
    z = lambda *s: torch.zeros(Fr, *s)
    return {"global_orient": z(3), "body_pose": z(21, 3), "betas": z(10), "left_hand_pose": z(15, 3),
            "right_hand_pose": z(15, 3), "jaw_pose": z(3), "leye_pose": z(3), "reye_pose": z(3), "expression": z(10)}


def test_lbs_rest_pose_is_template(model):
    v = O.lbs_vertices(_zero_params(2), model)
    assert (v - model["v_template"][None]).abs().max().item() < 2e-6


def test_lbs_global_orient_is_rigid_about_root(model):
    p = _zero_params(1)
    p["global_orient"][0] = torch.tensor([0.3, -0.7, 0.2])
    p["betas"][0] = torch.randn(10, generator=torch.Generator().manual_seed(3))
    v, A, J = O.lbs_vertices(p, model, return_transforms=True)
    q = _zero_params(1)
    q["betas"] = p["betas"]
    v0 = O.lbs_vertices(q, model)
    R = O.batch_rodrigues(p["global_orient"])[0]
    expect = (v0[0] - J[0, 0]) @ R.T + J[0, 0]
    assert (v[0] - expect).abs().max().item() < 5e-6


def test_subdivision_table_equals_levelwise(model):
    faces = model["faces"].numpy()
    V = model["v_template"].shape[0]
    levels = S.subdivision_levels(V, faces, 2)
    base = model["v_template"][None] + 0.01 * torch.randn(1, V, 3, generator=torch.Generator().manual_seed(5))
    dense = O.subdivide_vertices(base, levels)
    par, wgt = S.densify_table(V, faces, 2)
    flat = (base[0][torch.from_numpy(par).long()] * torch.from_numpy(wgt)[..., None]).sum(1)
    assert dense.shape[1] == par.shape[0] == 167420
    assert (dense[0] - flat).abs().max().item() < 1e-6


def test_exp_polynomial_close_to_libm():
    lib = O._lib()
    xs = np.concatenate([np.linspace(-30, 0, 20001), -np.logspace(-8, 1.5, 2000)]).astype(np.float32)
    got = np.array([lib.amav_oracle_exp(float(x)) for x in xs], dtype=np.float64)
    ref = np.exp(xs.astype(np.float64))
    assert np.max(np.abs(got - ref) / ref) < 2.5e-7


def _single_gaussian(W=64, H=48):
    means = torch.tensor([[0.05, -0.02, 2.0]])
    scales = torch.tensor([[0.03, 0.03, 0.03]])
    rots = torch.tensor([[1.0, 0, 0, 0]])
    opac = torch.tensor([[0.8]])
    col = torch.tensor([[0.2, 0.5, 0.9]])
    K = torch.tensor([[70.0, 0, W / 2], [0, 70.0, H / 2], [0, 0, 1]])
    view, full, tfx, tfy, _ = O.camera_setup(K, torch.eye(4), H, W)
    return means, scales, rots, opac, col, view, full, tfx, tfy, K


def test_single_gaussian_analytic_splat():
    W, H = 64, 48
    means, scales, rots, opac, col, view, full, tfx, tfy, K = _single_gaussian(W, H)
    pre, binned, out = O.rasterize(means, scales, rots, opac, col, view, full, tfx, tfy, (1, 1, 1), W, H)
    fx = float(K[0, 0])
    z = 2.0
    # isotropic: cov2D = (fx*s/z)^2 (+ small perspective terms from the off-axis Jacobian) + 0.3
    var = (fx * 0.03 / z) ** 2 + 0.3
    cx = fx * 0.05 / z + W / 2 - 0.5
    cy = fx * -0.02 / z + H / 2 - 0.5
    assert abs(pre["xy"][0, 0] - cx) < 1e-3 and abs(pre["xy"][0, 1] - cy) < 1e-3
    ys, xs = np.mgrid[0:H, 0:W]
    d2 = (xs - cx) ** 2 + (ys - cy) ** 2
    alpha = np.minimum(0.99, 0.8 * np.exp(-0.5 * d2 / var))
    inside = np.zeros((H, W), bool)
    r = pre["rect"][0]
    inside[r[1] * 16:r[3] * 16, r[0] * 16:r[2] * 16] = True
    alpha = np.where((alpha >= 1 / 255) & inside, alpha, 0.0)
    for ch in range(3):
        expect = float(col[0, ch]) * alpha + (1 - alpha)
        assert np.abs(out["color"][ch] - expect).max() < 2e-3
    assert np.abs(out["final_T"] - (1 - alpha)).max() < 2e-3


def test_weights_sum_and_sorted_lists():
    g = torch.Generator().manual_seed(11)
    N, W, H = 400, 80, 64
    means = torch.cat([torch.rand(N, 2, generator=g) * 1.2 - 0.6, 1.5 + torch.rand(N, 1, generator=g)], 1)
    means[:5, 2] = 0.1                                   # behind the near plane: culled
    scales = torch.rand(N, 3, generator=g) * 0.05 + 0.005
    rots = torch.nn.functional.normalize(torch.randn(N, 4, generator=g), dim=-1)
    opac = torch.rand(N, 1, generator=g)
    col = torch.rand(N, 3, generator=g)
    K = torch.tensor([[90.0, 0, W / 2], [0, 90.0, H / 2], [0, 0, 1]])
    view, full, tfx, tfy, _ = O.camera_setup(K, torch.eye(4), H, W)
    pre, binned, out = O.rasterize(means, scales, rots, opac, col, view, full, tfx, tfy, (0, 0, 0), W, H)
    assert (pre["radii"][:5] == 0).all() and binned["D"] == int(pre["tiles_touched"].sum())
    keys = binned["keys"]
    assert (keys[1:] >= keys[:-1]).all()
    # ties (same tile, same depth bits) keep ascending Gaussian index
    same = keys[1:] == keys[:-1]
    assert (binned["point_list"][1:][same] > binned["point_list"][:-1][same]).all()
    # sorted multiset equals emitted multiset
    assert np.array_equal(np.sort(binned["keys_unsorted"]), keys)
    # with white colours and black background, colour == 1 - final_T per pixel
    _, _, out1 = O.rasterize(means, scales, rots, opac, torch.ones(N, 3), view, full, tfx, tfy, (0, 0, 0), W, H)
    assert np.abs(out1["color"][0] - (1 - out1["final_T"])).max() < 1e-5
    r = binned["ranges"]
    assert ((r[:, 1] - r[:, 0]) >= out1["n_contrib"].reshape(H // 16, 16, W // 16, 16).max((1, 3)).reshape(-1)).all()


# ---- 8f-3 / 8f-4: restatements of pytorch3d's rotation conversions and of the demo's uint8 sink ----
def test_rotation_6d_and_axis_angle_round_trip():
    """rotation_6d_to_matrix gives proper rotations whose first two rows span the inputs; matrix_to_axis_angle is
    the inverse of smplx's Rodrigues (what the reference relies on between smplx_decoder.py:106 and lbs)."""
    g = torch.Generator().manual_seed(0)
    d6 = torch.randn(2000, 6, generator=g)
    R = O.rotation_6d_to_matrix(d6)
    assert (R @ R.transpose(-1, -2) - torch.eye(3)).abs().max().item() < 1e-5
    assert (torch.det(R) - 1).abs().max().item() < 1e-5
    assert (torch.nn.functional.normalize(d6[:, :3], dim=-1) - R[:, 0]).abs().max().item() < 1e-6
    back = O.batch_rodrigues(O.matrix_to_axis_angle(R))
    assert (back - R).abs().max().item() < 1e-5
    aa = 0.8 * torch.randn(500, 3, generator=g)                       # angles below pi: the map is one-to-one
    aa = aa * (aa.norm(dim=1, keepdim=True).clamp(max=3.0) / aa.norm(dim=1, keepdim=True))
    assert (O.matrix_to_axis_angle(O.batch_rodrigues(aa)) - aa).abs().max().item() < 2e-4
    assert torch.allclose(O.matrix_to_axis_angle(torch.eye(3)[None]), torch.zeros(1, 3), atol=1e-7)


def test_frames_to_u8_known_answers():
    x = np.array([[[[0.0, 1.0, 0.5], [1.0 / 255, 254.999 / 255, 0.25]]]], dtype=np.float32)
    assert O.frames_to_u8(x).tolist() == [[[[0, 255, 127], [1, 254, 63]]]]
    assert O.frames_to_u8(x, bgr=True).tolist() == [[[[127, 255, 0], [63, 254, 1]]]]
