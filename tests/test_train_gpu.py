"""Backward of stages 1+2 and of the whole Renderer on the GPU vs autograd through the oracle path
(torch CPU autograd for grid-sample / Linear / normalize / sigmoid exactly as the reference gets its gradients,
plus the C rasteriser's backward).  SURVEY.md 8c: gradients rel-err <= 1e-3 (of the gradient's max-abs)."""
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from audio_motion_avatar_b200 import synthetic as S
from oracle import ref_path as O

pytestmark = pytest.mark.gpu

NAMES = ("xyz", "rotation", "scaling", "opacity", "shs")


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


@pytest.mark.parametrize("C,R,N,F", [(64, 16, 1500, 2), (512, 64, 3000, 1)])
def test_decode_backward_matches_autograd(C, R, N, F):
    from audio_motion_avatar_b200.stages import decode_gaussians_autograd, decoder_matrix
    g = torch.Generator().manual_seed(C + N)
    tokens = torch.randn(F, C, 3 * R * R, generator=g)
    w = S.make_decoder_weights(C, seed=3)
    w["xyz_layer.weight"] = w["xyz_layer.weight"] * 30.0
    points = (torch.rand(F, N, 3, generator=g) * 2 - 1) * 1.5        # some outside radius 1.4: clamp path (zero gradient)
    transl = torch.randn(F, 3, generator=g)
    up = {k: torch.randn(F, N, d, generator=g) for k, d in (("xyz", 3), ("scale", 3), ("rot", 4), ("opacity", 1), ("color", 3))}
    # oracle: CPU autograd
    tok_r = tokens.clone().requires_grad_(True)
    pts_r = points.clone().requires_grad_(True)
    tr_r = transl.clone().requires_grad_(True)
    w_r = {k: v.clone().requires_grad_(True) for k, v in w.items()}
    tri = tok_r.view(F, C, 3, R, R).permute(0, 2, 1, 3, 4)
    ref = O.construct_gaussians(O.decode_heads(pts_r, O.sample_from_triplane_torch(tri, pts_r, 1.4), w_r), pts_r, tr_r)
    sum((ref[k] * up[k]).sum() for k in up).backward()
    # CUDA
    d = torch.device("cuda:0")
    tok = tokens.to(d).requires_grad_(True)
    pts = points.to(d).requires_grad_(True)
    tr = transl.to(d).requires_grad_(True)
    wd = {k: v.to(d).requires_grad_(True) for k, v in w.items()}
    W, b = decoder_matrix(wd)
    out = decode_gaussians_autograd(tok, pts, tr, W, b, C, R, 1.4)
    sum((out[k] * up[k].to(d)).sum() for k in up).backward()
    torch.cuda.synchronize()
    assert _rel(tok.grad.cpu(), tok_r.grad) <= 1e-3
    assert _rel(pts.grad.cpu(), pts_r.grad) <= 1e-3
    assert _rel(tr.grad.cpu(), tr_r.grad) <= 1e-3
    for n in NAMES:
        assert _rel(wd[f"{n}_layer.weight"].grad.cpu(), w_r[f"{n}_layer.weight"].grad) <= 1e-3, n
        assert _rel(wd[f"{n}_layer.bias"].grad.cpu(), w_r[f"{n}_layer.bias"].grad) <= 1e-3, n


def test_decode_backward_partial_grads():
    """Only some dict entries feed the loss (others get no gradient); tokens without requires_grad."""
    from audio_motion_avatar_b200.stages import decode_gaussians_autograd
    C, R, N = 64, 16, 500
    d = torch.device("cuda:0")
    g = torch.Generator().manual_seed(1)
    tokens = torch.randn(1, C, 3 * R * R, generator=g).to(d)
    W = torch.randn(14, 3 + 3 * C, generator=g).to(d).requires_grad_(True)
    b = torch.randn(14, generator=g).to(d).requires_grad_(True)
    pts = (torch.rand(1, N, 3, generator=g) - 0.5).to(d)
    out = decode_gaussians_autograd(tokens, pts, torch.zeros(1, 3, device=d), W, b, C, R, 1.4)
    out["opacity"].sum().backward()
    assert W.grad[10].abs().sum() > 0 and not W.grad[:10].any() and not W.grad[11:].any()
    assert torch.allclose(b.grad, torch.tensor([0.0] * 10 + [float(N)] + [0.0] * 3, device=d))


def _cfg(H, W, C, R):
    return SimpleNamespace(device="cuda:0", triplane_resolution=R, triplane_feature_dim=C, radius=1.4, smplx_model_path=None,
                           flat_hand_mean=True, num_expression_coeffs=10, upsample_triplane=False, densify_smplx_verts=False,
                           subdivide_steps=2, predict_smplx_params=False, no_point_refiner=True, image_size=(H, W), rgb=True,
                           sh_degree=3, num_gaussians=0)


def test_renderer_training_step_gradients():
    """Renderer.forward under autograd, then render_multi_view of the same Gaussians (the stage-1 training step,
    lightning_model_wrapper.py:51,132): gradients of an image loss w.r.t. the triplane tokens and every decoder
    parameter vs the oracle path."""
    from audio_motion_avatar_b200.renderer import Renderer, render_multi_view
    H, W, C, R, T = 64, 48, 64, 16, 2
    model = S.make_body_model(seed=0)
    w = S.make_decoder_weights(C, seed=0)
    tokens = S.make_triplane_tokens(T, C, R, seed=0)                     # [1,T,C,3RR]
    params = S.make_pose_sequence(1, T, seed=0)
    cams = S.make_cameras(1, T, H, W)
    gimg = torch.randn(1, T, H, W, 3, generator=torch.Generator().manual_seed(3))
    # oracle
    tok_r = tokens.clone().requires_grad_(True)
    w_r = {k: v.clone().requires_grad_(True) for k, v in w.items()}
    flat = {k: v.reshape(T, *v.shape[2:]) for k, v in params.items()}
    ref, ref_g = O.reference_frames(model, w_r, tok_r[0], flat, cams["intrinsic"][0], cams["extrinsic"][0], H, W, C, R,
                                    differentiable=True)
    # second term: frame 0's Gaussians seen from both cameras (render_multi_view)
    mv = torch.stack([O.render_one_autograd({k: v[0] for k, v in ref_g.items()}, cams["intrinsic"][0, t], cams["extrinsic"][0, t],
                                            H, W).permute(1, 2, 0) for t in range(T)])
    ((ref * gimg[0]).sum() + (mv * gimg[0]).sum() * 0.5).backward()
    # CUDA
    d = torch.device("cuda:0")
    r = Renderer(_cfg(H, W, C, R), body_model=model)
    r.gaussian_decoder.load_state_dict(w)
    tok = tokens.to(d).requires_grad_(True)
    cam_d = {k: v.to(d) for k, v in cams.items()}
    img, g = r(tok, cam_d, None, {k: v.to(d) for k, v in params.items()})
    assert img.requires_grad and img.shape == (1, T, H, W, 3)
    g0 = {k: v[:1] for k, v in g.items()}
    img2 = render_multi_view(g0, cam_d["intrinsic"], cam_d["extrinsic"], r.cfg)
    ((img * gimg.to(d)).sum() + (img2 * gimg.to(d)).sum() * 0.5).backward()
    torch.cuda.synchronize()
    assert (img.detach().cpu()[0] - ref.detach()).abs().max().item() <= 1.0 / 255.0 + 1e-3
    e_tok = _rel(tok.grad.cpu(), tok_r.grad)
    assert e_tok <= 2e-3, e_tok
    for n in NAMES:
        layer = getattr(r.gaussian_decoder, f"{n}_layer")
        ew = _rel(layer.weight.grad.cpu(), w_r[f"{n}_layer.weight"].grad)
        eb = _rel(layer.bias.grad.cpu(), w_r[f"{n}_layer.bias"].grad)
        assert ew <= 2e-3 and eb <= 2e-3, (n, ew, eb)


POSE_KEYS = ("global_orient", "body_pose", "jaw_pose", "leye_pose", "reye_pose", "left_hand_pose", "right_hand_pose",
             "betas", "expression")


def test_lbs_and_densify_backward_match_autograd():
    """d(verts)/d(pose, betas, expression) through amav_lbs_backward and the densify scatter vs torch autograd of the
    oracle's smplx restatement (19 frames: a ragged 16-frame group)."""
    from audio_motion_avatar_b200.stages import BodyModel, densify_sites, lbs_vertices
    model = S.make_body_model(seed=0)
    body = BodyModel(model, "cuda:0")
    Fr = 19
    params = S.make_pose_sequence(1, Fr, seed=4)
    flat = {k: v.reshape(Fr, *v.shape[2:]) for k, v in params.items()}
    V = model["v_template"].shape[0]
    par, wgt = S.densify_table(V, model["faces"].numpy(), 1)
    idx = torch.randperm(par.shape[0], generator=torch.Generator().manual_seed(1))[:20000]
    par_t, wgt_t = torch.from_numpy(par)[idx], torch.from_numpy(wgt)[idx]
    up = torch.randn(Fr, 20000, 3, generator=torch.Generator().manual_seed(2))
    # oracle
    leaves = {k: flat[k].clone().requires_grad_(True) for k in POSE_KEYS}
    verts = O.lbs_vertices({**flat, **leaves}, model)
    sites = (verts[:, par_t.long()] * wgt_t[None, :, :, None]).sum(2)
    (sites * up).sum().backward()
    # CUDA
    d = torch.device("cuda:0")
    dl = {k: flat[k].clone().to(d).requires_grad_(True) for k in POSE_KEYS}
    full, shape = body.full_pose(dl)
    out = densify_sites(lbs_vertices(body, full, shape), par_t.to(d).contiguous(), wgt_t.to(d).contiguous())
    assert (out.detach().cpu() - sites.detach()).abs().max().item() < 5e-5
    (out * up.to(d)).sum().backward()
    torch.cuda.synchronize()
    for k in POSE_KEYS:
        e = _rel(dl[k].grad.cpu(), leaves[k].grad)
        assert e <= 1e-3, (k, e)


def test_stage2_pose_gradients_end_to_end():
    """Image loss -> splat -> decode (position derivative of the triplane query) -> LBS -> pose parameters and transl:
    the gradient path of stage-2 training, where the SMPL-X parameters are predicted (triplane_audio_net.py:269)."""
    from audio_motion_avatar_b200.renderer import Renderer
    H, W, C, R, T = 64, 48, 64, 16, 2
    model = S.make_body_model(seed=0)
    w = S.make_decoder_weights(C, seed=0)
    tokens = S.make_triplane_tokens(T, C, R, seed=0)
    params = S.make_pose_sequence(1, T, seed=0)
    cams = S.make_cameras(1, T, H, W)
    gimg = torch.randn(1, T, H, W, 3, generator=torch.Generator().manual_seed(3))
    keys = POSE_KEYS + ("transl",)
    flat = {k: v.reshape(T, *v.shape[2:]) for k, v in params.items()}
    leaves = {k: flat[k].clone().requires_grad_(True) for k in keys}
    ref, _ = O.reference_frames(model, w, tokens[0], {**flat, **leaves}, cams["intrinsic"][0], cams["extrinsic"][0], H, W, C, R,
                                differentiable=True)
    (ref * gimg[0]).sum().backward()
    d = torch.device("cuda:0")
    r = Renderer(_cfg(H, W, C, R), body_model=model)
    r.gaussian_decoder.load_state_dict(w)
    dl = {k: params[k].clone().to(d).requires_grad_(True) for k in keys}
    img, _ = r(tokens.to(d), {k: v.to(d) for k, v in cams.items()}, None, {**{k: v.to(d) for k, v in params.items()}, **dl})
    (img * gimg.to(d)).sum().backward()
    torch.cuda.synchronize()
    for k in keys:
        e = _rel(dl[k].grad.cpu().reshape(leaves[k].grad.shape), leaves[k].grad)
        assert e <= 5e-3, (k, e)
