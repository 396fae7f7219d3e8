"""Stages 1-3 on the GPU vs the oracle (oracle/ref_path.py) on identical seeded inputs.
Floating-point stages: tolerance from north_star, max-abs <= 1e-3 on attributes (measured ~1e-5)."""
import numpy as np
import pytest
import torch

from audio_motion_avatar_b200 import synthetic as S
from oracle import ref_path as O

pytestmark = pytest.mark.gpu
TOL = 1e-3          # north_star: attributes max-abs <= 1e-3
TIGHT = 5e-5        # what fp32 reassociation actually costs; guards against silent regressions


@pytest.fixture(scope="module")
def world():
    model = S.make_body_model(seed=0)
    from audio_motion_avatar_b200.stages import BodyModel
    return model, BodyModel(model, "cuda:0")


def _flat(params):
    Fr = params["global_orient"].shape[0] * params["global_orient"].shape[1]
    return {k: v.reshape(Fr, *v.shape[2:]) for k, v in params.items()}


def test_lbs_matches_oracle(world):
    from audio_motion_avatar_b200.stages import lbs_forward
    model, body = world
    params = _flat(S.make_pose_sequence(1, 19, seed=4))          # 19: exercises a ragged 16-frame group
    ref, A_ref, J_ref = O.lbs_vertices(params, model, return_transforms=True)
    full, shape = body.full_pose({k: v.cuda() for k, v in params.items()})
    verts, aux = lbs_forward(body, full, shape, return_aux=True)
    torch.cuda.synchronize()
    assert (aux["joints"].cpu() - J_ref).abs().max().item() < TIGHT
    assert (aux["A"].cpu().view(19, 55, 3, 4) - A_ref[:, :, :3, :]).abs().max().item() < TIGHT
    err = (verts.cpu() - ref).abs().max().item()
    assert err < TIGHT, err


def test_lbs_rest_pose_and_pose_mean(world):
    from audio_motion_avatar_b200.stages import BodyModel, lbs_forward
    model, body = world
    z = {k: torch.zeros(2, *s) for k, s in dict(global_orient=(3,), body_pose=(21, 3), betas=(10,), left_hand_pose=(15, 3),
                                                 right_hand_pose=(15, 3), jaw_pose=(3,), leye_pose=(3,), reye_pose=(3,),
                                                 expression=(10,)).items()}
    full, shape = body.full_pose({k: v.cuda() for k, v in z.items()})
    v = lbs_forward(body, full, shape)
    assert (v.cpu() - model["v_template"][None]).abs().max().item() < 2e-6
    m2 = S.make_body_model(seed=0, flat_hand_mean=False)
    b2 = BodyModel(m2, "cuda:0")
    full2, shape2 = b2.full_pose({k: v.cuda() for k, v in z.items()})
    ref = O.lbs_vertices(z, m2)
    assert (lbs_forward(b2, full2, shape2).cpu() - ref).abs().max().item() < TIGHT


def test_densify_matches_levelwise_subdivision(world):
    from audio_motion_avatar_b200.stages import densify_gather
    model, body = world
    V = model["v_template"].shape[0]
    faces = model["faces"].numpy()
    levels = S.subdivision_levels(V, faces, 2)
    base = model["v_template"][None] + 0.01 * torch.randn(2, V, 3, generator=torch.Generator().manual_seed(5))
    dense = O.subdivide_vertices(base, levels)
    par, wgt = S.densify_table(V, faces, 2)
    idx = torch.randperm(par.shape[0], generator=torch.Generator().manual_seed(1))[:30000]
    sites = densify_gather(base.cuda(), torch.from_numpy(par)[idx].cuda().contiguous(), torch.from_numpy(wgt)[idx].cuda().contiguous())
    assert (sites.cpu() - dense[:, idx]).abs().max().item() < 1e-6
    # P = 1 plain gather
    one = densify_gather(base.cuda(), idx[:100].to(torch.int32).clamp(max=V - 1)[:, None].cuda().contiguous(),
                         torch.ones(100, 1).cuda())
    assert torch.equal(one.cpu(), base[:, idx[:100].clamp(max=V - 1)])


@pytest.mark.parametrize("C,R,N,F", [(64, 16, 1500, 2), (512, 64, 4000, 1)])
def test_project_sample_decode_matches_oracle(C, R, N, F):
    from audio_motion_avatar_b200.stages import decoder_matrix, sample_decode, triplane_project
    g = torch.Generator().manual_seed(C + N)
    tokens = torch.randn(F, C, 3 * R * R, generator=g)
    w = S.make_decoder_weights(C, seed=3)
    w["xyz_layer.weight"] = w["xyz_layer.weight"] * 100.0            # full-size offsets: a harder test
    points = (torch.rand(F, N, 3, generator=g) * 2 - 1) * 1.6       # some outside radius 1.4: clamp path
    points[:, :8] = torch.tensor([[1.4, -1.4, 0.0], [-1.4, 1.4, 1.4], [1.39999, 0, 0], [0, 0, 0], [5, 5, 5],
                                  [-5, 0.3, 0.2], [1.378, 1.378, -1.378], [-1.3999, -1.3999, -1.3999]])
    transl = torch.randn(F, 3, generator=g)
    tri = tokens.view(F, C, 3, R, R).permute(0, 2, 1, 3, 4).contiguous()                 # renderer.py:85-91
    feats = O.sample_from_triplane(tri, points, 1.4)
    ref = O.construct_gaussians(O.decode_heads(points, feats, w), points, transl)
    W, b = decoder_matrix(w)
    proj = triplane_project(tokens.cuda(), W.cuda(), C, R)
    # projected planes themselves: proj[f,p,texel,slot] = W[row(slot), 3+p*C:3+(p+1)*C] . tokens[f,:,p*R*R+texel]
    rows = [3, 4, 5, 6, 0, 1, 2, 10, 7, 8, 9, -1, 11, 12, 13, -1]
    pr = proj.cpu()
    for p in range(3):
        want = torch.einsum("oc,fct->fto", W[:, 3 + p * C:3 + (p + 1) * C].double(), tokens[:, :, p * R * R:(p + 1) * R * R].double())
        for s, r in enumerate(rows):
            if r < 0:
                assert torch.all(pr[:, p, :, s] == 0)
            else:
                assert (pr[:, p, :, s].double() - want[:, :, r]).abs().max().item() < TIGHT
    out = sample_decode(proj, points.cuda(), transl.cuda(), W.cuda(), b.cuda(), C, R, 1.4)
    torch.cuda.synchronize()
    assert out["shs"] is out["color"]
    for k in ("xyz", "scale", "rot", "opacity", "color"):
        err = (out[k].cpu() - ref[k]).abs().max().item()
        assert out[k].shape == ref[k].shape and err < TIGHT, (k, err)
        assert err <= TOL


def test_project_padded_frame_stride_and_empty():
    from audio_motion_avatar_b200.stages import triplane_project, sample_decode
    C, R = 64, 16
    g = torch.Generator().manual_seed(0)
    big = torch.randn(3, C + 5, 3 * R * R, generator=g).cuda()
    W = torch.randn(14, 3 + 3 * C, generator=g).cuda()
    a = triplane_project(big[:, :C], W, C, R)                         # frame stride (C+5)*3*R*R
    b = triplane_project(big[:, :C].contiguous(), W, C, R)
    assert torch.equal(a, b)
    e = triplane_project(torch.zeros(0, C, 3 * R * R).cuda(), W, C, R)
    assert e.shape[0] == 0
    out = sample_decode(a, torch.zeros(3, 0, 3).cuda(), torch.zeros(3, 3).cuda(), W, torch.zeros(14).cuda(), C, R, 1.4)
    assert out["xyz"].shape == (3, 0, 3)


@pytest.mark.parametrize("C,R,F", [(512, 64, 3), (64, 16, 2), (128, 32, 1)])
def test_tensor_core_projection_matches_float64_and_ffma(C, R, F):
    """amav_triplane_project_tc (tcgen05.mma kind::tf32, 3xTF32 split, TMEM accumulator) vs float64 W_feat . tokens and
    vs the FP32-FMA kernel: fp32-level accuracy (<= 2e-5 here; single-pass TF32 would be ~1e-3)."""
    from audio_motion_avatar_b200 import stages, synthetic as S
    g = torch.Generator().manual_seed(C + R)
    tokens = torch.randn(F, C, 3 * R * R, generator=g)
    W = torch.cat([v for k, v in S.make_decoder_weights(C, seed=3).items() if k.endswith("weight")], 0)     # [14, 3+3C]
    names = ("xyz", "rotation", "scaling", "opacity", "shs")
    wd = S.make_decoder_weights(C, seed=3)
    W = torch.cat([wd[f"{n}_layer.weight"] for n in names], 0)
    tok_d, W_d = tokens.cuda(), W.cuda()
    a = stages.triplane_project(tok_d, W_d, C, R, impl="tc")
    b = stages.triplane_project(tok_d, W_d, C, R, impl="ffma")
    torch.cuda.synchronize()
    slot_row = [3, 4, 5, 6, 0, 1, 2, 10, 7, 8, 9, -1, 11, 12, 13, -1]
    ref = torch.zeros(F, 3, R * R, 16, dtype=torch.float64)
    for p in range(3):
        Wp = W[:, 3 + p * C:3 + (p + 1) * C].double()                      # [14, C]
        T = tokens[:, :, p * R * R:(p + 1) * R * R].double()              # [F, C, R2]
        out = torch.einsum("oc,fct->fto", Wp, T)                          # [F, R2, 14]
        for s_, row in enumerate(slot_row):
            if row >= 0:
                ref[:, p, :, s_] = out[:, :, row]
    assert a.shape == b.shape == (F, 3, R * R, 16)
    err_tc, err_ffma = (a.cpu().double() - ref).abs().max().item(), (b.cpu().double() - ref).abs().max().item()
    assert err_tc <= 2e-5, err_tc
    assert err_ffma <= 2e-6, err_ffma           # (the tensor core accumulates with truncation: ~1e-5, the FMA kernel ~3e-7)
    assert float(a[..., 11].abs().max()) == 0.0 and float(a[..., 15].abs().max()) == 0.0


def test_locality_order_does_not_change_results():
    """amav_sample_decode / amav_densify_gather with a processing order (Morton order of the rest-pose sites) write the
    same bits at the same indices as without one."""
    from audio_motion_avatar_b200 import stages, synthetic as S
    g = torch.Generator().manual_seed(7)
    F, N, C, R, V = 3, 5000, 64, 16, 700
    verts = torch.randn(F, V, 3, generator=g).cuda() * 0.4
    par = torch.randint(0, V, (N, 4), generator=g, dtype=torch.int32).cuda()
    wgt = torch.rand(N, 4, generator=g).cuda()
    order = stages.locality_order(stages.densify_gather(verts[:1], par, wgt)[0], 1.4)
    assert sorted(order.cpu().tolist()) == list(range(N))
    a, b = stages.densify_gather(verts, par, wgt), stages.densify_gather(verts, par, wgt, order)
    assert torch.equal(a, b)
    wd = S.make_decoder_weights(C, seed=3)
    W = torch.cat([wd[f"{n}_layer.weight"] for n in ("xyz", "rotation", "scaling", "opacity", "shs")], 0).cuda()
    bias = torch.cat([wd[f"{n}_layer.bias"] for n in ("xyz", "rotation", "scaling", "opacity", "shs")], 0).cuda()
    proj = stages.triplane_project(torch.randn(F, C, 3 * R * R, generator=g).cuda(), W, C, R)
    tr = torch.randn(F, 3, generator=g).cuda()
    x, y = stages.sample_decode(proj, a, tr, W, bias, C, R, 1.4), stages.sample_decode(proj, a, tr, W, bias, C, R, 1.4, order=order)
    assert all(torch.equal(x[k], y[k]) for k in ("xyz", "scale", "rot", "opacity", "color"))
