"""SURVEY 8f-3 (6D pose straight into LBS) and 8f-4 (video sink) on the GPU vs the oracle's restatement of the
reference route (smplx_decoder.py:106-127 -> renderer.py:261-274; main2.py:351-353)."""
import numpy as np
import pytest
import torch

from audio_motion_avatar_b200 import synthetic as S
from oracle import ref_path as O
from tests.test_renderer_gpu import make_cfg

pytestmark = pytest.mark.gpu
TIGHT = 5e-5


def _pose_6d(Fr, seed):
    g = torch.Generator().manual_seed(seed)
    eye6 = torch.tensor([1.0, 0, 0, 0, 1.0, 0])
    return eye6[None, None] + 0.35 * torch.randn(Fr, 55, 6, generator=g)


def test_lbs_6d_matches_reference_route():
    """6D -> Gram-Schmidt -> LBS on the device == 6D -> matrix -> axis-angle -> Rodrigues -> LBS (oracle)."""
    from audio_motion_avatar_b200.stages import BodyModel, lbs_forward_6d, pose_6d_from_decoder_heads
    model = S.make_body_model(seed=0)
    body = BodyModel(model, "cuda:0")
    Fr = 19
    d6 = _pose_6d(Fr, 3)
    g = torch.Generator().manual_seed(4)
    betas, expr = torch.randn(Fr, 10, generator=g), 0.5 * torch.randn(Fr, 10, generator=g)
    ref = O.lbs_vertices(O.params_from_pose_6d(d6, betas, expr), model)
    # the decoder's six heads, packed the way SMPLXDecoder splits them
    packed = pose_6d_from_decoder_heads(d6[:, 0], d6[:, 1:22].reshape(Fr, 126), d6[:, 25:55].reshape(Fr, 180),
                                        d6[:, 22], d6[:, 23], d6[:, 24])
    assert torch.equal(packed, d6)
    verts = lbs_forward_6d(body, packed.cuda(), torch.cat([betas, expr], 1).cuda())
    err = (verts.cpu() - ref).abs().max().item()
    assert err < TIGHT, err


def test_lbs_6d_rejects_pose_mean():
    from audio_motion_avatar_b200.stages import BodyModel, lbs_forward_6d
    d6 = _pose_6d(2, 1).cuda()
    shape = torch.zeros(2, 20, device="cuda")
    with pytest.raises(ValueError):
        lbs_forward_6d(BodyModel(S.make_body_model(seed=0, flat_hand_mean=False), "cuda:0"), d6, shape)


def test_lbs_6d_backward_matches_autograd_of_reference_route():
    """dL/d(6D heads), dL/d(betas, expression) from amav_lbs_backward_6d vs torch CPU autograd through the oracle's
    6D -> matrix -> axis-angle -> Rodrigues -> LBS (what the reference's decoder receives in stage-2 training)."""
    from audio_motion_avatar_b200.stages import BodyModel, lbs_vertices_6d
    model = S.make_body_model(seed=0)
    body = BodyModel(model, "cuda:0")
    Fr = 5
    d6 = _pose_6d(Fr, 11)
    g = torch.Generator().manual_seed(12)
    shape = torch.cat([torch.randn(Fr, 10, generator=g), 0.5 * torch.randn(Fr, 10, generator=g)], 1)
    wgt = torch.randn(Fr, model["v_template"].shape[0], 3, generator=g)
    d6_ref, shape_ref = d6.clone().requires_grad_(True), shape.clone().requires_grad_(True)
    (O.lbs_vertices(O.params_from_pose_6d(d6_ref, shape_ref[:, :10], shape_ref[:, 10:]), model) * wgt).sum().backward()
    d6_dev, shape_dev = d6.cuda().requires_grad_(True), shape.cuda().requires_grad_(True)
    (lbs_vertices_6d(body, d6_dev, shape_dev) * wgt.cuda()).sum().backward()
    for got, ref, name in ((d6_dev.grad, d6_ref.grad, "pose_6d"), (shape_dev.grad, shape_ref.grad, "shape")):
        err = (got.cpu() - ref).abs().max().item() / ref.abs().max().item()
        assert err < 1e-3, (name, err)


def test_renderer_accepts_pose_6d():
    from audio_motion_avatar_b200.renderer import Renderer
    H, W, C, R = 64, 48, 64, 16
    model = S.make_body_model(seed=0)
    r = Renderer(make_cfg(H, W, C, R), body_model=model)
    d6 = _pose_6d(3, 7)
    betas, expr = torch.zeros(3, 10), torch.zeros(3, 10)
    pts = r.get_smpl_vertices({"pose_6d": d6.reshape(1, 3, 55, 6).cuda(), "betas": betas.cuda(), "expression": expr.cuda()})
    ref = O.lbs_vertices(O.params_from_pose_6d(d6, betas, expr), model)
    assert (pts.cpu() - ref).abs().max().item() < TIGHT


@pytest.mark.parametrize("H,W", [(64, 48), (70, 45), (1, 3)])     # 70x45 and 1x3: H*W % 4 != 0, scalar instantiation
@pytest.mark.parametrize("bgr", [False, True])
def test_frames_to_u8_bit_exact(H, W, bgr):
    from audio_motion_avatar_b200.stages import frames_to_u8
    g = torch.Generator().manual_seed(H * W)
    fr = torch.rand(5, H, W, 3, generator=g)
    fr.view(-1)[:6] = torch.tensor([0.0, 1.0, 0.5, 254.999 / 255, 1.0 / 255, 0.99999994])
    got = frames_to_u8(fr.cuda(), bgr=bgr).cpu().numpy()
    assert np.array_equal(got, O.frames_to_u8(fr.numpy(), bgr=bgr))


def test_frames_to_u8_interleaves_two_sequences():
    """combined_seq of main2.py:328-331: even_seq[i], odd_seq[i], even_seq[i+1], ..."""
    from audio_motion_avatar_b200.stages import frames_to_u8
    g = torch.Generator().manual_seed(0)
    a, b = torch.rand(3, 32, 40, 3, generator=g), torch.rand(3, 32, 40, 3, generator=g)
    out = torch.zeros(6, 32, 40, 3, dtype=torch.uint8, device="cuda")
    frames_to_u8(a.cuda(), out=out, out_first=0, out_step=2)
    frames_to_u8(b.cuda(), out=out, out_first=1, out_step=2)
    want = np.empty((6, 32, 40, 3), np.uint8)
    want[0::2], want[1::2] = O.frames_to_u8(a.numpy()), O.frames_to_u8(b.numpy())
    assert np.array_equal(out.cpu().numpy(), want)
    with pytest.raises(ValueError):
        frames_to_u8(a.cuda(), out=out, out_first=2, out_step=2)


def test_render_sequence_uint8_equals_sink_of_float_frames():
    from audio_motion_avatar_b200.renderer import Renderer
    H, W, C, R, T = 96, 80, 64, 16, 5
    model = S.make_body_model(seed=0)
    r = Renderer(make_cfg(H, W, C, R), body_model=model)
    r.gaussian_decoder.load_state_dict(S.make_decoder_weights(C, seed=1))
    tokens = S.make_triplane_tokens(T, C, R, seed=2).reshape(1, T, C, 3 * R * R).pin_memory()
    params, cams = S.make_pose_sequence(1, T, seed=2), S.make_cameras(1, T, H, W)
    with torch.no_grad():
        f32 = r.render_sequence(tokens, cams, params, chunk=2).clone()
        u8 = r.render_sequence(tokens, cams, params, chunk=2, out_dtype=torch.uint8, bgr=True)
        dev = r(tokens.cuda(), {k: v.cuda() for k, v in cams.items()}, None, {k: v.cuda() for k, v in params.items()})[0]
    assert torch.equal(f32, dev.cpu())                             # the chunked host path renders the same frames
    assert u8.dtype == torch.uint8 and u8.shape == (1, T, H, W, 3)
    assert np.array_equal(u8.numpy(), O.frames_to_u8(f32.numpy(), bgr=True))
    assert f32.min() < 0.5                                         # not a blank canvas


# ---- 8f-2: the projection from the producer's token-major layout ----
@pytest.mark.parametrize("C,R,F", [(64, 16, 3), (512, 64, 2), (192, 16, 1)])
def test_token_major_projection_matches_channel_major_and_oracle(C, R, F):
    """amav_triplane_project_tm on tokens^T == amav_triplane_project on tokens (fp32 summation order only), and both
    equal W_feat . tokens computed in float64."""
    from audio_motion_avatar_b200.stages import decoder_matrix, triplane_project
    g = torch.Generator().manual_seed(C + R)
    tokens = torch.randn(F, C, 3 * R * R, generator=g)
    Wd, _ = decoder_matrix(S.make_decoder_weights(C, seed=3))
    cm = triplane_project(tokens.cuda(), Wd.cuda(), C, R)
    tm_in = tokens.transpose(1, 2).contiguous().cuda()                       # [F, 3*R*R, C]
    tm = triplane_project(tm_in, Wd.cuda(), C, R, token_major=True)
    torch.cuda.synchronize()
    assert tm.shape == cm.shape == (F, 3, R * R, 16)
    slot_row = [3, 4, 5, 6, 0, 1, 2, 10, 7, 8, 9, -1, 11, 12, 13, -1]
    ref = torch.zeros(F, 3, R * R, 16, dtype=torch.float64)
    t64 = tokens.double().reshape(F, C, 3, R * R)
    for s, row in enumerate(slot_row):
        if row >= 0:
            for p in range(3):
                ref[:, p, :, s] = torch.einsum("c,fct->ft", Wd[row, 3 + p * C:3 + (p + 1) * C].double(), t64[:, :, p])
    assert (tm.cpu().double() - ref).abs().max().item() < 2e-5
    assert (tm - cm).abs().max().item() < 2e-5
    assert torch.equal(tm[..., 11], torch.zeros_like(tm[..., 11])) and torch.equal(tm[..., 15], torch.zeros_like(tm[..., 15]))
    # padded frame stride
    if F > 1:
        pad = torch.zeros(F, 3 * R * R + 8, C, device="cuda")
        pad[:, :3 * R * R] = tm_in
        assert torch.equal(triplane_project(pad[:, :3 * R * R], Wd.cuda(), C, R, token_major=True), tm)


def test_renderer_token_major_equals_channel_major():
    from audio_motion_avatar_b200.renderer import Renderer
    H, W, C, R, T = 96, 80, 64, 16, 2
    model = S.make_body_model(seed=0)
    w = S.make_decoder_weights(C, seed=1)
    tokens = S.make_triplane_tokens(T, C, R, seed=2)                          # [1,T,C,3RR]
    params = {k: v.cuda() for k, v in S.make_pose_sequence(1, T, seed=2).items()}
    cams = {k: v.cuda() for k, v in S.make_cameras(1, T, H, W).items()}
    imgs = []
    for tm in (False, True):
        cfg = make_cfg(H, W, C, R)
        cfg.triplane_token_major = tm
        r = Renderer(cfg, body_model=model)
        r.gaussian_decoder.load_state_dict(w)
        tk = tokens.transpose(2, 3).contiguous() if tm else tokens
        with torch.no_grad():
            imgs.append(r(tk.cuda(), cams, None, params)[0])
    d = (imgs[0] - imgs[1]).abs()
    # 1e-6 attribute differences can move a (pixel, Gaussian) pair across the rasteriser's alpha >= 1/255 cut
    assert (d > 1e-3).float().mean().item() <= 1e-4 and d.max().item() <= 1.0 / 255 + 1e-3
    from tests.helpers import psnr
    assert psnr(imgs[0].cpu(), imgs[1].cpu()) > 60


def test_token_major_decode_backward_equals_channel_major():
    """Training in the producer's layout: dL/dtokens (transposed), dL/dW, dL/db, dL/dpoints from the token-major
    autograd node equal the channel-major node's (itself checked against torch CPU autograd in test_train_gpu.py)."""
    from audio_motion_avatar_b200.stages import decode_gaussians_autograd, decoder_matrix
    C, R, F, N = 64, 16, 3, 3000
    g = torch.Generator().manual_seed(21)
    tokens = torch.randn(F, C, 3 * R * R, generator=g).cuda()
    pts = (torch.rand(F, N, 3, generator=g) * 2.4 - 1.2).cuda()
    transl = torch.randn(F, 3, generator=g).cuda()
    Wd, bd = decoder_matrix(S.make_decoder_weights(C, seed=3))
    cot = {k: torch.randn(F, N, n, generator=g).cuda() for k, n in (("xyz", 3), ("scale", 3), ("rot", 4), ("opacity", 1), ("color", 3))}
    grads = []
    for tm in (False, True):
        tk = (tokens.transpose(1, 2).contiguous() if tm else tokens.clone()).requires_grad_(True)
        p_, t_ = pts.clone().requires_grad_(True), transl.clone().requires_grad_(True)
        W_, b_ = Wd.cuda().clone().requires_grad_(True), bd.cuda().clone().requires_grad_(True)
        out = decode_gaussians_autograd(tk, p_, t_, W_, b_, C, R, 1.4, token_major=tm)
        sum((out[k] * cot[k]).sum() for k in cot).backward()
        grads.append((tk.grad.transpose(1, 2) if tm else tk.grad, p_.grad, t_.grad, W_.grad, b_.grad))
    for a, b, name in zip(grads[0], grads[1], ("tokens", "points", "transl", "weight", "bias")):
        err = (a - b).abs().max().item() / max(a.abs().max().item(), 1e-12)
        assert err < 1e-4, (name, err)
