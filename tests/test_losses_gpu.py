"""Image losses on the GPU vs the reference's golden vectors (tests/golden/ref_loss.npz, produced by
the reference's own l1_loss / ssim + autograd) and vs the pinned oracle on larger seeded frames.
Floating point: values within 2e-6 absolute, gradients within 1e-4 of their max-abs (separable vs 2-D window product
and fp32 summation order)."""
import os

import numpy as np
import pytest
import torch

from oracle import loss_oracle as L

pytestmark = pytest.mark.gpu


def test_losses_match_reference_golden():
    from audio_motion_avatar_b200 import losses
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_loss.npz"))
    dev = torch.device("cuda:0")
    img = torch.from_numpy(d["img"]).to(dev).requires_grad_(True)        # [1,2,37,45,3]: ragged tiles
    gt = torch.from_numpy(d["gt"]).to(dev)
    l1, s = losses.l1_loss(img, gt), losses.ssim(img, gt)
    assert abs(l1.item() - float(d["l1"])) <= 2e-6 and abs(s.item() - float(d["ssim"])) <= 2e-6
    per = losses.ssim(img, gt, size_average=False)
    assert np.abs(per.detach().cpu().numpy() - d["ssim_per_frame"]).max() <= 2e-6
    loss, l1b, dssim = losses.l1_ssim_loss(img, gt)
    assert abs(loss.item() - float(d["loss"])) <= 2e-6 and abs(dssim.item() - (1 - float(d["ssim"]))) <= 2e-6
    loss.backward()
    g, ref = img.grad.cpu().numpy(), d["grad"]
    assert g.shape == ref.shape
    assert np.abs(g - ref).max() <= 1e-4 * np.abs(ref).max()


@pytest.mark.parametrize("F,H,W", [(3, 70, 45), (2, 128, 160), (1, 16, 16)])
def test_losses_match_oracle(F, H, W):
    from audio_motion_avatar_b200 import losses
    g = torch.Generator().manual_seed(F * 1000 + H)
    gt = torch.rand(F, H, W, 3, generator=g)
    img = (gt + 0.2 * torch.randn(F, H, W, 3, generator=g)).clamp(0, 1)
    w = torch.rand(F, generator=g) + 0.5                                  # per-frame weights: size_average=False path
    x = img.clone().requires_grad_(True)
    ref = (L.l1_loss(x, gt) + (w * (1 - L.ssim(x, gt, size_average=False))).sum())
    ref.backward()
    dev = torch.device("cuda:0")
    xd = img.to(dev).requires_grad_(True)
    out = losses.l1_loss(xd, gt.to(dev)) + (w.to(dev) * (1 - losses.ssim(xd, gt.to(dev), size_average=False))).sum()
    out.backward()
    assert abs(out.item() - ref.item()) <= 1e-5 * max(1.0, abs(ref.item()))
    a, b = xd.grad.cpu().numpy(), x.grad.numpy()
    assert np.abs(a - b).max() <= 1e-4 * np.abs(b).max()


def test_losses_no_grad_and_errors():
    from audio_motion_avatar_b200 import _lib, losses
    dev = torch.device("cuda:0")
    a = torch.rand(2, 32, 32, 3, device=dev)
    with torch.no_grad():
        assert losses.l1_loss(a, a).item() == 0.0
        assert abs(losses.ssim(a, a).item() - 1.0) <= 1e-6
    with pytest.raises(_lib.AmavError):
        losses.l1_loss(a.cpu(), a.cpu())
    with pytest.raises(ValueError):
        losses.l1_loss(a, a[:, :16])


def test_training_step_with_fused_losses():
    """Renderer.forward -> l1 + 0.1 (1 - ssim) -> backward: the loss gradient feeds the splat backward directly."""
    from types import SimpleNamespace
    from audio_motion_avatar_b200 import losses, synthetic as S
    from audio_motion_avatar_b200.renderer import Renderer
    H, W, C, R, T = 64, 48, 64, 16, 2
    cfg = SimpleNamespace(device="cuda:0", triplane_resolution=R, triplane_feature_dim=C, radius=1.4, smplx_model_path=None,
                          flat_hand_mean=True, num_expression_coeffs=10, upsample_triplane=False, densify_smplx_verts=False,
                          subdivide_steps=2, predict_smplx_params=False, no_point_refiner=True, image_size=(H, W), rgb=True,
                          sh_degree=3, num_gaussians=0)
    dev = torch.device("cuda:0")
    r = Renderer(cfg, body_model=S.make_body_model(seed=0))
    r.gaussian_decoder.load_state_dict(S.make_decoder_weights(C, seed=0))
    tok = S.make_triplane_tokens(T, C, R, seed=0).to(dev).requires_grad_(True)
    params = {k: v.to(dev) for k, v in S.make_pose_sequence(1, T, seed=0).items()}
    cams = {k: v.to(dev) for k, v in S.make_cameras(1, T, H, W).items()}
    target = torch.rand(1, T, H, W, 3, device=dev)
    img, _ = r(tok, cams, None, params)
    loss, l1, ds = losses.l1_ssim_loss(img, target)
    loss.backward()
    ref = L.l1_ssim_loss(img.detach().cpu()[0], target.cpu()[0])
    assert abs(loss.item() - ref.item()) <= 1e-5
    assert tok.grad is not None and torch.isfinite(tok.grad).all() and tok.grad.abs().sum() > 0
