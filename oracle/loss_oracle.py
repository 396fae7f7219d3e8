"""oracle/loss_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement (torch CPU fp32) of the image losses the reference applies to the rendered frames:
``l1_loss`` (src/utils/loss_utils.py:18) and ``ssim`` / ``_ssim`` (:44-83: 11x11 Gaussian window, sigma 1.5,
``F.conv2d(padding=5, groups=C)`` = zero padding, C1 = 0.01^2, C2 = 0.03^2), combined as
``l1 + 0.1 * (1 - ssim)`` at src/models/lightning_model_wrapper.py:138-140.  Gradients come from torch autograd of
these same ops, which is how the reference gets them.

PINNED: tests/test_oracle_golden.py::test_image_losses_match_reference checks values AND the gradient against
tests/golden/ref_loss.npz, produced by the reference's own functions (tests/golden/make_golden_loss.py).
Only tests/ may import this module."""
from math import exp

import torch
import torch.nn.functional as F


def gaussian_window(window_size: int = 11, sigma: float = 1.5) -> torch.Tensor:
    """loss_utils.py:24-26."""
    g = torch.Tensor([exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)])
    return g / g.sum()


def l1_loss(img: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    return torch.abs(img - gt).mean()


def ssim_map(img: torch.Tensor, gt: torch.Tensor, window_size: int = 11) -> torch.Tensor:
    """[F,H,W,3] frames -> SSIM map [F,3,H,W] (loss_utils.py:50-79)."""
    x = img.permute(0, 3, 1, 2)
    y = gt.permute(0, 3, 1, 2)
    C = x.shape[1]
    w1 = gaussian_window(window_size).unsqueeze(1)
    w2 = w1.mm(w1.t()).float()[None, None].expand(C, 1, window_size, window_size).contiguous()
    conv = lambda t: F.conv2d(t, w2, padding=window_size // 2, groups=C)
    mu1, mu2 = conv(x), conv(y)
    mu1_sq, mu2_sq, mu12 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = conv(x * x) - mu1_sq
    s2 = conv(y * y) - mu2_sq
    s12 = conv(x * y) - mu12
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return ((2 * mu12 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))


def ssim(img: torch.Tensor, gt: torch.Tensor, size_average: bool = True) -> torch.Tensor:
    m = ssim_map(img.reshape(-1, *img.shape[-3:]), gt.reshape(-1, *gt.shape[-3:]))
    return m.mean() if size_average else m.mean(1).mean(1).mean(1)


def l1_ssim_loss(img: torch.Tensor, gt: torch.Tensor, ssim_weight: float = 0.1) -> torch.Tensor:
    return l1_loss(img, gt) + ssim_weight * (1 - ssim(img, gt))
