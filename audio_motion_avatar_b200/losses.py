"""Image losses of the training step on the B200 kernels: drop-in for ``l1_loss`` and ``ssim`` of the
reference's ``src/utils/loss_utils.py:18,44-83`` (same names, argument meaning and return values), differentiable
w.r.t. the first argument (the rendered frames), plus the fused combination the Lightning wrapper forms at
``lightning_model_wrapper.py:138-145``.  Frames are ``[B,T,H,W,3]`` (or ``[F,H,W,3]``) fp32 CUDA tensors in HWC, exactly
what ``render_batch`` returns; no permute, no CHW copy.  There is no CPU path."""
from __future__ import annotations

from typing import Tuple

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr


def _frames(t: torch.Tensor) -> torch.Tensor:
    if t.dim() < 4 or t.shape[-1] != 3:
        raise ValueError(f"expected [...,H,W,3] frames, got {tuple(t.shape)}")
    if not t.is_cuda:
        raise _lib.AmavError("image losses need CUDA tensors (no CPU path)")
    return t.reshape(-1, *t.shape[-3:]).to(torch.float32).contiguous()


class ImageLossFunction(torch.autograd.Function):
    """(img [F,H,W,3], gt [F,H,W,3]) -> (per-frame mean |img - gt| [F], per-frame mean SSIM [F]) in one pass;
    backward: one pass over the three SSIM derivative maps."""

    @staticmethod
    def forward(ctx, img, gt):
        Fr, H, W = int(img.shape[0]), int(img.shape[1]), int(img.shape[2])
        dev = img.device
        sums = torch.zeros(Fr, 2, device=dev)
        need = ctx.needs_input_grad[0]
        maps = torch.empty(Fr, 3, 3, H, W, device=dev) if need else None
        call("amav_image_loss_forward", Fr, H, W, ptr(img), ptr(gt), ptr(sums), ptr(maps), stream_ptr())
        ctx.save_for_backward(img, gt, maps)
        n = float(H * W * 3)
        ctx.n = n
        return sums[:, 0] / n, sums[:, 1] / n

    @staticmethod
    def backward(ctx, g_l1, g_ssim):
        img, gt, maps = ctx.saved_tensors
        Fr, H, W = int(img.shape[0]), int(img.shape[1]), int(img.shape[2])
        grad = torch.empty_like(img)
        gl = (g_l1 / ctx.n).to(torch.float32).contiguous()
        gs = (g_ssim / ctx.n).to(torch.float32).contiguous()
        call("amav_image_loss_backward", Fr, H, W, ptr(img), ptr(gt), ptr(maps), ptr(gl), ptr(gs), ptr(grad), stream_ptr())
        return grad, None


def image_loss_terms(img: torch.Tensor, gt: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per-frame (mean L1, mean SSIM), each [F]."""
    x, y = _frames(img), _frames(gt.detach())
    if x.shape != y.shape:
        raise ValueError("rendered and target frames differ in shape")
    return ImageLossFunction.apply(x, y)


def l1_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """loss_utils.py:18."""
    return image_loss_terms(network_output, gt)[0].mean()


def ssim(img1: torch.Tensor, img2: torch.Tensor, window_size: int = 11, size_average: bool = True) -> torch.Tensor:
    """loss_utils.py:44-83 ([B,T,H,W,C] frames, 11x11 Gaussian window, sigma 1.5, zero padding)."""
    if window_size != 11:
        raise NotImplementedError("only the reference's window_size=11 is built")
    s = image_loss_terms(img1, img2)[1]
    return s.mean() if size_average else s


def l1_ssim_loss(rendered: torch.Tensor, target: torch.Tensor, ssim_weight: float = 0.1):
    """lightning_model_wrapper.py:138-140: ``l1 + ssim_weight * (1 - ssim)``; returns (loss, l1, 1 - ssim)."""
    l1, s = image_loss_terms(rendered, target)
    l1, s = l1.mean(), s.mean()
    return l1 + ssim_weight * (1.0 - s), l1, 1.0 - s
