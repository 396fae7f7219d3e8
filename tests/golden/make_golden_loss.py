"""Generates tests/golden/ref_loss.npz by running the REFERENCE's own l1_loss / ssim (src/utils/loss_utils.py:18,44-83)
and torch autograd through them, on CPU.  Run in the build container only (it reads /root/reference):

    python tests/golden/make_golden_loss.py

loss_utils imports ``lpips`` (not installed offline); it is stubbed, the functions exercised here do not touch it."""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import REF, OUT, _StubFinder          # noqa: E402


def main():
    sys.meta_path.insert(0, _StubFinder())
    sys.path.insert(0, REF)
    lu = importlib.import_module("src.utils.loss_utils")
    g = torch.Generator().manual_seed(77)
    B, T, H, W = 1, 2, 37, 45                            # ragged: not a multiple of the 16-pixel tile
    gt = torch.rand(B, T, H, W, 3, generator=g)
    img = (gt + 0.15 * torch.randn(B, T, H, W, 3, generator=g)).clamp(0, 1)
    img[0, 0, :6, :7] = gt[0, 0, :6, :7]                 # exact zeros of |img - gt|: sign(0) = 0
    x = img.clone().requires_grad_(True)
    l1 = lu.l1_loss(x, gt)
    s = lu.ssim(x, gt)
    s_per = lu.ssim(x, gt, size_average=False)
    loss = l1 + 0.1 * (1 - s)                            # lightning_model_wrapper.py:138-140
    loss.backward()
    np.savez_compressed(os.path.join(OUT, "ref_loss.npz"), img=img.numpy(), gt=gt.numpy(), l1=l1.item(), ssim=s.item(),
                        ssim_per_frame=s_per.detach().numpy(), loss=loss.item(), grad=x.grad.numpy())
    print("l1", l1.item(), "ssim", s.item(), "written to", OUT)


if __name__ == "__main__":
    main()
