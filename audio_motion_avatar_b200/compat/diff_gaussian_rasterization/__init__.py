"""``diff_gaussian_rasterization``-compatible front end of the B200 splat kernels (secondary boundary).

Exports what src/models/renderer.py:422 imports -- ``GaussianRasterizationSettings`` (the 13 fields set at
renderer.py:516-530) and ``GaussianRasterizer`` whose call takes the keyword arguments of renderer.py:557-566 and
returns ``(color [3,H,W], radii [N] int32, invdepth [1,H,W])`` -- differentiable w.r.t. means3D, opacities, scales,
rotations and colors_precomp.  One frame per call, like upstream; the batched entry points of
``audio_motion_avatar_b200.renderer`` are the fast path.  Not supported (the reference does not use them): ``shs``
(SH colours), ``cov3D_precomp``, ``antialiasing=True``, gradients to ``means2D``."""
from typing import NamedTuple

import torch
import torch.nn as nn

from ...splat import SplatWorkspace, splat_backward, splat_forward


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    antialiasing: bool = False


class _Rasterize(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, scales, rotations, opacities, colors, settings):
        dev = means3D.device
        view = settings.viewmatrix.to(dev, torch.float32).reshape(1, 4, 4)
        proj = settings.projmatrix.to(dev, torch.float32).reshape(1, 4, 4)
        tan = torch.tensor([[float(settings.tanfovx), float(settings.tanfovy)]], dtype=torch.float32, device=dev)
        bg = settings.bg.to(dev, torch.float32).reshape(3)
        H, W = int(settings.image_height), int(settings.image_width)
        args = (means3D[None], scales[None], rotations[None], opacities.reshape(1, -1, 1), colors[None], view, proj, tan, bg)
        fwd = splat_forward(*args, H, W, want_aux=True, want_invdepth=True, scale_modifier=float(settings.scale_modifier),
                            workspace=SplatWorkspace())
        ctx.fwd, ctx.args, ctx.hw, ctx.smod = fwd, args, (H, W), float(settings.scale_modifier)
        ctx.opac_shape = opacities.shape
        radii = fwd["geom"][0, :, 11].to(torch.int32)
        invdepth = fwd["invdepth"][0]
        ctx.mark_non_differentiable(radii, invdepth)
        return fwd["color"][0], radii, invdepth

    @staticmethod
    def backward(ctx, g_color, _g_radii, _g_invdepth):
        H, W = ctx.hw
        g = splat_backward(ctx.fwd, g_color[None].contiguous(), *ctx.args, H, W, scale_modifier=ctx.smod)
        ctx.fwd = None
        return (g["means3d"][0], g["scales"][0], g["rots"][0], g["opacities"][0].reshape(ctx.opac_shape), g["colors"][0], None)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings: GaussianRasterizationSettings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D, means2D=None, opacities=None, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        if shs is not None or colors_precomp is None:
            raise NotImplementedError("SH colours are not supported: pass colors_precomp (renderer.py:547)")
        if cov3D_precomp is not None or scales is None or rotations is None:
            raise NotImplementedError("cov3D_precomp is not supported: pass scales and rotations")
        if getattr(self.raster_settings, "antialiasing", False):
            raise NotImplementedError("antialiasing=True is not supported (renderer.py:529 uses False)")
        return _Rasterize.apply(means3D.float(), scales.float(), rotations.float(), opacities.float(),
                                colors_precomp.float(), self.raster_settings)
