"""Generates tests/golden/ref_*.npz by running the REFERENCE's own functions on CPU.

Run in the build container only (it reads /root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

The reference module src/models/renderer.py imports packages that are not installed offline
(omegaconf, smplx, pytorch3d, diff_gaussian_rasterization, spconv, ...).  They are replaced by empty
stub modules so that the module body executes; the functions exercised here do not touch them:
  Renderer.sample_from_triplane (renderer.py:292-317), nn.Linear heads (:51-55,165-171),
  Renderer.construct_gaussians (:319-346), getWorld2View2_torch / getProjectionMatrix_torch /
  focal2fov_torch (graphic_utils.py:67-78,103-136,144-145) and the camera lines of render_one
  (renderer.py:489-510, re-executed here statement by statement because render_one itself hard-codes
  device="cuda").
"""
import importlib
import math
import os
import sys
import types
from types import SimpleNamespace

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


class _Anything:
    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        return _Anything()


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Anything


class _StubFinder:
    """Meta-path finder that fabricates a stub for any module that cannot be found."""
    NAMES = ("omegaconf", "smplx", "pytorch3d", "diff_gaussian_rasterization", "spconv", "torch_scatter",
             "addict", "timm", "pyrender", "trimesh", "matplotlib", "lightning", "diffusers", "kornia",
             "lpips", "sklearn", "cv2", "flash_attn", "PIL", "imageio", "av", "torchaudio", "torchvision")

    def find_spec(self, fullname, path=None, target=None):
        if fullname.split(".")[0] in self.NAMES:
            from importlib.machinery import ModuleSpec
            return ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        m = _StubModule(spec.name)
        m.__path__ = []
        return m

    def exec_module(self, module):
        pass


def import_reference():
    sys.meta_path.insert(0, _StubFinder())
    sys.path.insert(0, REF)
    renderer = importlib.import_module("src.models.renderer")
    gu = importlib.import_module("src.utils.graphic_utils")
    return renderer, gu


def main():
    renderer, gu = import_reference()
    torch.manual_seed(1234)
    g = torch.Generator().manual_seed(1234)

    # ---- query + decode + construct_gaussians ------------------------------------------------
    Fr, C, R, N, radius = 2, 24, 16, 301, 1.4
    tri = torch.randn(Fr, 3, C, R, R, generator=g)
    pts = (torch.rand(Fr, N, 3, generator=g) * 2 - 1) * 1.6        # some beyond +-radius: exercises clamp
    pts[0, :4] = torch.tensor([[1.4, -1.4, 0.0], [-1.4, 1.4, 1.4], [0.0, 0.0, 0.0], [1.3999, -1.3999, 0.7]])
    self_ = SimpleNamespace(cfg=SimpleNamespace(radius=radius))
    feats = renderer.Renderer.sample_from_triplane(self_, tri, pts)
    K_in = 3 * C + 3
    heads = {"xyz": 3, "rotation": 4, "scaling": 3, "opacity": 1, "shs": 3}
    weights, dec = {}, {}
    x = torch.cat([pts, feats], dim=-1)
    for name, k in heads.items():
        lin = torch.nn.Linear(K_in, k)
        with torch.no_grad():
            lin.weight.copy_((torch.rand(k, K_in, generator=g) * 2 - 1) / math.sqrt(K_in))
            lin.bias.copy_(torch.randn(k, generator=g) * 0.3)
        weights[f"{name}_layer.weight"] = lin.weight.detach().clone()
        weights[f"{name}_layer.bias"] = lin.bias.detach().clone()
        with torch.no_grad():
            dec[name] = lin(x)
    transl = torch.randn(Fr, 3, generator=g)
    gp = {"xyz_offset": dec["xyz"], "scaling": dec["scaling"], "rotation": dec["rotation"],
          "opacity": dec["opacity"], "shs": dec["shs"]}
    gauss = renderer.Renderer.construct_gaussians(self_, gp, pts, {"transl": transl})
    np.savez_compressed(
        os.path.join(OUT, "ref_query_decode.npz"),
        triplane=tri.numpy(), points=pts.numpy(), radius=np.float32(radius), feats=feats.numpy(),
        transl=transl.numpy(),
        **{f"w_{k}": v.numpy() for k, v in weights.items()},
        **{f"dec_{k}": v.numpy() for k, v in dec.items()},
        **{f"g_{k}": v.numpy() for k, v in gauss.items() if k != "shs"})

    # ---- camera set-up (renderer.py:489-510) -------------------------------------------------
    cams = []
    for i in range(6):
        H, W = [(512, 512), (1024, 1024), (896, 640), (1296, 2304), (64, 48), (128, 128)][i]
        ang = torch.randn(3, generator=g) * 0.4
        A = torch.zeros(3, 3)
        A[0, 1], A[0, 2], A[1, 2] = -ang[2], ang[1], -ang[0]
        A = A - A.T
        Rm = torch.linalg.matrix_exp(A)
        E = torch.eye(4)
        if i > 0:
            E[:3, :3] = Rm
            E[:3, 3] = torch.randn(3, generator=g) * 0.5 + torch.tensor([0.0, 0.0, 2.0])
        K = torch.tensor([[1.1 * W + i, 0.0, W / 2 + 0.37 * i], [0.0, 1.07 * W, H / 2 - 0.21 * i], [0.0, 0.0, 1.0]])
        R = E[:3, :3].reshape(3, 3).transpose(1, 0)
        T = E[:3, 3]
        FovY = gu.focal2fov_torch(K[1, 1], H)
        FovX = gu.focal2fov_torch(K[0, 0], W)
        tanfovx = math.tan(FovX * 0.5)
        tanfovy = math.tan(FovY * 0.5)
        wvt = gu.getWorld2View2_torch(R, T).transpose(0, 1)
        proj = gu.getProjectionMatrix_torch(znear=0.01, zfar=100.0, fovX=FovX, fovY=FovY, K=K, w=W, h=H).transpose(0, 1)
        full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)
        campos = wvt.inverse()[3, :3]
        cams.append(dict(K=K.numpy(), E=E.numpy(), H=H, W=W, view=wvt.numpy(), full=full.numpy(),
                         tanfovx=tanfovx, tanfovy=tanfovy, campos=campos.numpy()))
    np.savez_compressed(os.path.join(OUT, "ref_camera.npz"),
                        **{f"{k}_{i}": np.asarray(c[k]) for i, c in enumerate(cams) for k in c})

    # ---- activation lines of render_one (renderer.py:532-533,547) ----------------------------
    s = torch.randn(50, 3, generator=g) * 2
    o = torch.randn(50, 1, generator=g) * 2
    c = torch.rand(50, 3, generator=g) * 1.4 - 0.2
    scales = torch.min(torch.exp(s - renderer.SCALE_BIAS), torch.tensor(0.1))
    opac = torch.sigmoid(o - renderer.OPACITY_BIAS)
    col = torch.clamp(c, 0.0, 1.0)
    np.savez_compressed(os.path.join(OUT, "ref_activation.npz"), s=s.numpy(), o=o.numpy(), c=c.numpy(),
                        scales=scales.numpy(), opac=opac.numpy(), col=col.numpy(),
                        SCALE_BIAS=renderer.SCALE_BIAS, OPACITY_BIAS=renderer.OPACITY_BIAS)
    print("golden vectors written to", OUT)


if __name__ == "__main__":
    main()
