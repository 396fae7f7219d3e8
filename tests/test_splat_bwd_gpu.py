"""Stage 4 backward on the GPU vs the oracle's backward (oracle/raster_oracle.c, Appendix B) on identical inputs.
Upstream's own gradients are float atomicAdd sums in arbitrary order, so parity is a tolerance check:
max-abs error <= 1e-3 of the gradient's max-abs (rel-err <= 1e-3)."""
import numpy as np
import pytest
import torch

from oracle import ref_path as O
from tests.helpers import oracle_cameras, oracle_frames, random_scene

pytestmark = pytest.mark.gpu

KEYS = ("means3d", "scales", "rots", "opacities", "colors")


def _oracle_grads(sc, view, full, tan, outs, gimg, bg, H, W):
    """Sum over frames (all frames view the same Gaussian set)."""
    tot = None
    for f, (pre, binned, img) in enumerate(outs):
        g = O.raster_backward(sc["means"], sc["scales"], sc["rots"], sc["opac"], sc["col"], view[f], full[f],
                              float(tan[f, 0]), float(tan[f, 1]), bg, W, H, pre, binned, img, gimg[f])
        tot = g if tot is None else {k: tot[k] + g[k] for k in g}
    return tot


def _rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-20))


@pytest.mark.parametrize("N,W,H,frames,hwc", [(3000, 128, 96, 3, False), (20000, 256, 256, 2, True), (500, 70, 45, 1, False)])
def test_splat_backward_vs_oracle(N, W, H, frames, hwc):
    from audio_motion_avatar_b200.splat import splat_backward, splat_forward
    sc = random_scene(N, W, H, seed=N + 1, frames=frames)
    bg = (1.0, 0.5, 0.25)
    view, full, tan, outs = oracle_frames(sc, H, W, bg)
    gimg = torch.randn(frames, 3, H, W, generator=torch.Generator().manual_seed(N))
    ref = _oracle_grads(sc, view, full, tan, outs, gimg, bg, H, W)
    d = torch.device("cuda:0")
    c = lambda t: t.to(d)
    args = (c(sc["means"])[None], c(sc["scales"])[None], c(sc["rots"])[None], c(sc["opac"])[None], c(sc["col"])[None],
            c(view), c(full), c(tan), c(torch.tensor(bg)))
    fwd = splat_forward(*args, H, W, views_per_set=frames, want_aux=True, hwc=hwc)
    gdev = c(gimg.permute(0, 2, 3, 1).contiguous() if hwc else gimg)
    got = splat_backward(fwd, gdev, *args, H, W, views_per_set=frames, hwc=hwc)
    torch.cuda.synchronize()
    for k in KEYS:
        assert _rel(got[k][0].cpu().numpy().reshape(ref[k].shape), ref[k]) <= 1e-3, k
    # screen-space sums of one frame, before the per-Gaussian chain
    pre, binned, img = outs[0]
    g0 = O.raster_backward(sc["means"], sc["scales"], sc["rots"], sc["opac"], sc["col"], view[0], full[0], float(tan[0, 0]),
                           float(tan[0, 1]), bg, W, H, pre, binned, img, gimg[0])
    ds = got["dsum"][0].cpu().numpy()
    assert _rel(ds[:, 0:2], g0["mean2d"]) <= 1e-3 and _rel(ds[:, 2:5], g0["conic"]) <= 1e-3
    assert _rel(ds[:, 5], g0["opacities"][:, 0]) <= 1e-3 and _rel(ds[:, 6:9], g0["colors"]) <= 1e-3
    dead = pre["radii"] <= 0
    assert not ds[dead].any()


def test_splat_autograd_with_activations():
    """SplatFunction with render_one's activations fused (activate=True) vs the oracle's autograd node."""
    from audio_motion_avatar_b200.splat import splat_render
    N, W, H = 4000, 96, 80
    sc = random_scene(N, W, H, seed=11, frames=2)
    bg = (1.0, 1.0, 1.0)
    view, full, tan = oracle_cameras(sc, H, W)
    raw = {"xyz": sc["means"].clone(), "rot": sc["rots"].clone(), "scale": torch.log(sc["scales"]) + 3.9,
           "opacity": torch.logit(sc["opac"].clamp(0.02, 0.98)), "color": sc["col"].clone()}
    raw["scale"][:50] = 2.0                                                # exp(2 - 3.9) > 0.1: clamped, zero gradient
    gimg = torch.randn(2, 3, H, W, generator=torch.Generator().manual_seed(5))
    leaves = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
    loss = 0
    for f in range(2):
        img = O.render_one_autograd(leaves, sc["K"][f], sc["E"][f], H, W, bg)
        loss = loss + (img * gimg[f]).sum()
    loss.backward()
    d = torch.device("cuda:0")
    dl = {k: v.clone().to(d).requires_grad_(True) for k, v in raw.items()}
    out = splat_render(dl["xyz"][None], dl["scale"][None], dl["rot"][None], dl["opacity"][None], dl["color"][None],
                       view.to(d), full.to(d), tan.to(d), torch.tensor(bg, device=d), H, W, views_per_set=2,
                       activate=True, clamp01=True)
    (out * gimg.to(d)).sum().backward()
    for k in raw:
        a, b = dl[k].grad.cpu().numpy(), leaves[k].grad.numpy()
        assert _rel(a, b) <= 1e-3, k
    assert not dl["scale"].grad[:50].any()


def test_splat_backward_empty_and_culled():
    from audio_motion_avatar_b200.splat import splat_render
    d = torch.device("cuda:0")
    sc = random_scene(64, 64, 64, seed=2, frames=1)
    view, full, tan = oracle_cameras(sc, 64, 64)
    means = sc["means"].clone()
    means[:, 2] = -10.0                                                    # everything behind the camera
    leaves = [t.to(d).requires_grad_(True) for t in (means, sc["scales"], sc["rots"], sc["opac"], sc["col"])]
    out = splat_render(*[t[None] for t in leaves], view.to(d), full.to(d), tan.to(d), torch.ones(3, device=d), 64, 64)
    out.sum().backward()
    for t in leaves:
        assert t.grad is not None and not t.grad.any()


def test_clamp_masks_the_gradient_like_the_reference():
    """render_one clamps the blended colour to [0,1] (renderer.py:568); autograd zeroes dL/dimage wherever the colour left
    that range.  A background outside [0,1] makes that happen on most pixels: gradients through the public render_batch must
    match the oracle's autograd node followed by the same clamp."""
    from types import SimpleNamespace
    from audio_motion_avatar_b200.renderer import render_batch
    N, W, H = 3000, 96, 80
    sc = random_scene(N, W, H, seed=17, frames=1)
    bg = [1.5, -0.2, 1.0]
    raw = {"xyz": sc["means"].clone(), "rot": sc["rots"].clone(), "scale": torch.log(sc["scales"]) + 3.9,
           "opacity": torch.logit(sc["opac"].clamp(0.02, 0.98)), "color": sc["col"].clone()}
    gimg = torch.randn(3, H, W, generator=torch.Generator().manual_seed(6))
    leaves = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
    img = O.render_one_autograd(leaves, sc["K"][0], sc["E"][0], H, W, tuple(bg))
    (img * gimg).sum().backward()
    d = torch.device("cuda:0")
    dl = {k: v.clone()[None].to(d).requires_grad_(True) for k, v in raw.items()}
    args = SimpleNamespace(image_size=(H, W), rgb=True, sh_degree=3)
    out = render_batch(dl, sc["K"][None].to(d), sc["E"][None].to(d), args, bg_color=bg)          # [1,1,H,W,3]
    assert float(out.min()) >= 0.0 and float(out.max()) <= 1.0
    assert np.abs(out[0, 0].permute(2, 0, 1).detach().cpu().numpy() - img.detach().numpy()).max() <= 1e-3
    (out[0, 0].permute(2, 0, 1) * gimg.to(d)).sum().backward()
    clamped = float(((img.detach() <= 0) | (img.detach() >= 1)).float().mean())
    assert clamped > 0.2                                                   # the mask matters in this scene
    for k in raw:
        a, b = dl[k].grad[0].cpu().numpy(), leaves[k].grad.numpy()
        assert _rel(a, b) <= 1e-3, k
