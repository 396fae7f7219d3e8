"""Pins the C restatement of the rasteriser's backward pass (oracle/raster_oracle.c).

The reference has no tests or fixtures for it (the arithmetic lives in diff-gaussian-rasterization), so the pin is
autograd itself: the forward (Appendix A) is restated a second time in float64 torch ops, per tile and per list
entry with the forward's discrete decisions (power > 0, alpha < 1/255, saturation) as constant masks, and
torch.autograd differentiates that.  The C backward must agree to fp32 rounding."""
import math

import numpy as np
import torch

from oracle import ref_path as O
from tests.helpers import oracle_cameras, random_scene


def _torch_forward(means, scales, rots, opac, col, view, proj, tanx, tany, bg, W, H, binned):
    """float64, differentiable.  view/proj are the transposed matrices the reference passes (m[c*4+r])."""
    V = view.double().reshape(4, 4).T        # mathematical matrix: V[r, c] = view_flat[c*4 + r]
    P = proj.double().reshape(4, 4).T
    N = means.shape[0]
    hom = torch.cat([means, torch.ones(N, 1, dtype=torch.float64)], 1)
    pv = hom @ V.T
    ph = hom @ P.T
    pw = 1.0 / (ph[:, 3] + 1e-7)
    ppx, ppy = ph[:, 0] * pw, ph[:, 1] * pw
    r, x, y, z = rots.unbind(1)
    Rm = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                      2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                      2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).reshape(N, 3, 3)
    M = Rm * scales[:, None, :]
    Sig = M @ M.transpose(1, 2)
    fx, fy = W / (2 * tanx), H / (2 * tany)
    tz = pv[:, 2]
    tx, ty = pv[:, 0], pv[:, 1]              # test scenes stay inside the 1.3*tanfov clamp
    zero = torch.zeros_like(tz)
    J = torch.stack([fx / tz, zero, -fx * tx / tz ** 2, zero, fy / tz, -fy * ty / tz ** 2], 1).reshape(N, 2, 3)
    A = J @ V[:3, :3]
    cov = A @ Sig @ A.transpose(1, 2)
    a, b, c = cov[:, 0, 0] + 0.3, cov[:, 0, 1], cov[:, 1, 1] + 0.3
    det = a * c - b * b
    conx, cony, conz = c / det, -b / det, a / det
    px = ((ppx + 1) * W - 1) * 0.5
    py = ((ppy + 1) * H - 1) * 0.5
    gx = (W + 15) // 16
    img = torch.zeros(3, H, W, dtype=torch.float64)
    bgt = torch.tensor(bg, dtype=torch.float64)
    ranges, plist = binned["ranges"], binned["point_list"]
    for tile in range(ranges.shape[0]):
        ty0, tx0 = (tile // gx) * 16, (tile % gx) * 16
        ys = torch.arange(ty0, min(ty0 + 16, H), dtype=torch.float64)
        xs = torch.arange(tx0, min(tx0 + 16, W), dtype=torch.float64)
        PY, PX = torch.meshgrid(ys, xs, indexing="ij")
        T = torch.ones_like(PX)
        C = torch.zeros(3, *PX.shape, dtype=torch.float64)
        done = torch.zeros_like(PX, dtype=torch.bool)
        for k in range(int(ranges[tile, 0]), int(ranges[tile, 1])):
            g = int(plist[k])
            dx, dy = px[g] - PX, py[g] - PY
            power = -0.5 * (conx[g] * dx * dx + conz[g] * dy * dy) - cony[g] * dx * dy
            raw = opac[g, 0] * torch.exp(power)
            alpha = raw - (raw - 0.99).clamp_min(0).detach()       # min(0.99, .) with pass-through gradient
            skip = (power.detach() > 0) | (alpha.detach() < 1.0 / 255.0) | done
            Tn = T * (1 - alpha)
            newly = (~skip) & (Tn.detach() < 1e-4)
            done = done | newly
            use = (~skip) & (~newly)
            w = torch.where(use, alpha * T, torch.zeros_like(T))
            C = C + col[g][:, None, None] * w
            T = torch.where(use, Tn, T)
        img[:, ty0:ty0 + PX.shape[0], tx0:tx0 + PX.shape[1]] = C + T * bgt[:, None, None]
    return img


def _scene(seed, N=70, W=48, H=40):
    sc = random_scene(N, W, H, seed=seed, frames=1, spread=0.45, smin=0.01, smax=0.08)
    view, full, tan = oracle_cameras(sc, H, W)
    return sc, view[0], full[0], float(tan[0, 0]), float(tan[0, 1])


def test_backward_matches_float64_autograd():
    W, H = 48, 40
    bg = (1.0, 0.5, 0.25)
    for seed in (0, 1):
        sc, view, full, tx, ty = _scene(seed, W=W, H=H)
        pre, binned, out = O.rasterize(sc["means"], sc["scales"], sc["rots"], sc["opac"], sc["col"], view, full, tx, ty, bg, W, H)
        assert binned["D"] > 100 and out["n_contrib"].max() > 3
        gen = torch.Generator().manual_seed(100 + seed)
        gimg = torch.randn(3, H, W, generator=gen)
        leaves = [sc[k].double().requires_grad_(True) for k in ("means", "scales", "rots", "opac", "col")]
        img64 = _torch_forward(*leaves, view, full, tx, ty, bg, W, H, binned)
        assert np.abs(img64.detach().numpy() - out["color"]).max() < 2e-5      # the float64 restatement is the same forward
        (img64 * gimg.double()).sum().backward()
        got = O.raster_backward(sc["means"], sc["scales"], sc["rots"], sc["opac"], sc["col"], view, full, tx, ty, bg, W, H,
                                pre, binned, out, gimg)
        for leaf, key in zip(leaves, ("means3d", "scales", "rots", "opacities", "colors")):
            ref = leaf.grad.numpy()
            err = np.abs(got[key] - ref).max()
            scale = np.abs(ref).max()
            assert scale > 0
            assert err <= 2e-4 * scale + 1e-6, (key, err, scale)
        # culled Gaussians get exactly zero
        dead = pre["radii"] <= 0
        assert dead.any() and not got["means3d"][dead].any() and not got["scales"][dead].any()


def test_autograd_node_and_activations():
    """render_one_autograd chains the reference's activations (renderer.py:532-533,547) on top of the node."""
    W, H = 32, 32
    sc, view, full, tx, ty = _scene(3, N=40, W=W, H=H)
    g = {"xyz": sc["means"].clone().requires_grad_(True), "rot": sc["rots"].clone().requires_grad_(True),
         "scale": (torch.log(sc["scales"]) + 3.9).requires_grad_(True),
         "opacity": torch.logit(sc["opac"].clamp(0.02, 0.98)).requires_grad_(True),
         "color": sc["col"].clone().requires_grad_(True)}
    img = O.render_one_autograd(g, sc["K"][0], sc["E"][0], H, W)
    assert img.shape == (3, H, W)
    img.sum().backward()
    for k, v in g.items():
        assert v.grad is not None and torch.isfinite(v.grad).all(), k
    assert g["color"].grad.abs().sum() > 0 and g["xyz"].grad.abs().sum() > 0


def test_finite_difference_colour_and_opacity():
    """Independent of autograd: central differences of the C forward for the smooth inputs."""
    W, H = 32, 32
    bg = (0.2, 0.4, 0.9)
    sc, view, full, tx, ty = _scene(5, N=30, W=W, H=H)
    pre, binned, out = O.rasterize(sc["means"], sc["scales"], sc["rots"], sc["opac"], sc["col"], view, full, tx, ty, bg, W, H)
    gimg = torch.randn(3, H, W, generator=torch.Generator().manual_seed(7))
    got = O.raster_backward(sc["means"], sc["scales"], sc["rots"], sc["opac"], sc["col"], view, full, tx, ty, bg, W, H,
                            pre, binned, out, gimg)
    loss = lambda col: float((torch.from_numpy(O.rasterize(sc["means"], sc["scales"], sc["rots"], sc["opac"], col, view, full,
                                                           tx, ty, bg, W, H)[2]["color"]).double() * gimg.double()).sum())
    live = np.nonzero(np.abs(got["colors"]).sum(1) > 1e-3)[0][:4]
    assert len(live) > 0
    for i in live:
        for ch in range(3):
            hi, lo = sc["col"].clone(), sc["col"].clone()
            hi[i, ch] += 0.05; lo[i, ch] -= 0.05                  # the image is linear in colour: a wide step is exact
            fd = (loss(hi) - loss(lo)) / 0.1
            assert math.isclose(fd, got["colors"][i, ch], rel_tol=2e-3, abs_tol=2e-4), (i, ch, fd, got["colors"][i, ch])
