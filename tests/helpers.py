"""Shared builders for the parity tests."""
import math

import numpy as np
import torch

from oracle import ref_path as O


def random_scene(N, W, H, seed=0, frames=1, spread=0.6, smin=0.005, smax=0.05):
    """Activated Gaussians (one set shared by all frames) + per-frame cameras on an arc."""
    g = torch.Generator().manual_seed(seed)
    means = torch.cat([torch.rand(N, 2, generator=g) * 2 * spread - spread,
                       torch.rand(N, 1, generator=g) * 0.6 - 0.3], 1)
    if N >= 8:
        means[:3, 2] = -5.0          # behind every camera: near-culled
        means[3, 0] = 40.0           # far outside the frustum
    scales = torch.rand(N, 3, generator=g) * (smax - smin) + smin
    rots = torch.nn.functional.normalize(torch.randn(N, 4, generator=g), dim=-1)
    opac = torch.rand(N, 1, generator=g)
    col = torch.rand(N, 3, generator=g)
    Ks, Es = [], []
    for f in range(frames):
        ang = 0.5 * f
        c, s = math.cos(ang), math.sin(ang)
        E = torch.eye(4)
        E[:3, :3] = torch.tensor([[c, 0.0, -s], [0.0, 1.0, 0.0], [s, 0.0, c]])
        E[:3, 3] = torch.tensor([0.02 * f, -0.01 * f, 2.0])
        Ks.append(torch.tensor([[1.1 * W, 0, W / 2 + 0.3], [0, 1.05 * W, H / 2 - 0.2], [0, 0, 1.0]]))
        Es.append(E)
    return dict(means=means, scales=scales, rots=rots, opac=opac, col=col, K=torch.stack(Ks), E=torch.stack(Es))


def oracle_cameras(sc, H, W):
    views, fulls, tans = [], [], []
    for K, E in zip(sc["K"], sc["E"]):
        v, fu, tx, ty, _ = O.camera_setup(K, E, H, W)
        views.append(v); fulls.append(fu); tans.append([tx, ty])
    return torch.stack(views), torch.stack(fulls), torch.tensor(tans, dtype=torch.float32)


def oracle_frames(sc, H, W, bg=(1.0, 1.0, 1.0)):
    view, full, tan = oracle_cameras(sc, H, W)
    outs = []
    for f in range(view.shape[0]):
        outs.append(O.rasterize(sc["means"], sc["scales"], sc["rots"], sc["opac"], sc["col"], view[f], full[f],
                                float(tan[f, 0]), float(tan[f, 1]), bg, W, H))
    return view, full, tan, outs


def psnr(a, b):
    mse = float(np.mean((np.asarray(a, np.float64) - np.asarray(b, np.float64)) ** 2))
    return 999.0 if mse == 0 else 10.0 * math.log10(1.0 / mse)
