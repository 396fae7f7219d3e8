"""BASELINE.json's full sizes (configs[1]: 100k Gaussians at 1024x1024; configs[4]: 500k Gaussians at 2048x2048, level-3
densification) checked through size-independent properties of the path -- the CPU oracle cannot finish these sizes in
seconds, so nothing here calls it:
  * binning: D = sum of tiles_touched = tile_start[-1]; every list entry's tile lies inside its Gaussian's tile
    rectangle; every tile's list is strictly ascending in (depth bits, Gaussian index) = upstream's sorted key order;
  * blending: white Gaussians on a white background give a white image (sum of weights + final T = 1); halving the colours
    on a black background halves the image bit for bit (every product and FMA scales exactly by a power of two);
    n_contrib never exceeds the tile's list length; 0 <= final_T <= 1;
  * determinism and frame independence: a second run, and a frame rendered alone, are bit-identical."""
import pytest
import torch

from audio_motion_avatar_b200 import synthetic as S
from tests.test_renderer_gpu import make_cfg

pytestmark = pytest.mark.gpu
C, R = 512, 64


def _gaussians(N, size, steps, F):
    from audio_motion_avatar_b200.renderer import Renderer, camera_matrices
    r = Renderer(make_cfg(size, size, C, R, densify=True, steps=steps, num_gaussians=N), body_model=S.make_body_model(seed=0))
    r.gaussian_decoder.load_state_dict(S.make_decoder_weights(C, seed=0))
    r.fix_site_subset(0)
    tokens = S.make_triplane_tokens(F, C, R, seed=0)[0].cuda()
    params = {k: v.cuda() for k, v in S.make_pose_sequence(1, F, seed=0).items()}
    cams = S.make_cameras(1, F, size, size)
    with torch.no_grad():
        pts = r.get_smpl_vertices(params)
        g = r.decode_gaussians(tokens, pts, params["transl"].reshape(F, 3))
    view, full, tan = camera_matrices(cams["intrinsic"].reshape(F, 3, 3).cuda(), cams["extrinsic"].reshape(F, 4, 4).cuda(), size, size)
    assert g["xyz"].shape == (F, N, 3)
    return g, view, full, tan


def _splat(g, cam, size, color, bg, frames=None):
    from audio_motion_avatar_b200.splat import SplatWorkspace, splat_forward
    view, full, tan = cam
    sel = (lambda t: t) if frames is None else (lambda t: t[frames].contiguous())
    return splat_forward(sel(g["xyz"]), sel(g["scale"]), sel(g["rot"]), sel(g["opacity"]), sel(color), sel(view), sel(full),
                         sel(tan), torch.tensor(bg, device="cuda"), size, size, activate=True, hwc=True, want_aux=True,
                         workspace=SplatWorkspace())


@pytest.mark.parametrize("N,size,steps,F", [(100_000, 1024, 2, 3), (500_000, 2048, 3, 2)])
def test_full_size_properties(N, size, steps, F):
    g, view, full, tan = _gaussians(N, size, steps, F)
    cam = (view, full, tan)
    res = _splat(g, cam, size, g["color"], (0.0, 0.0, 0.0))
    torch.cuda.synchronize()
    gx = gy = (size + 15) // 16
    tiles = gx * gy
    D = res["D"]
    ts = res["tile_start"][:F * tiles + 1].long()
    assert D == int(res["tiles_touched"].long().sum().item()) == int(ts[-1].item())
    assert D > 3 * N * F and bool((ts[1:] >= ts[:-1]).all())

    # ---- binning ----
    pl = res["point_list"][:D].long()
    assert int(pl.min()) >= 0 and int(pl.max()) < N
    tile = torch.searchsorted(ts[1:].contiguous(), torch.arange(D, device="cuda"), right=True)   # tile of each entry
    fr, tin = tile // tiles, tile % tiles
    tx, ty = tin % gx, tin // gx
    rect = res["rect"].view(F, N, 4)[fr, pl].long() & 0xFFFF
    assert bool(((rect[:, 0] <= tx) & (tx < rect[:, 2]) & (rect[:, 1] <= ty) & (ty < rect[:, 3])).all())
    depth_bits = res["geom"].view(F, N, 12)[fr, pl, 10].view(torch.int32).long()               # depths are positive floats
    key = depth_bits * (1 << 20) + pl                                                            # N < 2^20
    same_tile = tile[1:] == tile[:-1]
    assert bool((key[1:][same_tile] > key[:-1][same_tile]).all()), "a tile's list is not in (depth, index) order"
    counts = ts[1:] - ts[:-1]
    nc = res["n_contrib"].view(F, gy, 16, gx, 16).permute(0, 1, 3, 2, 4).reshape(F * tiles, 256).long()
    assert bool((nc.max(dim=1).values <= counts).all())
    assert float(res["final_T"].min()) >= 0.0 and float(res["final_T"].max()) <= 1.0
    assert float(res["final_T"].min()) < 0.5 and float(res["final_T"].max()) == 1.0             # a body and empty background

    # ---- blending ----
    half = _splat(g, cam, size, g["color"] * 0.5, (0.0, 0.0, 0.0))
    assert torch.equal(half["color"], res["color"] * 0.5)
    assert torch.equal(half["point_list"][:D], res["point_list"][:D]) and torch.equal(half["n_contrib"], res["n_contrib"])
    white = _splat(g, cam, size, torch.ones_like(g["color"]), (1.0, 1.0, 1.0))
    assert float((white["color"] - 1.0).abs().max()) <= 2e-4

    # ---- determinism, frame independence ----
    again = _splat(g, cam, size, g["color"], (0.0, 0.0, 0.0))
    assert torch.equal(again["color"], res["color"]) and torch.equal(again["final_T"], res["final_T"])
    alone = _splat(g, cam, size, g["color"], (0.0, 0.0, 0.0), frames=slice(F - 1, F))
    assert torch.equal(alone["color"][0], res["color"][F - 1]) and torch.equal(alone["n_contrib"][0], res["n_contrib"][F - 1])
