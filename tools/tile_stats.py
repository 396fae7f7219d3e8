"""Tile-list statistics of the two bench workloads (run on the GPU box): how many tiles are non-empty, how long the
lists are, how much of the blend work sits in the largest tiles.  Used to size the work decomposition of the blend
kernels."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import bench
from audio_motion_avatar_b200 import splat, synthetic as S
from audio_motion_avatar_b200.renderer import Renderer, render_multi_view

captured = []
_orig = splat.splat_forward


def _hook(*a, **k):
    k["want_aux"] = True
    out = _orig(*a, **k)
    captured.append({"tile_start": out["tile_start"].clone(), "n_contrib": out["n_contrib"].clone()})
    return out


splat.splat_forward = _hook
import audio_motion_avatar_b200.renderer as R
if hasattr(R, "splat_forward"):
    R.splat_forward = _hook


def report(name, cap, F, size):
    ts = cap["tile_start"].cpu().numpy().astype(np.int64) & 0xFFFFFFFF
    cnt = np.diff(ts)
    tiles = cnt.size // F
    nc = cap["n_contrib"].cpu().numpy().reshape(F, size // 16, 16, size // 16, 16).max(axis=(2, 4)).reshape(-1)
    print(f"{name}: F={F} tiles/frame={tiles} entries/frame={cnt.sum() / F:.0f} non-empty/frame={np.count_nonzero(cnt) / F:.0f}")
    nz = cnt[cnt > 0]
    print("  list length percentiles 50/90/99/max:", [int(np.percentile(nz, q)) for q in (50, 90, 99, 100)])
    print("  deepest-contributor (walk length) percentiles 50/90/99/max:", [int(np.percentile(nc[cnt > 0], q)) for q in (50, 90, 99, 100)],
          " sum walk / sum list = %.3f" % (nc.sum() / cnt.sum()))
    order = np.sort(nc)[::-1]
    print("  share of total walk in the top 1%% / 5%% of tiles: %.3f / %.3f; max tile = %.4f of total, total/max = %.0f" % (
        order[:max(1, order.size // 100)].sum() / order.sum(), order[:max(1, order.size // 20)].sum() / order.sum(),
        order[0] / order.sum(), order.sum() / order[0]))


def main():
    dev = torch.device("cuda:0")
    C, Rr = bench.C, bench.R
    for name, size, B, Tv in (("train c4", 512, 4, 4), ("sequence c2 (8 frames)", 1024, 8, 1)):
        N = 100000
        r = Renderer(bench.make_cfg(size, N, "cuda:0"), body_model=S.make_body_model(seed=0))
        r.gaussian_decoder.load_state_dict(S.make_decoder_weights(C, seed=0))
        r.fix_site_subset(seed=0)
        g = torch.Generator(device=dev).manual_seed(0)
        captured.clear()
        with torch.no_grad():
            if Tv > 1:
                tokens = torch.randn(B, 1, C, 3 * Rr * Rr, device=dev, generator=g)
                params = {k: v.reshape(B, 1, *v.shape[2:]).to(dev) for k, v in S.make_pose_sequence(1, B, seed=0).items()}
                params["transl"] = torch.tensor([0.0, 0.35, 0.0], device=dev).expand(B, 1, 3).contiguous()
                cams = {k: v.to(dev) for k, v in S.make_cameras(B, Tv, size, size, orbit=True).items()}
                img, gs = r(tokens, {k: v[:, :1] for k, v in cams.items()}, None, params)
                render_multi_view(gs, cams["intrinsic"][:, 1:], cams["extrinsic"][:, 1:], r.cfg)
                report(name + " view 0", captured[0], B, size)
                report(name + " views 1..", captured[1], B * (Tv - 1), size)
            else:
                tokens = torch.randn(1, B, C, 3 * Rr * Rr, device=dev, generator=g)
                params = {k: v.to(dev) for k, v in S.make_pose_sequence(1, B, seed=0).items()}
                cams = {k: v.to(dev) for k, v in S.make_cameras(1, B, size, size).items()}
                r(tokens, cams, None, params)
                report(name, captured[0], B, size)


if __name__ == "__main__":
    main()
