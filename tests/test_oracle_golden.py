"""The oracle's restatements vs outputs of the reference's own functions (tests/golden/make_golden.py)."""
import os

import numpy as np
import torch

from oracle import ref_path as O


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_sample_from_triplane_matches_reference(golden_dir):
    g = _load(golden_dir, "ref_query_decode.npz")
    feats = O.sample_from_triplane(torch.from_numpy(g["triplane"]), torch.from_numpy(g["points"]), float(g["radius"]))
    ref = torch.from_numpy(g["feats"])
    assert feats.shape == ref.shape
    assert (feats - ref).abs().max().item() <= 2e-6      # grid_sample vs explicit bilinear: fp32 rounding only


def test_decode_and_construct_match_reference(golden_dir):
    g = _load(golden_dir, "ref_query_decode.npz")
    w = {k[2:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("w_")}
    pts = torch.from_numpy(g["points"])
    dec = O.decode_heads(pts, torch.from_numpy(g["feats"]), w)
    for name, _ in O.HEADS:
        assert (dec[name] - torch.from_numpy(g[f"dec_{name}"])).abs().max().item() <= 5e-6, name
    gs = O.construct_gaussians(dec, pts, torch.from_numpy(g["transl"]))
    for k in ("xyz", "scale", "rot", "opacity", "color"):
        assert (gs[k] - torch.from_numpy(g[f"g_{k}"])).abs().max().item() <= 5e-6, k
    assert gs["shs"] is gs["color"]


def test_camera_setup_matches_reference(golden_dir):
    g = _load(golden_dir, "ref_camera.npz")
    for i in range(6):
        H, W = int(g[f"H_{i}"]), int(g[f"W_{i}"])
        view, full, tfx, tfy, campos = O.camera_setup(torch.from_numpy(g[f"K_{i}"]), torch.from_numpy(g[f"E_{i}"]), H, W)
        # same torch ops on the same CPU: bit-exact
        assert np.array_equal(view.numpy(), g[f"view_{i}"])
        assert np.array_equal(full.numpy(), g[f"full_{i}"])
        assert tfx == float(g[f"tanfovx_{i}"]) and tfy == float(g[f"tanfovy_{i}"])
        assert np.array_equal(campos.numpy(), g[f"campos_{i}"])


def test_activation_matches_reference(golden_dir):
    g = _load(golden_dir, "ref_activation.npz")
    assert float(g["SCALE_BIAS"]) == O.SCALE_BIAS and float(g["OPACITY_BIAS"]) == O.OPACITY_BIAS
    s, o, c = O.activate_for_raster({"scale": torch.from_numpy(g["s"]), "opacity": torch.from_numpy(g["o"]),
                                     "color": torch.from_numpy(g["c"])})
    assert np.array_equal(s.numpy(), g["scales"])
    assert np.array_equal(o.numpy(), g["opac"])
    assert np.array_equal(c.numpy(), g["col"])


def test_grid_sample_baseline_equals_explicit_restatement(golden_dir):
    g = _load(golden_dir, "ref_query_decode.npz")
    tri, pts = torch.from_numpy(g["triplane"]), torch.from_numpy(g["points"])
    a = O.sample_from_triplane_torch(tri, pts, float(g["radius"]))
    assert (a - torch.from_numpy(g["feats"])).abs().max().item() <= 1e-6
    assert (a - O.sample_from_triplane(tri, pts, float(g["radius"]))).abs().max().item() <= 2e-6


def test_image_losses_match_reference():
    """oracle/loss_oracle.py vs the reference's own l1_loss / ssim and autograd through them (ref_loss.npz)."""
    import os
    from oracle import loss_oracle as L
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_loss.npz"))
    img = torch.from_numpy(d["img"]).clone().requires_grad_(True)
    gt = torch.from_numpy(d["gt"])
    l1, s = L.l1_loss(img, gt), L.ssim(img, gt)
    assert abs(l1.item() - float(d["l1"])) <= 1e-7 and abs(s.item() - float(d["ssim"])) <= 1e-6
    per = L.ssim(img, gt, size_average=False)
    assert np.abs(per.detach().numpy() - d["ssim_per_frame"]).max() <= 1e-6
    loss = L.l1_ssim_loss(img, gt)
    assert abs(loss.item() - float(d["loss"])) <= 1e-6
    loss.backward()
    assert np.abs(img.grad.numpy() - d["grad"]).max() <= 1e-9 + 1e-5 * np.abs(d["grad"]).max()
