"""Distance between the bit-exact oracle (polynomial exp + FMA nesting shared with the CUDA kernels) and upstream's literal
blend arithmetic (libm expf, `-0.5f*(A dx^2 + C dy^2) - B dx dy` without FMA; oracle/Makefile: libamav_oracle_upstream.so),
counted in blend decisions on the same sorted lists.  CPU only."""
import numpy as np
import pytest

from oracle import ref_path as O
from tests.helpers import oracle_frames, random_scene


@pytest.mark.parametrize("N,W,H", [(3000, 128, 96), (20000, 256, 256)])
def test_upstream_arithmetic_takes_the_same_blend_decisions(N, W, H):
    sc = random_scene(N, W, H, seed=N, frames=1)
    bg = (1.0, 0.5, 0.25)
    _, _, _, outs = oracle_frames(sc, H, W, bg)
    pre, binned, img = outs[0]
    with O.oracle_variant("upstream"):
        up = O.raster_blend(pre, binned, sc["col"], bg, W, H)
    flips = int((up["n_contrib"] != img["n_contrib"]).sum())
    dcol = np.abs(up["color"] - img["color"])
    # a pair on the alpha >= 1/255 or T < 1e-4 cut may flip (2 ulp of exp): bounded count, bounded effect
    assert flips <= max(2, H * W // 5000), flips
    assert (dcol.max(axis=0) > 1e-5).sum() <= max(2, H * W // 5000)
    assert dcol.max() <= 1.0 / 255.0 + 1e-3
    mse = float((dcol.astype(np.float64) ** 2).mean())
    assert mse <= 1e-9                                   # PSNR >= 90 dB


def test_fast_build_is_bit_identical_to_the_parity_build():
    """The -O3 -march=native build that bench.py times (cpu_baseline, --impl reference) computes the same bits."""
    sc = random_scene(4000, 160, 128, seed=3, frames=1)
    _, _, _, a = oracle_frames(sc, 128, 160)
    with O.oracle_variant("fast"):
        _, _, _, b = oracle_frames(sc, 128, 160)
    for x, y in zip(a[0], b[0]):
        for k in x:
            if isinstance(x[k], np.ndarray):
                assert np.array_equal(x[k], y[k]), k
