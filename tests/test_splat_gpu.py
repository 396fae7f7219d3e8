"""Stage 4 parity on the GPU: CUDA splat path vs oracle/raster_oracle.c on identical inputs.
Bit-exact: depth bits, radii, tile rectangles, tile ranges, sorted per-tile lists (= upstream keys
and sort order), n_contrib, final_T and colours (the blend arithmetic is FMA-for-FMA the oracle's)."""
import numpy as np
import pytest
import torch

from tests.helpers import oracle_cameras, oracle_frames, random_scene, psnr

pytestmark = pytest.mark.gpu


def _run_cuda(sc, view, full, tan, H, W, bg, **kw):
    from audio_motion_avatar_b200.splat import splat_forward
    d = torch.device("cuda:0")
    c = lambda t: t.to(d)
    out = splat_forward(c(sc["means"])[None], c(sc["scales"])[None], c(sc["rots"])[None], c(sc["opac"])[None],
                        c(sc["col"])[None], c(view), c(full), c(tan), c(torch.tensor(bg, dtype=torch.float32)), H, W,
                        views_per_set=view.shape[0], want_aux=True, want_invdepth=True, **kw)
    torch.cuda.synchronize()
    return out


def _check_frame(out, f, pre, binned, img, N, H, W):
    geom = out["geom"][f].cpu().numpy()
    live = pre["radii"] > 0
    assert np.array_equal(out["tiles_touched"][f].cpu().numpy(), pre["tiles_touched"])
    assert np.array_equal(out["rect"][f].cpu().numpy().astype(np.int32)[live], pre["rect"][live])
    assert np.array_equal(geom[live, 10].view(np.uint32), pre["depth"][live].view(np.uint32))       # depth bits
    assert np.array_equal(geom[live, 11].astype(np.int32), pre["radii"][live])
    assert np.array_equal(geom[live, 0:2].view(np.uint32), pre["xy"][live].view(np.uint32))
    con = pre["conic_opacity"][live]
    assert np.array_equal(geom[live, 2], -0.5 * con[:, 0]) and np.array_equal(geom[live, 3], -con[:, 1])
    assert np.array_equal(geom[live, 4], -0.5 * con[:, 2]) and np.array_equal(geom[live, 5], con[:, 3])
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    ts = out["tile_start"][f * tiles:(f + 1) * tiles + 1].cpu().numpy().astype(np.int64)
    base = ts[0]
    cnt = binned["ranges"][:, 1] - binned["ranges"][:, 0]
    assert np.array_equal(np.diff(ts), cnt)
    nonempty = cnt > 0
    assert np.array_equal((ts[:-1] - base)[nonempty], binned["ranges"][nonempty, 0])
    pl = out["point_list"][base:base + binned["D"]].cpu().numpy()
    assert np.array_equal(pl, binned["point_list"])                                                # sort order
    # upstream keys rebuilt from (tile, depth bits of the listed Gaussian)
    tile_of = np.repeat(np.arange(tiles, dtype=np.uint64), cnt)
    keys = (tile_of << np.uint64(32)) | geom[pl, 10].view(np.uint32).astype(np.uint64)
    assert np.array_equal(keys, binned["keys"])
    assert np.array_equal(out["n_contrib"][f].cpu().numpy(), img["n_contrib"])
    assert np.array_equal(out["final_T"][f].cpu().numpy(), img["final_T"])
    col = out["color"][f].cpu().numpy()
    assert np.array_equal(col, img["color"]), f"max diff {np.abs(col - img['color']).max()}"
    assert np.array_equal(out["invdepth"][f, 0].cpu().numpy(), img["invdepth"])


@pytest.mark.parametrize("N,W,H,frames", [(3000, 128, 96, 3), (20000, 256, 256, 2), (500, 70, 45, 1)])
def test_splat_bit_exact_vs_oracle(N, W, H, frames):
    sc = random_scene(N, W, H, seed=N, frames=frames)
    bg = (1.0, 0.5, 0.25)
    view, full, tan, outs = oracle_frames(sc, H, W, bg)
    out = _run_cuda(sc, view, full, tan, H, W, bg)
    assert out["D"] == sum(o[1]["D"] for o in outs)
    for f, (pre, binned, img) in enumerate(outs):
        _check_frame(out, f, pre, binned, img, N, H, W)


def test_splat_empty_and_degenerate():
    from audio_motion_avatar_b200.splat import splat_forward
    d = torch.device("cuda:0")
    H, W = 40, 56
    sc = random_scene(16, W, H, seed=1)
    sc["means"][:, 2] = -9.0                                   # everything culled
    view, full, tan, outs = oracle_frames(sc, H, W)
    out = _run_cuda(sc, view, full, tan, H, W, (1.0, 1.0, 1.0))
    assert out["D"] == 0 and torch.all(out["color"] == 1.0)
    z = torch.zeros(1, 0, 3, device=d)
    o0 = splat_forward(z, z, torch.zeros(1, 0, 4, device=d), torch.zeros(1, 0, 1, device=d), z, view.to(d), full.to(d),
                       tan.to(d), torch.zeros(3, device=d), H, W)
    assert torch.all(o0["color"] == 0.0)


@pytest.mark.parametrize("N,min_largest", [(12000, 8192), (24000, 16384), (50000, 32768)])
def test_splat_large_buckets(N, min_largest):
    # Thousands of Gaussians stacked on a few tiles.  12000: buckets of 4097..16384 entries (the 1024-thread
    # shared-memory instantiation of the per-tile sort) next to small ones; 24000: a bucket above 16384 entries
    # (bitonic network in place in global memory, strides below 16384 on chunks in shared memory: one merge level
    # across two chunks); 50000: above 32768 (two cross-chunk merge levels incl. a cross-chunk half-cleaner).
    W, H = 32, 32
    sc = random_scene(N, W, H, seed=5, spread=0.05, smin=0.002, smax=0.004)
    sc["opac"] *= 0.02
    view, full, tan, outs = oracle_frames(sc, H, W)
    pre, binned, img = outs[0]
    sizes = binned["ranges"][:, 1] - binned["ranges"][:, 0]
    assert sizes.max() > min_largest and (N > 16384 or sizes.max() <= 16384)
    out = _run_cuda(sc, view, full, tan, H, W, (1.0, 1.0, 1.0))
    _check_frame(out, 0, pre, binned, img, N, H, W)


def test_hwc_clamp_layout_and_activation():
    N, W, H = 2000, 96, 64
    sc = random_scene(N, W, H, seed=9)
    view, full, tan, outs = oracle_frames(sc, H, W)
    a = _run_cuda(sc, view, full, tan, H, W, (1.0, 1.0, 1.0))["color"].clone()
    b = _run_cuda(sc, view, full, tan, H, W, (1.0, 1.0, 1.0), hwc=True, clamp01=True)["color"]
    assert torch.equal(a.clamp(0, 1).permute(0, 2, 3, 1), b)
    # raw -> activated inside the kernel (render_one, renderer.py:532-533,547): tolerance, not bits
    raw = dict(sc)
    raw["scales"] = torch.log(sc["scales"]) + 3.9
    raw["opac"] = torch.logit(sc["opac"].clamp(1e-4, 1 - 1e-4))
    sc2 = dict(sc); sc2["opac"] = sc["opac"].clamp(1e-4, 1 - 1e-4)
    _, _, _, outs2 = oracle_frames(sc2, H, W)
    c = _run_cuda(raw, view, full, tan, H, W, (1.0, 1.0, 1.0), activate=True)["color"][0].cpu().numpy()
    assert np.abs(c - outs2[0][2]["color"]).max() <= 1e-3 and psnr(c, outs2[0][2]["color"]) >= 60.0


def test_bucket_overflow_is_flagged_not_fatal_and_renderer_recovers():
    """sync_free mode trusts the cached entry capacity; a scene that needs more sets the overflow flag, reads nothing
    out of bounds, and render_batch re-renders with the sizing read."""
    from types import SimpleNamespace
    from audio_motion_avatar_b200.splat import SplatWorkspace, splat_forward
    from audio_motion_avatar_b200 import renderer as Rn
    d = torch.device("cuda:0")
    H, W = 128, 128
    small, big = random_scene(200, W, H, seed=1), random_scene(20000, W, H, seed=2)
    ws = SplatWorkspace()

    def run(sc, **kw):
        view, full, tan = oracle_cameras(sc, H, W)
        c = lambda t: t.to(d)
        return splat_forward(c(sc["means"])[None], c(sc["scales"])[None], c(sc["rots"])[None], c(sc["opac"])[None],
                             c(sc["col"])[None], c(view), c(full), c(tan), torch.ones(3, device=d), H, W, workspace=ws, **kw)

    run(small)
    cap = ws.capacity
    out = run(big, sync_free=True)
    torch.cuda.synchronize()
    assert out["D"] == -1 and ws.capacity == cap and int(ws.overflow.item()) == 1
    with pytest.raises(Exception):
        ws.check()
    ref = run(big)["color"].clone()
    # through the public API: first a small call to seed the per-device workspace, then the big one
    args = SimpleNamespace(image_size=(H, W), rgb=True, sh_degree=3)
    Rn._WORKSPACES.clear()

    def gs(sc):
        return {"xyz": sc["means"][None].to(d), "rot": sc["rots"][None].to(d), "scale": (torch.log(sc["scales"]) + 3.9)[None].to(d),
                "opacity": torch.logit(sc["opac"].clamp(1e-4, 1 - 1e-4))[None].to(d), "color": sc["col"][None].to(d)}

    K, E = small["K"][None].to(d), small["E"][None].to(d)
    Rn.render_batch(gs(small), K, E, args)
    a = Rn.render_batch(gs(big), K, E, args)
    Rn._WORKSPACES.clear()
    b = Rn.render_batch(gs(big), K, E, args)
    assert torch.equal(a, b)


def test_equal_depths_fall_back_to_the_sorting_network():
    """A fronto-parallel sheet: every Gaussian has the same depth bits, so the depth-bin sort cannot separate them and
    the tile is sorted by the bitonic network on (depth, index); ties must come out in ascending Gaussian index."""
    N, W, H = 1500, 64, 48
    sc = random_scene(N, W, H, seed=21, spread=0.3, smin=0.004, smax=0.01)
    sc["means"][:, 2] = 0.125
    sc["E"][0, :3, :3] = torch.eye(3)
    view, full, tan, outs = oracle_frames(sc, H, W)
    pre, binned, img = outs[0]
    live = pre["radii"] > 0
    assert len(np.unique(pre["depth"][live].view(np.uint32))) == 1
    out = _run_cuda(sc, view, full, tan, H, W, (1.0, 1.0, 1.0))
    _check_frame(out, 0, pre, binned, img, N, H, W)


@pytest.mark.parametrize("W,H", [(96, 72), (64, 40), (1024, 1024)])
def test_row_store_variant_equals_per_pixel_stores(W, H):
    """The blend instantiation used when the frames are written into another GPU's buffer (tile pairs, colour leaves
    as 16 rows of 384 bytes) produces the same HWC frames as the per-pixel stores, including the rows below the image
    edge in the last tile row.  row_stores=True forces it on a local buffer (AMAV_LAYOUT_HWC_ROWS; the library otherwise
    picks it by cudaPointerGetAttributes); W must be a multiple of 32, other widths are rejected."""
    from audio_motion_avatar_b200._lib import AmavError
    from audio_motion_avatar_b200.splat import splat_forward
    N = 4000 if W < 1024 else 60000
    sc = random_scene(N, W, H, seed=21)
    view, full, tan = oracle_cameras(sc, H, W)
    d = torch.device("cuda:0")
    c = lambda t: t.to(d)

    def run(rows, width=W):
        out = splat_forward(c(sc["means"])[None], c(sc["scales"])[None], c(sc["rots"])[None], c(sc["opac"])[None],
                            c(sc["col"])[None], c(view), c(full), c(tan), torch.tensor([0.2, 0.5, 0.9], device=d), H, width,
                            views_per_set=view.shape[0], hwc=True, clamp01=True, row_stores=rows)["color"].clone()
        torch.cuda.synchronize()
        return out

    ref = run(False)
    rows = run(True)
    auto = run(None)                                   # a local buffer: the library keeps the per-pixel stores
    assert ref.shape == (view.shape[0], H, W, 3) and torch.equal(ref, rows) and torch.equal(ref, auto)
    assert (ref != torch.tensor([0.2, 0.5, 0.9], device=d)).any()          # something was actually rendered
    with pytest.raises(AmavError):
        run(True, width=W + 8)


def test_scan_flags_a_wrapped_entry_count():
    """tile offsets are 32-bit: a call whose entry count reaches 2^32 must not wrap silently.  The scan writes
    the sentinel D = 0xffffffff, on which splat_forward raises and the scatter kernel sets the overflow flag."""
    from audio_motion_avatar_b200 import _lib
    from audio_motion_avatar_b200._lib import SplatDims, call, ptr, stream_ptr
    d = torch.device("cuda:0")
    W = H = 64                                                  # 16 tiles per frame, 2 frames
    dims = SplatDims(2, 1, 1, W, H, 0, 1.0)
    nt = 32
    count = torch.full((nt,), 150_000_000, dtype=torch.int32, device=d)          # 32 x 1.5e8 = 4.8e9 > 2^32
    start = torch.zeros(nt + 1, dtype=torch.int32, device=d)
    tmp = torch.zeros(int(_lib.lib().amav_splat_scan_tmp_elems(nt)), dtype=torch.int32, device=d)
    call("amav_splat_scan", dims, ptr(count), ptr(start), ptr(tmp), stream_ptr())
    torch.cuda.synchronize()
    assert int(start[nt].item()) & 0xFFFFFFFF == 0xFFFFFFFF
    count.fill_(1000)
    call("amav_splat_scan", dims, ptr(count), ptr(start), ptr(tmp), stream_ptr())
    torch.cuda.synchronize()
    assert int(start[nt].item()) == 32000 and int(start[5].item()) == 5000
