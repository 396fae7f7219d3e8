"""Host side of stage 4 (tile-based Gaussian splatting) over the C ABI of libamav_b200.so.

Replaces the per-frame ``GaussianRasterizer`` calls of the reference (src/models/renderer.py:475-477,
516-566) by one batched pass over all frames of a call: preprocess -> tile scan -> bucket scatter ->
per-tile sort -> blend, with at most ONE device->host read per call (the entry count D, to size the
bucket buffer; skipped when the cached workspace is already large enough... see ``SplatWorkspace``).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional

import torch

from . import _lib
from ._lib import SplatDims, call, ptr, stream_ptr

TILE = 16


def tile_grid(height: int, width: int):
    return (width + TILE - 1) // TILE, (height + TILE - 1) // TILE


@dataclass
class SplatWorkspace:
    """Reusable device buffers.  ``capacity`` (entries) only grows; a call whose D fits does not
    re-allocate, and with ``sync_free=True`` does not read D back either: it trusts the capacity
    and raises at the next ``check()`` if the overflow flag was set."""
    capacity: int = 0
    entries: Optional[torch.Tensor] = None
    point_list: Optional[torch.Tensor] = None
    bufs: Dict[str, torch.Tensor] = field(default_factory=dict)
    overflow: Optional[torch.Tensor] = None
    last_D: int = -1
    # entries per frame seen at the last sizing read, per (N, H, W, views_per_set): a call of another shape (or with
    # more frames than the cached capacity covers) must not trust the capacity a different workload left behind
    per_frame: Dict[tuple, float] = field(default_factory=dict)

    def covers(self, key: tuple, frames: int) -> bool:
        est = self.per_frame.get(key)
        return est is not None and self.capacity > 0 and self.capacity >= est * frames * 1.1

    def get(self, name: str, numel: int, dtype, device, zero: bool = False) -> torch.Tensor:
        t = self.bufs.get(name)
        if t is None or t.numel() < numel or t.dtype != dtype or t.device != device:
            t = torch.empty(max(numel, 1), dtype=dtype, device=device)
            self.bufs[name] = t
        t = t[:numel]
        if zero:
            t.zero_()
        return t

    def ensure_capacity(self, D: int, device) -> None:
        if self.entries is None or self.capacity < D or self.entries.device != device:
            self.capacity = max(int(D * 1.25) + 1024, 1024)
            self.entries = torch.empty(self.capacity, dtype=torch.int64, device=device)
            self.point_list = torch.empty(self.capacity, dtype=torch.int32, device=device)

    def check(self) -> None:
        if self.overflow is not None and int(self.overflow.item()) != 0:
            raise _lib.AmavError("splat bucket capacity overflow: rendered frames are incomplete; "
                                 "call again with sync_free=False")


def splat_forward(means3d: torch.Tensor, scales: torch.Tensor, rots: torch.Tensor, opacities: torch.Tensor,
                  colors: torch.Tensor, view: torch.Tensor, proj: torch.Tensor, tanfov: torch.Tensor,
                  bg: torch.Tensor, height: int, width: int, *, views_per_set: int = 1, activate: bool = False,
                  hwc: bool = False, clamp01: bool = False, want_aux: bool = False, want_invdepth: bool = False,
                  scale_modifier: float = 1.0, workspace: Optional[SplatWorkspace] = None,
                  sync_free: bool = False, out: Optional[torch.Tensor] = None,
                  row_stores: Optional[bool] = None) -> Dict[str, torch.Tensor]:
    """Render F frames.

    means3d/scales/colors [G,N,3], rots [G,N,4], opacities [G,N,1] with G = F // views_per_set
    Gaussian sets; view/proj [F,4,4] (already transposed as the reference passes them), tanfov [F,2],
    bg [3].  Returns a dict with ``color`` ([F,3,H,W], or [F,H,W,3] when ``hwc``) plus the
    intermediate buffers (``geom``, ``rect``, ``tiles_touched``, ``tile_start``, ``point_list``, ``D``
    and, with ``want_aux``, ``final_T`` / ``n_contrib``) used by the backward pass and the parity tests.
    ``row_stores`` (HWC inference only): None lets the library decide (row stores when ``out`` is another GPU's memory),
    True / False force the choice (include/amav.h: AMAV_LAYOUT_HWC_ROWS / _PIXELS).
    """
    dev = means3d.device
    if dev.type != "cuda":
        raise _lib.AmavError("splat_forward needs CUDA tensors (no CPU path)")
    F = int(view.shape[0])
    N = int(means3d.shape[-2])
    G = means3d.numel() // max(N * 3, 1) if N > 0 else F // views_per_set
    if F != G * views_per_set:
        raise ValueError(f"frames ({F}) != gaussian sets ({G}) * views_per_set ({views_per_set})")
    f32 = lambda t: t.detach().to(torch.float32).contiguous()
    means3d, scales, rots, opacities, colors = map(f32, (means3d, scales, rots, opacities, colors))
    view, proj, tanfov, bg = map(f32, (view, proj, tanfov, bg))
    ws = workspace if workspace is not None else SplatWorkspace()
    gx, gy = tile_grid(height, width)
    nt = F * gx * gy
    dims = SplatDims(F, N, views_per_set, width, height, int(activate), float(scale_modifier))
    st = stream_ptr()

    geom = ws.get("geom", F * N * 12, torch.float32, dev)
    rect = ws.get("rect", F * N * 4, torch.int16, dev)
    touched = ws.get("tiles_touched", F * N, torch.int32, dev)
    counters = ws.get("counters", 2 * nt + 3, torch.int32, dev, zero=True)  # tile_count | tile_cursor | overflow, sort queue | blend queue
    tile_count, tile_cursor, overflow = counters[:nt], counters[nt:2 * nt], counters[2 * nt:2 * nt + 2]
    work_counter = counters[2 * nt + 2:2 * nt + 3]
    tile_start = ws.get("tile_start", nt + 1, torch.int32, dev)
    scan_tmp = ws.get("scan_tmp", int(_lib.lib().amav_splat_scan_tmp_elems(nt)), torch.int32, dev)
    tile_order = scan_tmp[int(_lib.lib().amav_splat_scan_order_offset(nt)):]      # deep tiles first (written by the scan)
    ws.overflow = overflow[:1]

    call("amav_splat_preprocess", dims, ptr(means3d), ptr(scales), ptr(rots), ptr(opacities), ptr(colors),
         ptr(view), ptr(proj), ptr(tanfov), ptr(geom), ptr(rect), ptr(touched), ptr(tile_count), st)
    call("amav_splat_scan", dims, ptr(tile_count), ptr(tile_start), ptr(scan_tmp), st)
    if sync_free and ws.capacity > 0:
        D = -1
    else:
        D = int(tile_start[nt].item()) & 0xFFFFFFFF
        if D == 0xFFFFFFFF:
            raise _lib.AmavError("more than 2^32 - 2 (Gaussian, tile) entries in one call: the 32-bit tile offsets cannot "
                                 "address them; render fewer frames per call")
        ws.ensure_capacity(D, dev)
        ws.last_D = D
        ws.per_frame[(N, height, width, views_per_set)] = D / max(F, 1)
    call("amav_splat_bin", dims, ptr(geom), ptr(rect), ptr(tile_start), ptr(tile_cursor), ptr(ws.entries),
         ptr(ws.point_list), ws.capacity, ptr(overflow), st)

    shape = (F, height, width, 3) if hwc else (F, 3, height, width)
    if out is None:
        out = torch.empty(shape, dtype=torch.float32, device=dev)
    elif tuple(out.shape) != shape or not out.is_contiguous():
        raise ValueError("`out` has the wrong shape")
    final_T = torch.empty(F, height, width, dtype=torch.float32, device=dev) if want_aux else None
    n_contrib = torch.empty(F, height, width, dtype=torch.int32, device=dev) if want_aux else None
    invdepth = torch.empty(F, 1, height, width, dtype=torch.float32, device=dev) if want_invdepth else None
    layout = 0 if not hwc else 1 if row_stores is None else 2 if row_stores else 3
    call("amav_splat_blend", dims, ptr(geom), ptr(tile_start), ptr(ws.point_list), ws.capacity, ptr(bg), layout, int(clamp01),
         ptr(out, peer=True), ptr(final_T), ptr(n_contrib), ptr(invdepth), ptr(work_counter), ptr(tile_order), st)
    return {"color": out, "tile_order": tile_order, "geom": geom.view(F, N, 12), "rect": rect.view(F, N, 4), "tiles_touched": touched.view(F, N),
            "tile_start": tile_start, "point_list": ws.point_list, "D": D, "final_T": final_T,
            "n_contrib": n_contrib, "invdepth": invdepth, "workspace": ws}


def splat_backward(fwd: Dict[str, torch.Tensor], grad_color: torch.Tensor, means3d, scales, rots, opacities, colors,
                   view, proj, tanfov, bg, height: int, width: int, *, views_per_set: int = 1, activate: bool = False,
                   hwc: bool = False, scale_modifier: float = 1.0) -> Dict[str, torch.Tensor]:
    """Gradients of the splat inputs from dL/d(colour output): blend backward (per-(frame, Gaussian) screen-space
    sums) then the per-Gaussian chain.  ``fwd`` is the dict ``splat_forward(..., want_aux=True)`` returned (its
    buffers must still hold that call's data).  Replaces _RasterizeGaussians.backward + the autograd of render_one's
    activations (renderer.py:532-533,547,555-568)."""
    dev = grad_color.device
    F = int(view.shape[0])
    N = int(means3d.shape[-2])
    G = F // views_per_set
    f32 = lambda t: t.detach().to(torch.float32).contiguous()
    means3d, scales, rots, opacities, colors = map(f32, (means3d, scales, rots, opacities, colors))
    view, proj, tanfov, bg, grad_color = map(f32, (view, proj, tanfov, bg, grad_color))
    dims = SplatDims(F, N, views_per_set, width, height, int(activate), float(scale_modifier))
    st = stream_ptr()
    dsum = torch.zeros(F, N, 12, dtype=torch.float32, device=dev)
    counter = torch.zeros(1, dtype=torch.int32, device=dev)
    ws = fwd["workspace"]
    geom = fwd["geom"].reshape(-1)
    call("amav_splat_blend_backward", dims, ptr(geom), ptr(fwd["tile_start"]), ptr(fwd["point_list"]), ws.capacity,
         ptr(bg), int(hwc), ptr(fwd["final_T"]), ptr(fwd["n_contrib"]), ptr(grad_color), ptr(dsum), ptr(counter), ptr(fwd.get("tile_order")), st)
    d_means = torch.empty(G, N, 3, device=dev)
    d_scales = torch.empty(G, N, 3, device=dev)
    d_rots = torch.empty(G, N, 4, device=dev)
    d_opac = torch.empty(G, N, 1, device=dev)
    d_colors = torch.empty(G, N, 3, device=dev)
    call("amav_splat_preprocess_backward", dims, ptr(means3d), ptr(scales), ptr(rots), ptr(opacities), ptr(colors),
         ptr(view), ptr(proj), ptr(tanfov), ptr(geom), ptr(dsum), ptr(d_means), ptr(d_scales), ptr(d_rots),
         ptr(d_opac), ptr(d_colors), st)
    return {"means3d": d_means, "scales": d_scales, "rots": d_rots, "opacities": d_opac, "colors": d_colors, "dsum": dsum}


# Training under a CUDA graph: SplatFunction normally sizes its bucket buffer with one host read of D per call, which a
# stream capture cannot contain.  Eager steps record the largest D seen per call shape; inside ``fixed_training_capacity()``
# the buffers are sized from that record (x1.25), D is not read, and every call's overflow flag is kept for
# ``training_overflow_flags()`` to be read after the replays (an overflow means: run eager steps again, recapture).
_TRAIN_D: Dict[tuple, int] = {}
_TRAIN_FIXED = False
_TRAIN_FLAGS: list = []


class fixed_training_capacity:
    def __enter__(self):
        global _TRAIN_FIXED
        self._prev, _TRAIN_FIXED = _TRAIN_FIXED, True
        _TRAIN_FLAGS.clear()
        return self

    def __exit__(self, *exc):
        global _TRAIN_FIXED
        _TRAIN_FIXED = self._prev


def training_overflow_flags() -> list:
    """Overflow flags (device tensors) of the SplatFunction calls made in the last ``fixed_training_capacity()`` block; for
    a captured step they are the graph's own static tensors, valid after every replay."""
    return list(_TRAIN_FLAGS)


class SplatFunction(torch.autograd.Function):
    """Differentiable batched splat: forward = splat_forward (own workspace, aux kept), backward = splat_backward.
    Cameras, background and the image size carry no gradient (as in the reference, where they are plain settings)."""

    @staticmethod
    def forward(ctx, means3d, scales, rots, opacities, colors, view, proj, tanfov, bg, height, width, views_per_set,
                activate, hwc, clamp01, scale_modifier):
        ws = SplatWorkspace()
        key = (int(view.shape[0]), int(means3d.shape[-2]), int(height), int(width), int(views_per_set), means3d.device.index)
        fixed = _TRAIN_FIXED and key in _TRAIN_D
        if fixed:
            ws.ensure_capacity(_TRAIN_D[key], means3d.device)
        fwd = splat_forward(means3d, scales, rots, opacities, colors, view, proj, tanfov, bg, height, width,
                            views_per_set=views_per_set, activate=activate, hwc=hwc, clamp01=clamp01, want_aux=True,
                            scale_modifier=scale_modifier, workspace=ws, sync_free=fixed)
        if fixed:
            _TRAIN_FLAGS.append(ws.overflow.clone())
        else:
            _TRAIN_D[key] = max(_TRAIN_D.get(key, 0), int(fwd["D"]))
        ctx.fwd = fwd
        ctx.args = (height, width, views_per_set, activate, hwc, scale_modifier)
        ctx.save_for_backward(means3d, scales, rots, opacities, colors, view, proj, tanfov, bg)
        ctx.shapes = tuple(t.shape for t in (means3d, scales, rots, opacities, colors))
        return fwd["color"]

    @staticmethod
    def backward(ctx, grad_color):
        means3d, scales, rots, opacities, colors, view, proj, tanfov, bg = ctx.saved_tensors
        height, width, vps, activate, hwc, smod = ctx.args
        g = splat_backward(ctx.fwd, grad_color, means3d, scales, rots, opacities, colors, view, proj, tanfov, bg,
                           height, width, views_per_set=vps, activate=activate, hwc=hwc, scale_modifier=smod)
        ctx.fwd = None
        outs = [g[k].reshape(s) for k, s in zip(("means3d", "scales", "rots", "opacities", "colors"), ctx.shapes)]
        return (*outs, None, None, None, None, None, None, None, None, None, None, None)


def splat_render(means3d, scales, rots, opacities, colors, view, proj, tanfov, bg, height, width, *,
                 views_per_set=1, activate=False, hwc=False, clamp01=False, scale_modifier=1.0) -> torch.Tensor:
    """Autograd-aware colour output of the splat ([F,3,H,W] or [F,H,W,3])."""
    return SplatFunction.apply(means3d, scales, rots, opacities, colors, view, proj, tanfov, bg, int(height), int(width),
                               int(views_per_set), bool(activate), bool(hwc), bool(clamp01), float(scale_modifier))
