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
                  sync_free: bool = False, out: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
    """Render F frames.

    means3d/scales/colors [G,N,3], rots [G,N,4], opacities [G,N,1] with G = F // views_per_set
    Gaussian sets; view/proj [F,4,4] (already transposed as the reference passes them), tanfov [F,2],
    bg [3].  Returns a dict with ``color`` ([F,3,H,W], or [F,H,W,3] when ``hwc``) plus the
    intermediate buffers (``geom``, ``rect``, ``tiles_touched``, ``tile_start``, ``point_list``, ``D``
    and, with ``want_aux``, ``final_T`` / ``n_contrib``) used by the backward pass and the parity tests.
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
    ws.overflow = overflow[:1]

    call("amav_splat_preprocess", dims, ptr(means3d), ptr(scales), ptr(rots), ptr(opacities), ptr(colors),
         ptr(view), ptr(proj), ptr(tanfov), ptr(geom), ptr(rect), ptr(touched), ptr(tile_count), st)
    call("amav_splat_scan", dims, ptr(tile_count), ptr(tile_start), ptr(scan_tmp), st)
    if sync_free and ws.capacity > 0:
        D = -1
    else:
        D = int(tile_start[nt].item()) & 0xFFFFFFFF
        ws.ensure_capacity(D, dev)
        ws.last_D = D
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
    call("amav_splat_blend", dims, ptr(geom), ptr(tile_start), ptr(ws.point_list), ws.capacity, ptr(bg), int(hwc), int(clamp01),
         ptr(out), ptr(final_T), ptr(n_contrib), ptr(invdepth), ptr(work_counter), st)
    return {"color": out, "geom": geom.view(F, N, 12), "rect": rect.view(F, N, 4), "tiles_touched": touched.view(F, N),
            "tile_start": tile_start, "point_list": ws.point_list, "D": D, "final_T": final_T,
            "n_contrib": n_contrib, "invdepth": invdepth, "workspace": ws}
