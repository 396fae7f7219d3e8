"""ctypes binding of libamav_b200.so (C ABI declared in include/amav.h).

There is no CPU fallback: if the shared library is missing or a call fails, this module raises."""
from __future__ import annotations

import ctypes
import os
import threading
from ctypes import c_double, c_float, c_int, c_int32, c_int64, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libamav_b200.so")


class AmavError(RuntimeError):
    pass


class SplatDims(ctypes.Structure):
    _fields_ = [("frames", c_int32), ("num_gaussians", c_int32), ("views_per_set", c_int32),
                ("width", c_int32), ("height", c_int32), ("activate", c_int32), ("scale_modifier", c_float)]


_SIGS = {
    "amav_triplane_project": [c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p, c_void_p],
    "amav_triplane_project_tc": [c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p],
    "amav_triplane_project_tc_static": [c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p, c_void_p],
    "amav_triplane_project_tm": [c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p, c_void_p],
    "amav_sample_decode": [c_int, c_int, c_int, c_int, c_float] + [c_void_p] * 11 + [c_void_p],
    "amav_lbs_precompute": [c_int, c_int, c_int] + [c_void_p] * 7 + [c_void_p],
    "amav_lbs_forward": [c_int, c_int, c_int, c_int] + [c_void_p] * 14 + [c_void_p],
    "amav_lbs_forward_6d": [c_int, c_int, c_int, c_int] + [c_void_p] * 14 + [c_void_p],
    "amav_frames_to_u8": [c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_int, c_void_p],
    "amav_densify_gather": [c_int, c_int, c_int, c_int] + [c_void_p] * 5 + [c_void_p],
    "amav_splat_preprocess": [ctypes.POINTER(SplatDims)] + [c_void_p] * 12 + [c_void_p],
    "amav_splat_scan": [ctypes.POINTER(SplatDims)] + [c_void_p] * 3 + [c_void_p],
    "amav_splat_bin": [ctypes.POINTER(SplatDims)] + [c_void_p] * 6 + [c_int64, c_void_p, c_void_p],
    "amav_sample_decode_backward": [c_int, c_int, c_int, c_int, c_float] + [c_void_p] * 13 + [c_void_p],
    "amav_triplane_project_backward": [c_int, c_int, c_int, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p],
    "amav_triplane_project_backward_tm": [c_int, c_int, c_int, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p],
    "amav_densify_gather_backward": [c_int, c_int, c_int, c_int] + [c_void_p] * 4 + [c_void_p],
    "amav_lbs_backward": [c_int, c_int, c_int, c_int] + [c_void_p] * 16 + [c_void_p],
    "amav_lbs_backward_6d": [c_int, c_int, c_int, c_int] + [c_void_p] * 16 + [c_void_p],
    "amav_camera_setup": [c_int, c_void_p, c_void_p, c_int, c_int, c_double, c_double, c_void_p, c_void_p, c_void_p, c_void_p],
    "amav_image_loss_forward": [c_int, c_int, c_int] + [c_void_p] * 4 + [c_void_p],
    "amav_image_loss_backward": [c_int, c_int, c_int] + [c_void_p] * 6 + [c_void_p],
    "amav_splat_blend_backward": [ctypes.POINTER(SplatDims)] + [c_void_p] * 3 + [c_int64, c_void_p, c_int] + [c_void_p] * 6 + [c_void_p],
    "amav_splat_preprocess_backward": [ctypes.POINTER(SplatDims)] + [c_void_p] * 15 + [c_void_p],
    "amav_splat_blend": [ctypes.POINTER(SplatDims)] + [c_void_p] * 3 + [c_int64, c_void_p, c_int, c_int] + [c_void_p] * 6 + [c_void_p],
}

_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise AmavError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                            "(there is no CPU or PyTorch fallback for the render path)")
        L = ctypes.CDLL(LIB_PATH)
        L.amav_last_error.restype = ctypes.c_char_p
        L.amav_version.restype = c_int
        L.amav_launch_count.restype = c_int64
        L.amav_splat_scan_tmp_elems.restype = c_int64
        L.amav_splat_scan_tmp_elems.argtypes = [c_int64]
        L.amav_splat_scan_order_offset.restype = c_int64
        L.amav_splat_scan_order_offset.argtypes = [c_int64]
        for name, sig in _SIGS.items():
            if hasattr(L, name):
                fn = getattr(L, name)
                fn.argtypes = sig
                fn.restype = c_int
        _lib = L
    return _lib


_tls = threading.local()


def ptr(t, peer: bool = False):
    """Device pointer of a contiguous CUDA tensor (None -> NULL); ``peer``: the buffer may be another GPU's memory
    mapped into this process (amav_splat_blend's out_color) and does not take part in the device check.  The tensor's device is remembered until the next
    ``call``: the C ABI launches on the calling thread's CURRENT device (include/amav.h, device contract), so ``call``
    makes the arguments' device current and refuses arguments that live on different devices."""
    if t is None:
        return None
    if not t.is_cuda:
        raise AmavError("libamav_b200 takes CUDA tensors only (no CPU path)")
    if not t.is_contiguous():
        raise AmavError("tensor must be contiguous")
    if not peer:
        devs = getattr(_tls, "devs", None)
        if devs is None:
            devs = _tls.devs = set()
        devs.add(t.device.index)
    return c_void_p(t.data_ptr())


class _StreamOfArguments:
    """Placeholder for the `stream` argument: ``call`` replaces it by the current stream of the device the pointer
    arguments live on (not of whatever device happens to be current)."""


_STREAM = _StreamOfArguments()


def stream_ptr():
    return _STREAM


class StageTimer:
    """CUDA-event timing of every C-ABI call made while it is installed (``with StageTimer() as t``); events are
    recorded on the current stream around each call, resolved by ``summary()`` after a synchronize."""

    def __init__(self):
        self.events = []

    def __enter__(self):
        global _TIMER
        self._prev, _TIMER = _TIMER, self
        return self

    def __exit__(self, *exc):
        global _TIMER
        _TIMER = self._prev

    def summary(self):
        torch.cuda.synchronize()
        out = {}
        for name, a, b in self.events:
            ms, n = out.get(name, (0.0, 0))
            out[name] = (ms + a.elapsed_time(b), n + 1)
        return out


_TIMER = None


def launch_count() -> int:
    return int(lib().amav_launch_count())


def call(name: str, *args) -> None:
    fn = getattr(lib(), name, None)
    if fn is None:
        raise AmavError(f"{name} is not exported by {LIB_PATH}")
    devs = getattr(_tls, "devs", None)
    _tls.devs = set()
    if devs and len(devs) > 1:
        raise AmavError(f"{name}: arguments live on different CUDA devices {sorted(devs)}; every buffer of a call must "
                        "belong to one device (include/amav.h, device contract)")
    cur = torch.cuda.current_device()
    idx = next(iter(devs)) if devs else cur
    st = c_void_p(torch.cuda.current_stream(idx).cuda_stream)
    args = [st if a is _STREAM else a for a in args]
    if idx != cur:
        with torch.cuda.device(idx):
            return _call_on_current_device(name, fn, args)
    return _call_on_current_device(name, fn, args)


def _call_on_current_device(name, fn, args) -> None:
    if _TIMER is not None:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        rc = fn(*args)
        b.record()
        _TIMER.events.append((name, a, b))
    else:
        rc = fn(*args)
    if rc != 0:
        raise AmavError(f"{name} failed ({rc}): {lib().amav_last_error().decode()}")
