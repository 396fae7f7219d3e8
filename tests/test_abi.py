"""The C-ABI library loads and exports every symbol include/amav.h declares (no compute: runs without a GPU)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "amav.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(amav_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_path():
    syms = declared_symbols()
    for want in ("amav_triplane_project", "amav_sample_decode", "amav_lbs_precompute", "amav_lbs_forward",
                 "amav_densify_gather", "amav_splat_preprocess", "amav_splat_scan", "amav_splat_bin", "amav_splat_blend",
                 "amav_last_error", "amav_version", "amav_launch_count"):
        assert want in syms


def test_library_exports_every_declared_symbol():
    from audio_motion_avatar_b200 import build
    lib = ctypes.CDLL(build.build())
    for name in declared_symbols():
        assert hasattr(lib, name), f"{name} declared in include/amav.h but not exported"
    lib.amav_version.restype = ctypes.c_int
    assert lib.amav_version() >= 100


def test_ctypes_signatures_cover_the_header():
    from audio_motion_avatar_b200 import _lib
    bound = set(_lib._SIGS) | {"amav_last_error", "amav_version", "amav_launch_count", "amav_splat_scan_tmp_elems",
             "amav_splat_scan_order_offset", "amav_triplane_project_tc_scratch_elems"}
    assert set(declared_symbols()) <= bound, set(declared_symbols()) - bound


def test_no_cpu_fallback_in_product_path():
    """The host mirror refuses CPU tensors instead of falling back, and never imports the oracle."""
    import pytest
    import torch
    from audio_motion_avatar_b200 import _lib
    with pytest.raises(_lib.AmavError):
        _lib.ptr(torch.zeros(4))
    pkg = os.path.join(ROOT, "audio_motion_avatar_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            text = open(os.path.join(pkg, fn)).read()
            assert "import oracle" not in text and "from oracle" not in text, fn
