"""Import-compatible stand-ins for third-party modules the reference's render path imports."""
import importlib
import sys


def install_diff_gaussian_rasterization() -> None:
    """Register the shim as ``diff_gaussian_rasterization`` so the reference's own ``render_one``
    (src/models/renderer.py:422,516-566) runs unmodified on libamav_b200.so."""
    mod = importlib.import_module("audio_motion_avatar_b200.compat.diff_gaussian_rasterization")
    sys.modules.setdefault("diff_gaussian_rasterization", mod)
