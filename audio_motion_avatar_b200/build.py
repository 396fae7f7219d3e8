"""In-tree build of libamav_b200.so (explicit nvcc, sm_100a only).  No JIT cache: the .so sits next
to this file so it travels to the GPU box with the repo snapshot."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
LIB = os.path.join(HERE, "libamav_b200.so")

# (source, extra flags).  -fmad=false: FMAs only where fmaf() is written (bit-exact vs the oracle).
SOURCES = [
    ("capi.cu", []),
    ("splat_preprocess.cu", ["-fmad=false"]),
    ("splat_bin.cu", []),
    ("splat_blend.cu", ["-fmad=false"]),
    ("splat_blend_bwd.cu", ["-fmad=false"]),
    ("splat_preprocess_bwd.cu", []),
    ("triplane.cu", []),
    ("triplane_bwd.cu", []),
    ("triplane_tm.cu", []),
    ("triplane_tc.cu", []),
    ("lbs.cu", []),
    ("lbs_tc.cu", []),
    ("lbs_bwd.cu", []),
    ("image_loss.cu", []),
    ("camera.cu", ["-fmad=false"]),
    ("video_sink.cu", []),
]
COMMON = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
          "-Xcompiler", "-fPIC", "-I", INCLUDE, "-I", CSRC]


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(verbose: bool = False, force: bool = False) -> str:
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, h) for h in os.listdir(CSRC) if h.endswith((".cuh", ".h"))]
    headers.append(os.path.join(INCLUDE, "amav.h"))
    objs, rebuilt = [], False
    for src, extra in SOURCES:
        path = os.path.join(CSRC, src)
        if not os.path.exists(path):
            continue
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [path] + headers):
            cmd = [nvcc] + COMMON + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", path, "-o", obj]
            if verbose:
                print(" ".join(cmd), file=sys.stderr)
            subprocess.check_call(cmd)
            rebuilt = True
    if rebuilt or not os.path.exists(LIB):
        subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB] + objs + ["-lcudart"])
    return LIB


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
