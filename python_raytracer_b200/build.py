"""Builds libvoxtrace.so in-tree with nvcc for sm_100a (no JIT cache, no torch extension machinery)."""
from __future__ import annotations

import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libvoxtrace.so")
SOURCES = ["voxtrace.cu", "goxel.cpp"]
HEADERS = [os.path.join(CSRC, "vt_device.cuh"), os.path.join(CSRC, "vt_march.cuh"), os.path.join(CSRC, "vt_scene.cuh"), os.path.join(CSRC, "vt_postfx.cuh"), os.path.join(CSRC, "vt_physics.cuh"), os.path.join(CSRC, "vt_pow_cr.h"), os.path.join(os.path.dirname(PKG_DIR), "include", "voxtrace.h")]

# -fmad=false: the reference computes in Python floats, every a*b+c rounds twice; FMA contraction would change
# the sampled positions.  -lineinfo: ncu source pages map to our lines.
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC", "-lz"]


RENDER_KERNEL_FILES = ["vt_march.cuh", "vt_device.cuh", "vt_pow_cr.h"]


def kernel_source_sha256() -> str:
    """sha256 over the sources of the render kernel (vt_march.cuh with its launch bounds, vt_device.cuh, vt_pow_cr.h) and the
    nvcc flags: ties a committed ncu capture of render_persistent to the kernel it was taken from (profiles/r*_traffic.json
    carry it, bench.py reports `traffic_stale` when it differs).  Host code and the other kernels are not part of it."""
    import hashlib
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    for name in RENDER_KERNEL_FILES:
        with open(os.path.join(CSRC, name), "rb") as f:
            h.update(name.encode() + b"\0" + f.read())
    return h.hexdigest()


def _nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found; libvoxtrace.so cannot be built (there is no CPU fallback)")
    return exe


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + HEADERS
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False, defines=(), out: str | None = None) -> str:
    """defines / out: tuning variants (-DNAME=VALUE) built next to the product library, for A/B runs only."""
    if out is None and not force and not needs_build():
        return LIB_PATH
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-D" + d for d in defines] + \
          ["-o", out or LIB_PATH] + [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return out or LIB_PATH


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
