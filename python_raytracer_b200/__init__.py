"""voxtrace-b200: B200-native replacement for python_raytracer's per-pixel voxel ray-march hot path.

Public surface (mirrors the reference's Camera protocol, reference init.py:13-150):
    GpuCamera        -- pos / rot / lens / chunks / chunk_set / chunk_get / tile, plus render()
    Settings         -- the `data.settings` subset the path reads
There is no CPU fallback: everything that renders goes through libvoxtrace.so (include/voxtrace.h).
"""
from .abi import VoxtraceError  # noqa: F401
from .settings import Settings  # noqa: F401
from .camera import GpuCamera  # noqa: F401

__all__ = ["GpuCamera", "Settings", "VoxtraceError"]
