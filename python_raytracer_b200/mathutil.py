"""Host-side helpers that follow the reference's own conventions (reference lib.py)."""
import math


def euler_to_quaternion(x: float, y: float, z: float):
    """vec3.quaternion (lib.py:322-338): euler degrees (roll x, yaw y, pitch z) -> (x, y, z, w)."""
    rx, ry, rz = math.radians(x), math.radians(y), math.radians(z)
    sx, cx = math.sin(rx / 2), math.cos(rx / 2)
    sy, cy = math.sin(ry / 2), math.cos(ry / 2)
    sz, cz = math.sin(rz / 2), math.cos(rz / 2)
    return (sx * cy * cz - cx * sy * sz,
            cx * sy * cz - sx * cy * sz,
            cx * cy * sz + sx * sy * cz,
            cx * cy * cz + sx * sy * sz)


def samples_per_pixel(width, height, samples, lod_edge):
    """Traced samples per pixel (init.py:131-134) as an int array [height, width] -- rays/s counts these."""
    import numpy as np
    x = np.arange(width, dtype=np.float64)
    y = np.arange(height, dtype=np.float64)
    dx = -1 + (x / width) * 2
    dy = -1 + (y / height) * 2
    detail = 1 - np.abs(dx[None, :] * dy[:, None]) * lod_edge
    return np.maximum(1, np.rint(samples * detail)).astype(np.int64)
