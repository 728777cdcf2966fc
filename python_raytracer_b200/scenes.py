"""Synthetic worlds of the named BASELINE.json shapes: C4 terrain, C5 large Goxel-text model.

Pure numpy generators shared by bench.py and the tests; deterministic (fixed seeds).  Material rows are
(albedo r, g, b, roughness, absorption, ior, energy), the values of the default mod's castle materials
(reference mods/default/init.py:6-173) where the survey names them.
"""
from __future__ import annotations

import numpy as np

STONE_DARK = (63.0, 63.0, 63.0, 0.5, 2.0, 1.0, 0.0)
STONE_GRAY = (127.0, 127.0, 127.0, 0.5, 1.5, 1.0, 0.0)
STONE_LIGHT = (191.0, 191.0, 191.0, 0.5, 1.0, 1.0, 0.0)
MARBLE = (255.0, 255.0, 255.0, 0.0, 1.0, 1.0, 0.0)
LIGHT = (255.0, 255.0, 0.0, 0.5, 1.0, 1.0, 2.0)
METAL = (0.0, 0.0, 0.0, 0.1, 0.5, 1.0, 0.0)
GLASS = (0.0, 255.0, 255.0, 0.0, 0.25, 0.25, 0.0)
MIST = (255.0, 0.0, 255.0, 0.0, 0.25, 0.0, 0.0)


def _value_noise(rng, n, cells):
    """Bilinear value noise on an n x n grid with `cells` lattice cells per side."""
    lat = rng.random((cells + 2, cells + 2))
    t = np.arange(n) * (cells / n)
    i = np.floor(t).astype(int)
    f = t - i
    f = f * f * (3 - 2 * f)
    a = lat[i][:, i]; b = lat[i + 1][:, i]; c = lat[i][:, i + 1]; d = lat[i + 1][:, i + 1]
    fx, fz = f[:, None], f[None, :]
    return (a * (1 - fx) + b * fx) * (1 - fz) + (c * (1 - fx) + d * fx) * fz


def terrain_world(n=512, seed=1234):
    """C4: n^3 cells, height field h = 0.375 n + 0.1875 n * fbm (4 octaves of value noise, period n/4 at the base),
    solid below h; four stone materials by height band, 1 % of the surface cells emissive.
    Returns (origin[3], cells u8 [n, n, n], materials [6, 7]); x and z are centred on 0, y starts at 0."""
    rng = np.random.default_rng(seed)
    fbm = np.zeros((n, n))
    amp, tot = 1.0, 0.0
    for o in range(4):
        fbm += amp * (_value_noise(rng, n, 4 << o) * 2 - 1)
        tot += amp
        amp *= 0.5
    h = np.clip(np.rint(0.375 * n + 0.1875 * n * fbm / tot), 1, n - 1).astype(np.int32)       # [x, z]
    y = np.arange(n, dtype=np.int32)[None, :, None]
    solid = y < h[:, None, :]
    band = np.clip((y * 4) // max(1, int(0.5625 * n)), 0, 3).astype(np.uint8) + 1               # 1..4 by height
    cells = np.where(solid, band, 0).astype(np.uint8)
    top = np.zeros_like(solid)
    xi, zi = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    lit = rng.random((n, n)) < 0.01
    cells[xi[lit], h[lit] - 1, zi[lit]] = 5
    del top
    mats = np.array([(0,) * 7, STONE_DARK, STONE_GRAY, STONE_LIGHT, MARBLE, LIGHT], np.float64)
    origin = np.array([-n // 2, 0, -n // 2], np.int32)
    return origin, cells, mats


C5_KEYS = ["000000", "3f3f3f", "7f7f7f", "bfbfbf", "ffffff", "ffff00", "00ffff", "ff00ff"]
C5_MATERIALS = np.array([(0,) * 7, METAL, STONE_DARK, STONE_GRAY, STONE_LIGHT, MARBLE, LIGHT, GLASS, MIST], np.float64)


def goxel_model_grid(size=(512, 256, 512), shapes=80, seed=5678):
    """The C5 model as the dense engine-space grid [sx, sy, sz] of colour indices 1..8 (0 = empty): what loading the file
    written from ``goxel_model`` gives back (every listed voxel lands inside the sprite box)."""
    return _goxel_grid(size, shapes, np.random.default_rng(seed))


def goxel_model(size=(512, 256, 512), shapes=80, seed=5678, target_voxels=None):
    """C5: shell-filled union of random boxes and spheres (engine axes: size = x, y, z), eight colours.
    Returns (xyz_file int32 [n, 3] in Goxel FILE coordinates, colour strings [n]) -- what a Goxel export of the model
    would list; engine position = (size.x - x_file, z_file, y_file) (data.py:285)."""
    rng = np.random.default_rng(seed)
    sx = size[0]
    grid = _goxel_grid(size, shapes, rng)
    ex, ey, ez = np.nonzero(grid)
    if target_voxels is not None and len(ex) > target_voxels:
        keep = np.sort(rng.choice(len(ex), target_voxels, replace=False))
        ex, ey, ez = ex[keep], ey[keep], ez[keep]
    col = grid[ex, ey, ez]
    # engine (x, y, z) -> file (size.x - x, z, y)
    xyz_file = np.stack([sx - ex, ez, ey], axis=1).astype(np.int32)
    return xyz_file, np.asarray(C5_KEYS)[col - 1]


def _goxel_grid(size, shapes, rng):
    sx, sy, sz = size
    grid = np.zeros((sx, sy, sz), np.uint8)
    for k in range(shapes):
        col = int(rng.integers(1, 9))
        r = int(rng.integers(max(4, sx // 40), max(6, sx // 8)))
        c = [int(rng.integers(r + 1, s - r - 1)) if s > 2 * r + 3 else s // 2 for s in size]
        r3 = [min(r, (s - 3) // 2) for s in size]
        lo = [cc - rr for cc, rr in zip(c, r3)]
        hi = [cc + rr + 1 for cc, rr in zip(c, r3)]
        sub = grid[lo[0]:hi[0], lo[1]:hi[1], lo[2]:hi[2]]
        gx, gy, gz = np.ogrid[lo[0]:hi[0], lo[1]:hi[1], lo[2]:hi[2]]
        thick = 3 + k % 3
        if k % 2 == 0:      # box shell
            d = np.maximum(np.maximum(np.abs(gx - c[0]) / max(1, r3[0]), np.abs(gy - c[1]) / max(1, r3[1])), np.abs(gz - c[2]) / max(1, r3[2]))
            shell = d >= 1.0 - thick / max(1, min(r3))
        else:               # ellipsoid shell
            d = np.sqrt(((gx - c[0]) / max(1, r3[0])) ** 2 + ((gy - c[1]) / max(1, r3[1])) ** 2 + ((gz - c[2]) / max(1, r3[2])) ** 2)
            shell = (d <= 1.0) & (d >= 1.0 - thick / max(1, min(r3)))
        sub[shell] = col
    return grid


# ---- the named configurations (BASELINE.json configs[3], configs[4]), shared by tests and bench.py
C4_SETTINGS = dict(width=3840, height=2160, samples=16, static=True, shutter=0.25, falloff=0.25, dof=0.5, dist_min=0,
                   dist_max=768, max_light=1, max_bounces=4, lod_bounces=0.5, lod_samples=0, lod_random=0.25,
                   lod_edge=0, chunk_size=16, chunk_lod=2, fov=90, culling=False)
C5_SETTINGS = dict(width=1920, height=1080, samples=8, static=True, shutter=0.25, falloff=0.25, dof=0.5, dist_min=0,
                   dist_max=384, max_light=1, max_bounces=4, lod_bounces=0.5, lod_samples=0, lod_random=0.25,
                   lod_edge=0, chunk_size=16, chunk_lod=2, fov=90, culling=True)
C4_POS, C4_YAWS, C4_PITCH = (0.0, 300.0, 0.0), (0.0, 90.0, 135.0, 270.0), 30.0      # camera 30 degrees down, four yaw angles
C5_SIZE, C5_POS, C5_EULER = (512, 256, 512), (-300.0, 40.0, -300.0), (0.0, 45.0, 10.0)


def c4_workload(n=512):
    """C4: (origin, cells, materials, chunk_res, poses) -- the per-chunk LOD table depends on the camera position only,
    which the four poses share."""
    from .mathutil import euler_to_quaternion
    from .world import chunk_presence, lod_table
    sd = C4_SETTINGS
    origin, cells, mats = terrain_world(n)
    present = chunk_presence(cells, sd["chunk_size"])
    res = lod_table(origin, present, C4_POS, sd["chunk_size"], sd["chunk_lod"], sd["dist_max"])
    poses = [dict(pos=C4_POS, rot=euler_to_quaternion(0.0, yaw, C4_PITCH)) for yaw in C4_YAWS]
    return origin, cells, mats, res, poses
