"""CPU restatement of the physics slice test (reference data.py:529-542) -- TEST INFRASTRUCTURE ONLY.

Plain Python loops in the reference's cell order, on the same inputs as the CUDA kernels (dense sprite grids of material
ids, Sprite.pos_rotated per data.py:338-371).  Used by tests to check the host mirror python_raytracer_b200.physics on
CPU against tests/golden/physics_steps.npz and as the checker of vt_slice_collide on the GPU.  Solidity 0 / 1 only."""
from __future__ import annotations

import numpy as np


def pos_rotated(size, angle, x, y, z):
    ex, ey, ez = size[0] - 1, size[1] - 1, size[2] - 1
    ax, ay, az = angle
    if ax and size[1] == size[2]:
        x, y, z = ((x, ez - z, y), (x, ey - y, ez - z), (x, z, ey - y))[ax - 1]
    if ay and size[0] == size[2]:
        x, y, z = ((z, y, ex - x), (ex - x, y, ez - z), (ez - z, y, x))[ay - 1]
    if az and size[0] == size[1]:
        x, y, z = ((ey - y, x, z), (ex - x, ey - y, z), (y, ex - x, z))[az - 1]
    return x, y, z


def sprite_voxel(obj, x, y, z):
    mins, _ = obj.box()
    size = obj.size
    x, y, z = pos_rotated(size, obj.angles(), x - mins[0], y - mins[1], z - mins[2])
    grid = np.asarray(obj.grid if obj.grid is not None else obj.sprite)
    x, y, z = x - obj.grid_lo[0], y - obj.grid_lo[1], z - obj.grid_lo[2]
    if 0 <= x < grid.shape[0] and 0 <= y < grid.shape[1] and 0 <= z < grid.shape[2]:
        return int(grid[x, y, z])
    return 0


def slice_collide(props, scale, obj, other, post6, vel_dir, friction, elasticity):
    blocked = False
    for x in range(post6[0], post6[3] + 1):
        for y in range(post6[1], post6[4] + 1):
            for z in range(post6[2], post6[5] + 1):
                om = sprite_voxel(other, x, y, z)
                if om and props[om][0] >= 1:
                    sm = sprite_voxel(obj, x - vel_dir[0], y - vel_dir[1], z - vel_dir[2])
                    if sm and props[sm][0] >= 1:
                        friction += props[om][1] * props[sm][1] * scale
                        elasticity += props[om][2] * props[sm][2] * scale
                        blocked = True
    return friction, elasticity, blocked
