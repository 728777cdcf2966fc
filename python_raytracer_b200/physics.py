"""Object physics (reference ``Object.update_physics``, data.py:495-560) with the voxel slice test on the GPU.

The reference moves a physical object one cell at a time along its largest velocity component; before each move it walks
the one-cell-thick slice the object would enter and tests it, cell by cell, against every other object that touches the
slice (data.py:520-542).  That voxel loop is the data-parallel part and runs as CUDA kernels (``vt_slice_collide``,
csrc/vt_physics.cuh); everything around it -- direction choice, velocity transfer between physical objects, the move,
gravity / friction / terminal velocity -- is a handful of scalar operations per step and stays on the host, written with
the same Python float operations in the same order as the reference so that positions and velocities come out bit for
bit.  There is no CPU fallback for the slice test: ``PhysicsWorld`` needs a GpuCamera's context.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass

import numpy as np
import torch

from .abi import VoxtraceError, VtPhysObject
from .builder import SceneObject


@dataclass
class PhysObject(SceneObject):
    """A ``data.Object`` as the physics step sees it: SceneObject (sprite grid, pos, rot, visible) + velocity, the
    ``physics`` flag and the weight ``Object.set_weight`` computed (data.py:607-611)."""
    vel: tuple = (0.0, 0.0, 0.0)
    physics: bool = False
    weight: float = 0.0
    redraw: bool = False
    # The voxel dictionary of the sprite over the bounding box of its keys, when that is larger than [0, size): Sprite.load
    # maps file column x = 0 to x = size.x (data.py:285) -- never drawn (init.py:417-421), but the slice test asks for it.
    grid: object = None              # [gx, gy, gz] u8, default: ``sprite``
    grid_lo: tuple = (0, 0, 0)       # sprite-local coordinate of grid[0, 0, 0]

    def move(self, pos):
        """Object.move (data.py:478-486)."""
        if tuple(pos) != tuple(self.pos):
            self.pos = tuple(pos)
            self.redraw = True


def _sub(a, b):
    return (a[0] - b[0], a[1] - b[1], a[2] - b[2])


def _add(a, b):
    return (a[0] + b[0], a[1] + b[1], a[2] + b[2])


def _mul(a, b):
    if isinstance(b, tuple):
        return (a[0] * b[0], a[1] * b[1], a[2] * b[2])
    return (a[0] * b, a[1] * b, a[2] * b)


class PhysicsWorld:
    def __init__(self, cam, phys_materials, gravity=0.0, friction=0.0, friction_air=0.0, max_velocity=0.0, min_velocity=0.0,
                 seed=0):
        """phys_materials: [n + 1, 3] solidity, friction, elasticity per material id (row 0 unused); the other arguments
        are the PHYSICS settings of config.cfg (data.py:52-59)."""
        self.cam = cam
        self.gravity, self.friction, self.friction_air = gravity, friction, friction_air
        self.max_velocity, self.min_velocity = max_velocity, min_velocity
        self.seed = int(seed)
        t = np.zeros((len(phys_materials), 4), np.float64)
        t[:, :3] = np.asarray(phys_materials, np.float64)
        cam._check(cam._lib.vt_set_phys_materials(cam._ctx, t.ctypes.data, len(t) - 1))
        self.slice_tests = 0

    # ---- the GPU part ------------------------------------------------------------------------------------------------
    def _vt_object(self, o):
        if getattr(o, "_phys_dev", None) is None:
            g = o.grid if o.grid is not None else o.sprite
            t = g if isinstance(g, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(g, np.uint8))
            o._phys_dev = t.to(self.cam.device, dtype=torch.uint8).contiguous()
        v = VtPhysObject()
        mins, _ = o.box()
        v.dev_grid = o._phys_dev.data_ptr()
        v.grid_lo[:] = [int(c) for c in o.grid_lo]
        v.grid_dims[:] = [int(c) for c in o._phys_dev.shape]
        v.size[:] = list(o.size)
        v.mins[:] = mins
        v.angle[:] = o.angles()
        return v

    def slice_collide(self, obj, other, post6, vel_dir, friction, elasticity):
        """data.py:529-542 for one (self, other) pair -> (friction, elasticity, blocked)."""
        cam = self.cam
        a, b = self._vt_object(obj), self._vt_object(other)
        lo = (C.c_int32 * 3)(*[int(v) for v in post6[:3]])
        hi = (C.c_int32 * 3)(*[int(v) for v in post6[3:]])
        d = (C.c_int32 * 3)(*[int(v) for v in vel_dir])
        sums = (C.c_double * 2)(float(friction), float(elasticity))
        blocked = C.c_int32(0)
        cam._check(cam._lib.vt_slice_collide(cam._ctx, C.byref(a), C.byref(b), lo, hi, d, float(self.friction),
                                             self.seed + self.slice_tests, sums, C.byref(blocked)))
        self.slice_tests += 1
        return sums[0], sums[1], bool(blocked.value)

    # ---- the host part: data.py:495-560, same operations in the same order -------------------------------------------
    def update_physics(self, obj, objects):
        friction = elasticity = 0
        vel_apply = tuple(obj.vel)
        while vel_apply[0] != 0 or vel_apply[1] != 0 or vel_apply[2] != 0:
            ref = max(abs(vel_apply[0]), abs(vel_apply[1]), abs(vel_apply[2]))            # vec3.normalize, lib.py:310-314
            n = (vel_apply[0] / ref, vel_apply[1] / ref, vel_apply[2] / ref) if ref and ref != 1 else vel_apply
            vel_dir = (math.trunc(n[0]), math.trunc(n[1]), math.trunc(n[2]))
            blocked = False
            mins, maxs = obj.box()
            if vel_dir[0] < 0:
                post6 = (mins[0] - 1, mins[1], mins[2], mins[0], maxs[1], maxs[2])
            elif vel_dir[0] > 0:
                post6 = (maxs[0], mins[1], mins[2], maxs[0] + 1, maxs[1], maxs[2])
            elif vel_dir[1] < 0:
                post6 = (mins[0], mins[1] - 1, mins[2], maxs[0], mins[1], maxs[2])
            elif vel_dir[1] > 0:
                post6 = (mins[0], maxs[1], mins[2], maxs[0], maxs[1] + 1, maxs[2])
            elif vel_dir[2] < 0:
                post6 = (mins[0], mins[1], mins[2] - 1, maxs[0], maxs[1], mins[2])
            elif vel_dir[2] > 0:
                post6 = (mins[0], mins[1], maxs[2], maxs[0], maxs[1], maxs[2] + 1)
            else:
                raise VoxtraceError("update_physics: velocity %r has no direction" % (vel_apply,))
            for other in objects:
                if other is obj or not other.visible:
                    continue
                omins, omaxs = other.box()
                if not (all(post6[a] <= omaxs[a] for a in range(3)) and all(post6[3 + a] >= omins[a] for a in range(3))):
                    continue                                                            # Object.intersects, data.py:463-464
                if other.physics:                                                       # data.py:523-527
                    av = max(abs(vel_apply[0]), abs(vel_apply[1]), abs(vel_apply[2]))
                    vel_transfer = _mul(vel_apply, max(0, min(1, av * obj.weight - other.weight)))
                    other.vel = _add(tuple(other.vel), vel_transfer)
                    obj.vel = _sub(tuple(obj.vel), vel_transfer)
                    vel_apply = _sub(vel_apply, vel_transfer)
                friction, elasticity, hit = self.slice_collide(obj, other, post6, vel_dir, friction, elasticity)
                blocked = blocked or hit
            av3 = (abs(vel_apply[0]), abs(vel_apply[1]), abs(vel_apply[2]))
            vel_step = _mul(vel_dir, av3) if blocked else _mul(vel_dir, (min(av3[0], 1), min(av3[1], 1), min(av3[2], 1)))
            vel_apply = _sub(vel_apply, vel_step)
            if not blocked:
                obj.move(_add(tuple(obj.pos), vel_step))
        v = _sub(tuple(obj.vel), (0, obj.weight * self.gravity, 0))                      # data.py:551-560
        v = _sub(v, _mul(v, elasticity))
        d = 1 + max(0, friction + self.friction_air)
        v = (v[0] / d, v[1] / d, v[2] / d)
        v = tuple(max(min(c, +self.max_velocity), -self.max_velocity) for c in v)
        obj.vel = tuple(0 if abs(c) < self.min_velocity else c for c in v)
