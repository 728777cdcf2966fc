"""tests/golden/physics_steps.npz: the reference's own Object.update_physics (data.py:495-560) run for 40 steps on the
default mod's objects (castle + six material cubes + player), some of them given a push, one turned by 90 degrees and
one placed at a fractional position.  Inputs: the objects' sprite grids (tests/golden/default_objects.npz, material ids =
index into default_world.npz `names`), physical material fields, weights, the PHYSICS settings.  Outputs per step:
every object's pos and vel.   python -m oracle.gen_physics_golden      TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import os
import sys

import numpy as np

from oracle import refharness as rh

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
STEPS = 40


def main():
    ref = rh.load_reference()
    data, lib = ref.data, ref.lib
    vec3 = lib.vec3
    rh.restore_settings()
    world = np.load(os.path.join(OUT, "default_world.npz"))
    names = [str(n) for n in world["names"]]
    mod = sys.modules["mods.default.init"]
    props = np.zeros((len(names), 3), np.float64)
    for k, nm in enumerate(names):
        if nm and hasattr(mod, nm):
            m = getattr(mod, nm)
            props[k] = (m.solidity, m.friction, m.elasticity)
    objs = list(data.objects.values())
    n = len(objs)
    for o in objs:
        o.visible = bool(o.sprite)                      # what Object.update sets for objects within dist_max (data.py:569)
        o.set_weight()
    # a scenario with contacts: cubes pushed sideways and upwards, one turned, one at a fractional position
    pushes = {1: (0.75, 0.0, 0.0), 2: (0.0, 1.5, -0.5), 3: (-2.25, 0.5, 0.0), 5: (0.3, 0.0, 2.0), n - 1: (0.0, 0.0, -1.25)}
    for k, v in pushes.items():
        if 0 <= k < n and objs[k].physics:
            objs[k].vel = vec3(*v)
    if n > 4 and objs[4].physics:
        objs[4].rot = vec3(0, 90, 0)
        objs[4].move(objs[4].pos + vec3(0.5, 0.25, 0))
    ident = {id(v): k for k, v in vars(mod).items() if isinstance(v, data.Material)}
    grids = {}
    for k, o in enumerate(objs):                          # the sprite's voxel dictionary over the bounding box of its keys
        vox = o.get_sprite().get_voxels(None)
        keys = np.array(list(vox.keys()), np.int64)
        lo, hi = keys.min(axis=0), keys.max(axis=0) + 1
        lo = np.minimum(lo, 0); hi = np.maximum(hi, [int(o.sprite.size.x), int(o.sprite.size.y), int(o.sprite.size.z)])
        g = np.zeros(tuple(hi - lo), np.uint8)
        for key, m in vox.items():
            g[tuple(np.array(key) - lo)] = names.index(ident[id(m)])
        grids["grid_%d" % k], grids["gridlo_%d" % k] = g, lo.astype(np.int32)
    s = data.settings
    init = np.array([[o.pos.x, o.pos.y, o.pos.z, o.rot.x, o.rot.y, o.rot.z, o.vel.x, o.vel.y, o.vel.z, float(o.physics), o.weight,
                      float(o.visible)] for o in objs], np.float64)
    steps = np.zeros((STEPS, n, 6), np.float64)
    for t in range(STEPS):
        for o in objs:                                   # the frame loop's order (init.py:470-471, data.py:575-583)
            if o.physics and o.visible:
                o.update_physics()
        steps[t] = [[o.pos.x, o.pos.y, o.pos.z, o.vel.x, o.vel.y, o.vel.z] for o in objs]
    moved = int((np.abs(steps[-1, :, :3] - init[:, :3]).max(axis=1) > 0).sum())
    print("physics_steps: %d objects (%d physical), %d steps, %d moved; final y of physical objects: %s" % (
        n, int(init[:, 9].sum()), STEPS, moved, np.round(steps[-1, init[:, 9] > 0, 1], 3).tolist()))
    np.savez_compressed(os.path.join(OUT, "physics_steps.npz"), init=init, steps=steps, props=props, **grids,
                        settings=np.array([s.gravity, s.friction, s.friction_air, s.max_velocity, s.min_velocity], np.float64))
    rh.restore_settings()


if __name__ == "__main__":
    main()
