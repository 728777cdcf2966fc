"""Host logic of the physics row on CPU: python_raytracer_b200.physics.PhysicsWorld.update_physics with the slice test
swapped for its plain-Python restatement (oracle/physics_np.py) must reproduce, bit for bit, the positions and velocities
the reference's own Object.update_physics produced (tests/golden/physics_steps.npz, oracle/gen_physics_golden.py)."""
import os

import numpy as np

from tests.golden_util import GOLDEN_DIR


def load_scene():
    from python_raytracer_b200.physics import PhysObject
    g = np.load(os.path.join(GOLDEN_DIR, "physics_steps.npz"))
    o = np.load(os.path.join(GOLDEN_DIR, "default_objects.npz"))
    objs = []
    for k, row in enumerate(g["init"]):
        objs.append(PhysObject(sprite=o["sprite_%d" % k], pos=tuple(row[0:3]), rot=tuple(row[3:6]), vel=tuple(row[6:9]),
                               physics=bool(row[9]), weight=float(row[10]), visible=bool(row[11]),
                               grid=g["grid_%d" % k], grid_lo=tuple(int(v) for v in g["gridlo_%d" % k])))
    return g, objs


def replay(world, g, objs):
    worst = 0.0
    for t in range(g["steps"].shape[0]):
        for ob in objs:
            if ob.physics and ob.visible:
                world.update_physics(ob, objs)
        got = np.array([list(ob.pos) + list(ob.vel) for ob in objs], np.float64)
        assert np.array_equal(got, g["steps"][t]), "step %d differs by %g" % (t, np.abs(got - g["steps"][t]).max())
    return worst


def test_update_physics_host_logic_matches_reference():
    from oracle import physics_np
    from python_raytracer_b200.physics import PhysicsWorld
    g, objs = load_scene()
    gravity, friction, friction_air, vmax, vmin = (float(v) for v in g["settings"])
    world = PhysicsWorld.__new__(PhysicsWorld)               # no GPU context: only the host part is under test here
    world.gravity, world.friction, world.friction_air, world.max_velocity, world.min_velocity = gravity, friction, friction_air, vmax, vmin
    calls = []

    def slice_fn(obj, other, post6, vel_dir, fr, el):
        r = physics_np.slice_collide(g["props"], friction, obj, other, post6, vel_dir, fr, el)
        calls.append(r[2])
        return r

    world.slice_collide = slice_fn
    replay(world, g, objs)
    assert sum(calls) > 50 and len(calls) > sum(calls)        # the scenario has blocked and free moves
