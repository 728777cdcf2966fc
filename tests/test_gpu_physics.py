"""GPU tests of the physics row (Object.update_physics, reference data.py:495-560): the CUDA slice test behind
vt_slice_collide against its plain-Python restatement on random configurations (rotations, partial overlaps, sprites
whose dictionary extends past size.x), and the whole step loop against positions / velocities the reference itself
produced over 40 steps (tests/golden/physics_steps.npz)."""
import numpy as np
import pytest

from tests.test_physics_cpu import load_scene, replay

pytestmark = pytest.mark.gpu


def _bg(ray, settings):
    raise AssertionError("never called")


def _world(props, settings):
    from python_raytracer_b200 import GpuCamera, Settings
    from python_raytracer_b200.physics import PhysicsWorld
    cam = GpuCamera(Settings.from_object(dict(width=8, height=8, chunk_size=16)), background_function=_bg, background=_bg)
    gravity, friction, friction_air, vmax, vmin = (float(v) for v in settings)
    return cam, PhysicsWorld(cam, props, gravity=gravity, friction=friction, friction_air=friction_air, max_velocity=vmax, min_velocity=vmin)


def test_physics_steps_match_the_reference():
    g, objs = load_scene()
    cam, world = _world(g["props"], g["settings"])
    replay(world, g, objs)
    assert world.slice_tests > 100
    cam.close()


def test_slice_kernel_equals_restatement_on_random_configurations():
    from oracle import physics_np
    from python_raytracer_b200.physics import PhysObject
    rng = np.random.default_rng(11)
    props = np.array([[0, 0, 0], [1, 0.5, 0.25], [1, 0.125, 0.75], [0, 1.0, 1.0], [1, 0.3, 0.1]], np.float64)   # id 3: solidity 0
    cam, world = _world(props, (1.0, 0.7, 0.1, 10.0, 0.01))
    dirs = [(-1, 0, 0), (1, 0, 0), (0, -1, 0), (0, 1, 0), (0, 0, -1), (0, 0, 1)]
    blocked_seen = 0
    for trial in range(60):
        sa = tuple(int(v) for v in rng.choice([4, 6, 8], 3)) if trial % 3 else (6, 6, 6)
        sb = tuple(int(v) for v in rng.choice([6, 10, 16], 3))
        ga = (rng.random((sa[0] + 1,) + sa[1:]) < 0.5) * rng.integers(1, 5, (sa[0] + 1,) + sa[1:])     # keys up to x == size.x
        gb = (rng.random(sb) < 0.4) * rng.integers(1, 5, sb)
        a = PhysObject(sprite=np.zeros(sa, np.uint8), grid=ga.astype(np.uint8), grid_lo=(0, 0, 0), pos=tuple(rng.integers(-4, 5, 3) + rng.choice([0.0, 0.5], 3)),
                       rot=tuple(float(v) for v in rng.choice([0, 90, 180, 270], 3)), physics=True)
        b = PhysObject(sprite=gb.astype(np.uint8), pos=tuple(float(v) for v in rng.integers(-6, 7, 3)), rot=tuple(float(v) for v in rng.choice([0, 90, 180, 270], 3)))
        d = dirs[trial % 6]
        mins, maxs = a.box()
        k = [i for i in range(3) if d[i]][0]
        post6 = list(mins) + list(maxs)
        if d[k] < 0:
            post6[k], post6[3 + k] = mins[k] - 1, mins[k]
        else:
            post6[k], post6[3 + k] = maxs[k], maxs[k] + 1
        want = physics_np.slice_collide(props, world.friction, a, b, post6, d, 0.25, 1.5)
        got = world.slice_collide(a, b, post6, d, 0.25, 1.5)
        assert got == (float(want[0]), float(want[1]), bool(want[2])), (trial, got, want)
        blocked_seen += bool(want[2])
    assert 5 < blocked_seen < 60
    cam.close()
