"""Host side of the scene-building row: geometry helpers and the reference-object adapter (no GPU needed)."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_object_boxes_and_angles_follow_the_reference():
    from python_raytracer_b200.builder import SceneObject, world_box
    objs = np.load(os.path.join(GOLD, "default_objects.npz"))
    scene = [SceneObject(sprite=objs["sprite_%d" % k], pos=tuple(objs["pos_%d" % k]), rot=tuple(objs["rot_%d" % k]))
             for k in range(int(objs["count"]))]
    for k, o in enumerate(scene):
        mins, maxs = o.box()                          # data.py:480-484, 592-594 as the reference computed them
        assert mins == objs["mins_%d" % k].tolist() and maxs == objs["maxs_%d" % k].tolist()
    origin, dims = world_box(scene, 16)
    world = np.load(os.path.join(GOLD, "default_world.npz"))
    assert (origin <= world["origin"]).all() and (origin + dims >= world["origin"] + np.array(world["cells"].shape)).all()
    assert (origin % 16 == 0).all() and (dims % 16 == 0).all()
    # round(rot / 90) % 4 with Python's banker's rounding (data.py:339-342): 45 -> 0, 135 -> 2, 225 -> 2, -90 -> 3
    o = SceneObject(sprite=np.zeros((2, 2, 2), np.uint8), rot=(45.0, 135.0, 225.0))
    assert o.angles() == [0, 2, 2]
    assert SceneObject(sprite=np.zeros((2, 2, 2), np.uint8), rot=(-90.0, 450.0, 314.0)).angles() == [3, 1, 3]
    # fractional positions: mins = ceil(pos) - half, maxs = floor(pos) + half
    assert SceneObject(sprite=np.zeros((4, 4, 4), np.uint8), pos=(1.5, 0.0, -0.5)).box() == ([0, -2, -2], [3, 2, 1])


def test_scene_from_reference_objects():
    """Adapter over live reference objects, where the reference is importable (the build container only)."""
    from oracle import refharness as rh
    if not rh.available():
        pytest.skip("reference tree not present")
    from python_raytracer_b200.builder import scene_from_reference
    ref = rh.load_reference()
    world = np.load(os.path.join(GOLD, "default_world.npz"))
    names = [str(n) for n in world["names"]]
    import sys
    mod = sys.modules["mods.default.init"]
    ident = {id(v): k for k, v in vars(mod).items() if isinstance(v, ref.data.Material)}
    for o in ref.data.objects.values():
        o.visible = True
    scene = scene_from_reference(ref.data.objects.values(), lambda m: names.index(ident[id(m)]))
    objs = np.load(os.path.join(GOLD, "default_objects.npz"))
    assert len(scene) == int(objs["count"])
    for k, o in enumerate(scene):
        assert np.array_equal(o.sprite, objs["sprite_%d" % k])
        assert list(o.pos) == objs["pos_%d" % k].tolist() and list(o.rot) == objs["rot_%d" % k].tolist()


def test_chunk_box_matches_reference_chunk_loop():
    """chunk_box = the chunks Window.chunk_update loops over for an object (init.py:406-411: mins.snapped ..
    maxs.snapped inclusive), checked on the object states the reference recorded in default_moves.npz."""
    import os
    import numpy as np
    from python_raytracer_b200.builder import SceneObject, chunk_box
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    objs, mv = np.load(os.path.join(gold, "default_objects.npz")), np.load(os.path.join(gold, "default_moves.npz"))
    for s in range(int(mv["steps"])):
        st = mv["state_%d" % s]
        for k in range(int(mv["count"])):
            o = SceneObject(sprite=objs["sprite_%d" % k], pos=tuple(st[k, 1:4]), rot=tuple(st[k, 4:7]))
            mins, maxs = o.box()
            lo, hi = chunk_box(o, 16)
            for a in range(3):
                assert lo[a] % 16 == 0 and hi[a] % 16 == 0
                assert lo[a] <= mins[a] < lo[a] + 16 and hi[a] - 16 <= maxs[a] < hi[a]
