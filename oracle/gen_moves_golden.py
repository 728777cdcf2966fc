"""Fixture for incremental scene updates: the UNMODIFIED reference's Window.chunk_update run several
times while objects move, turn, disappear and leave the initial box.  TEST INFRASTRUCTURE ONLY.
    python -m oracle.gen_moves_golden   ->  tests/golden/default_moves.npz

Per step: the state of every object of the default mod right before chunk_update ran (present in data.objects,
pos, rot, visible, redraw flag) and the merged full-resolution world chunk_update left behind (cells by absolute
position, material ids as in tests/golden/default_world.npz).  Sprites are those of tests/golden/default_objects.npz.
What the fixture pins: which chunks are rebuilt (old and new boxes of redrawn objects, init.py:397-428) and the
override order after a redraw (a redrawn object is re-inserted at the end of chunks_objects, init.py:437-439)."""
from __future__ import annotations

import os
import types

import numpy as np

from oracle import refharness as rh

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def main():
    ref = rh.load_reference()
    data, lib = ref.data, ref.lib
    vec3 = lib.vec3
    names = [str(n) for n in np.load(os.path.join(OUT, "default_world.npz"))["names"]]
    cam = rh.make_camera((-12.0, 4.0, 4.0), (0.0, 0.0, 0.0))
    objs = list(data.objects.values())
    for o in objs:
        o.redraw = True
    ws = types.SimpleNamespace(cam=cam, chunks={}, chunks_objects={}, timer=0.0, traversed=[[]])
    cube = [k for k, o in enumerate(objs) if o.sprite and max(o.sprite.size.x, o.sprite.size.y, o.sprite.size.z) <= 16]
    print("objects", len(objs), "small ones", cube)
    a, b = objs[cube[0]], objs[cube[1]]

    def act0():
        pass

    def act1():                                  # a jumps into b: a is redrawn later than b -> a wins the overlap
        a.move(vec3(b.pos.x + 3, b.pos.y + 1, b.pos.z - 2))

    def act2():                                  # b moves one cell: now b is the last one redrawn
        b.move(vec3(b.pos.x + 1, b.pos.y, b.pos.z))

    def act3():                                  # a turns a quarter around y and z
        a.rotate(vec3(0, 90, 90))

    def act4():                                  # b is removed from the scene
        b.remove()

    def act5():                                  # a leaves the box the scene started with
        a.move(vec3(a.pos.x + 150, a.pos.y + 40, a.pos.z))

    out = {"count": np.int32(len(objs))}
    for s, act in enumerate((act0, act1, act2, act3, act4, act5)):
        act()
        for o in list(data.objects.values()):
            o.update(cam.pos)
        state = np.zeros((len(objs), 9), np.float64)
        for k, o in enumerate(objs):
            present = o.id in data.objects
            state[k] = (present, o.pos.x, o.pos.y, o.pos.z, o.rot.x, o.rot.y, o.rot.z, bool(o.visible), bool(o.redraw))
        ws.timer = 0.0
        ref.Window.chunk_update(ws, max(1.0, data.settings.chunk_time + 1.0))
        lo, cells, _, _, mat_names = rh.extract_world(ws, cam)
        remap = np.zeros(len(mat_names), np.uint8)
        for i, nm in enumerate(mat_names):
            if i:
                remap[i] = names.index(nm)
        out["state_%d" % s] = state
        out["origin_%d" % s] = lo.astype(np.int32)
        out["cells_%d" % s] = remap[cells]
        print("step", s, "redraw", [k for k in range(len(objs)) if state[k, 8]], "origin", lo, "shape", cells.shape,
              "voxels", int((cells > 0).sum()))
    out["steps"] = np.int32(6)
    np.savez_compressed(os.path.join(OUT, "default_moves.npz"), **out)


if __name__ == "__main__":
    main()
