"""Fixtures for the scene-building rows, made by running the UNMODIFIED reference here.
TEST INFRASTRUCTURE ONLY.   python -m oracle.gen_objects_golden

  tests/golden/default_objects.npz   every object of the default mod as the reference holds it after import
      (data.py:430-624): dense sprite grid of material ids (Sprite.get_voxels of the active frame), pos, rot, sprite
      size, mins, maxs -- the inputs of Window.chunk_update (init.py:389-452).  The expected output is
      tests/golden/default_world.npz (the chunk frames chunk_update built, flattened cell by cell).
  tests/golden/goxel_quirks.txt(.gz) + goxel_quirks.npz   a small hand-written Goxel text file exercising the
      parser rules of Sprite.load (data.py:271-289) and the voxel dict the reference's own loader produced.
"""
from __future__ import annotations

import gzip
import os
import sys

import numpy as np

from oracle import refharness as rh

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

QUIRKS = """# Goxel 0.14.0
# One line per voxel
# X Y Z RRGGBB
4 2 0 7f7f7f
5 2 0 7f7f7f
0 0 0 ffffff
12 11 10 ffffff
4 2 0 ffffff
3 3 1 7F7F7F
3 3 2 123456
-1 2 3 7f7f7f
6  2 3 7f7f7f
7 2 3 7f7f7f extra
  8 2 3 ffffff  

9 9 9 7f7f7f\r
10 1 1 ffffff\t
1.5 2 3 7f7f7f
+2 2 3 7f7f7f
007 008 009 ffffff
11 5 5 7f7f7f
"""


def main():
    ref = rh.load_reference()
    data, lib = ref.data, ref.lib
    world = np.load(os.path.join(OUT, "default_world.npz"))
    names = [str(n) for n in world["names"]]
    mod = sys.modules["mods.default.init"]
    ident = {id(v): k for k, v in vars(mod).items() if isinstance(v, data.Material)}
    objs = {}
    for k, obj in enumerate(data.objects.values()):
        spr = obj.get_sprite()
        size = (int(spr.size.x), int(spr.size.y), int(spr.size.z))
        dense = np.zeros(size, np.uint8)
        extra = 0
        for (x, y, z), m in spr.get_voxels(None).items():
            if 0 <= x < size[0] and 0 <= y < size[1] and 0 <= z < size[2]:
                nm = ident[id(m)]
                dense[x, y, z] = names.index(nm) if nm in names else 255
            else:
                extra += 1
        objs["sprite_%d" % k] = dense
        objs["pos_%d" % k] = np.array([obj.pos.x, obj.pos.y, obj.pos.z], np.float64)
        objs["rot_%d" % k] = np.array([obj.rot.x, obj.rot.y, obj.rot.z], np.float64)
        objs["mins_%d" % k] = np.array([obj.mins.x, obj.mins.y, obj.mins.z], np.int32)
        objs["maxs_%d" % k] = np.array([obj.maxs.x, obj.maxs.y, obj.maxs.z], np.int32)
        print("object %d size %s pos %s rot %s voxels %d (outside the drawn box: %d)" % (
            k, size, objs["pos_%d" % k], objs["rot_%d" % k], int((dense > 0).sum()), extra))
    objs["count"] = np.int32(len(data.objects))
    np.savez_compressed(os.path.join(OUT, "default_objects.npz"), **objs)

    # ---- Sprite.load on the quirks file, plain and gzip ----
    txt = os.path.join(OUT, "goxel_quirks.txt")
    with open(txt, "wt", newline="") as f:
        f.write(QUIRKS)
    with gzip.open(txt + ".gz", "wt", newline="") as f:
        f.write(QUIRKS)
    mats = {"7f7f7f": data.Material(albedo=lib.rgb(1, 1, 1)), "ffffff": data.Material(albedo=lib.rgb(2, 2, 2))}
    key_id = {id(mats["7f7f7f"]): 1, id(mats["ffffff"]): 2}
    res = {}
    for tag, path in (("txt", txt), ("gz", txt + ".gz")):
        spr = data.Sprite(size=lib.vec3(12, 12, 12), frames=1, lod=0)
        spr.load([path], mats)
        vox = spr.get_voxels(0)
        arr = np.array(sorted((x, y, z, key_id[id(m)]) for (x, y, z), m in vox.items()), np.int32).reshape(-1, 4)
        res[tag] = arr
        print(tag, "voxels", len(arr))
    assert np.array_equal(res["txt"], res["gz"])
    np.savez_compressed(os.path.join(OUT, "goxel_quirks.npz"), voxels=res["txt"], size=np.array([12, 12, 12], np.int32),
                        keys=np.array(["7f7f7f", "ffffff"]))
    print(res["txt"])

    # ---- Sprite.get_voxel under rotation (data.py:338-371, 417-419): cubic and non-cubic sprites ----
    rng = np.random.default_rng(99)
    rot_cases = {}
    rots = [(0, 0, 0), (90, 0, 0), (0, 90, 0), (0, 0, 90), (180, 270, 90), (270, 180, 0), (-90, 0, 450), (45, 135, 225),
            (100, 44.9, 314), (90, 90, 90)]
    for tag, size in (("cube", (6, 6, 6)), ("slab", (4, 6, 6)), ("bar", (4, 6, 8)), ("xz", (6, 4, 6))):
        spr = data.Sprite(size=lib.vec3(*size), frames=1, lod=0)
        grid = (rng.random(size) < 0.35) * rng.integers(1, 4, size)
        mats3 = [None, data.Material(albedo=lib.rgb(1, 1, 1)), data.Material(albedo=lib.rgb(2, 2, 2)), data.Material(albedo=lib.rgb(3, 3, 3))]
        mid = {id(m): k for k, m in enumerate(mats3) if m is not None}
        vox = {(int(x), int(y), int(z)): mats3[int(grid[x, y, z])] for x, y, z in zip(*np.nonzero(grid))}
        spr.get_frame(0).set_voxels(vox, True)
        outs = []
        for rot in rots:
            r = lib.vec3(*rot)
            o = np.zeros(size, np.uint8)
            for x in range(size[0]):
                for y in range(size[1]):
                    for z in range(size[2]):
                        m = spr.get_voxel(None, lib.vec3(x, y, z), r)
                        o[x, y, z] = mid[id(m)] if m is not None else 0
            outs.append(o)
        rot_cases["sprite_" + tag] = grid.astype(np.uint8)
        rot_cases["drawn_" + tag] = np.stack(outs)
    rot_cases["rots"] = np.array(rots, np.float64)
    np.savez_compressed(os.path.join(OUT, "sprite_rotations.npz"), **rot_cases)
    print("rotation cases written", {k: v.shape for k, v in rot_cases.items()})


if __name__ == "__main__":
    main()
