#!/usr/bin/env python
"""A/B of the world-build kernels (dev tool): vt_build_world on the C5 world (one 512 x 256 x 512 sprite -> 528 x 272 x 528
cells) and on a many-object world (the default mod's 8 unrotated objects tiled 5 x 5: 200 objects), for buildrows = 0 (build_cells_kernel),
2 and 4 (build_cells16_kernel).  CUDA events on the library's stream, L2 flushed between calls, median of 20; the
worlds of all variants are compared cell for cell.  Prints one JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from python_raytracer_b200 import GpuCamera, Settings
from python_raytracer_b200.builder import SceneObject, WorldBuilder
from python_raytracer_b200.scenes import C5_MATERIALS, C5_SIZE, goxel_model_grid

PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6500.0


def scene(which):
    if which == "c5":
        return [SceneObject(sprite=goxel_model_grid(C5_SIZE), pos=(0.0, 0.0, 0.0))], C5_MATERIALS
    o = np.load(os.path.join(ROOT, "tests", "golden", "default_objects.npz"))
    w = np.load(os.path.join(ROOT, "tests", "golden", "default_world.npz"))
    objs = []
    for tx in range(5):
        for tz in range(5):
            for k in range(int(o["count"])):
                p = o["pos_%d" % k]
                objs.append(SceneObject(sprite=o["sprite_%d" % k], pos=(float(p[0]) + 144 * tx, float(p[1]), float(p[2]) + 144 * tz),
                                        rot=tuple(float(v) for v in o["rot_%d" % k])))
    return objs, w["materials"]


def main():
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    res = {"peak_gbs": PEAK}
    for which in ("c5", "tiled_default"):
        objs, mats = scene(which)
        ref = None
        for rows in (0, 2, 4):
            cam = GpuCamera(Settings.from_object(dict(width=64, height=48, chunk_size=16, chunk_lod=2, dist_max=768, threads=1)), background=None)
            cam._check(cam._lib.vt_set_tuning(cam._ctx, b"buildrows", rows))
            cam._check(cam._lib.vt_set_stream(cam._ctx, torch.cuda.current_stream(cam.device).cuda_stream))
            wb = WorldBuilder(cam)
            world = wb.build(objs, mats)
            arr = wb._object_array(wb.order)
            ts = []
            for k in range(22):
                flush.zero_()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); cam._check(cam._lib.vt_build_world(cam._ctx, arr, len(wb.order), world.data_ptr())); e1.record()
                torch.cuda.synchronize()
                if k >= 2:
                    ts.append(e0.elapsed_time(e1))
            if ref is None:
                ref = world.clone()
            same = bool(torch.equal(world, ref))
            ms = float(np.median(ts))
            nbytes = int(sum(int(np.prod(o.size)) for o in objs if True) + world.numel())
            res["%s/rows%d" % (which, rows)] = dict(ms=round(ms, 4), gbs=round(nbytes / ms / 1e6, 1), frac_of_hbm_peak=round(nbytes / ms / 1e6 / PEAK, 4),
                                                     bytes=nbytes, world=list(world.shape), objects=len(objs), same_cells_as_rows0=same)
            cam.close()
    print(json.dumps(res))


if __name__ == "__main__":
    main()
