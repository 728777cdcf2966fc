#!/usr/bin/env python
"""Measurement of the physics row: one slice test of a large object against a large object (the 512 x 256 x 512 C5 model
resting on a copy of itself: a 513 x 513-cell slice, 263 169 cells) on the GPU, and the same test through the plain-Python
restatement of the reference's loop on a strip of the slice (the reference visits cells at ~1 us each in CPython; the
restatement does the same lookups on numpy grids).  Prints one JSON line."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from oracle import physics_np
from python_raytracer_b200 import GpuCamera, Settings
from python_raytracer_b200.physics import PhysObject, PhysicsWorld
from python_raytracer_b200.scenes import C5_SIZE, goxel_model_grid


def main():
    grid = goxel_model_grid(C5_SIZE)
    props = np.array([[0, 0, 0]] + [[1, 0.1 * k, 0.05 * k] for k in range(1, 9)], np.float64)
    cam = GpuCamera(Settings.from_object(dict(width=8, height=8, chunk_size=16)), background=None)
    world = PhysicsWorld(cam, props, gravity=1.0, friction=1.0, friction_air=0.1, max_velocity=10.0, min_velocity=0.01)
    cut = np.ascontiguousarray(grid[:, 40:, :])                       # the model without its empty bottom margin: its lowest layer holds voxels
    a = PhysObject(sprite=cut, pos=(3.0, 150.0, -2.0), physics=True)  # sunk into the copy below: thousands of contacts
    b = PhysObject(sprite=grid, pos=(0.0, 0.0, 0.0))
    mins, maxs = a.box()
    post6 = (mins[0], mins[1] - 1, mins[2], maxs[0], mins[1], maxs[2])
    d = (0, -1, 0)
    world.slice_collide(a, b, post6, d, 0.0, 0.0)                     # warm-up (uploads the grids)
    ts = []
    for _ in range(20):
        t0 = time.perf_counter()
        got = world.slice_collide(a, b, post6, d, 0.0, 0.0)
        ts.append(time.perf_counter() - t0)
    cells = (post6[3] - post6[0] + 1) * (post6[4] - post6[1] + 1) * (post6[5] - post6[2] + 1)
    strip = (post6[0], post6[1], post6[2], post6[0] + 15, post6[4], post6[5])   # 16 of the 513 x-rows
    t0 = time.perf_counter()
    physics_np.slice_collide(props, 1.0, a, b, strip, d, 0.0, 0.0)
    cpu_s = time.perf_counter() - t0
    strip_cells = 16 * 2 * (post6[5] - post6[2] + 1)
    print(json.dumps({"what": "vt_slice_collide, C5 model on a copy of itself", "cells": cells, "blocked": got[2], "friction_sum": got[0],
                      "gpu_call_ms_median": 1000 * float(np.median(ts)), "gpu_cells_per_s": cells / float(np.median(ts)),
                      "cpu_restatement_cells_per_s": strip_cells / cpu_s, "cpu_sample": "%d cells of the same slice, plain Python" % strip_cells}))
    cam.close()


if __name__ == "__main__":
    main()
