#!/usr/bin/env python
"""Frame time vs image size on the bench scene (dev tool): shows the fixed cost of a launch (start-up + tail)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch, bench
from python_raytracer_b200 import GpuCamera, Settings
origin, cells, mats, poses = bench.load_workload()
for w, h in ((1920, 1080), (1920, 540), (1920, 272), (1920, 136), (960, 136), (480, 136), (240, 136), (64, 48)):
    sd = dict(bench.C3_SETTINGS); sd.update(width=w, height=h)
    cam = GpuCamera(Settings.from_object(sd), device=0, background_function=bench.BUILTIN_BG, background=bench.BUILTIN_BG)
    cam.set_world_dense(origin, cells, poses[0]["chunk_res"], mats)
    out, res = {}, []
    for p in poses:
        cam.pos, cam.rot = p["pos"], p["rot"]; cam.set_chunk_res(p["chunk_res"]); cam._prepare()
        for _ in range(3):
            cam.render(want=("rgba8",), prepare=False, out=out, clear=False)
        ts = []
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); cam.render(want=("rgba8",), prepare=False, out=out, clear=False); e1.record(); e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        res.append(float(np.median(ts)))
    rays = w * h * 4
    print("%4dx%-4d rays %8d  ms/pose %s  mean %.4f  ns/ray %.3f" % (w, h, rays, [round(v, 4) for v in res], np.mean(res), 1e6 * np.mean(res) / rays), flush=True)
    cam.close()
