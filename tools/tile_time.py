import time, sys, os
sys.path.insert(0, "/root/repo")
import numpy as np, torch, bench
from python_raytracer_b200 import GpuCamera, Settings
origin, cells, mats, poses = bench.load_workload()
sd = dict(bench.C3_SETTINGS)
cam = GpuCamera(Settings.from_object(sd), device=0, background_function=bench.BUILTIN_BG, background=bench.BUILTIN_BG)
cam.set_world_dense(origin, cells, poses[0]["chunk_res"], mats)
for i in range(3):
    cam.tile(0, 0)
ts = []
for i in range(20):
    p = poses[i % 4]
    cam.pos, cam.rot = p["pos"], p["rot"]; cam.set_chunk_res(p["chunk_res"])
    t0 = time.perf_counter(); img, trav, th = cam.tile(0, 0); ts.append(time.perf_counter() - t0)
print("tile() ms: median %.3f min %.3f  bytes %d traversed %d" % (1e3 * np.median(ts), 1e3 * min(ts), len(img), len(trav)))
