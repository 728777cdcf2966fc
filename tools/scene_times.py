#!/usr/bin/env python
"""Frame times of the large named configurations (dev tool): C4 terrain 4K 16 spp, C5 Goxel model 1080p 8 spp + culling,
world build / brick statistics.  python tools/scene_times.py"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from python_raytracer_b200 import GpuCamera, Settings
from python_raytracer_b200.builder import SceneObject, WorldBuilder
from python_raytracer_b200.goxel import load_goxel, sprite_grid, write_goxel
from python_raytracer_b200.mathutil import euler_to_quaternion
from python_raytracer_b200.scenes import C5_KEYS, C5_MATERIALS, goxel_model, terrain_world
from python_raytracer_b200.world import chunk_presence, lod_table
from tests.test_gpu_scenes import C4_SETTINGS, C5_SETTINGS


def bg(r, s):
    raise AssertionError


def timed(cam, frames=5, want=("rgba8",)):
    out = {}
    for _ in range(2):
        cam.render(want=want, out=out)
    ts = []
    for _ in range(frames):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); cam.render(want=want, out=out, prepare=False); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


res = {}
# ---- C4
sd = dict(C4_SETTINGS); sd["threads"] = 1
origin, cells, mats = terrain_world(512)
cam = GpuCamera(Settings.from_object(sd), background_function=bg, background=bg)
pos = (0.0, 300.0, 0.0)
present = chunk_presence(cells, 16)
t0 = time.perf_counter()
dev_cells = torch.from_numpy(cells).cuda(); torch.cuda.synchronize()
t1 = time.perf_counter()
cam.set_world_dense(origin, dev_cells, lod_table(origin, present, pos, 16, sd["chunk_lod"], sd["dist_max"]), mats)
torch.cuda.synchronize(); t2 = time.perf_counter()
st = cam.world_stats()
c4 = []
for yaw in (0.0, 90.0, 180.0, 270.0):
    cam.pos, cam.rot = pos, euler_to_quaternion(0.0, yaw, 30.0)
    ms = timed(cam)
    cnt = cam.render(want=("rgba8", "counters"))["counters"].cpu().numpy().astype(float)
    c4.append(dict(yaw=yaw, ms=round(ms, 3), rays=int(cnt[0]), grays_per_s=round(cnt[0] / ms / 1e6, 3), samples_per_ray=round((cnt[1] + cnt[2]) / cnt[0], 2),
                   hits_per_ray=round(cnt[3] / cnt[0], 3)))
res["c4"] = dict(frames=c4, bricks=st["bricks"], world_bytes=st["bytes"], dense_bytes=int(cells.size), h2d_s=round(t1 - t0, 4), brick_s=round(t2 - t1, 4))
cam.close(); del dev_cells
if os.environ.get("ST_ONLY") == "c4":                    # quick A/B of kernel variants on the terrain (VT_LIB=...)
    print(json.dumps(res)); sys.exit(0)
# ---- C5
sd = dict(C5_SETTINGS); sd["threads"] = 1
size = (512, 256, 512)
xyz_file, cols = goxel_model(size)
path = "/tmp/vt_c5_model.txt"
t0 = time.perf_counter(); write_goxel(path, xyz_file, cols); t1 = time.perf_counter()
xyz, m = load_goxel(path, size[0], C5_KEYS); t2 = time.perf_counter()
grid = sprite_grid(xyz, m, size); t3 = time.perf_counter()
cam = GpuCamera(Settings.from_object(sd), background_function=bg, background=bg)
cam.pos, cam.rot = (-300.0, 40.0, -300.0), euler_to_quaternion(0.0, 45.0, 10.0)
wb = WorldBuilder(cam)
obj = SceneObject(sprite=grid)
wb.build([obj], C5_MATERIALS); torch.cuda.synchronize()
t4 = time.perf_counter(); wb.build([obj], C5_MATERIALS); torch.cuda.synchronize(); t5 = time.perf_counter()
res_all = wb.update_lod(None)
ms_all = timed(cam)
out = cam.render(want=("rgba8", "visited"))
vis = cam.visited_slots(out["visited"]).copy()
res_c = wb.update_lod(visible=vis)
ms_cull = timed(cam)
cnt = cam.render(want=("rgba8", "counters"))["counters"].cpu().numpy().astype(float)
res["c5"] = dict(voxels=int(len(xyz)), file_mb=round(os.path.getsize(path) / 1e6, 1), write_s=round(t1 - t0, 2), parse_s=round(t2 - t1, 3),
                 grid_s=round(t3 - t2, 3), build_and_brick_s=round(t5 - t4, 4), ms_all_chunks=round(ms_all, 3), ms_culled=round(ms_cull, 3),
                 chunks_all=int((res_all > 0).sum()), chunks_culled=int((res_c > 0).sum()), rays=int(cnt[0]),
                 grays_per_s=round(cnt[0] / ms_cull / 1e6, 3), bricks=cam.world_stats()["bricks"])
# ---- a small object moving through the C5 world: incremental update (dirty chunks only) vs full rebuild
cube = SceneObject(sprite=np.full((16, 16, 16), 2, np.uint8), pos=(40.0, 150.0, 40.0))
wb.build([obj, cube], C5_MATERIALS, margin=1); torch.cuda.synchronize()
ups = []
for k in range(20):
    cube.pos = (40.0 + 3.0 * k, 150.0, 40.0 + 2.0 * k)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    n_chunks = wb.update([obj, cube], [cube]); torch.cuda.synchronize()
    ups.append(time.perf_counter() - t0)
t0 = time.perf_counter(); wb.build([obj, cube], C5_MATERIALS, margin=1); torch.cuda.synchronize(); t_full = time.perf_counter() - t0
res["c5_moving_object"] = dict(update_ms=round(1e3 * float(np.median(ups)), 3), chunks_redrawn=int(n_chunks), full_rebuild_ms=round(1e3 * t_full, 2))
os.remove(path)
print(json.dumps(res, indent=1))
