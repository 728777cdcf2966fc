#!/usr/bin/env python
"""Scene-building row measured against the HBM roofline: world build from objects, dense -> bricks,
chunk presence, in-place update of a moving object.  Each call is timed with CUDA events on the stream the library
launches on (vt_set_stream -> torch's current stream) and by the host clock; bytes are ALGORITHMIC (cells read once,
cells / bricks written once).  python tools/scene_build_bench.py [--quick] [--reps N] > gpurun_out/scene_build.json"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from python_raytracer_b200 import GpuCamera, Settings
from python_raytracer_b200.builder import SceneObject, WorldBuilder
from python_raytracer_b200.scenes import C5_KEYS, C5_MATERIALS, goxel_model, terrain_world
from python_raytracer_b200.goxel import sprite_grid

ap = argparse.ArgumentParser()
ap.add_argument("--quick", action="store_true", help="one repetition of every call (for an ncu pass)")
ap.add_argument("--reps", type=int, default=20)
args = ap.parse_args()
REPS = 1 if args.quick else args.reps
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    PEAK_SRC = "MEASURED_PEAKS.json hbm_gbs"
except Exception:
    PEAK, PEAK_SRC = 6500.0, "fallback (B200_PROFILING.md)"


def bg(r, s):
    raise AssertionError


def cam_for(cs=16):
    cam = GpuCamera(Settings.from_object(dict(width=64, height=48, chunk_size=cs, chunk_lod=2, dist_max=768, threads=1)),
                    background_function=bg, background=bg)
    cam._check(cam._lib.vt_set_stream(cam._ctx, torch.cuda.current_stream(cam.device).cuda_stream))
    return cam


def timed(fn, reps=REPS, flush=None):
    """median / min of the event time (ms) and of the host wall time (ms) of fn()"""
    ev, wall = [], []
    for k in range(reps + (0 if args.quick else 2)):
        if flush is not None:
            flush.zero_()                                    # 256 MiB memset: L2 does not hold the grid between calls
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        if args.quick or k >= 2:
            ev.append(e0.elapsed_time(e1)); wall.append(1e3 * (t1 - t0))
    return dict(ms=round(float(np.median(ev)), 4), ms_min=round(float(np.min(ev)), 4), wall_ms=round(float(np.median(wall)), 4))


def rate(entry, nbytes):
    entry["bytes"] = int(nbytes)
    entry["gbs"] = round(nbytes / entry["ms"] / 1e6, 1)
    entry["frac_of_hbm_peak"] = round(entry["gbs"] / PEAK, 4)
    return entry


res = {"peak_gbs": PEAK, "peak_source": PEAK_SRC, "reps": REPS, "l2_flush_between_calls": True}
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
import ctypes as C

# ---- C4: 512^3 terrain (128 MiB dense) -> bricks
origin, cells, mats = terrain_world(512)
dev = torch.from_numpy(cells).cuda()
cam = cam_for()
res_tab = np.ones((32, 32, 32), np.uint8)
cam.set_world_dense(origin, dev, res_tab, mats)
cam._check(cam._lib.vt_set_stream(cam._ctx, torch.cuda.current_stream(cam.device).cuda_stream))
st = cam.world_stats()
e = timed(lambda: cam._check(cam._lib.vt_upload_dense(cam._ctx, dev.data_ptr())), flush=flush)
res["c4_upload_dense"] = rate(e, cells.size + st["bricks"] * 512 + (cells.size // 512) * 4)
res["c4_upload_dense"].update(bricks=st["bricks"], dense_bytes=int(cells.size), note="reads the grid twice (mask pass, pack pass); algorithmic bytes count it once")
out = np.empty((32, 32, 32), np.uint8)
e = timed(lambda: cam._check(cam._lib.vt_chunk_presence(cam._ctx, out.ctypes.data, out.size)))
res["c4_chunk_presence"] = e
back = torch.empty_like(dev)
e = timed(lambda: cam._check(cam._lib.vt_download_dense(cam._ctx, back.data_ptr())), flush=flush)
res["c4_download_dense"] = rate(e, cells.size + st["bricks"] * 512)
assert torch.equal(back, dev)
cam.close(); del dev, back

# ---- C5: one 512 x 256 x 512 sprite (9.9 M voxels) -> world -> bricks
size = (512, 256, 512)
xyz_file, cols = goxel_model(size)
# the model as Sprite.load would place it (data.py:283-288): (size.x - x, z, y)
xyz = np.stack([size[0] - xyz_file[:, 0], xyz_file[:, 2], xyz_file[:, 1]], axis=1).astype(np.int32)
uniq, inv = np.unique(cols, return_inverse=True)
lut = np.array([C5_KEYS.index(str(u)) + 1 for u in uniq], np.uint8)
grid = sprite_grid(xyz, lut[inv], size)
cam = cam_for()
wb = WorldBuilder(cam)
obj = SceneObject(sprite=grid)
world = wb.build([obj], C5_MATERIALS)
cam._check(cam._lib.vt_set_stream(cam._ctx, torch.cuda.current_stream(cam.device).cuda_stream))
arr = wb._object_array([obj])
n_cells = int(np.prod(cam._world_shape))
e = timed(lambda: cam._check(cam._lib.vt_build_world(cam._ctx, arr, 1, world.data_ptr())), flush=flush)
res["c5_build_world"] = rate(e, grid.size + n_cells)
res["c5_build_world"].update(world_shape=list(cam._world_shape), sprite_bytes=int(grid.size))
e = timed(lambda: cam._check(cam._lib.vt_upload_dense(cam._ctx, world.data_ptr())), flush=flush)
st = cam.world_stats()
res["c5_upload_dense"] = rate(e, n_cells + st["bricks"] * 512 + (n_cells // 512) * 4)
res["c5_upload_dense"].update(bricks=st["bricks"])
t = timed(lambda: wb.build([obj], C5_MATERIALS), reps=max(1, REPS // 4))
res["c5_worldbuilder_build_call"] = dict(wall_ms=t["wall_ms"], note="WorldBuilder.build: set_world + build + brick + presence + LOD table, host clock")
# a small object moving through the model: in-place update of the dirty chunks
cube = SceneObject(sprite=np.full((16, 16, 16), 2, np.uint8), pos=(40.0, 150.0, 40.0))
wb.build([obj, cube], C5_MATERIALS, margin=1)
ups, k = [], 0
for k in range(10 if not args.quick else 1):
    cube.pos = (40.0 + 3.0 * k, 150.0, 40.0 + 2.0 * k)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    n_chunks = wb.update([obj, cube], [cube]); torch.cuda.synchronize()
    ups.append(1e3 * (time.perf_counter() - t0))
res["c5_moving_object_update"] = dict(wall_ms=round(float(np.median(ups)), 4), chunks_redrawn=int(n_chunks))
cam.close()
print(json.dumps(res, indent=1))
