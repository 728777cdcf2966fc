"""Times the REFERENCE'S OWN Python trace loop (Camera.tile, init.py:126-150) on the host cores -- bench.py's
cpu_baseline of kind "reference".  TEST / MEASUREMENT INFRASTRUCTURE ONLY.

Runs from /root/reference when the tree exists (build container) and from the bytecode build oracle/_ref/
(oracle/build_ref.py) otherwise (GPU box).  The workload is C1: the default mod exactly as
shipped (64 x 48, 1 spp, dof 0.5, max_bounces 2, dist_max 192, chunk_lod 2), culling off so that all 198 chunks are
resident.  Like the reference's frame loop (init.py:162, 193-205) the frame is fanned out over a persistent
multiprocessing pool with one pixel list per worker; wall time from dispatch to the last result.  The scene build
(Window.chunk_update, ~5 s of Python) is timed separately and excluded.
"""
from __future__ import annotations

import multiprocessing as mp
import os
import time

from oracle import refharness as rh


def time_reference(seconds: float = 8.0, procs: int | None = None, max_frames: int = 400):
    """-> dict(rays_per_sec, frames, rays_per_frame, seconds, procs, build_seconds) or raises RuntimeError."""
    ref = rh.load_reference()
    data = ref.data
    rh.restore_settings()
    rh.override_settings(culling=False, dist_move=0)
    s = data.settings
    cam = ref.Camera()
    cam.pos, cam.rot = data.player.cam_pos, data.player.cam_rot
    t0 = time.perf_counter()
    rh.build_chunks(cam)
    t_build = time.perf_counter() - t0
    procs = procs or (os.cpu_count() or 1)
    # data.py:70-77: pixel (x, y) belongs to worker (x ^ y) % threads, x-major order
    lists = [[] for _ in range(procs)]
    for x in range(s.width):
        for y in range(s.height):
            lists[(x ^ y) % procs].append((x, y))
    s.pixels = lists
    rh._WORK["cam"] = cam
    rays_frame = s.width * s.height * max(1, s.samples)
    frames, t_total = 0, 0.0
    try:
        ctx = mp.get_context("fork")
        with ctx.Pool(procs) as pool:
            pool.map(_tile, range(procs))                          # warm-up frame
            while t_total < seconds and frames < max_frames:
                t0 = time.perf_counter()
                pool.map(_tile, range(procs))
                t_total += time.perf_counter() - t0
                frames += 1
    finally:
        rh._WORK.clear()
        rh.restore_settings()
    return dict(rays_per_sec=rays_frame * frames / t_total, frames=frames, rays_per_frame=rays_frame, seconds=t_total,
                procs=procs, build_seconds=t_build, source="tree" if rh.available() else "bytecode (oracle/_ref)")


class C3Reference:
    """The reference's Camera.tile on bench.py's C3 workload (default scene, 1920 x 1080, 4 spp, 4 bounces, dof 0.5, four
    poses), on a strided pixel sample: a full frame is 8.3 M rays, minutes of Python.  One camera per pose, each holding
    the chunk frames the reference's own chunk_update picked for it; one persistent fork pool."""

    def __init__(self, settings: dict, poses_euler, stride: int, procs: int | None = None):
        ref = rh.load_reference()
        data = ref.data
        rh.restore_settings()
        keys = ("width", "height", "samples", "static", "shutter", "falloff", "dof", "dist_min", "dist_max", "max_light",
                "max_bounces", "lod_bounces", "lod_samples", "lod_random", "lod_edge", "chunk_size", "chunk_lod", "fov")
        rh.override_settings(culling=False, dist_move=0, **{k: settings[k] for k in keys})
        s = data.settings
        self.procs = procs or (os.cpu_count() or 1)
        t0 = time.perf_counter()
        self.cams, st = [], None
        for pos, eul in poses_euler:
            cam = rh.make_camera(pos, eul)
            cam.lens = s.fov * 3.141592653589793 / 8
            st = rh.build_chunks(cam, st)
            self.cams.append(cam)
        self.build_seconds = time.perf_counter() - t0
        px = [(x, y) for x in range(stride // 2, s.width, stride) for y in range(stride // 2, s.height, stride)]
        lists = [[] for _ in range(self.procs)]
        for x, y in px:
            lists[(x ^ y) % self.procs].append((x, y))
        s.pixels = lists
        self.rays_per_step = len(px) * max(1, s.samples)
        rh._WORK["cams"] = self.cams
        self.pool = mp.get_context("fork").Pool(self.procs)
        self.source = "tree" if rh.available() else "bytecode (oracle/_ref)"

    def step(self, pose_index: int) -> float:
        t0 = time.perf_counter()
        self.pool.map(_tile_pose, [(pose_index % len(self.cams), t) for t in range(self.procs)])
        return time.perf_counter() - t0

    def close(self):
        self.pool.close()
        self.pool.join()
        rh._WORK.clear()
        rh.restore_settings()


def _tile_pose(args):
    k, t = args
    image, traversed, thread = rh._WORK["cams"][k].tile(t, 0)
    return len(image)


def _tile(t):
    cam = rh._WORK["cam"]
    image, traversed, thread = cam.tile(t, 0)
    return len(image)


if __name__ == "__main__":
    print(time_reference(4.0))
