#!/usr/bin/env python
"""Kernel A/B timing on the three named workloads (dev tool):  python tools/cbench.py [variant ...]

A variant is a comma-separated list of key=value pairs for vt_set_tuning (voidruns=0, runs=0, refill=n, ...), optionally
LIB=<alternative libvoxtrace.so> and WL=c3c4c5 (which workloads; default all).  Prints median ms/frame per pose and the
frame CRCs (a variant must not change them)."""
import json
import os
import sys
import zlib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import bench
from python_raytracer_b200 import GpuCamera, Settings, abi

_CACHE = {}


def workloads(which):
    out = []
    if "c3" in which:
        if "c3" not in _CACHE:
            _CACHE["c3"] = bench.load_workload()
        origin, cells, mats, poses = _CACHE["c3"]
        out.append(("c3", bench.C3_SETTINGS, origin, cells, mats, poses))
    if "c4" in which:
        from python_raytracer_b200.scenes import C4_SETTINGS, c4_workload
        if "c4" not in _CACHE:
            _CACHE["c4"] = c4_workload()
        origin, cells, mats, res, poses = _CACHE["c4"]
        out.append(("c4", C4_SETTINGS, origin, cells, mats, [dict(p, chunk_res=res) for p in poses]))
    return out


def time_poses(cam, poses, frames, flush, want=("rgba8",)):
    ms, crcs, out = [], [], {}
    for p in poses:
        cam.pos, cam.rot = p["pos"], p["rot"]
        if p.get("chunk_res") is not None:
            cam.set_chunk_res(p["chunk_res"])
        cam._prepare()
        for _ in range(2):
            cam.render(want=want, prepare=False, out=out)
        ts = []
        for _ in range(frames):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); cam.render(want=want, prepare=False, out=out); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms.append(float(np.median(ts)))
        crcs.append("%08x" % (zlib.crc32(out["rgba8"].cpu().numpy().tobytes()) & 0xffffffff))
    return ms, crcs


def run_variant(spec):
    env = dict(kv.split("=", 1) for kv in spec.split(",") if kv)
    lib = env.pop("LIB", None)
    which = env.pop("WL", "c3c4c5")
    abi._lib = abi.load(os.path.join(ROOT, lib)) if lib else None
    if not lib:
        abi.load()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    res = {"variant": spec}

    def tune(cam):
        for k, v in env.items():
            cam._check(cam._lib.vt_set_tuning(cam._ctx, k.encode(), int(v)))

    for name, sd, origin, cells, mats, poses in workloads(which):
        sd = dict(sd); sd["threads"] = 1
        cam = GpuCamera(Settings.from_object(sd), device=0, background_function=bench.BUILTIN_BG, background=bench.BUILTIN_BG)
        tune(cam)
        cam.set_world_dense(origin, torch.from_numpy(np.ascontiguousarray(cells)), poses[0]["chunk_res"], mats)
        ms, crcs = time_poses(cam, poses, 8 if name == "c3" else 3, flush)
        res[name] = {"ms": [round(v, 3) for v in ms], "mean": round(float(np.mean(ms)), 3), "crc": crcs}
        cam.close()
    if "c5" in which:
        from python_raytracer_b200.builder import SceneObject, WorldBuilder
        from python_raytracer_b200.mathutil import euler_to_quaternion
        from python_raytracer_b200.scenes import C5_EULER, C5_MATERIALS, C5_POS, C5_SETTINGS, C5_SIZE, goxel_model_grid
        if "c5" not in _CACHE:
            _CACHE["c5"] = goxel_model_grid(C5_SIZE)
        sd = dict(C5_SETTINGS); sd["threads"] = 1
        cam = GpuCamera(Settings.from_object(sd), device=0, background_function=bench.BUILTIN_BG, background=bench.BUILTIN_BG)
        tune(cam)
        cam.pos, cam.rot = C5_POS, euler_to_quaternion(*C5_EULER)
        wb = WorldBuilder(cam)
        wb.build([SceneObject(sprite=_CACHE["c5"], pos=(0.0, 0.0, 0.0))], C5_MATERIALS)
        wb.update_lod(visible=None)
        for _ in range(4):
            out = cam.render(want=("visited",))
            wb.update_lod(visible=cam.visited_slots(out["visited"]).copy())
        ms, crcs = time_poses(cam, [dict(pos=cam.pos, rot=cam.rot)], 8, flush, want=("rgba8", "visited"))
        res["c5"] = {"ms": [round(v, 3) for v in ms], "mean": round(ms[0], 3), "crc": crcs}
        cam.close()
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    for spec in (sys.argv[1:] or [""]):
        run_variant(spec)
