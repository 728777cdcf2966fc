#!/usr/bin/env python
"""Kernel A/B timing on the bench workload (dev tool, not a test):  python tools/kbench.py [variant ...]

A variant is a comma-separated list of key=value pairs handed to vt_set_tuning (refill=n, runs=0, powtab=0, group=n) plus
optionally LIB=<path to an alternative libvoxtrace.so> (tools/build_variants.py).  Prints ms/frame per camera pose.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import bench
from python_raytracer_b200 import GpuCamera, Settings, abi


def run_variant(spec, frames=8, settings=None, workload=None):
    env = dict(kv.split("=", 1) for kv in spec.split(",") if kv)
    lib = env.pop("LIB", None)
    if lib:
        abi._lib = abi.load(os.path.join(ROOT, lib))
    else:
        abi._lib = None
        abi.load()
    origin, cells, mats, poses = workload or bench.load_workload()
    if os.environ.get("KB_POSES"):
        poses = [poses[int(i)] for i in os.environ["KB_POSES"].split(",")]
    frames = int(os.environ.get("KB_FRAMES", frames))
    s = Settings.from_object(settings or bench.C3_SETTINGS)
    cam = GpuCamera(s, device=0, background_function=bench.BUILTIN_BG, background=bench.BUILTIN_BG)
    for k, v in env.items():
        cam._check(cam._lib.vt_set_tuning(cam._ctx, k.encode(), int(v)))
    cam.set_world_dense(origin, cells, poses[0]["chunk_res"], mats)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    out, ms = {}, []
    sums = []
    for p in poses:
        cam.pos, cam.rot = p["pos"], p["rot"]
        cam.set_chunk_res(p["chunk_res"])
        cam._prepare()
        for _ in range(2):
            cam.render(want=("rgba8",), prepare=False, out=out)
        ts = []
        for _ in range(frames):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); cam.render(want=("rgba8",), prepare=False, out=out); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms.append(float(np.median(ts)))
        if os.environ.get("KB_COUNTERS"):
            o2 = cam.render(want=("rgba8", "counters"), prepare=False)
            cn = o2["counters"].cpu().numpy().astype(float)
            print("   pose counters per ray: samples_in %.2f void %.2f hits %.3f probes %.2f rechunk %.2f rays %d" %
                  (cn[1] / cn[0], cn[2] / cn[0], cn[3] / cn[0], cn[4] / cn[0], cn[5] / cn[0], int(cn[0])), flush=True)
            tr = o2["trips"].cpu().numpy().astype(float)
            cn = np.concatenate([cn, tr[:6], [0, 0]])
            if tr[6:10].sum() > 0:                      # cycles per warp phase (VT_TRIPS build)
                tot = tr[6] + tr[7] + tr[8]
                print("   warp cycles: refill %.1f %%  shade %.1f %%  march %.1f %% (of which chunk re-evaluation block %.1f %%)" %
                      (100 * tr[6] / tot, 100 * tr[7] / tot, 100 * tr[8] / tot, 100 * tr[9] / tot), flush=True)
                b = tr[10:18]
                print("   march trips by marching lanes <=4 / <=8 / <=16 / <=32: %s %% of trips, %s %% of march cycles" % (
                    " / ".join("%.0f" % (100 * v / max(1, b[:4].sum())) for v in b[:4]),
                    " / ".join("%.0f" % (100 * v / max(1, b[4:].sum())) for v in b[4:])), flush=True)
            if cn[8] > 0:                                # VT_TRIPS build: phase trips per 32 rays and lanes per trip
                k = cn[0] / 32.0
                print("   trips per 32 rays: march %.1f (lanes %.1f, %.1f phases) shade %.2f (lanes %.1f) refill %.2f (lanes %.1f)" %
                      (cn[8] / k, (cn[1] + cn[2]) / cn[8], cn[13] / k, cn[9] / k, cn[10] / max(1, cn[9]), cn[11] / k, cn[12] / max(1, cn[11])), flush=True)
        sums.append(int(out["rgba8"].to(torch.int64).sum()))
    cam.close()
    print(json.dumps({"variant": spec, "ms_per_pose": [round(v, 3) for v in ms], "ms_mean": round(float(np.mean(ms)), 3),
                      "checksums": sums}), flush=True)


if __name__ == "__main__":
    for spec in (sys.argv[1:] or [""]):
        run_variant(spec)
