"""Whole-frame parity on the benchmark workloads (run on the B200 box): EVERY pixel of a frame through the C ABI against
the C oracle (oracle/vt_oracle.c, pinned bit for bit to the reference on the 19 goldens) under the same Philox streams.

The committed goldens cover every 20th / 24th pixel of a 1080p frame; here all 2 073 600 pixels are compared and the
number of rays / pixels that differ is COUNTED and written to gpurun_out/r2_parity_counts.json (copied to profiles/),
not hidden in a tolerance.  What can differ at all: the device evaluates pow / sin / cos with libdevice
and its own polynomial (<= 1-2 ulp from glibc), and `q ** 2` -- libm pow(q, 2.0) in CPython -- as q * q; a last-bit
difference in a direction or an energy can put one floor() on the other side of a cell face.
"""
import json
import os
import time

import numpy as np
import pytest
import torch

from tests.golden_util import case_names, load_case

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COUNTS = os.path.join(ROOT, "gpurun_out", "r2_parity_counts.json")
INTS = [0, 1, 2, 6, 7, 8, 9, 10]            # colour r g b | hits | first-hit voxel x y z | hit flag
F64 = [3, 4, 5, 11, 12, 13, 14]             # energy bounces step first_step vel x y z


def _bg(ray, settings):
    raise AssertionError("never called")


def _record(name, entry):
    os.makedirs(os.path.dirname(COUNTS), exist_ok=True)
    try:
        data = json.load(open(COUNTS))
    except Exception:
        data = {}
    data[name] = entry
    with open(COUNTS, "w") as f:
        json.dump(data, f, indent=1, sort_keys=True)
    print("parity counts %s: %s" % (name, json.dumps(entry)))


def _row_major_pixels(w, h, step=1):
    ys, xs = np.meshgrid(np.arange(0, h, step, dtype=np.int32), np.arange(0, w, step, dtype=np.int32), indexing="ij")
    return np.ascontiguousarray(np.stack([xs.ravel(), ys.ravel()], axis=1))


def _camera(sd, origin, cells, chunk_res, mats, pos, rot, lens=None):
    from python_raytracer_b200 import GpuCamera, Settings
    sd = dict(sd); sd["threads"] = 1
    bg = _bg if sd.get("background", 1) else None
    cam = GpuCamera(Settings.from_object(sd), background_function=_bg, background=bg)
    cam.set_world_dense(origin, cells, chunk_res, mats)
    cam.pos, cam.rot = pos, rot
    if lens is not None:
        cam.lens = float(lens)
    return cam


def _compare(g_ray, g_mean, g_rgba, o, n_px):
    """-> dict of counts: sample-0 rays whose integer outputs differ, pixels whose RGBA8 differs, max |delta mean|,
    sample-0 rays whose fp64 state is not bit-identical and the largest such difference."""
    ray_bad = (g_ray[:, INTS] != o["ray0"][:, INTS]).any(axis=1)
    pix_bad = (g_rgba != o["rgba8"]).any(axis=1)
    ok = ~ray_bad
    return dict(pixels=int(n_px), rays_differ=int(ray_bad.sum()), pixels_differ=int(pix_bad.sum()),
                max_abs_dmean=float(np.abs(g_mean.astype(np.float64) - o["mean"]).max()), **_f64_stats(g_ray[ok], o["ray0"][ok]))


def _f64_stats(g, o):
    """fp64 ray state (energy, bounces, step, first-hit step, velocity) of rays whose integer outputs agree: how many are
    not bit-identical, and the largest difference relative to max(1, |value|), per field."""
    a, b = g[:, F64], o[:, F64]
    d = np.abs(a - b) / np.maximum(1.0, np.abs(b))
    names = ("energy", "bounces", "step", "first_step", "vel_x", "vel_y", "vel_z")
    return dict(f64_not_bit_identical=int((d != 0).any(axis=1).sum()), f64_max_rel_diff=float(d.max()) if d.size else 0.0,
                f64_max_rel_diff_by_field={n: float(d[:, k].max()) if d.size else 0.0 for k, n in enumerate(names)},
                f64_not_bit_identical_by_field={n: int((d[:, k] != 0).sum()) for k, n in enumerate(names)})


C2_CASES = [n for n in case_names() if n.startswith("c2_")]


@pytest.mark.parametrize("name", C2_CASES)
def test_c2_every_pixel_identical(name):
    """The seven RNG-free 1080p poses of oracle/gen_golden.py, all 1920 x 1080 pixels: first-hit voxel, hit count, 8-bit
    colour, RGBA8 frame and visited set identical on EVERY pixel; also through the production instantiation (no ray
    record, early exit)."""
    from oracle import cport
    c = load_case(name)
    sd = c["settings"]
    w, h = int(sd["width"]), int(sd["height"])
    assert (w, h) == (1920, 1080)
    cam = _camera(sd, c["origin"], c["cells"], c["chunk_res"], c["materials"], c["cam_pos"], c["cam_rot"], c["cam_lens"])
    out = cam.render(want=("rgba8", "mean", "ray0", "visited"))
    torch.cuda.synchronize()
    g_ray = out["ray0"].cpu().numpy().reshape(-1, 16)
    g_mean = out["mean"].cpu().numpy().reshape(-1, 4)
    g_rgba = out["rgba8"].cpu().numpy().reshape(-1, 4)
    g_vis = cam.visited_slots(out["visited"]).copy()
    prod = cam.render(want=("rgba8", "visited"))                       # FAST / no REC / early exit: what bench.py times
    torch.cuda.synchronize()
    p_rgba = prod["rgba8"].cpu().numpy().reshape(-1, 4)
    p_vis = cam.visited_slots(prod["visited"]).copy()
    cam.close()
    world = cport.World(c["origin"], c["cells"], c["chunk_res"], c["materials"])
    px = _row_major_pixels(w, h)
    t0 = time.time()
    o = cport.render(sd, c["cam_pos"], c["cam_rot"], float(c["cam_lens"]), world, pixels=px, rng=cport.RNG_PHILOX, seed=0)
    e = _compare(g_ray, g_mean, g_rgba, o, w * h)
    e["oracle_seconds"] = round(time.time() - t0, 2)
    e["visited_differ"] = int((g_vis != o["visited"]).sum())
    e["production_rgba8_differ"] = int((p_rgba != o["rgba8"]).any(axis=1).sum())
    e["production_visited_differ"] = int((p_vis != o["visited"]).sum())
    # the committed golden (the reference itself, every 20th pixel) is a subset of this frame
    gp = c["pixels"]
    idx = gp[:, 1].astype(np.int64) * w + gp[:, 0]
    e["golden_subset_differ"] = int((g_ray[idx][:, INTS] != c["ray0"][:, INTS]).any(axis=1).sum())
    _record("c2/" + name, e)
    assert e["rays_differ"] == 0 and e["pixels_differ"] == 0 and e["visited_differ"] == 0, e
    assert e["production_rgba8_differ"] == 0 and e["production_visited_differ"] == 0, e
    assert e["golden_subset_differ"] == 0 and e["max_abs_dmean"] == 0.0, e
    assert e["f64_max_rel_diff"] <= 1e-11, e


def test_c3_bench_poses_every_pixel():
    """bench.py's four C3 poses (1080p, 4 spp, dof 0.5, rough / glass / mist): every pixel, all samples.  Lens jitter
    makes the direction's sin / cos device-computed (<= 1 ulp from libm), so some of the 8.3 M rays could take another
    path: they are counted and written down.  Observed: none."""
    import bench
    from oracle import cport
    origin, cells, mats, poses = bench.load_workload()
    sd = dict(bench.C3_SETTINGS)
    w, h = sd["width"], sd["height"]
    px = _row_major_pixels(w, h)
    lens = sd["fov"] * np.pi / 8
    total_rays = 0
    for k, p in enumerate(poses):
        cam = _camera(sd, origin, cells, p["chunk_res"], mats, p["pos"], p["rot"])
        out = cam.render(want=("rgba8", "mean", "ray0", "visited", "counters"))
        torch.cuda.synchronize()
        g_ray = out["ray0"].cpu().numpy().reshape(-1, 16)
        g_mean = out["mean"].cpu().numpy().reshape(-1, 4)
        g_rgba = out["rgba8"].cpu().numpy().reshape(-1, 4)
        g_vis = cam.visited_slots(out["visited"]).copy()
        g_cnt = [int(v) for v in out["counters"].cpu().numpy()[:6]]
        prod = cam.render(want=("rgba8",))["rgba8"].cpu().numpy().reshape(-1, 4)
        cam.close()
        world = cport.World(origin, cells, p["chunk_res"], mats)
        t0 = time.time()
        o = cport.render(sd, p["pos"], p["rot"], lens, world, pixels=px, rng=cport.RNG_PHILOX, seed=0)
        e = _compare(g_ray, g_mean, g_rgba, o, w * h)
        e["oracle_seconds"] = round(time.time() - t0, 2)
        e["rays"] = g_cnt[0]
        e["visited_differ"] = int((g_vis != o["visited"]).sum())
        e["counters_gpu"] = g_cnt
        e["counters_oracle"] = [int(v) for v in o["counters"][:6]]
        e["production_vs_rec_rgba8_differ"] = int((prod != g_rgba).any(axis=1).sum())
        _record("c3/bench_pose%d" % k, e)
        total_rays += g_cnt[0]
        assert g_cnt[0] == w * h * sd["samples"] == int(o["counters"][0])
        assert e["production_vs_rec_rgba8_differ"] == 0, e
        # observed on B200 (profiles/r2_parity_counts.json): 0 rays, 0 pixels, identical counters -- so that is the bound
        assert e["rays_differ"] == 0 and e["pixels_differ"] == 0 and e["visited_differ"] == 0, e
        assert e["counters_gpu"] == e["counters_oracle"], e
        assert e["f64_max_rel_diff"] <= 1e-11, e


def test_c5_whole_frame_after_culling_warmup():
    """BASELINE.json configs[4] at full size: the 9.9 M-voxel model, 1080p, 8 spp, culling on; after the warm-up frames
    every pixel (16.6 M rays) is compared with the oracle run on the SAME culled chunk table."""
    from oracle import cport
    from python_raytracer_b200.builder import SceneObject, WorldBuilder
    from python_raytracer_b200.mathutil import euler_to_quaternion
    from python_raytracer_b200.scenes import C5_EULER, C5_MATERIALS, C5_POS, C5_SETTINGS, C5_SIZE, goxel_model_grid
    from python_raytracer_b200 import GpuCamera, Settings
    sd = dict(C5_SETTINGS); sd["threads"] = 1
    w, h = sd["width"], sd["height"]
    grid = goxel_model_grid(C5_SIZE)
    cam = GpuCamera(Settings.from_object(sd), background_function=_bg, background=_bg)
    rot = euler_to_quaternion(*C5_EULER)
    cam.pos, cam.rot = C5_POS, rot
    wb = WorldBuilder(cam)
    cells = wb.build([SceneObject(sprite=grid, pos=(0.0, 0.0, 0.0))], C5_MATERIALS)
    res = wb.update_lod(visible=None)
    for _ in range(3):                                                    # init.py:446-452: traversed -> resident set
        out = cam.render(want=("visited",))
        torch.cuda.synchronize()
        res = wb.update_lod(visible=cam.visited_slots(out["visited"]).copy())
    out = cam.render(want=("rgba8", "mean", "ray0", "visited", "counters"))
    torch.cuda.synchronize()
    g_ray = out["ray0"].cpu().numpy().reshape(-1, 16)
    g_mean = out["mean"].cpu().numpy().reshape(-1, 4)
    g_rgba = out["rgba8"].cpu().numpy().reshape(-1, 4)
    g_vis = cam.visited_slots(out["visited"]).copy()
    rays = int(out["counters"][0])
    cells_np = cells.cpu().numpy()
    origin = wb.origin.astype(np.int32)
    cam.close()
    world = cport.World(origin, cells_np, res, C5_MATERIALS)
    t0 = time.time()
    o = cport.render(sd, C5_POS, rot, sd["fov"] * np.pi / 8, world, pixels=_row_major_pixels(w, h), rng=cport.RNG_PHILOX, seed=0)
    e = _compare(g_ray, g_mean, g_rgba, o, w * h)
    e.update(oracle_seconds=round(time.time() - t0, 2), rays=rays, resident_chunks=int((res > 0).sum()),
             visited_differ=int((g_vis != o["visited"]).sum()))
    _record("c5/whole_frame_culled", e)
    assert rays == w * h * sd["samples"]
    assert e["rays_differ"] == 0 and e["pixels_differ"] == 0 and e["visited_differ"] == 0, e
    assert e["f64_max_rel_diff"] <= 1e-11, e


def test_c4_whole_4k_frame():
    """BASELINE.json configs[3] at full size (512^3 terrain, 3840 x 2160, 16 spp = 132.7 M rays on the GPU).  The oracle
    does ~3 x 10^5 rays/s/8 cores on this scene, so the comparison is: the sample-0 ray of EVERY one of the 8.3 M pixels
    (oracle run with samples=1: sample 0 of a pixel does not depend on the sample count when lod_samples == 0), and all
    16 samples (mean, RGBA8) on every 4th pixel in x and y (518 400 pixels, 8.3 M rays)."""
    from oracle import cport
    from python_raytracer_b200.scenes import C4_SETTINGS, c4_workload
    sd = dict(C4_SETTINGS)
    w, h = sd["width"], sd["height"]
    origin, cells, mats, res, poses = c4_workload()
    p = poses[2]
    cam = _camera(sd, origin, torch.from_numpy(cells), res, mats, p["pos"], p["rot"])
    out = cam.render(want=("rgba8", "mean", "ray0", "visited", "counters"))
    torch.cuda.synchronize()
    rays = int(out["counters"][0])
    g_ray = out["ray0"].cpu().numpy().reshape(-1, 16)
    g_mean = out["mean"].cpu().numpy().reshape(h, w, 4)
    g_rgba = out["rgba8"].cpu().numpy().reshape(h, w, 4)
    cam.close()
    del out
    torch.cuda.empty_cache()
    world = cport.World(origin, cells, res, mats)
    lens = sd["fov"] * np.pi / 8
    sd1 = dict(sd); sd1["samples"] = 1
    t0 = time.time()
    o1 = cport.render(sd1, p["pos"], p["rot"], lens, world, pixels=_row_major_pixels(w, h), rng=cport.RNG_PHILOX, seed=0)
    ray_bad = (g_ray[:, INTS] != o1["ray0"][:, INTS]).any(axis=1)
    ok = ~ray_bad
    f64 = _f64_stats(g_ray[ok], o1["ray0"][ok])
    t1 = time.time()
    sub = _row_major_pixels(w, h, 4)
    o16 = cport.render(sd, p["pos"], p["rot"], lens, world, pixels=sub, rng=cport.RNG_PHILOX, seed=0)
    s_rgba, s_mean = g_rgba[sub[:, 1], sub[:, 0]], g_mean[sub[:, 1], sub[:, 0]]
    e = dict(pixels=w * h, rays=rays, sample0_rays_compared=w * h, rays_differ=int(ray_bad.sum()),
             subset_pixels=int(len(sub)), subset_rays=int(o16["counters"][0]),
             pixels_differ=int((s_rgba != o16["rgba8"]).any(axis=1).sum()),
             max_abs_dmean=float(np.abs(s_mean.astype(np.float64) - o16["mean"]).max()),
             oracle_seconds=[round(t1 - t0, 2), round(time.time() - t1, 2)], **f64)
    _record("c4/whole_4k_frame_yaw135", e)
    assert rays == w * h * sd["samples"]
    assert e["rays_differ"] == 0 and e["pixels_differ"] == 0, e
    assert e["f64_max_rel_diff"] <= 1e-10, e
