"""Generates tests/golden/*.npz by running the UNMODIFIED reference (oracle/refharness.py).

Run in the build container only (needs /root/reference):  python -m oracle.gen_golden
The fixtures travel to the GPU box; the reference does not.  TEST INFRASTRUCTURE ONLY.

Files written:
  default_world.npz        the default mod's voxel world, flattened cell by cell with Frame.get_voxel
                           (origin, cells[x,y,z] u8, materials[n+1,7] f64, material names)
  case_<name>.npz          one per case: settings (json), camera, the world as the reference built it for that
                           camera (cells, per-case material table, the chunk resolution table the
                           reference's chunk_update chose), the pixel list, and the
                           reference outputs (mean RGBA, nsamp, sample-0 ray record, visited chunk slots)
"""
from __future__ import annotations

import json
import os
import sys
import time

import numpy as np

from oracle import refharness as rh

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

RNG_FREE = dict(dof=0, lod_random=0, lod_edge=0, samples=1)


def strided(w, h, step, off=0):
    return [(x, y) for x in range(off, w, step) for y in range(off, h, step)]


# name -> dict(settings overrides, pos, rot (euler deg: roll, yaw, pitch), mirror (all roughness 0), pixels)
CASES = {
    # C1: the shipped config, full frame, stochastic (static MT seeding)
    "c1_shipped": dict(settings={}, pos=None, rot=None, mirror=False, pixels="all"),
    # shipped config from other view points (LOD 2/3 chunks, sky, long traversals)
    "c1_yaw135": dict(settings={}, pos=(40, 20, -90), rot=(0, 135, 0), mirror=False, pixels="all"),
    "c1_far": dict(settings={}, pos=(-150, 50, -150), rot=(0, 45, 20), mirror=False, pixels="all"),
    # C2: 1080p, RNG-free mirror materials, six poses, strided pixel subset
    "c2_pose0": dict(settings=dict(width=1920, height=1080, **RNG_FREE), pos=None, rot=None, mirror=True, pixels=20),
    "c2_yaw90": dict(settings=dict(width=1920, height=1080, **RNG_FREE), pos=(-12, 4, 4), rot=(0, 90, 0), mirror=True, pixels=20),
    "c2_yaw180": dict(settings=dict(width=1920, height=1080, **RNG_FREE), pos=(-12, 4, 4), rot=(0, 180, 0), mirror=True, pixels=20),
    "c2_yaw270_roll": dict(settings=dict(width=1920, height=1080, **RNG_FREE), pos=(-12.5, 4.25, 3.75), rot=(20, 270, 0), mirror=True, pixels=20),
    "c2_lookdown": dict(settings=dict(width=1920, height=1080, **RNG_FREE), pos=(0, 70, 0), rot=(0, 30, 60), mirror=True, pixels=20),
    "c2_outside": dict(settings=dict(width=1920, height=1080, **RNG_FREE), pos=(-120, 40, -130), rot=(0, 40, 15), mirror=True, pixels=20),
    "c2_onboundary": dict(settings=dict(width=1920, height=1080, **RNG_FREE), pos=(-16, 0, 16), rot=(0, 315, 350), mirror=True, pixels=20),
    # C3: 1080p 4 spp, rough PBR + glass + mist, 4 bounces (stochastic, static MT seeding)
    "c3_cubes": dict(settings=dict(width=1920, height=1080, samples=4, max_bounces=4, lod_edge=0, lod_samples=0,
                                   lod_random=0.25, dof=0.5), pos=(-30, 20, -10), rot=(0, 20, 0), mirror=False, pixels=24),
    "c3_glass": dict(settings=dict(width=1920, height=1080, samples=4, max_bounces=4, lod_edge=0, lod_samples=0,
                                   lod_random=0.25, dof=0.5), pos=(-4, 18, -6), rot=(0, 0, 0), mirror=False, pixels=24),
    "c3_mist": dict(settings=dict(width=1920, height=1080, samples=4, max_bounces=4, lod_edge=0, lod_samples=0,
                                  lod_random=0.25, dof=0.5), pos=(-36, 19, -60), rot=(0, 0, 0), mirror=False, pixels=24),
    # per-pixel sample count / edge LOD / sample LOD (init.py:133-139)
    "lod_edge_samples": dict(settings=dict(width=160, height=90, samples=5, lod_edge=0.6, lod_samples=0.5),
                             pos=(-40, 10, 30), rot=(0, 200, 10), mirror=False, pixels="all"),
    # setting corner cases
    "no_background": dict(settings=dict(width=96, height=64), pos=None, rot=None, mirror=False, pixels="all", background=False),
    "falloff0_distmin": dict(settings=dict(width=96, height=64, falloff=0, dist_min=3, max_light=0.5, shutter=0),
                             pos=(10, 6, -20), rot=(0, 250, 5), mirror=False, pixels="all"),
    "bounces8_long": dict(settings=dict(width=96, height=64, max_bounces=8, lod_bounces=0, dist_max=320, **RNG_FREE),
                          pos=(-12, 4, 4), rot=(0, 0, 0), mirror=True, pixels="all"),
    # non power-of-two chunk size + deeper LOD chain (float // and % semantics, init.py:68, 114)
    "chunk12_lod3": dict(settings=dict(width=96, height=64, chunk_size=12, chunk_lod=3, **RNG_FREE),
                         pos=(-100, 30, -90), rot=(0, 50, 10), mirror=True, pixels="all"),
    # culling on: chunks the previous frames did not traverse are withheld from the camera (init.py:447-452)
    "culling_warm": dict(settings=dict(width=96, height=64, culling=True, **RNG_FREE), pos=None, rot=None, mirror=True,
                         pixels="all", warm_frames=4),
}


def main(only=None):
    ref = rh.load_reference()
    data = ref.data
    os.makedirs(OUT, exist_ok=True)
    shipped_rough = None
    world_saved = False
    for name, case in CASES.items():
        if only and name not in only:
            continue
        t0 = time.time()
        rh.restore_settings()
        rh.override_settings(culling=False, dist_move=0)
        rh.override_settings(**case["settings"])
        s = data.settings
        player = data.player
        if case["pos"] is None:
            cam = ref.Camera()
            cam.pos, cam.rot = player.cam_pos, player.cam_rot
            rot_euler = (player.rot.x, player.rot.y, player.rot.z)
        else:
            cam = rh.make_camera(case["pos"], case["rot"])
            rot_euler = case["rot"]
        warm = case.get("warm_frames", 0)
        if warm:
            # culling needs history: frame 0 sees nothing (init.py:18), so feed traversed lists forward
            rh.override_settings(culling=False)
            st = rh.build_chunks(cam)
            rh.override_settings(culling=True)
            mats = rh.all_materials(cam)
            trav = None
            for _ in range(warm):
                r = rh.render(cam, pixels=strided(s.width, s.height, 1))
                trav = r["traversed"]
                rh.build_chunks(cam, st, traversed=trav)
        else:
            st = rh.build_chunks(cam)      # fresh state per case: object redraw flags are global (data.py:444)
            mats = rh.all_materials(cam)
        saved = [(m, m.roughness) for m in mats]
        if case["mirror"]:
            for m in mats:
                m.roughness = 0
        bg_saved = data.background
        if case.get("background", True) is False:
            data.background = None
        try:
            origin, cells, chunk_res, table, names = rh.extract_world(st, cam)
            if case["pixels"] == "all":
                pixels = strided(s.width, s.height, 1)
            else:
                pixels = strided(s.width, s.height, case["pixels"], off=case["pixels"] // 2)
            out = rh.render(cam, pixels=pixels)
        finally:
            for m, r in saved:
                m.roughness = r
            data.background = bg_saved
        if not world_saved and s.chunk_size == 16:
            np.savez_compressed(os.path.join(OUT, "default_world.npz"), origin=origin, cells=cells,
                                materials=table, names=np.array(names))
            world_saved = True
        visited = np.zeros(chunk_res.shape, np.uint8)
        for post in out["traversed"]:
            c = (np.array(post) - origin) // s.chunk_size
            if np.all(c >= 0) and np.all(c < np.array(chunk_res.shape)):
                visited[tuple(int(v) for v in c)] = 1
        sd = rh.settings_dict()
        sd["background"] = 0 if data.background is None or case.get("background", True) is False else 1
        np.savez_compressed(
            os.path.join(OUT, "case_%s.npz" % name),
            settings=json.dumps(sd), cam_pos=np.array([cam.pos.x, cam.pos.y, cam.pos.z], np.float64),
            cam_rot=np.array([cam.rot.x, cam.rot.y, cam.rot.z, cam.rot.w], np.float64), cam_lens=np.float64(cam.lens),
            rot_euler=np.array(rot_euler, np.float64), origin=origin, world_shape=np.array(cells.shape, np.int32),
            cells=cells, chunk_res=chunk_res, materials=table, pixels=out["pixels"], mean=out["mean"],
            nsamp=out["nsamp"].astype(np.int32), ray0=out["ray0"], visited=visited, rays=np.int64(out["rays"]))
        print("%-18s %6d px %7d rays  chunks %3d  res-hist %s  hit-frac %.3f  %.1fs" % (
            name, len(pixels), out["rays"], int((chunk_res > 0).sum()),
            np.bincount(chunk_res.ravel(), minlength=5)[1:].tolist(), out["ray0"][:, 10].mean(), time.time() - t0))
    rh.restore_settings()


if __name__ == "__main__":
    main(sys.argv[1:] or None)
