"""GPU tests of GpuCamera.trace / vt_trace_rays (Camera.trace, reference init.py:37-121) and of the culling cold start
on the chunk_set path (init.py:447-452, 455-471), against fixtures written by oracle/gen_r2_golden.py from the reference."""
import json
import os

import numpy as np
import pytest
import torch

from tests.golden_util import GOLDEN_DIR
from tests.test_gpu_camera_api import Frame, Mat, material, material_background

pytestmark = pytest.mark.gpu
INTS = [0, 1, 2, 6, 7, 8, 9, 10]


def _cam(sd, **kw):
    from python_raytracer_b200 import GpuCamera, Settings
    sd = dict(sd); sd["threads"] = 1
    return GpuCamera(Settings.from_object(sd), material_function=material, background_function=material_background,
                     background=material_background, **kw)


def test_trace_rays_match_reference_and_oracle():
    from oracle import cport
    g = np.load(os.path.join(GOLDEN_DIR, "trace_rays.npz"))
    for tag in ("free", "stoch"):
        sd = json.loads(str(g["settings_" + tag]))
        cam = _cam(sd)
        cam.set_world_dense(g["origin"], g["cells"], g["chunk_res"], g["materials_" + tag])
        cam.pos, cam.rot, cam.lens = g["cam_pos"], g["cam_rot"], float(g["cam_lens"])
        out = cam.trace_rays(g["dirs"], want=("visited", "counters"))
        rec = out["records"].cpu().numpy()
        vis = cam.visited_slots(out["visited"]).copy()
        cnt = [int(v) for v in out["counters"].cpu().numpy()[:6]]
        world = cport.World(g["origin"], g["cells"], g["chunk_res"], g["materials_" + tag])
        o = cport.trace_rays(sd, g["cam_pos"], g["cam_rot"], float(g["cam_lens"]), world, g["dirs"], rng=cport.RNG_PHILOX, seed=0)
        assert np.array_equal(rec[:, INTS], o["records"][:, INTS]), tag
        np.testing.assert_allclose(rec, o["records"], rtol=1e-11, atol=1e-11)
        assert np.array_equal(vis, o["visited"]) and cnt == [int(v) for v in o["counters"][:6]]
        if tag == "free":                                  # no random draw anywhere: the reference's own records
            ref = g["records_free"]
            assert np.array_equal(rec[:, INTS], ref[:, INTS])
            np.testing.assert_allclose(rec, ref, rtol=1e-11, atol=1e-11)
            assert np.array_equal(vis, g["visited_free"])
            # host-libm half-angle tables: the start direction is the reference's except where pow(q, 2) != q * q
            d0 = g["dirs"][3]                               # detail 0: life 0, the ray never moves, vel is the primary direction
            assert d0[2] == 0.0 and np.array_equal(rec[3, 12:15], ref[3, 12:15])
            one = cam.trace(*g["dirs"][0])
            assert one.color == tuple(int(v) for v in ref[0, 0:3]) and one.hits == int(ref[0, 6])
            assert one.first_hit == (tuple(int(v) for v in ref[0, 7:10]) if ref[0, 10] else None)
            assert abs(one.life - ref[0, 15]) == 0.0 and len(one.traversed) > 0
        cam.close()


def _window_frames(world, posts, mats, chunk_lod):
    """Window.chunks of the default scene: per chunk position one stand-in Frame per LOD holding the cells whose
    coordinates are multiples of the resolution (init.py:441-444, data.py:165-169)."""
    origin, cells = world["origin"].astype(int), world["cells"]
    out = {}
    for post in posts:
        lo = np.array(post) - origin
        blk = cells[lo[0]:lo[0] + 16, lo[1]:lo[1] + 16, lo[2]:lo[2] + 16]
        frames = []
        for lod in range(chunk_lod + 1):
            r = lod + 1
            f = Frame(r)
            for x, y, z in np.argwhere(blk > 0):
                wx, wy, wz = post[0] + int(x), post[1] + int(y), post[2] + int(z)
                if wx % r == 0 and wy % r == 0 and wz % r == 0:
                    f.data3[(wx // r, wy // r, wz // r)] = mats[int(blk[x, y, z])]
            frames.append(f)
        out[tuple(post)] = frames
    return out


def test_culling_cold_start_follows_the_reference():
    """culling=true is the shipped default and the frame loop starts with cam.chunks == {} (init.py:18).
    The reference's own loop from a cold start (tests/golden/culling_cold.npz): frame 0 is all sky but reports every chunk
    position its rays snapped into, chunk_update hands over the chunks among them, and the picture converges.  The same
    loop through GpuCamera.tile / chunk_set must report the same traversed sets, end up with the same resident chunks
    and draw the same frames."""
    g = np.load(os.path.join(GOLDEN_DIR, "culling_cold.npz"))
    world = np.load(os.path.join(GOLDEN_DIR, "default_world.npz"))
    sd = json.loads(str(g["settings"]))
    assert sd["culling"] is True
    table = world["materials"].copy()
    table[:, 3] = 0                                                     # the fixture is RNG-free: every roughness 0
    mats = [None] + [Mat(row) for row in table[1:]]
    posts = [tuple(int(v) for v in p) for p in g["window_posts"]]
    window = _window_frames(world, posts, mats, int(sd["chunk_lod"]))
    cam = _cam(sd)
    cam.pos, cam.rot, cam.lens = tuple(g["cam_pos"]), tuple(g["cam_rot"]), float(g["cam_lens"])
    W, H = int(sd["width"]), int(sd["height"])
    px = g["pixels"]
    cs, radius = int(sd["chunk_size"]), round(int(sd["chunk_size"]) / 2)
    for f in range(int(g["frames"])):
        image, traversed, _ = cam.tile(0, 0)                            # Window.draw
        frame = np.frombuffer(image, np.uint8).reshape(H, W, 4)
        want_trav = {tuple(int(v) for v in t) for t in g["traversed_%d" % f]}
        assert set(traversed) == want_trav, "frame %d: traversed differs (%d vs %d)" % (f, len(traversed), len(want_trav))
        assert np.array_equal(frame[px[:, 1], px[:, 0]], np.trunc(g["mean_%d" % f]).astype(np.uint8)), "frame %d" % f
        for post, frames in window.items():                            # Window.chunk_update, init.py:446-452
            if post in want_trav:
                d = np.sqrt(sum((post[a] + radius - cam.pos[a]) ** 2 for a in range(3)))
                lod = min(int(d / (sd["dist_max"] / (1 + sd["chunk_lod"]))), sd["chunk_lod"])
                cam.chunk_set(post, frames[lod])
            else:
                cam.chunk_set(post, None)
        resident = sorted((p[0], p[1], p[2], int(fr.resolution)) for p, fr in cam.chunks.items())
        assert resident == [tuple(int(v) for v in r) for r in g["resident_%d" % f]], "frame %d: resident set" % f
    assert len(cam.chunks) > 0
    cam.close()


def test_empty_stripe_publishes_its_flag_and_chunk_size_can_change():
    """a rank whose stripe set is empty (64 x 48 canvas, 8 ranks of 8-row stripes: ranks 6, 7) must still raise
    its arrival flag; and an engine that edits settings.chunk_size in place gets a rebuilt world, not an error."""
    import ctypes as C
    from tests.golden_util import load_case
    from python_raytracer_b200 import GpuCamera, Settings
    c = load_case("c1_shipped")
    sd = dict(c["settings"]); sd["threads"] = 1
    cam = GpuCamera(Settings.from_object(sd), background_function=material_background, background=material_background)
    cam.set_world_dense(c["origin"], c["cells"], c["chunk_res"], c["materials"])
    cam.pos, cam.rot, cam.lens = c["cam_pos"], c["cam_rot"], float(c["cam_lens"])
    frame = torch.zeros((48, 64, 4), dtype=torch.uint8, device=cam.device)
    flag = torch.zeros(4, dtype=torch.int32, device=cam.device)
    cam._prepare()
    cam._check(cam._lib.vt_set_frame_target(cam._ctx, frame.data_ptr(), flag.data_ptr(), 7))
    cam.render(stripes=(8, 8, 7), want=(), prepare=False)               # rows of rank 7: none
    cam._check(cam._lib.vt_stream_wait_flags(cam._ctx, flag.data_ptr(), 1, 7, 5.0))
    st = C.c_int(-1)
    cam._check(cam._lib.vt_wait_status(cam._ctx, C.byref(st)))
    assert st.value == 0 and int(flag[0]) == 7 and not frame.any()
    cam._check(cam._lib.vt_set_frame_target(cam._ctx, None, None, 0))
    cam.close()
    # chunk_size edited in place on the chunk_set path
    s = Settings.from_object(sd)
    cam = GpuCamera(s, material_function=material, background_function=material_background, background=material_background)
    cam.pos, cam.rot, cam.lens = c["cam_pos"], c["cam_rot"], float(c["cam_lens"])
    m = Mat(c["materials"][1])
    f = Frame(1); f.data3[(-12, 4, 8)] = m
    cam.chunk_set((-16, 0, 0), f)
    a = cam.render(want=("rgba8",))["rgba8"].clone()
    s.chunk_size = 8
    cam.chunks = {}
    f2 = Frame(1); f2.data3[(-12, 4, 8)] = m
    cam.chunk_set((-16, 0, 8), f2)
    b = cam.render(want=("rgba8",))["rgba8"]
    assert a.any() and b.any()
    cam.close()


def test_device_pow_equals_host_pow():
    """vt_pow_cr.h compiles to the same arithmetic on the device (explicit FMAs, -fmad=false) as on the host: bit-identical
    results, so the CPU test's proof of correct rounding carries over."""
    import ctypes as C
    from tests.golden_util import load_case
    from python_raytracer_b200 import GpuCamera, Settings
    rng = np.random.default_rng(3)
    x = np.concatenate([1.0 + rng.random(20000) * rng.choice([1.0, 8.0, 60.0], 20000), [1.0, 2.0, 0.5, 1e3, 1e-3]])
    y = np.concatenate([1.0 + rng.random(20000) * 2.5, [1.25, 2.0, 1.25, 1.25, 2.5]])
    sd = dict(load_case("c1_shipped")["settings"]); sd["threads"] = 1
    cam = GpuCamera(Settings.from_object(sd), background=None)
    dx, dy = torch.from_numpy(x).to(cam.device), torch.from_numpy(y).to(cam.device)
    out = torch.empty_like(dx)
    cam._check(cam._lib.vt_set_stream(cam._ctx, torch.cuda.current_stream(cam.device).cuda_stream))
    cam._check(cam._lib.vt_selftest_pow(cam._ctx, dx.data_ptr(), dy.data_ptr(), len(x), out.data_ptr()))
    torch.cuda.synchronize()
    host = np.array([cam._lib.vt_pow_host(float(a), float(b)) for a, b in zip(x, y)])
    assert np.array_equal(out.cpu().numpy(), host)
    cam.close()
