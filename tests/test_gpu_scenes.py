"""GPU tests of the rows either side of the trace loop and of the large named configurations (run on the B200 box):
scene building (vt_build_world vs what the reference's chunk_update produced), C4 (512^3 terrain, 3840x2160, 16 spp)
and C5 (10 M-voxel Goxel text model, 1080p 8 spp, culling on) at FULL size through size-independent properties plus the
C oracle on a pixel sub-sample."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _bg(ray, settings):
    raise AssertionError("never called")


def _camera(sd):
    from python_raytracer_b200 import GpuCamera, Settings
    sd = dict(sd); sd["threads"] = 1
    return GpuCamera(Settings.from_object(sd), background_function=_bg, background=_bg)


# ------------------------------------------------------------------------------------------------ scene building
def test_builder_reproduces_reference_chunk_update():
    """Objects of the default mod as the reference holds them -> device world == the chunk frames the reference's
    own Window.chunk_update built (tests/golden/default_world.npz), cell for cell."""
    from python_raytracer_b200.builder import SceneObject, WorldBuilder
    objs = np.load(os.path.join(GOLD, "default_objects.npz"))
    world = np.load(os.path.join(GOLD, "default_world.npz"))
    scene = [SceneObject(sprite=objs["sprite_%d" % k], pos=tuple(objs["pos_%d" % k]), rot=tuple(objs["rot_%d" % k]))
             for k in range(int(objs["count"]))]
    for k, o in enumerate(scene):                       # Object.mins / maxs (data.py:480-484)
        mins, maxs = o.box()
        assert mins == objs["mins_%d" % k].tolist() and maxs == objs["maxs_%d" % k].tolist()
    cam = _camera(dict(width=64, height=48, chunk_size=16, chunk_lod=2, dist_max=192))
    cam.pos = (-12.0, 4.0, 4.0)
    wb = WorldBuilder(cam)
    cells = wb.build(scene, world["materials"]).cpu().numpy()
    off = world["origin"].astype(np.int64) - wb.origin
    ref = np.zeros_like(cells)
    sx, sy, sz = world["cells"].shape
    ref[off[0]:off[0] + sx, off[1]:off[1] + sy, off[2]:off[2] + sz] = world["cells"]
    assert np.array_equal(cells, ref)
    # and the bricked world renders like the dense-upload path: shipped 64x48 golden frame
    from tests.golden_util import load_case
    from tests.test_gpu_parity import gpu_case
    c = load_case("c1_shipped")
    g = gpu_case(c, want=("rgba8",))
    cam2 = _camera(c["settings"])
    cam2.pos, cam2.rot, cam2.lens = c["cam_pos"], c["cam_rot"], float(c["cam_lens"])
    wb2 = WorldBuilder(cam2)
    wb2.build(scene, c["materials"])
    out = cam2.render(pixels=c["pixels"], want=("rgba8",))
    assert np.array_equal(out["rgba8"].cpu().numpy(), g["rgba8"])
    cam.close(); cam2.close()


def test_builder_rotations_match_reference_sprite_get_voxel():
    """Sprite.pos_rotated (data.py:338-371) for cubic and non-cubic sprites, ten rotations incl. non-multiples of 90,
    against grids produced by the reference's own Sprite.get_voxel."""
    from python_raytracer_b200.builder import SceneObject, WorldBuilder
    g = np.load(os.path.join(GOLD, "sprite_rotations.npz"))
    cam = _camera(dict(width=8, height=8, chunk_size=16))
    mats = np.zeros((4, 7)); mats[1:, 4] = 1.0
    for tag in ("cube", "slab", "bar", "xz"):
        spr = g["sprite_" + tag]
        for r, rot in enumerate(g["rots"]):
            o = SceneObject(sprite=spr, pos=(0.0, 0.0, 0.0), rot=tuple(rot))
            wb = WorldBuilder(cam)
            cells = wb.build([o], mats).cpu().numpy()
            mins, maxs = o.box()
            lo = np.array(mins) - wb.origin
            sx, sy, sz = spr.shape
            got = cells[lo[0]:lo[0] + sx, lo[1]:lo[1] + sy, lo[2]:lo[2] + sz]
            assert np.array_equal(got, g["drawn_" + tag][r]), (tag, rot)
            assert int((cells > 0).sum()) == int((got > 0).sum())
    cam.close()


def test_builder_object_order_and_offsets():
    """Later objects override earlier ones (init.py:437-439); fractional positions use ceil / floor (data.py:480-484)."""
    from python_raytracer_b200.builder import SceneObject, WorldBuilder
    a = np.full((4, 4, 4), 1, np.uint8)
    b = np.full((4, 4, 4), 2, np.uint8)
    b[0, 0, 0] = 0
    cam = _camera(dict(width=8, height=8, chunk_size=16))
    mats = np.zeros((3, 7)); mats[1:, 4] = 1.0
    wb = WorldBuilder(cam)
    objs = [SceneObject(sprite=a, pos=(0.0, 0.0, 0.0)), SceneObject(sprite=b, pos=(1.5, 0.0, -0.5))]
    cells = wb.build(objs, mats).cpu().numpy()
    exp = np.zeros_like(cells)
    for o, grid in zip(objs, (a, b)):
        mins, maxs = o.box()
        for x in range(mins[0], maxs[0]):
            for y in range(mins[1], maxs[1]):
                for z in range(mins[2], maxs[2]):
                    m = grid[x - mins[0], y - mins[1], z - mins[2]]
                    if m:
                        exp[x - wb.origin[0], y - wb.origin[1], z - wb.origin[2]] = m
    assert np.array_equal(cells, exp)
    assert objs[1].box() == ([0, -2, -2], [3, 2, 1])
    cam.close()


# ------------------------------------------------------------------------------------------- C4 / C5 at full size
from python_raytracer_b200.scenes import C4_SETTINGS, C5_SETTINGS  # noqa: E402


def _sample_pixels(w, h, n, seed):
    rng = np.random.default_rng(seed)
    return np.stack([rng.integers(0, w, n), rng.integers(0, h, n)], axis=1).astype(np.int32)


def _check_against_oracle(cam, sd, origin, cells_np, chunk_res, mats, pos, rot, px, frame_rgba8):
    """List-mode render of `px` vs the C oracle with the same Philox streams, and vs the full frame's pixels."""
    from oracle import cport
    out = cam.render(pixels=px, want=("rgba8", "mean", "ray0"))
    torch.cuda.synchronize()
    g_ray, g_rgba = out["ray0"].cpu().numpy(), out["rgba8"].cpu().numpy()
    world = cport.World(origin, cells_np, chunk_res, mats)
    o = cport.render(sd, pos, rot, float(cam.lens), world, pixels=px, rng=cport.RNG_PHILOX, seed=0)
    ints = [0, 1, 2, 6, 7, 8, 9, 10]
    bad = (g_ray[:, ints] != o["ray0"][:, ints]).any(axis=1)
    # a 1-ulp difference of device pow / sincos in a jittered direction can move one sample across a cell face
    assert bad.sum() <= max(1, int(5e-4 * len(px))), "%d / %d sample-0 rays differ from the oracle" % (bad.sum(), len(px))
    pix_bad = (g_rgba != o["rgba8"]).any(axis=1)
    assert pix_bad.sum() <= max(2, int(4e-3 * len(px))), "%d / %d pixels differ from the oracle" % (pix_bad.sum(), len(px))
    # rectangle mode and list mode are the same rays (streams are keyed by the global pixel index)
    assert np.array_equal(frame_rgba8[px[:, 1], px[:, 0]], g_rgba)
    return int(bad.sum()), int(pix_bad.sum())


def test_c4_terrain_4k_16spp_full_size():
    """BASELINE.json configs[3]: 512^3 bricked terrain at 3840x2160, 16 spp -- the full frame (132.7 M rays) on the
    GPU; oracle on 1500 random pixels; idempotence; 2- and 8-way stripe split reassembles to the same frame."""
    from python_raytracer_b200.mathutil import euler_to_quaternion
    from python_raytracer_b200.scenes import terrain_world
    from python_raytracer_b200.world import chunk_presence, lod_table
    sd = C4_SETTINGS
    origin, cells, mats = terrain_world(512)
    present = chunk_presence(cells, sd["chunk_size"])
    cam = _camera(sd)
    pos = (0.0, 300.0, 0.0)
    res0 = lod_table(origin, present, pos, sd["chunk_size"], sd["chunk_lod"], sd["dist_max"])
    cam.set_world_dense(origin, torch.from_numpy(cells), res0, mats)
    st = cam.world_stats()
    assert 0 < st["bricks"] < (512 // 8) ** 3
    px = _sample_pixels(sd["width"], sd["height"], 1500, 4)
    for yaw in (0.0, 135.0):
        rot = euler_to_quaternion(0.0, yaw, 30.0)
        cam.pos, cam.rot = pos, rot
        out = cam.render(want=("rgba8", "visited", "counters"))
        torch.cuda.synchronize()
        frame = out["rgba8"].cpu().numpy()
        cnt = out["counters"].cpu().numpy()
        assert frame.shape == (2160, 3840, 4)
        assert int(cnt[0]) == 3840 * 2160 * 16                                  # every sample traced
        assert int(cnt[3]) > int(cnt[0]) // 4                                   # the camera looks at terrain
        visited = cam.visited_slots(out["visited"]).copy()
        again = cam.render(want=("rgba8", "visited"))
        assert torch.equal(again["rgba8"], out["rgba8"])                        # static streams: frames repeat
        assert np.array_equal(cam.visited_slots(again["visited"]), visited)
        for n in (2, 8):                                                        # stripe split == whole frame
            rank = n - 1
            band = cam.render(stripes=(8, n, rank), want=("rgba8",))["rgba8"].cpu().numpy()
            rows = np.arange(2160)[(np.arange(2160) // 8) % n == rank]
            assert np.array_equal(band, frame[rows])
        _check_against_oracle(cam, sd, origin, cells, res0, mats, pos, rot, px, frame)
    cam.close()


def test_c5_goxel_model_culling_full_size(tmp_path):
    """BASELINE.json configs[4]: a ~10 M-voxel Goxel text model written to disk, parsed by the C++ loader, built on
    the device, rendered at 1080p 8 spp with occlusion culling fed by the previous frames' traversed chunks
    (init.py:446-452)."""
    from python_raytracer_b200.builder import SceneObject, WorldBuilder
    from python_raytracer_b200.goxel import load_goxel, sprite_grid, write_goxel
    from python_raytracer_b200.mathutil import euler_to_quaternion
    from python_raytracer_b200.scenes import C5_KEYS, C5_MATERIALS, goxel_model
    sd = C5_SETTINGS
    size = (512, 256, 512)
    xyz_file, cols = goxel_model(size)
    path = str(tmp_path / "model.txt")
    write_goxel(path, xyz_file, cols)
    xyz, m = load_goxel(path, size[0], C5_KEYS)
    assert len(xyz) == len(xyz_file) > 9_000_000
    assert np.array_equal(xyz[:, 0], size[0] - xyz_file[:, 0]) and np.array_equal(xyz[:, 1], xyz_file[:, 2])
    grid = sprite_grid(xyz, m, size)
    cam = _camera(sd)
    pos = (-300.0, 40.0, -300.0)
    rot = euler_to_quaternion(0.0, 45.0, 10.0)
    cam.pos, cam.rot = pos, rot
    wb = WorldBuilder(cam)
    cells = wb.build([SceneObject(sprite=grid, pos=(0.0, 0.0, 0.0))], C5_MATERIALS)
    # file column x == 0 maps to size.x, outside the drawn box: everything else is drawn
    assert int((cells > 0).sum().item()) == int((grid > 0).sum())
    # culling: frame 0 sees every chunk; afterwards only chunks some ray traversed stay resident
    res_all = wb.update_lod(visible=None)
    visible = None
    frames = []
    for f in range(4):
        out = cam.render(want=("rgba8", "visited"))
        torch.cuda.synchronize()
        frames.append(out["rgba8"].cpu().numpy().copy())
        visible = cam.visited_slots(out["visited"]).copy()
        res = wb.update_lod(visible=visible)
        assert ((res > 0) <= (res_all > 0)).all() and ((res > 0) == ((res_all > 0) & (visible > 0))).all()
    # culled chunks were never traversed, so withholding them cannot change the picture: the frames converge at once
    assert np.array_equal(frames[1], frames[2]) and np.array_equal(frames[2], frames[3])
    assert np.array_equal(frames[0], frames[1])
    assert (res > 0).sum() < (res_all > 0).sum()
    px = _sample_pixels(sd["width"], sd["height"], 1500, 5)
    cells_np = cells.cpu().numpy()
    _check_against_oracle(cam, sd, wb.origin.astype(np.int32), cells_np, res, C5_MATERIALS, pos, rot, px, frames[3])
    cam.close()


# ------------------------------------------------------------------------------------------------ incremental updates
def _embed(cells, origin, lo, hi):
    out = np.zeros(tuple(int(v) for v in (hi - lo)), np.uint8)
    o = (np.asarray(origin, np.int64) - lo).astype(int)
    s = cells.shape
    out[o[0]:o[0] + s[0], o[1]:o[1] + s[1], o[2]:o[2] + s[2]] = cells
    return out


def test_builder_update_follows_reference_chunk_update():
    """tests/golden/default_moves.npz: the reference's chunk_update run six times while objects move into each other,
    turn, get removed and leave the initial box.  WorldBuilder.update must leave the same merged world after every
    step (dirty chunks only, override order = re-insertion order), and the bricked world must render like a world
    built from scratch from the same cells."""
    from python_raytracer_b200.builder import SceneObject, WorldBuilder
    objs = np.load(os.path.join(GOLD, "default_objects.npz"))
    world = np.load(os.path.join(GOLD, "default_world.npz"))
    mv = np.load(os.path.join(GOLD, "default_moves.npz"))
    n = int(mv["count"])
    assert n == int(objs["count"])
    cam = _camera(dict(width=96, height=64, chunk_size=16, chunk_lod=0, dist_max=192, samples=1, max_bounces=2, static=True))
    cam.pos = (-12.0, 4.0, 4.0)
    wb = WorldBuilder(cam)
    scene = [SceneObject(sprite=objs["sprite_%d" % k]) for k in range(n)]
    full_rebuilds = 0
    for s in range(int(mv["steps"])):
        st = mv["state_%d" % s]
        live, redraw = [], []
        for k, o in enumerate(scene):
            present, dirty = bool(st[k, 0]), bool(st[k, 8])
            o.pos, o.rot, o.visible = tuple(st[k, 1:4]), tuple(st[k, 4:7]), bool(st[k, 7])
            if present:
                live.append(o)
            if dirty or (not present and id(o) in wb.drawn):
                redraw.append(o)
        if s == 0:
            wb.build(live, world["materials"], margin=1)
        else:
            before = wb.cells.data_ptr()
            wb.update(live, redraw)
            full_rebuilds += wb.cells.data_ptr() != before
        got, want = wb.cells.cpu().numpy(), mv["cells_%d" % s]
        lo = np.minimum(wb.origin, mv["origin_%d" % s].astype(np.int64))
        hi = np.maximum(wb.origin + np.array(got.shape), mv["origin_%d" % s].astype(np.int64) + np.array(want.shape))
        a, b = _embed(got, wb.origin, lo, hi), _embed(want, mv["origin_%d" % s], lo, hi)
        assert np.array_equal(a, b), "step %d: %d cells differ" % (s, int((a != b).sum()))
        # the incrementally patched bricks render like bricks made from scratch out of the same cells
        cam2 = _camera(dict(width=96, height=64, chunk_size=16, chunk_lod=0, dist_max=192, samples=1, max_bounces=2, static=True))
        cam2.pos = cam.pos
        from python_raytracer_b200.world import lod_table
        res = lod_table(wb.origin, wb.present, list(cam.pos), 16, 0, 192)
        cam2.set_world_dense(wb.origin, got, res, world["materials"])
        wb.update_lod()
        frame = cam.render(want=("rgba8",))["rgba8"].cpu().numpy()
        frame2 = cam2.render(want=("rgba8",))["rgba8"].cpu().numpy()
        assert np.array_equal(frame, frame2), "step %d: patched bricks render differently" % s
        cam2.close()
    assert full_rebuilds == 1                     # only the step that leaves the box rebuilds everything
    cam.close()
