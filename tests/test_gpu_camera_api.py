"""GPU tests of the reference-facing protocol: chunk_set/chunk_get/tile, dirty-chunk re-upload, rejection of custom
material callbacks, rect/stripe regions.  Frames are stand-ins with the reference Frame's attributes
(resolution, data3, data6 -- reference data.py:96-106)."""
import numpy as np
import pytest
import torch

from tests.golden_util import load_case

pytestmark = pytest.mark.gpu


def material(ray, mat, settings):          # stands in for lib.material: only its identity matters on the GPU path
    raise AssertionError("never called")


def material_background(ray, settings):
    raise AssertionError("never called")


class Albedo:
    def __init__(self, r, g, b):
        self.r, self.g, self.b = r, g, b


class Mat:
    def __init__(self, row, function=material):
        self.albedo = Albedo(*row[:3])
        self.roughness, self.absorption, self.ior, self.energy = row[3:7]
        self.function = function


class Frame:
    def __init__(self, resolution=1):
        self.resolution, self.data3, self.data6 = resolution, {}, {}


def frames_from_case(c, mats):
    """Golden world -> {post: Frame} the way chunk_update hands them to Camera.chunk_set (init.py:441-450):
    one frame per chunk at the chunk's LOD, holding only cells whose coordinates are multiples of the resolution."""
    cs = c["settings"]["chunk_size"]
    origin, cells, res = c["origin"], c["cells"], c["chunk_res"]
    frames = {}
    for cx, cy, cz in np.argwhere(res > 0):
        r = int(res[cx, cy, cz])
        f = Frame(r)
        blk = cells[cx * cs:(cx + 1) * cs, cy * cs:(cy + 1) * cs, cz * cs:(cz + 1) * cs]
        post = (int(origin[0] + cx * cs), int(origin[1] + cy * cs), int(origin[2] + cz * cs))
        for x, y, z in np.argwhere(blk > 0):
            wx, wy, wz = post[0] + int(x), post[1] + int(y), post[2] + int(z)
            if wx % r == 0 and wy % r == 0 and wz % r == 0:
                f.data3[(wx // r, wy // r, wz // r)] = mats[int(blk[x, y, z])]
        frames[post] = f
    return frames


def make_cam(c, **over):
    from python_raytracer_b200 import GpuCamera, Settings
    sd = dict(c["settings"]); sd.update(over)
    sd.setdefault("threads", 1)
    cam = GpuCamera(Settings.from_object(sd), material_function=material, background_function=material_background,
                    background=material_background if sd.get("background", 1) else None)
    cam.pos, cam.rot, cam.lens = c["cam_pos"], c["cam_rot"], float(c["cam_lens"])
    return cam


@pytest.mark.parametrize("name", ["c1_shipped", "c2_outside", "chunk12_lod3"])
def test_chunk_set_path_equals_dense_path_and_reference(name):
    c = load_case(name)
    mats = [None] + [Mat(row) for row in c["materials"][1:]]
    cam = make_cam(c)
    for post, f in frames_from_case(c, mats).items():
        cam.chunk_set(post, f)
    out = cam.render(pixels=c["pixels"], want=("rgba8", "ray0"))
    ray0 = out["ray0"].cpu().numpy()
    from tests.test_gpu_parity import gpu_case
    dense = gpu_case(c, want=("rgba8", "ray0", "visited"))
    ints = [0, 1, 2, 6, 7, 8, 9, 10]
    assert np.array_equal(ray0[:, ints], dense["ray0"][:, ints])
    assert np.array_equal(out["rgba8"].cpu().numpy(), dense["rgba8"])
    # chunk_get mirrors init.py:28-33
    some = next(iter(cam.chunks))
    assert cam.chunk_get((some[0] + 0.5, some[1] + 1, some[2] + 2.25)) is cam.chunks[some]
    cam.close()


def test_dirty_chunk_reupload_and_removal():
    c = load_case("c2_pose0")
    mats = [None] + [Mat(row) for row in c["materials"][1:]]
    frames = frames_from_case(c, mats)
    cam = make_cam(c)
    for post, f in frames.items():
        cam.chunk_set(post, f)
    a = cam.render(pixels=c["pixels"], want=("rgba8", "ray0"))["ray0"].cpu().numpy()
    # remove the chunk most first hits land in: those rays must now travel further (or hit nothing)
    hit_chunk = tuple(int(v) for v in (np.floor(a[a[:, 10] > 0][:, 7:10] / 16) * 16)[0])
    launches = cam.launch_count
    cam.chunk_set(hit_chunk, None)
    b = cam.render(pixels=c["pixels"], want=("rgba8", "ray0"))["ray0"].cpu().numpy()
    in_removed = (np.floor(a[:, 7:10] / 16) * 16 == np.array(hit_chunk)).all(axis=1) & (a[:, 10] > 0)
    assert in_removed.any()
    still = (np.floor(b[:, 7:10] / 16) * 16 == np.array(hit_chunk)).all(axis=1) & (b[:, 10] > 0)
    assert not still.any()
    assert cam.launch_count - launches == 2          # one brick scatter + one render: only the dirty chunk moved
    # put it back: identical to the first render
    cam.chunk_set(hit_chunk, frames[hit_chunk])
    d = cam.render(pixels=c["pixels"], want=("rgba8", "ray0"))["ray0"].cpu().numpy()
    assert np.array_equal(a, d)
    cam.close()


def test_custom_material_callback_is_rejected():
    from python_raytracer_b200 import VoxtraceError
    c = load_case("c1_shipped")

    def my_shader(ray, mat, settings):
        return 1.0

    mats = [None] + [Mat(row) for row in c["materials"][1:]]
    mats[1].function = my_shader
    cam = make_cam(c)
    for post, f in frames_from_case(c, mats).items():
        cam.chunk_set(post, f)
    with pytest.raises(VoxtraceError, match="custom Python function"):
        cam.render(pixels=c["pixels"])
    cam2 = make_cam(c)
    cam2.background = my_shader
    with pytest.raises(VoxtraceError, match="background"):
        cam2.render(pixels=c["pixels"])
    cam.close(); cam2.close()


def test_tile_protocol_threads_and_stripes():
    """tile(thread, t) returns (W*H*4 RGBA bytes, traversed, thread); pixels of other threads stay (0,0,0,0)
    (init.py:127, 149-150; data.py:70-77).  The union over threads equals the single-launch frame, and so does the
    union of row stripes (the multi-GPU partition)."""
    c = load_case("c1_shipped")
    s = c["settings"]
    W, H = s["width"], s["height"]
    from tests.test_gpu_parity import _builtin_background
    from python_raytracer_b200 import GpuCamera, Settings

    def cam_for(threads):
        sd = dict(s); sd["threads"] = threads
        cam = GpuCamera(Settings.from_object(sd), background_function=_builtin_background, background=_builtin_background)
        cam.set_world_dense(c["origin"], c["cells"], c["chunk_res"], c["materials"])
        cam.pos, cam.rot, cam.lens = c["cam_pos"], c["cam_rot"], float(c["cam_lens"])
        return cam

    cam1 = cam_for(1)
    img, trav, th = cam1.tile(0, 0)
    full = np.frombuffer(img, np.uint8).reshape(H, W, 4)
    assert th == 0 and len(img) == W * H * 4
    # traversed: chunk-origin int tuples, comparable with the keys chunk_update looks up (init.py:447)
    assert len(trav) > 0 and all(len(t) == 3 and all(isinstance(v, int) and v % 16 == 0 for v in t) for t in trav)
    assert (-16, 0, 0) in trav                              # the chunk the default camera (-12, 4, 4) starts in
    cam3 = cam_for(3)
    acc = np.zeros_like(full)
    for t in range(3):
        img_t, trav_t, th_t = cam3.tile(t, 0)
        part = np.frombuffer(img_t, np.uint8).reshape(H, W, 4)
        ys, xs = np.mgrid[0:H, 0:W]
        mine = ((xs ^ ys) % 3) == t
        assert th_t == t and not part[~mine].any()
        assert np.array_equal(part[mine], full[mine])
        acc[mine] = part[mine]
    assert np.array_equal(acc, full)
    # stripes: 3 "ranks" of 8-row stripes reassemble the same frame bit for bit
    stitched = np.zeros_like(full)
    for r in range(3):
        out = cam1.render(stripes=(8, 3, r), want=("rgba8",))["rgba8"].cpu().numpy()
        rows = [y for y in range(H) if (y // 8) % 3 == r]
        stitched[rows] = out
    assert np.array_equal(stitched, full)
    # rect sub-region
    sub = cam1.render(rect=(5, 7, 41, 30), want=("rgba8",))["rgba8"].cpu().numpy()
    assert np.array_equal(sub, full[7:30, 5:41])
    cam1.close(); cam3.close()


def test_static_off_changes_noise_per_frame():
    c = load_case("c1_shipped")
    from tests.test_gpu_parity import gpu_case
    a = gpu_case(c, want=("rgba8",))
    b = gpu_case(c, want=("rgba8",))
    assert np.array_equal(a["rgba8"], b["rgba8"])            # static: identical frames (README "static")
    from tests.test_gpu_parity import _builtin_background
    from python_raytracer_b200 import GpuCamera, Settings
    sd = dict(c["settings"]); sd["static"] = False; sd["threads"] = 1
    cam = GpuCamera(Settings.from_object(sd), background_function=_builtin_background, background=_builtin_background)
    cam.set_world_dense(c["origin"], c["cells"], c["chunk_res"], c["materials"])
    cam.pos, cam.rot, cam.lens = c["cam_pos"], c["cam_rot"], float(c["cam_lens"])
    f0 = cam.render(want=("rgba8",))["rgba8"].clone()
    f1 = cam.render(want=("rgba8",))["rgba8"].clone()
    assert not torch.equal(f0, f1)
    cam.close()
