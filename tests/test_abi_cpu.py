"""CPU-side checks: the C-ABI library loads and exports every symbol include/voxtrace.h declares, struct layouts
agree with the header, host logic (settings parsing, frame flattening) behaves like the reference."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_header_symbols():
    from python_raytracer_b200 import abi
    lib = abi.load()
    hdr = open(os.path.join(ROOT, "include", "voxtrace.h")).read()
    declared = set(re.findall(r"\b(vt_[a-z_]+)\s*\(", hdr))
    assert declared == set(abi.EXPORTS), declared ^ set(abi.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.vt_abi_version() == 1


def test_struct_sizes_match_header():
    from python_raytracer_b200 import abi
    assert C.sizeof(abi.VtSettings) == 8 * 4 + 12 * 8
    assert C.sizeof(abi.VtMaterial) == 8 * 8
    assert C.sizeof(abi.VtRegion) == 8 * 4 + 8 + 8
    assert C.sizeof(abi.VtObject) == 8 + 13 * 4 + 4               # pointer, size / mins / maxs / angle, reserved (+ padding)
    assert C.sizeof(abi.VtPhysObject) == 8 + 16 * 4               # pointer, grid_lo / grid_dims / size / mins / angle, reserved
    assert C.sizeof(abi.VtFx) == 6 * 4 + 6 * 8


def test_no_cuda_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from python_raytracer_b200 import GpuCamera, Settings, VoxtraceError, abi
    with pytest.raises(VoxtraceError):
        GpuCamera(Settings())
    ctx = C.c_void_p()
    lib = abi.load()
    assert lib.vt_create(0, C.byref(ctx)) != 0
    assert b"no CPU fallback" in lib.vt_last_error(None)


def test_region_pixels():
    from python_raytracer_b200 import abi
    lib = abi.load()
    r = abi.VtRegion()
    r.x0, r.y0, r.x1, r.y1 = 0, 0, 1920, 1080
    r.stripe_rows, r.stripe_count, r.stripe_index = 8, 4, 1
    rows = sum(1 for y in range(1080) if (y // 8) % 4 == 1)
    assert lib.vt_region_pixels(C.byref(r)) == rows * 1920
    tot = 0
    for i in range(4):
        r.stripe_index = i
        tot += lib.vt_region_pixels(C.byref(r))
    assert tot == 1920 * 1080


def test_settings_parse_rules(tmp_path):
    from python_raytracer_b200 import Settings
    cfg = tmp_path / "config.cfg"
    cfg.write_text("[WINDOW]\nwidth = 64\nheight = 48\n[RENDER]\nculling = true\nstatic = true\nsamples = 0\n"
                   "shutter = 0.25\nfov = 0\nfalloff = 0.25\nchunk_size = 0\nchunk_lod = 2\ndof = 0.5\ndist_min = 0\n"
                   "dist_max = 0\nmax_light = 1\nmax_bounces = 2\nlod_bounces = 0.5\nlod_samples = 0.5\n"
                   "lod_random = 0.25\nlod_edge = 0.25\nthreads = 0\n")
    s = Settings.from_config(str(cfg))
    # `value or default` replaces falsy values (data.py:29, 36, 39, 43)
    assert (s.samples, s.fov, s.chunk_size, s.dist_max, s.threads) == (1, 90, 16, 32, 1)
    assert s.proportions == ((64 + 48) / 2) / 64 and s.chunk_radius == 8
    bad = tmp_path / "bad.cfg"
    bad.write_text("[WINDOW]\nwidth = 64\n")
    with pytest.raises(Exception):
        Settings.from_config(str(bad))


class _Frame:
    def __init__(self, res):
        self.resolution, self.data3, self.data6 = res, {}, {}


def test_flatten_frame_precedence_and_lod():
    from python_raytracer_b200.world import flatten_frame
    ids = {"a": 1, "b": 2, "c": 3}
    f = _Frame(1)
    f.data6[(16, 0, 0, 19, 3, 3)] = "a"
    f.data6[(18, 0, 0, 31, 0, 0)] = "b"          # overlaps the first box: first box wins (data.py:142-144)
    f.data3[(17, 1, 1)] = "c"                    # a point wins over any box (data.py:139-140)
    out = flatten_frame(f, (16, 0, 0), 16, ids.__getitem__)
    assert out[0, 0, 0] == 1 and out[3, 0, 0] == 1 and out[4, 0, 0] == 2 and out[1, 1, 1] == 3
    assert out[15, 0, 0] == 2 and out.sum() == 1 * (64 - 1) + 3 + 2 * 12
    # LOD 3 frame of the chunk at -16: keys are cell // 3, stored at cell = key * 3 (data.py:165-169)
    g = _Frame(3)
    g.data3[(-5, 0, -1)] = "a"                   # cell (-15, 0, -3)
    g.data6[(-4, 1, -5, -3, 2, -4)] = "b"        # cells x in {-12,-9}, y in {3,6}, z in {-15,-12}
    out = flatten_frame(g, (-16, 0, -16), 16, ids.__getitem__)
    assert out[1, 0, 13] == 1
    assert out[4, 3, 1] == 2 and out[7, 6, 4] == 2 and (out == 2).sum() == 8 and (out > 0).sum() == 9


def test_lod_table_matches_reference_chunk_update():
    """The LOD the reference's own chunk_update chose (stored in every golden case) is reproduced by lod_table."""
    from python_raytracer_b200.world import chunk_presence, lod_table
    from tests.golden_util import case_names, load_case
    for name in case_names():
        c = load_case(name)
        s = c["settings"]
        if s["culling"]:
            continue
        present = chunk_presence(c["cells"], s["chunk_size"])
        res = lod_table(c["origin"], present, c["cam_pos"], s["chunk_size"], s["chunk_lod"], s["dist_max"])
        assert np.array_equal(res, c["chunk_res"]), name


def test_goxel_loader_matches_reference_sprite_load():
    """C++ loader vs the voxel dict the reference's own Sprite.load produced for the quirks file (generated by
    oracle/gen_objects_golden.py): header lines, double spaces, signs, unknown colours, trailing text, CRLF,
    duplicates (last wins), x = 0 landing on size.x; .txt and .gz."""
    import os
    import numpy as np
    from python_raytracer_b200.goxel import load_goxel, sprite_grid
    gdir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    g = np.load(os.path.join(gdir, "goxel_quirks.npz"))
    for ext in ("txt", "txt.gz"):
        xyz, m = load_goxel(os.path.join(gdir, "goxel_quirks." + ext), 12, [str(k) for k in g["keys"]])
        d = {}
        for p, k in zip(xyz.tolist(), m.tolist()):
            d[tuple(p)] = k
        got = np.array(sorted((x, y, z, k) for (x, y, z), k in d.items()), np.int32)
        assert np.array_equal(got, g["voxels"])
    grid = sprite_grid(xyz, m, (12, 12, 12))
    inside = g["voxels"][(g["voxels"][:, :3] < 12).all(axis=1)]
    assert int((grid > 0).sum()) == len(inside)
    for x, y, z, k in inside:
        assert grid[x, y, z] == k


def test_goxel_loader_errors(tmp_path):
    import pytest
    from python_raytracer_b200 import VoxtraceError
    from python_raytracer_b200.goxel import load_goxel
    with pytest.raises(VoxtraceError):
        load_goxel(str(tmp_path / "missing.txt"), 8, ["ffffff"])
    p = tmp_path / "short.txt"
    p.write_text("1 2 3 ffffff\n4 5\n")
    with pytest.raises(VoxtraceError, match="IndexError"):          # the reference raises on this line
        load_goxel(str(p), 8, ["ffffff"])
    q = tmp_path / "model.vox"
    q.write_text("1 2 3 ffffff\n")
    with pytest.raises(VoxtraceError, match="extension"):
        load_goxel(str(q), 8, ["ffffff"])
    e = tmp_path / "empty.txt"
    e.write_text("")
    xyz, m = load_goxel(str(e), 8, ["ffffff"])
    assert len(xyz) == 0 and len(m) == 0


def test_product_library_is_not_an_instrumented_build():
    from python_raytracer_b200 import abi
    assert abi.load().vt_build_flags() == 0
