"""ctypes binding of the C oracle (oracle/vt_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_DIR, "libvt_oracle.so")
_lib = None

RNG_MT19937, RNG_PHILOX = 0, 1

SETTINGS_INT = ("width", "height", "samples", "static_noise", "chunk_size", "chunk_radius", "background", "pad_")
SETTINGS_F64 = ("shutter", "falloff", "dof", "dist_min", "dist_max", "max_light", "max_bounces",
                "lod_bounces", "lod_samples", "lod_random", "lod_edge", "proportions")


class Settings(C.Structure):
    _fields_ = [(n, C.c_int32) for n in SETTINGS_INT] + [(n, C.c_double) for n in SETTINGS_F64]


class Camera(C.Structure):
    _fields_ = [("pos", C.c_double * 3), ("rot", C.c_double * 4), ("lens", C.c_double)]


class Scene(C.Structure):
    _fields_ = [("origin", C.c_int32 * 3), ("dims", C.c_int32 * 3), ("cells", C.c_void_p),
                ("chunk_res", C.c_void_p), ("materials", C.c_void_p), ("n_materials", C.c_int32),
                ("pad_", C.c_int32)]


def build(force: bool = False) -> str:
    """Compile libvt_oracle.so with the committed Makefile (gcc, seconds)."""
    src = os.path.join(_DIR, "vt_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < max(
            os.path.getmtime(src), os.path.getmtime(os.path.join(_DIR, "vt_oracle.h"))):
        subprocess.run(["make", "-C", _DIR, "-B", "libvt_oracle.so"], check=True, capture_output=True)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.vto_render.restype = C.c_int
        _lib.vto_render.argtypes = [C.POINTER(Settings), C.POINTER(Camera), C.POINTER(Scene), C.c_void_p,
                                    C.c_int64, C.c_int, C.c_uint64, C.c_int] + [C.c_void_p] * 6
        _lib.vto_trace_rays.restype = C.c_int
        _lib.vto_trace_rays.argtypes = [C.POINTER(Settings), C.POINTER(Camera), C.POINTER(Scene), C.c_void_p, C.c_int64,
                                        C.c_int, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.vto_py_floordiv.restype = C.c_double
        _lib.vto_py_floordiv.argtypes = [C.c_double, C.c_double]
        _lib.vto_py_mod.restype = C.c_double
        _lib.vto_py_mod.argtypes = [C.c_double, C.c_double]
        _lib.vto_mt_random.restype = C.c_double
        _lib.vto_mt_random.argtypes = [C.c_uint64, C.c_int]
        _lib.vto_philox.restype = None
        _lib.vto_philox.argtypes = [C.c_uint32] * 6 + [C.POINTER(C.c_uint32)]
        _lib.vto_ray_dir.restype = None
        _lib.vto_ray_dir.argtypes = [C.POINTER(Camera), C.c_double, C.c_double, C.POINTER(C.c_double)]
    return _lib


def make_settings(d: dict) -> Settings:
    """d: plain dict of the reference's settings names (static -> static_noise)."""
    s = Settings()
    cs = int(d["chunk_size"])
    vals = dict(d)
    vals.setdefault("static_noise", int(bool(d.get("static", True))))
    vals.setdefault("chunk_radius", round(cs / 2))
    vals.setdefault("background", 1)
    vals.setdefault("proportions", ((d["width"] + d["height"]) / 2) / max(d["width"], d["height"]))
    vals["pad_"] = 0
    for n in SETTINGS_INT:
        setattr(s, n, int(vals[n]))
    for n in SETTINGS_F64:
        setattr(s, n, float(vals[n]))
    return s


def make_camera(pos, rot_xyzw, lens) -> Camera:
    c = Camera()
    c.pos[:] = [float(v) for v in pos]
    c.rot[:] = [float(v) for v in rot_xyzw]
    c.lens = float(lens)
    return c


class World:
    """Dense world for the oracle: cells[x,y,z] u8, chunk_res[cx,cy,cz] u8, materials[n+1,7] f64."""

    def __init__(self, origin, cells, chunk_res, materials):
        self.origin = np.asarray(origin, dtype=np.int32)
        self.cells = np.ascontiguousarray(cells, dtype=np.uint8)
        self.chunk_res = np.ascontiguousarray(chunk_res, dtype=np.uint8)
        self.materials = np.ascontiguousarray(materials, dtype=np.float64)
        assert self.materials.ndim == 2 and self.materials.shape[1] == 7

    def struct(self) -> Scene:
        sc = Scene()
        sc.origin[:] = [int(v) for v in self.origin]
        sc.dims[:] = [int(v) for v in self.cells.shape]
        sc.cells = self.cells.ctypes.data
        sc.chunk_res = self.chunk_res.ctypes.data
        sc.materials = self.materials.ctypes.data
        sc.n_materials = self.materials.shape[0] - 1
        return sc


def render(settings: dict, cam_pos, cam_rot, lens, world: World, pixels=None, rng=RNG_PHILOX, seed=0,
           threads=None):
    """Runs the oracle over ``pixels`` ([n,2] int32 (x,y); default every pixel in x-major order like
    data.py:74-77).  Returns dict(mean, rgba8, nsamp, ray0, visited, counters, pixels)."""
    L = lib()
    s = make_settings(settings)
    cs = s.chunk_size
    assert all(d % cs == 0 for d in world.cells.shape), "world dims must be multiples of chunk_size"
    assert tuple(world.chunk_res.shape) == tuple(d // cs for d in world.cells.shape)
    if pixels is None:
        xs, ys = np.meshgrid(np.arange(s.width, dtype=np.int32), np.arange(s.height, dtype=np.int32), indexing="ij")
        pixels = np.stack([xs.ravel(), ys.ravel()], axis=1)
    pixels = np.ascontiguousarray(pixels, dtype=np.int32)
    n = pixels.shape[0]
    cam = make_camera(cam_pos, cam_rot, lens)
    sc = world.struct()
    mean = np.zeros((n, 4), np.float64)
    rgba8 = np.zeros((n, 4), np.uint8)
    nsamp = np.zeros(n, np.int32)
    ray0 = np.zeros((n, 15), np.float64)
    visited = np.zeros(world.chunk_res.shape, np.uint8)
    counters = np.zeros(8, np.uint64)
    threads = threads or (os.cpu_count() or 1)
    rc = L.vto_render(C.byref(s), C.byref(cam), C.byref(sc), pixels.ctypes.data, n, int(rng), int(seed),
                      int(threads), mean.ctypes.data, rgba8.ctypes.data, nsamp.ctypes.data, ray0.ctypes.data,
                      visited.ctypes.data, counters.ctypes.data)
    if rc != 0:
        raise RuntimeError("vto_render failed: %d" % rc)
    return dict(mean=mean, rgba8=rgba8, nsamp=nsamp, ray0=ray0, visited=visited, counters=counters, pixels=pixels)


def trace_rays(settings: dict, cam_pos, cam_rot, lens, world: World, dirs, rng=RNG_PHILOX, seed=0):
    """Camera.trace (init.py:37-121) for ``dirs`` [n,3] = (dir_x, dir_y, detail).  Returns dict(records [n,20], visited,
    counters); record columns: r g b energy bounces step hits fx fy fz fflag first_step vx vy vz life px py pz 0."""
    L = lib()
    s = make_settings(settings)
    dirs = np.ascontiguousarray(dirs, dtype=np.float64).reshape(-1, 3)
    n = dirs.shape[0]
    cam = make_camera(cam_pos, cam_rot, lens)
    sc = world.struct()
    rec = np.zeros((n, 20), np.float64)
    visited = np.zeros(world.chunk_res.shape, np.uint8)
    counters = np.zeros(8, np.uint64)
    rc = L.vto_trace_rays(C.byref(s), C.byref(cam), C.byref(sc), dirs.ctypes.data, n, int(rng), int(seed), rec.ctypes.data,
                          visited.ctypes.data, counters.ctypes.data)
    if rc != 0:
        raise RuntimeError("vto_trace_rays failed: %d" % rc)
    return dict(records=rec, visited=visited, counters=counters)
