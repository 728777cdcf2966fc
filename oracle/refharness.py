"""Oracle harness: runs the UNMODIFIED reference trace loop headless.  TEST INFRASTRUCTURE ONLY.

Nothing here is imported by the product package.  It only works where ``/root/reference`` exists
(the build container); it is used to (a) generate the golden fixtures under ``tests/golden/``
(``oracle/gen_golden.py``) and (b) validate the C restatement ``oracle/vt_oracle.c``.

How the reference is loaded, nothing is copied into this repo:
  * ``oracle/pygame_stub`` + ``/root/reference`` are put on ``sys.path``; cwd is switched to the
    reference root because its paths are cwd-relative (data.py:17, data.py:627,
    mods/default/init.py:176);
  * ``import data`` executes the mod (data.py:627);
  * ``init.py`` cannot be imported (its last line constructs ``Window()``, init.py:474), so its AST is
    parsed at run time and only the imports + ``class Camera`` + ``class Window`` are exec'd into a
    synthetic module.
"""
from __future__ import annotations

import ast
import math
import multiprocessing as mp
import os
import sys
import types

import numpy as np

REFERENCE_ROOT = os.environ.get("VT_REFERENCE_ROOT", "/root/reference")
_STUB_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "pygame_stub")
BYTECODE_ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")     # oracle/build_ref.py

_REF = None


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "init.py"))


def available_bytecode() -> bool:
    """oracle/_ref/ (bytecode of the reference built by oracle/build_ref.py) is present and was compiled by this
    interpreter version."""
    import importlib.util
    stamp = os.path.join(BYTECODE_ROOT, "BUILT_WITH")
    if not os.path.isfile(stamp) or not os.path.isfile(os.path.join(BYTECODE_ROOT, "vt_reference_init.vtbc")) \
            or not os.path.isfile(os.path.join(BYTECODE_ROOT, "assets.zip")):
        return False
    return open(stamp).read().strip().endswith(importlib.util.MAGIC_NUMBER.hex())


class _BytecodeFinder:
    """Imports the reference's modules from the bytecode files of oracle/_ref/ (oracle/build_ref.py)."""
    FILES = {"lib": "lib.vtbc", "data": "data.vtbc", "vt_reference_init": "vt_reference_init.vtbc",
             "mods.default.init": os.path.join("mods", "default", "init.vtbc")}

    def find_spec(self, fullname, path=None, target=None):
        import importlib.util
        rel = self.FILES.get(fullname)
        if rel is None or not os.path.isfile(os.path.join(BYTECODE_ROOT, rel)):
            return None
        return importlib.util.spec_from_loader(fullname, self, origin=os.path.join(BYTECODE_ROOT, rel))

    def create_module(self, spec):
        return None

    def exec_module(self, module):
        import marshal
        path = os.path.join(BYTECODE_ROOT, self.FILES[module.__name__])
        with open(path, "rb") as f:
            code = marshal.loads(f.read()[16:])          # pyc: 16-byte header, then the marshalled code object
        module.__file__ = path
        exec(code, module.__dict__)


def load_reference(mod: str = "default"):
    """Import the reference once per process; returns a namespace (data, lib, Camera, Window, pg)."""
    global _REF
    if _REF is not None:
        return _REF
    from_source = available()
    if not from_source and not available_bytecode():
        raise RuntimeError("reference tree not present at %s (and no bytecode build under %s)" % (REFERENCE_ROOT, BYTECODE_ROOT))
    root = REFERENCE_ROOT if from_source else BYTECODE_ROOT
    if not from_source:
        if not any(isinstance(f, _BytecodeFinder) for f in sys.meta_path):
            sys.meta_path.insert(0, _BytecodeFinder())
        # the mod's data files (config, voxel models) are read relative to the working directory while `data` is imported
        import tempfile
        import zipfile
        import atexit
        import shutil
        root = tempfile.mkdtemp(prefix="vt_ref_")
        atexit.register(shutil.rmtree, root, True)
        with zipfile.ZipFile(os.path.join(BYTECODE_ROOT, "assets.zip")) as z:
            z.extractall(root)
    sys.dont_write_bytecode = True
    for p in (_STUB_DIR, root):
        if p not in sys.path:
            sys.path.insert(0, p)
    old_cwd, old_argv = os.getcwd(), sys.argv
    os.chdir(root)
    sys.argv = ["init.py", mod]
    try:
        import pygame as pg  # the stub
        import lib
        import data
        if from_source:
            with open(os.path.join(REFERENCE_ROOT, "init.py"), "rt") as f:
                tree = ast.parse(f.read())
            keep = [n for n in tree.body
                    if isinstance(n, (ast.Import, ast.ImportFrom))
                    or (isinstance(n, ast.ClassDef) and n.name in ("Camera", "Window"))]
            module = types.ModuleType("vt_reference_init")
            module.__file__ = os.path.join(REFERENCE_ROOT, "init.py")
            sys.modules["vt_reference_init"] = module
            exec(compile(ast.Module(body=keep, type_ignores=[]), module.__file__, "exec"), module.__dict__)
        else:                                   # the same extraction, done at build time by oracle/build_ref.py
            import vt_reference_init as module
    finally:
        os.chdir(old_cwd)
        sys.argv = old_argv
    _REF = types.SimpleNamespace(data=data, lib=lib, pg=pg, Camera=module.Camera, Window=module.Window,
                                 module=module, shipped=dict(vars(data.settings)))
    return _REF


def override_settings(**kw):
    """Mutate ``data.settings`` and recompute the derived fields exactly as data.py:64-68 does.
    ``pixels`` is rebuilt lazily by ``render`` (the full x-major list of data.py:70-77 is 2 M tuples at 1080p)."""
    ref = load_reference()
    s = ref.data.settings
    for k, v in kw.items():
        if not hasattr(s, k):
            raise KeyError(k)
        setattr(s, k, v)
    s.window = s.width, s.height
    s.window_scaled = s.window[0] * s.scale, s.window[1] * s.scale
    s.proportions = ((s.width + s.height) / 2) / max(s.width, s.height)
    s.chunk_time = s.chunk_rate / 1000
    s.chunk_radius = round(s.chunk_size / 2)
    return s


def restore_settings():
    ref = load_reference()
    for k, v in ref.shipped.items():
        setattr(ref.data.settings, k, v)


def all_materials(cam):
    """Every distinct Material object referenced by the camera's chunk frames (identity-unique)."""
    seen, out = set(), []
    for frame in cam.chunks.values():
        for d in (frame.data3, frame.data6):
            for m in d.values():
                if id(m) not in seen:
                    seen.add(id(m))
                    out.append(m)
    return out


def make_camera(pos, rot_euler):
    """Camera placed like Window.update does (init.py:464-465): pos vec3, rot = euler.quaternion()."""
    ref = load_reference()
    cam = ref.Camera()
    cam.pos = ref.lib.vec3(float(pos[0]), float(pos[1]), float(pos[2]))
    cam.rot = ref.lib.vec3(rot_euler[0], rot_euler[1], rot_euler[2]).quaternion()
    return cam


def build_chunks(cam, window_state=None, traversed=None):
    """One pass of the reference's own ``Window.chunk_update`` (init.py:389-452), called unbound on a
    namespace that carries the attributes it touches.  Returns the state so later passes are incremental."""
    ref = load_reference()
    if window_state is None:
        for obj in ref.data.objects.values():
            obj.redraw = True
        window_state = types.SimpleNamespace(cam=cam, chunks={}, chunks_objects={}, timer=0.0,
                                             traversed=[[]])
    window_state.cam = cam
    if traversed is not None:
        window_state.traversed = [list(traversed)]
    for obj in ref.data.objects.values():
        obj.update(cam.pos)
    window_state.timer = 0.0
    ref.Window.chunk_update(window_state, max(1.0, ref.data.settings.chunk_time + 1.0))
    return window_state


def _wrap_materials(cam):
    """First-hit recorder through the reference's own plug-in point (init.py:81): the wrapper sees the live
    ray on every material call and stamps the first one."""
    ref = load_reference()
    originals = []
    for m in all_materials(cam):
        fn = m.function
        if getattr(fn, "_vt_wrapped", False):
            continue

        def wrapper(ray, mat, settings, _fn=fn):
            if not hasattr(ray, "first_hit"):
                p = math.floor(ray.pos)
                ray.first_hit = (p.x, p.y, p.z)
                ray.first_step = ray.step
                ray.hits = 0
            ray.hits += 1
            return _fn(ray, mat, settings)

        wrapper._vt_wrapped = True
        originals.append((m, fn))
        m.function = wrapper
    return originals


def _unwrap_materials(originals):
    for m, fn in originals:
        m.function = fn


_WORK = {}


def _tile_worker(t):
    """Runs the reference's own Camera.tile on pixel list ``t`` with a recording shim around Camera.trace."""
    ref = load_reference()
    cam = _WORK["cam"]
    rays = []
    real_trace = ref.Camera.trace

    def trace(dir_x, dir_y, detail):
        ray = real_trace(cam, dir_x, dir_y, detail)
        fh = getattr(ray, "first_hit", None)
        rays.append((ray.color.r, ray.color.g, ray.color.b, ray.energy, ray.bounces, ray.step,
                     getattr(ray, "hits", 0),
                     fh[0] if fh else 0, fh[1] if fh else 0, fh[2] if fh else 0, 1 if fh else 0,
                     getattr(ray, "first_step", -1), ray.vel.x, ray.vel.y, ray.vel.z))
        return ray

    cam.trace = trace
    try:
        image, traversed, thread = cam.tile(t, 0)
    finally:
        del cam.trace
    surf = ref.pg.LAST_SURFACE
    px = ref.data.settings.pixels[t]
    means = np.array([surf.px.get((x, y), (0.0, 0.0, 0.0, 0.0)) for (x, y) in px], dtype=np.float64).reshape(-1, 4)
    return t, means, np.array(rays, dtype=np.float64).reshape(-1, 15), [tuple(c) for c in traversed]


def render(cam, pixels=None, procs=None):
    """Render ``pixels`` (list of (x, y); default = every pixel, x-major like data.py:74-77) through the
    reference's ``Camera.tile``.  Returns a dict of arrays aligned with ``pixels``:
      mean[n,4]   float64 per-pixel RGBA mean handed to Surface.set_at (pre-quantisation)
      nsamp[n]    samples traced for the pixel
      ray0[n,15]  sample-0 ray record: r,g,b,energy,bounces,step,hits,fx,fy,fz,fflag,first_step,vx,vy,vz
      rays        total rays;  traversed: set of chunk-origin tuples
    """
    ref = load_reference()
    s = ref.data.settings
    if pixels is None:
        pixels = [(x, y) for x in range(s.width) for y in range(s.height)]
    pixels = [(int(x), int(y)) for x, y in pixels]
    procs = procs or min(os.cpu_count() or 1, 16)
    procs = max(1, min(procs, len(pixels)))
    # contiguous split keeps per-worker order == caller order (tile() itself is order-agnostic per pixel)
    bounds = np.linspace(0, len(pixels), procs + 1).astype(int)
    s.pixels = [pixels[bounds[i]:bounds[i + 1]] for i in range(procs)]
    originals = _wrap_materials(cam)
    _WORK["cam"] = cam
    try:
        if procs == 1:
            results = [_tile_worker(0)]
        else:
            ctx = mp.get_context("fork")
            with ctx.Pool(procs) as pool:
                results = pool.map(_tile_worker, range(procs))
    finally:
        _unwrap_materials(originals)
        _WORK.clear()
    results.sort(key=lambda r: r[0])
    mean = np.concatenate([r[1] for r in results], axis=0)
    # per-pixel sample count per init.py:133-134
    nsamp = np.empty(len(pixels), dtype=np.int64)
    for i, (x, y) in enumerate(pixels):
        dx = -1 + (x / s.width) * 2
        dy = -1 + (y / s.height) * 2
        nsamp[i] = max(1, round(s.samples * (1 - abs(dx * dy) * s.lod_edge)))
    allrays = np.concatenate([r[2] for r in results], axis=0)
    assert allrays.shape[0] == int(nsamp.sum())
    first = np.concatenate([[0], np.cumsum(nsamp)[:-1]])
    traversed = set()
    for r in results:
        traversed.update(r[3])
    return dict(pixels=np.array(pixels, dtype=np.int32), mean=mean, nsamp=nsamp, ray0=allrays[first],
                rays=int(nsamp.sum()), traversed=traversed)


# ---------------------------------------------------------------------------------------------------
# Scene extraction: the reference's chunk frames -> the dense world the C oracle and the fixtures use.
# Deliberately brute force (one Frame.get_voxel call per cell) so it cannot share a bug with the
# product's own flattening code.

MATERIAL_FIELDS = ("albedo_r", "albedo_g", "albedo_b", "roughness", "absorption", "ior", "energy")


def extract_world(window_state, cam):
    """Returns (origin[3], cells[x,y,z] u8, chunk_res[cx,cy,cz] u8, materials[n+1,7] f64, names).
    cells come from the full-resolution frames ``window_state.chunks[post][0]`` (init.py:441-444);
    chunk_res is the resolution of the frame the camera currently holds (init.py:448-450), 0 if culled/absent."""
    ref = load_reference()
    cs = ref.data.settings.chunk_size
    posts = [p for p, v in window_state.chunks.items() if v]
    lo = np.array(posts).min(axis=0)
    hi = np.array(posts).max(axis=0) + cs
    dims = (hi - lo).astype(int)
    cells = np.zeros(tuple(dims), np.uint8)
    chunk_res = np.zeros(tuple(dims // cs), np.uint8)
    mats, names = {}, {}
    mod = sys.modules.get("mods.default.init")
    if mod is not None:
        for k, v in vars(mod).items():
            if isinstance(v, ref.data.Material):
                names[id(v)] = k
    vec3 = ref.lib.vec3
    for post in posts:
        frame = window_state.chunks[post][0]
        assert frame.resolution == 1
        for x in range(post[0], post[0] + cs):
            for y in range(post[1], post[1] + cs):
                for z in range(post[2], post[2] + cs):
                    m = frame.get_voxel(vec3(x, y, z))
                    if m is not None:
                        if id(m) not in mats:
                            mats[id(m)] = (len(mats) + 1, m)
                        cells[x - lo[0], y - lo[1], z - lo[2]] = mats[id(m)][0]
        if post in cam.chunks:
            c = tuple((np.array(post) - lo) // cs)
            chunk_res[c] = cam.chunks[post].resolution
    table = np.zeros((len(mats) + 1, 7), np.float64)
    mat_names = [""] * (len(mats) + 1)
    for mid, m in mats.values():
        table[mid] = (m.albedo.r, m.albedo.g, m.albedo.b, m.roughness, m.absorption, m.ior, m.energy)
        mat_names[mid] = names.get(id(m), "mat%d" % mid)
    return lo.astype(np.int32), cells, chunk_res, table, mat_names


def chunk_res_table(cam, origin, shape):
    """Resolution of each chunk frame the camera holds, on the dense chunk-slot grid."""
    ref = load_reference()
    cs = ref.data.settings.chunk_size
    out = np.zeros(shape, np.uint8)
    for post, frame in cam.chunks.items():
        c = tuple((np.array(post) - np.asarray(origin)) // cs)
        out[c] = frame.resolution
    return out


def settings_dict():
    """The settings subset the path reads, as a plain dict (data.py:18-68)."""
    s = load_reference().data.settings
    keys = ("width", "height", "samples", "static", "shutter", "falloff", "dof", "dist_min", "dist_max",
            "max_light", "max_bounces", "lod_bounces", "lod_samples", "lod_random", "lod_edge",
            "chunk_size", "chunk_lod", "fov", "culling")
    return {k: getattr(s, k) for k in keys}
