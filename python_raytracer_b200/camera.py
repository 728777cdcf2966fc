"""GpuCamera: the reference's ``Camera`` protocol (reference init.py:13-150) on top of libvoxtrace.so.

Same attribute and method names as the reference class so ``Window``-style callers can swap it in:
``pos``, ``rot``, ``lens``, ``chunks``, ``chunk_set``, ``chunk_get``, ``tile``, ``trace``.  The per-tile Python thread pool
(init.py:162, 193-205) is replaced by one kernel launch per call; ``tile(thread, t)`` still returns
``(rgba_bytes, traversed, thread)`` like init.py:149-150.

No CPU fallback: constructing a GpuCamera without a CUDA device raises VoxtraceError.
User-defined Python material callbacks cannot run inside a CUDA kernel and are rejected with a clear error
(only ``lib.material`` / ``lib.material_background`` are compiled device functions).
"""
from __future__ import annotations

import ctypes as C
import math
import sys

import numpy as np
import torch

from . import abi
from .abi import VoxtraceError, VtMaterial, VtRegion
from .settings import Settings
from .world import WorldMirror, flatten_frame

_UNSET = object()


def _xyz(v):
    if hasattr(v, "x"):
        return float(v.x), float(v.y), float(v.z)
    return float(v[0]), float(v[1]), float(v[2])


def _xyzw(q):
    if hasattr(q, "w"):
        return float(q.x), float(q.y), float(q.z), float(q.w)
    return float(q[0]), float(q[1]), float(q[2]), float(q[3])


def _is_builtin(fn, name):
    return callable(fn) and getattr(fn, "__name__", None) == name and getattr(fn, "__module__", None) == "lib"


class GpuCamera:
    def __init__(self, settings, device: int = 0, material_function=None, background_function=None,
                 background=_UNSET):
        """settings: reference ``data.settings`` store, dict or Settings.
        material_function / background_function: the callables accepted as "the built-in" (pass ``lib.material``
        and ``lib.material_background`` of the running engine); when omitted, functions named ``material`` /
        ``material_background`` from a module called ``lib`` are accepted.
        background: what ``data.background`` is set to (init.py:119): the built-in function or None."""
        self._lib = abi.load()
        if not torch.cuda.is_available():
            raise VoxtraceError("GpuCamera needs a CUDA device; the GPU path has no CPU fallback")
        self.settings = settings if isinstance(settings, Settings) else Settings.from_object(settings)
        self._live_settings = None if isinstance(settings, (Settings, dict)) else settings
        self.device = torch.device("cuda", device)
        self._material_function = material_function
        self._background_function = background_function
        ctx = C.c_void_p()
        if self._lib.vt_create(int(device), C.byref(ctx)) != 0:
            raise VoxtraceError(self._lib.vt_last_error(None).decode())
        self._ctx = ctx
        # reference attributes (init.py:15-18)
        self.pos = (0.0, 0.0, 0.0)
        self.rot = (0.0, 0.0, 0.0, 0.0)
        self.lens = float(self.settings.fov) * math.pi / 8
        self.chunks = {}
        self.background = background_function if background is _UNSET else background
        # device-side state
        self._known_posts = set()      # every chunk position chunk_set was ever called with (frame or None)
        self._mirror = None
        self._uploaded = {}            # post -> (frame object, resolution) last flattened
        self._mat_ids = {}             # id(Material) -> (index, Material)
        self._dense_mode = False
        self._settings_key = None
        self._frame_seed = 0
        self._visited = None
        self._counters = torch.zeros(32, dtype=torch.int64, device=self.device)
        self._pix_cache = {}
        self._static_materials = None
        self._mats_sent = None
        self._world_shape = None
        self._world_origin = None

    # ------------------------------------------------------------------ plumbing
    def _check(self, rc):
        if rc != 0:
            raise VoxtraceError(self._lib.vt_last_error(self._ctx).decode())

    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.vt_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launch_count(self) -> int:
        return int(self._lib.vt_launch_count(self._ctx))

    def _push_settings(self):
        if self._live_settings is not None:      # the engine mutates data.settings in place; re-read it
            self.settings = Settings.from_object(self._live_settings)
        if self.background is not None:
            ok = (self.background is self._background_function) if self._background_function is not None \
                else _is_builtin(self.background, "material_background")
            if not ok:
                raise VoxtraceError("custom background callbacks cannot run on the GPU path; use "
                                    "lib.material_background or None (got %r)" % (self.background,))
        key = (self.settings.key(), self.background is not None)
        if key != self._settings_key:
            st = self.settings.to_struct(background=self.background is not None)
            if self._mirror is not None and self._mirror.cs != int(self.settings.chunk_size):
                # data.settings.chunk_size was edited in place: the chunk dictionary is keyed by the old size, so the
                # device world is dropped and rebuilt from whatever chunk_set holds now (vt_set_settings refuses a
                # chunk_size change while a world is loaded)
                self._check(self._lib.vt_clear_world(self._ctx))
                self._mirror, self._uploaded, self._visited = None, {}, None
                self._known_posts = set(self.chunks)
            self._check(self._lib.vt_set_settings(self._ctx, C.byref(st)))
            self._settings_key = key

    # ------------------------------------------------------------------ reference protocol: chunks
    def chunk_set(self, post: tuple, chunk):
        """init.py:21-25."""
        if self._dense_mode:
            raise VoxtraceError("chunk_set after set_world_dense: pick one world path per camera")
        # Window.chunk_update calls this for EVERY chunk it knows, with None for the culled ones (init.py:447-452):
        # the positions are remembered so that the visited-chunk grid covers them (see _sync_world)
        self._known_posts.add(tuple(int(v) for v in post))
        if chunk:
            self.chunks[post] = chunk
        elif post in self.chunks:
            del self.chunks[post]

    def chunk_get(self, pos):
        """init.py:28-33."""
        cs = self.settings.chunk_size
        x, y, z = _xyz(pos)
        post = ((x // cs) * cs, (y // cs) * cs, (z // cs) * cs)
        return self.chunks.get(post)

    def _material_id(self, mat):
        ent = self._mat_ids.get(id(mat))
        if ent is None:
            fn = getattr(mat, "function", None)
            ok = (fn is self._material_function) if self._material_function is not None else _is_builtin(fn, "material")
            if not ok:
                raise VoxtraceError(
                    "material %r uses a custom Python function (%r); user-defined material callbacks cannot run "
                    "on the GPU path -- only the built-in lib.material is compiled" % (mat, fn))
            if len(self._mat_ids) >= 255:
                raise VoxtraceError("more than 255 distinct materials")
            ent = (len(self._mat_ids) + 1, mat)
            self._mat_ids[id(mat)] = ent
        return ent[0]

    def _push_materials(self):
        """Material rows -> device.  Mods may mutate Material attributes in place (README.md:77), so the table is
        rebuilt every frame, but it is sent only when a value changed."""
        if self._static_materials is not None:
            table = self._static_materials
        else:
            table = np.zeros((len(self._mat_ids) + 1, 7), np.float64)
            for idx, m in self._mat_ids.values():
                a = m.albedo
                rgb = (a.r, a.g, a.b) if hasattr(a, "r") else tuple(a)
                table[idx] = (rgb[0], rgb[1], rgb[2], m.roughness, m.absorption, m.ior, m.energy)
        if self._mats_sent is not None and self._mats_sent.shape == table.shape and np.array_equal(self._mats_sent, table):
            return
        n = table.shape[0] - 1
        rows = np.zeros((n + 1, 8), np.float64)          # vt_material: albedo[3], roughness, absorption, ior, energy, reserved
        rows[:, :7] = table
        self._check(self._lib.vt_set_materials(self._ctx, rows.ctypes.data, n))
        self._mats_sent = np.array(table, copy=True)

    def _reach_box(self):
        """Chunk-aligned box no ray of this camera can leave: a ray travels less than dist_min + life + one step per axis
        (|vel| <= 1 per axis: primary rays are not longer than 1, lib.py:372-376, and a bounce normalises the largest
        axis to 1, lib.py:310-314); life <= dist_max * detail with detail <= 1 + max(0, -lod_edge) (init.py:56, 133)."""
        s = self.settings
        cs = int(s.chunk_size)
        reach = float(s.dist_max) * (1.0 + max(0.0, -float(s.lod_edge))) + cs + 2
        p = np.array(_xyz(self.pos), np.float64)
        lo = (np.floor((p - reach) / cs) * cs).astype(np.int64)
        hi = (np.floor((p + reach) / cs) * cs).astype(np.int64) + cs
        return lo, hi

    def _sync_world(self):
        """Flatten chunks whose Frame changed since the last call and upload the bricks they touch.

        The device box (brick table, chunk table, visited grid) spans every chunk position chunk_set was ever called
        with.  With settings.culling on it also spans the camera's reach: the reference appends EVERY chunk position a
        ray snaps into to ray.traversed, empty and culled ones included (init.py:67-73), and chunk_update tests
        `post_chunk in traversed` against Window.chunks, not cam.chunks (init.py:447) -- on the first frame cam.chunks is
        empty and the picture only ever appears because those positions are reported."""
        if self._dense_mode:
            return
        s = self.settings
        cs = int(s.chunk_size)
        if self._mirror is None or self._mirror.cs != cs:
            self._mirror, self._uploaded = WorldMirror(cs), {}
        posts = set(self._known_posts) | {tuple(int(v) for v in p) for p in self.chunks}
        lo = hi = None
        if posts:
            lo, dims = self._mirror.box_for(posts)
            hi = lo + dims
        if bool(getattr(s, "culling", False)) or lo is None:
            rlo, rhi = self._reach_box() if bool(getattr(s, "culling", False)) else (np.zeros(3, np.int64), np.full(3, cs, np.int64))
            lo = rlo if lo is None else np.minimum(lo, rlo)
            hi = rhi if hi is None else np.maximum(hi, rhi)
        dims = hi - lo
        rebuilt = False
        if not self._mirror.covers(lo, dims):
            if bool(getattr(s, "culling", False)):       # room to move before the next rebuild
                lo, hi = lo - 4 * cs, hi + 4 * cs
            if self._mirror.origin is not None:
                # grow: keep the old box inside the new one, unless that makes it more than twice the size needed
                # (a camera that walked far away from where it started)
                old_lo, old_hi = self._mirror.origin, self._mirror.origin + self._mirror.dims
                glo, ghi = np.minimum(lo, old_lo), np.maximum(hi, old_hi)
                if np.all(ghi - glo <= 2 * (hi - lo)):
                    lo, hi = glo, ghi
            dims = hi - lo
            self._mirror.reset(lo, dims)
            self._uploaded = {}
            o = (C.c_int32 * 3)(*[int(v) for v in lo])
            d = (C.c_int32 * 3)(*[int(v) for v in dims])
            self._check(self._lib.vt_set_world(self._ctx, o, d))
            rebuilt = True
        dirty = set()
        table_changed = rebuilt
        for post in list(self._uploaded):
            if post not in self.chunks:
                dirty |= self._mirror.write_chunk(post, None, 0)
                del self._uploaded[post]
                table_changed = True
        for post, frame in self.chunks.items():
            key = tuple(int(v) for v in post)
            prev = self._uploaded.get(key)
            if prev is not None and prev[0] is frame and prev[1] == frame.resolution:
                continue
            cells = flatten_frame(frame, key, cs, self._material_id)
            dirty |= self._mirror.write_chunk(key, cells, int(frame.resolution))
            self._uploaded[key] = (frame, frame.resolution)
            table_changed = True
        if dirty:
            coords, cells = self._mirror.gather_bricks(dirty)
            self._check(self._lib.vt_upload_bricks(self._ctx, coords.ctypes.data, cells.ctypes.data, len(coords)))
        if table_changed:
            t = np.ascontiguousarray(self._mirror.chunk_res)
            self._check(self._lib.vt_set_chunk_table(self._ctx, t.ctypes.data, t.size))
            self._visited = None

    def mark_dirty(self, post=None):
        """Force re-flattening of one chunk (or all) after an in-place Frame edit."""
        if post is None:
            self._uploaded = {}
        else:
            self._uploaded.pop(tuple(int(v) for v in post), None)

    # ------------------------------------------------------------------ bulk world path (synthetic / imported)
    def set_world_dense(self, origin, cells, chunk_res, materials):
        """Load a whole world at once: ``cells`` [nx, ny, nz] u8 material ids (numpy or CUDA tensor), world cell of
        its corner ``origin`` (multiples of chunk_size), ``chunk_res`` [nx/cs, ny/cs, nz/cs] u8 (0 = no frame,
        else Frame.resolution, what chunk_update would pick -- see lod_table) and the material table
        [n+1, 7] = albedo r,g,b, roughness, absorption, ior, energy (row 0 unused)."""
        self._push_settings()
        self._dense_mode = True
        self.chunks = {}
        t = cells if isinstance(cells, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(cells, dtype=np.uint8))
        t = t.to(self.device, dtype=torch.uint8).contiguous()
        o = (C.c_int32 * 3)(*[int(v) for v in origin])
        d = (C.c_int32 * 3)(*[int(v) for v in t.shape])
        self._check(self._lib.vt_set_world(self._ctx, o, d))
        torch.cuda.synchronize(self.device)
        self._check(self._lib.vt_upload_dense(self._ctx, t.data_ptr()))
        self._world_origin = np.asarray([int(v) for v in origin], np.int64)
        self._world_shape = tuple(int(v) for v in t.shape)
        self.set_chunk_res(chunk_res)
        self._static_materials = np.ascontiguousarray(materials, dtype=np.float64)
        assert self._static_materials.ndim == 2 and self._static_materials.shape[1] == 7

    def set_chunk_res(self, chunk_res):
        """Replace the per-chunk resolution table (LOD / culling state) of a dense world."""
        cs = int(self.settings.chunk_size)
        t = np.ascontiguousarray(chunk_res, dtype=np.uint8)
        if tuple(t.shape) != tuple(v // cs for v in self._world_shape):
            raise VoxtraceError("chunk_res shape %s does not match world %s / chunk_size %d" % (t.shape, self._world_shape, cs))
        self._check(self._lib.vt_set_chunk_table(self._ctx, t.ctypes.data, t.size))
        if getattr(self, "_chunk_res_shape", None) != t.shape:       # same grid: the visited words keep their size (cleared per render)
            self._visited = None
        self._chunk_res_shape = t.shape

    def download_dense(self):
        """The world held in bricks as a dense [nx, ny, nz] u8 CUDA tensor (inverse of ``set_world_dense``)."""
        if self._dense_mode:
            shape = self._world_shape
        else:
            self._sync_world()
            shape = tuple(int(v) for v in self._mirror.dims)
        if shape is None:
            raise VoxtraceError("download_dense: no world")
        out = torch.empty(shape, dtype=torch.uint8, device=self.device)
        torch.cuda.synchronize(self.device)
        self._check(self._lib.vt_download_dense(self._ctx, out.data_ptr()))
        return out

    def chunk_presence(self):
        """[ncx, ncy, ncz] bool: chunk slots that hold a voxel (``post_chunk in self.chunks`` after chunk_update,
        init.py:441-452), derived on the device from the brick table.  chunk_size % 8 == 0 only."""
        cs = int(self.settings.chunk_size)
        if not self._dense_mode:
            self._sync_world()
        world = self._world_shape if self._dense_mode else tuple(int(v) for v in self._mirror.dims)
        shape = tuple(v // cs for v in world)
        out = np.empty(shape, np.uint8)
        self._check(self._lib.vt_chunk_presence(self._ctx, out.ctypes.data, out.size))
        return out.astype(bool)

    def world_stats(self):
        nb, by = C.c_int64(), C.c_int64()
        self._check(self._lib.vt_world_stats(self._ctx, C.byref(nb), C.byref(by)))
        return dict(bricks=nb.value, bytes=by.value)

    # ------------------------------------------------------------------ render
    def _chunk_grid(self):
        if self._dense_mode:
            return self._world_origin, self._chunk_res_shape
        return self._mirror.origin, self._mirror.chunk_res.shape

    def _prepare(self):
        self._push_settings()
        self._sync_world()
        self._push_materials()
        px, py, pz = _xyz(self.pos)
        pos = (C.c_double * 3)(px, py, pz)
        rot = (C.c_double * 4)(*_xyzw(self.rot))
        self._check(self._lib.vt_set_camera(self._ctx, pos, rot, float(self.lens)))
        self._check(self._lib.vt_set_stream(self._ctx, torch.cuda.current_stream(self.device).cuda_stream))
        _, shape = self._chunk_grid()
        nwords = (int(np.prod(shape)) + 31) // 32
        if self._visited is None or self._visited.numel() != nwords:
            self._visited = torch.zeros(nwords, dtype=torch.int32, device=self.device)

    def render(self, pixels=None, rect=None, stripes=(1, 1, 0), want=("rgba8",), prepare=True, out=None,
               clear=True, visited_out=None):
        """One pass of the hot path.  ``pixels``: [n,2] int32 (x,y) list (numpy / tensor) or None for the rectangle
        ``rect`` = (x0, y0, x1, y1) (default: whole canvas) restricted to rows whose stripe
        (y // stripes[0]) % stripes[1] == stripes[2].  ``want`` picks outputs among rgba8, mean, ray0, visited,
        counters.  ``visited_out``: an int32 word tensor to OR the visited bits into instead of the camera's own (the
        caller clears it).  Returns a dict of CUDA tensors (nothing is copied to the host here)."""
        if prepare:
            self._prepare()
        s = self.settings
        reg = VtRegion()
        reg.stripe_rows, reg.stripe_count, reg.stripe_index = (int(v) for v in stripes)
        if pixels is not None:
            p = pixels if isinstance(pixels, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(pixels, dtype=np.int32))
            p = p.to(self.device, dtype=torch.int32).contiguous()
            reg.dev_pixels, reg.n_pixels = p.data_ptr(), p.shape[0]
            shape = (p.shape[0],)
        else:
            x0, y0, x1, y1 = rect if rect is not None else (0, 0, int(s.width), int(s.height))
            reg.x0, reg.y0, reg.x1, reg.y1 = int(x0), int(y0), int(x1), int(y1)
            reg.dev_pixels, reg.n_pixels = None, 0
            n = int(self._lib.vt_region_pixels(C.byref(reg)))
            shape = (n // max(1, x1 - x0), x1 - x0)
        res = out if out is not None else {}
        ptr = {}
        for name, dt, tail in (("rgba8", torch.uint8, (4,)), ("mean", torch.float32, (4,)),
                               ("ray0", torch.float64, (abi.RAY_RECORD,))):
            if name in want:
                if name not in res:
                    res[name] = torch.empty(shape + tail, dtype=dt, device=self.device)
                ptr[name] = res[name].data_ptr()
        vis = self._visited if visited_out is None else visited_out
        if clear:                      # only what this call hands out: each zero_() is a kernel launch inside the frame
            if visited_out is None and "visited" in want:
                self._visited.zero_()
            if "counters" in want:
                self._counters.zero_()
        self._check(self._lib.vt_render(self._ctx, C.byref(reg), ptr.get("rgba8"), ptr.get("mean"), ptr.get("ray0"),
                                        vis.data_ptr() if "visited" in want else None,
                                        self._counters.data_ptr() if "counters" in want else None,
                                        int(self._frame_seed)))
        if not s.static:
            self._frame_seed += 1
        if "visited" in want:
            res["visited"] = vis
        if "counters" in want:
            res["counters"] = self._counters[:8]
            res["trips"] = self._counters[8:26]          # phase trip counts and cycles, only filled by a VT_TRIPS build (tools/kbench.py)
        return res

    def visited_slots(self, visited_bits) -> np.ndarray:
        """Bitmask -> u8 array over the chunk-slot grid."""
        _, shape = self._chunk_grid()
        words = visited_bits.cpu().numpy().view(np.uint32)
        bits = np.unpackbits(words.view(np.uint8), bitorder="little")[:int(np.prod(shape))]
        return bits.reshape(shape)

    def traversed(self, visited_bits):
        """Chunk-origin tuples of the slots a frame touched -- what tile() returns as ``traversed`` (init.py:143, 150).
        The grid spans every chunk position chunk_update ever named and, with culling on, the camera's reach (see
        _sync_world), i.e. every position init.py:447 can look up."""
        origin, _ = self._chunk_grid()
        cs = int(self.settings.chunk_size)
        idx = np.argwhere(self.visited_slots(visited_bits))
        return [tuple(int(v) for v in (i * cs + origin)) for i in idx]

    # ------------------------------------------------------------------ reference protocol: trace
    TRACE_FIELDS = ("r", "g", "b", "energy", "bounces", "step", "hits", "hit_x", "hit_y", "hit_z", "hit", "hit_step",
                    "vel_x", "vel_y", "vel_z", "life", "pos_x", "pos_y", "pos_z")

    def trace_rays(self, dirs, want=(), prepare=True):
        """Camera.trace (init.py:37-121) for a batch: ``dirs`` [n, 3] = (dir_x, dir_y, detail) rows (numpy or tensor).
        Returns a dict with ``records`` [n, 20] f64 on the device (columns: TRACE_FIELDS, then 0) and, when asked for in
        ``want``, ``visited`` / ``counters`` like render().  With dof == 0 the half-angle sines and cosines are evaluated
        here with the interpreter's own libm (math.sin / math.cos, lib.py:327-332), so the primary directions are the
        reference's bit for bit; with dof != 0 the angles carry the per-ray lens jitter and the device evaluates them.
        Ray i draws its jitter from the Philox stream keyed i."""
        if prepare:
            self._prepare()
        s = self.settings
        d = dirs.detach().cpu().numpy() if isinstance(dirs, torch.Tensor) else np.asarray(dirs)
        d = np.ascontiguousarray(d, dtype=np.float64).reshape(-1, 3)
        n = d.shape[0]
        rec = torch.empty((n, abi.TRACE_RECORD), dtype=torch.float64, device=self.device)
        res = {"records": rec}
        if n == 0:
            return res
        sc_ptr = None
        if float(s.dof) == 0.0:
            prop, lens = float(s.proportions), float(self.lens)
            sc = np.empty((n, 4), np.float64)
            for i in range(n):                      # scalar math on purpose: numpy's vector sin / cos are not glibc's
                rad_y = math.radians(-((d[i, 0] / prop) * lens))
                rad_z = math.radians(+((d[i, 1] * prop) * lens))
                sc[i] = (math.sin(rad_y / 2), math.cos(rad_y / 2), math.sin(rad_z / 2), math.cos(rad_z / 2))
            sc_dev = torch.from_numpy(sc).to(self.device)
            sc_ptr = sc_dev.data_ptr()
        d_dev = torch.from_numpy(d).to(self.device)
        self._visited.zero_()
        self._counters.zero_()
        self._check(self._lib.vt_trace_rays(self._ctx, d_dev.data_ptr(), sc_ptr, n, rec.data_ptr(),
                                            self._visited.data_ptr() if "visited" in want else None,
                                            self._counters.data_ptr() if "counters" in want else None,
                                            int(self._frame_seed)))
        torch.cuda.current_stream(self.device).synchronize()      # d_dev / sc_dev may be released on return
        if "visited" in want:
            res["visited"] = self._visited
        if "counters" in want:
            res["counters"] = self._counters[:8]
        return res

    def trace(self, dir_x: float, dir_y: float, detail: float):
        """Camera.trace (init.py:37-121) for one ray; returns the ray as an attribute bag with the reference's names
        (color, energy, pos, vel, step, life, bounces, traversed -- init.py:50-59) plus hits / first_hit.  One launch per
        call: use trace_rays for more than a handful."""
        from types import SimpleNamespace
        out = self.trace_rays([(float(dir_x), float(dir_y), float(detail))], want=("visited",))
        r = out["records"][0].cpu().numpy()
        return SimpleNamespace(color=(int(r[0]), int(r[1]), int(r[2])), energy=float(r[3]), bounces=float(r[4]), step=float(r[5]),
                               hits=int(r[6]), first_hit=(int(r[7]), int(r[8]), int(r[9])) if r[10] else None,
                               vel=(float(r[12]), float(r[13]), float(r[14])), life=float(r[15]),
                               pos=(float(r[16]), float(r[17]), float(r[18])), traversed=self.traversed(out["visited"]))

    # ------------------------------------------------------------------ reference protocol: tile
    def _thread_pixels(self, thread, threads):
        key = (thread, threads, int(self.settings.width), int(self.settings.height))
        if key not in self._pix_cache:
            w, h = key[2], key[3]
            xs = torch.arange(w, dtype=torch.int32, device=self.device).view(-1, 1).expand(w, h)
            ys = torch.arange(h, dtype=torch.int32, device=self.device).view(1, -1).expand(w, h)
            sel = ((xs ^ ys) % threads) == thread               # data.py:76, x-major order like data.py:74-75
            self._pix_cache = {key: torch.stack([xs[sel], ys[sel]], dim=1).contiguous()}
        return self._pix_cache[key]

    def tile(self, thread: int, t: int = 0):
        """Camera.tile (init.py:126-150): returns (RGBA bytes of the full window, traversed, thread).  Pixels that
        belong to other threads stay (0,0,0,0) like a fresh SRCALPHA surface.  With settings.threads == 1 the
        whole frame is one launch."""
        self._prepare()
        s = self.settings
        threads = max(1, int(getattr(s, "threads", 1) or 1))
        if threads == 1:
            out = self.render(want=("rgba8", "visited"), prepare=False)
            frame = out["rgba8"]
        else:
            px = self._thread_pixels(int(thread), threads)
            out = self.render(pixels=px, want=("rgba8", "visited"), prepare=False)
            frame = torch.zeros((int(s.height), int(s.width), 4), dtype=torch.uint8, device=self.device)
            frame[px[:, 1].long(), px[:, 0].long()] = out["rgba8"]
        # device -> host through pinned staging buffers (frame and visited bits in flight together), then the one
        # copy `bytes` needs: pg.image.frombytes wants a real bytes object (init.py:186)
        key = (tuple(frame.shape), int(out["visited"].numel()))
        if getattr(self, "_tile_stage", (None,))[0] != key:
            self._tile_stage = (key, torch.empty(frame.shape, dtype=torch.uint8).pin_memory(),
                                torch.empty(out["visited"].shape, dtype=out["visited"].dtype).pin_memory())
        _, h_frame, h_vis = self._tile_stage
        h_frame.copy_(frame, non_blocking=True)
        h_vis.copy_(out["visited"], non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        image = h_frame.numpy().tobytes()
        return image, self.traversed(h_vis), thread
