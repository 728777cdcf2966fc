"""Device-side scene building: what ``Window.chunk_update`` (reference init.py:389-452) does before the trace loop.

The reference visits every cell of every dirty object in Python (5.1 s for the default scene),
merges the per-object chunk frames in object order and builds ``chunk_lod + 1`` LOD frames per chunk.  Here the
objects' sprite grids live on the GPU and one kernel (``vt_build_world``) writes the merged world, which is then
bricked on the device (``vt_upload_dense``); LOD needs no extra storage and the per-chunk LOD /
culling table is ``world.lod_table``.

Object geometry follows data.py: ``size = trunc(sprite.size / 2)``, ``mins = ceil(pos) - size``,
``maxs = floor(pos) + size`` (data.py:480-484, 592-594); drawn cells are ``mins <= p < maxs`` (init.py:413-419);
the sprite is sampled at ``pos_rotated(p - mins)`` with 90-degree steps ``round(rot / 90) % 4`` (data.py:338-371).
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, field

import numpy as np
import torch

from .abi import VoxtraceError, VtObject
from .world import lod_table


@dataclass
class SceneObject:
    """One ``data.Object`` with its active sprite frame as a dense grid of material ids (0 = empty)."""
    sprite: np.ndarray                 # [sx, sy, sz] u8 (numpy or CUDA tensor)
    pos: tuple = (0.0, 0.0, 0.0)       # Object.pos
    rot: tuple = (0.0, 0.0, 0.0)       # Object.rot, euler degrees
    visible: bool = True               # Object.visible (data.py:569)
    _dev: object = field(default=None, repr=False)

    @property
    def size(self):
        return tuple(int(v) for v in self.sprite.shape)

    def box(self):
        half = [math.trunc(s / 2) for s in self.size]
        mins = [math.ceil(p) - h for p, h in zip(self.pos, half)]
        maxs = [math.floor(p) + h for p, h in zip(self.pos, half)]
        return mins, maxs

    def angles(self):
        return [round(r / 90) % 4 for r in self.rot]


def world_box(objects, chunk_size):
    """Chunk-aligned box covering every chunk chunk_update would loop over (init.py:409-414)."""
    cs = int(chunk_size)
    lo, hi = None, None
    for o in objects:
        if not o.visible:
            continue
        mins, maxs = o.box()
        cmin = [(m // cs) * cs for m in mins]
        cmax = [(m // cs) * cs + cs for m in maxs]
        lo = cmin if lo is None else [min(a, b) for a, b in zip(lo, cmin)]
        hi = cmax if hi is None else [max(a, b) for a, b in zip(hi, cmax)]
    if lo is None:
        lo, hi = [0, 0, 0], [cs, cs, cs]
    return np.array(lo, np.int64), np.array(hi, np.int64) - np.array(lo, np.int64)


def chunk_box(obj, chunk_size):
    """Cells of the chunks chunk_update loops over for one object (init.py:406-411): [lo, hi)."""
    cs = int(chunk_size)
    mins, maxs = obj.box()
    return [(m // cs) * cs for m in mins], [(m // cs) * cs + cs for m in maxs]


class WorldBuilder:
    """Builds the camera's device world from a list of objects and keeps it up to date when objects move.

    Override order: the reference merges the per-object chunk frames in the insertion order of its
    ``chunks_objects`` dict (init.py:437-439); an object that is redrawn is deleted from that dict and inserted
    again (init.py:399-401, 426-428), i.e. it moves to the END of the order.  ``build`` starts with the list order
    (every object is drawn for the first time, in ``data.objects`` order), ``update`` reproduces the re-insertion."""

    def __init__(self, cam):
        self.cam = cam
        self.cells = None              # CUDA tensor [nx, ny, nz] u8, the merged world
        self.origin = None
        self.present = None            # numpy bool [ncx, ncy, ncz]
        self.order = []                # drawn objects in override order (later ones win)
        self.drawn = {}                # id(object) -> chunk box it was last drawn into
        self.margin = 0

    def _object_array(self, objects):
        cam = self.cam
        arr = (VtObject * max(1, len(objects)))()
        for k, o in enumerate(objects):
            if o._dev is None:
                t = o.sprite if isinstance(o.sprite, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(o.sprite, np.uint8))
                o._dev = t.to(cam.device, dtype=torch.uint8).contiguous()
            mins, maxs = o.box()
            arr[k].dev_sprite = o._dev.data_ptr()
            arr[k].size[:] = list(o.size)
            arr[k].mins[:] = mins
            arr[k].maxs[:] = maxs
            arr[k].angle[:] = o.angles()
        return arr

    def build(self, objects, materials, margin=0):
        """Full build.  ``margin``: extra chunks around the objects' box, room for objects to move into without a
        full rebuild."""
        cam = self.cam
        cam._push_settings()
        cs = int(cam.settings.chunk_size)
        objects = [o for o in objects if o.visible]
        origin, dims = world_box(objects, cs)
        self.margin = int(margin)
        origin = origin - self.margin * cs
        dims = dims + 2 * self.margin * cs
        lib, ctx = cam._lib, cam._ctx
        o3 = (C.c_int32 * 3)(*[int(v) for v in origin])
        d3 = (C.c_int32 * 3)(*[int(v) for v in dims])
        cam._check(lib.vt_set_world(ctx, o3, d3))
        arr = self._object_array(objects)
        cells = torch.empty(tuple(int(v) for v in dims), dtype=torch.uint8, device=cam.device)
        torch.cuda.synchronize(cam.device)
        cam._check(lib.vt_build_world(ctx, arr, len(objects), cells.data_ptr()))
        cam._check(lib.vt_upload_dense(ctx, cells.data_ptr()))
        cam._dense_mode = True
        cam.chunks = {}
        cam._world_origin = origin
        cam._world_shape = tuple(int(v) for v in dims)
        if cs % 8 == 0:
            self.present = cam.chunk_presence()                # from the brick table, on the device
        else:
            ncx, ncy, ncz = (int(v) // cs for v in dims)
            self.present = cells.view(ncx, cs, ncy, cs, ncz, cs).amax(dim=(1, 3, 5)).ne(0).cpu().numpy()
        self.cells, self.origin = cells, origin
        self.order = list(objects)
        self.drawn = {id(o): chunk_box(o, cs) for o in objects}
        cam._static_materials = np.ascontiguousarray(materials, dtype=np.float64)
        self.update_lod(visible=None)
        return cells

    def update(self, objects, redraw):
        """Objects moved, turned, changed frame, appeared or disappeared (``Object.redraw``, init.py:397-404).

        ``objects``: the current object list (``data.objects`` order); ``redraw``: the objects among them -- and any
        that were removed -- whose chunks must be rebuilt.  Only the chunks those objects covered before and cover
        now are redrawn and re-bricked; an object that leaves the world box triggers a full rebuild (with the same
        override order).  Returns the number of chunks redrawn (0 after a full rebuild)."""
        cam = self.cam
        cs = int(cam.settings.chunk_size)
        if self.cells is None:
            raise VoxtraceError("WorldBuilder.update before build")
        current = {id(o) for o in objects}
        todo = [o for o in objects if any(o is r for r in redraw)] + [r for r in redraw if id(r) not in current]
        boxes = []
        for o in todo:
            old = self.drawn.pop(id(o), None)
            if old is not None:
                boxes.append(old)
            self.order = [q for q in self.order if q is not o]
            o._dev = None if not isinstance(o.sprite, torch.Tensor) else o._dev    # the sprite frame may have changed
            if id(o) in current and o.visible:
                self.order.append(o)
                box = chunk_box(o, cs)
                self.drawn[id(o)] = box
                boxes.append(box)
        if not boxes:
            return 0
        lo_w = [int(v) for v in self.origin]
        hi_w = [int(a + b) for a, b in zip(self.origin, self.cam._world_shape)]
        if any(b[0][a] < lo_w[a] or b[1][a] > hi_w[a] for b in boxes for a in range(3)):
            self.build(self.order, cam._static_materials, margin=self.margin)
            return 0
        # boxes that touch or overlap are merged (an object's old and new boxes almost always do)
        merged = []
        for lo, hi in boxes:
            lo, hi = list(lo), list(hi)
            again = True
            while again:
                again = False
                for k, (mlo, mhi) in enumerate(merged):
                    if all(lo[a] <= mhi[a] and mlo[a] <= hi[a] for a in range(3)):
                        lo = [min(lo[a], mlo[a]) for a in range(3)]
                        hi = [max(hi[a], mhi[a]) for a in range(3)]
                        del merged[k]
                        again = True
                        break
            merged.append((lo, hi))
        arr = self._object_array(self.order)
        n_chunks = 0
        cam._check(cam._lib.vt_set_stream(cam._ctx, torch.cuda.current_stream(cam.device).cuda_stream))
        for lo, hi in merged:
            l3, h3 = (C.c_int32 * 3)(*lo), (C.c_int32 * 3)(*hi)
            cam._check(cam._lib.vt_update_region(cam._ctx, arr, len(self.order), self.cells.data_ptr(), l3, h3))
            c0 = [(lo[a] - lo_w[a]) // cs for a in range(3)]
            c1 = [(hi[a] - lo_w[a]) // cs for a in range(3)]
            if cs % 8 != 0:
                sub = self.cells[c0[0] * cs:c1[0] * cs, c0[1] * cs:c1[1] * cs, c0[2] * cs:c1[2] * cs]
                pres = sub.reshape(c1[0] - c0[0], cs, c1[1] - c0[1], cs, c1[2] - c0[2], cs).amax(dim=(1, 3, 5)).ne(0).cpu().numpy()
                self.present[c0[0]:c1[0], c0[1]:c1[1], c0[2]:c1[2]] = pres
            n_chunks += int(np.prod([c1[a] - c0[a] for a in range(3)]))
        if cs % 8 == 0:
            self.present = cam.chunk_presence()                # one small kernel over the brick table + 1 byte per chunk D2H
        return n_chunks

    def update_lod(self, visible=None):
        """Per-chunk LOD by camera distance and, when ``visible`` is given, occlusion culling by the previous
        frames' traversed set (init.py:446-452)."""
        s = self.cam.settings
        pos = [float(v) for v in (self.cam.pos.x, self.cam.pos.y, self.cam.pos.z)] if hasattr(self.cam.pos, "x") else list(self.cam.pos)
        res = lod_table(self.origin, self.present, pos, s.chunk_size, s.chunk_lod, s.dist_max, visible=visible)
        self.cam.set_chunk_res(res)
        return res


def scene_from_reference(objects, material_id):
    """Live reference objects -> SceneObjects, for callers that keep the engine's ``data.objects`` (data.py:430-624).

    ``objects``: iterable of reference ``data.Object`` in ``data.objects`` order; ``material_id``: callable mapping a
    reference ``Material`` to its device id (``GpuCamera._material_id`` -- it also rejects custom material functions).
    Only visible objects with a sprite are drawn (init.py:404, data.py:569).  The active frame's voxels are taken
    through the reference's own ``Sprite.get_voxels`` (which expands sprites stored at a coarser resolution), so the
    grid answers exactly what ``Sprite.get_voxel`` would for positions inside the sprite box."""
    out = []
    for obj in objects:
        spr = obj.get_sprite() if hasattr(obj, "get_sprite") else getattr(obj, "sprite", None)
        if not spr or not getattr(obj, "visible", True):
            continue
        size = (int(spr.size.x), int(spr.size.y), int(spr.size.z))
        grid = np.zeros(size, np.uint8)
        for (x, y, z), mat in spr.get_voxels(None).items():
            if 0 <= x < size[0] and 0 <= y < size[1] and 0 <= z < size[2]:
                grid[x, y, z] = material_id(mat)
        out.append(SceneObject(sprite=grid, pos=(float(obj.pos.x), float(obj.pos.y), float(obj.pos.z)),
                               rot=(float(obj.rot.x), float(obj.rot.y), float(obj.rot.z))))
    return out
