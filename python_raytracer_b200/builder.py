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


class WorldBuilder:
    """Builds and rebuilds the camera's device world from a list of objects (object order = override order)."""

    def __init__(self, cam):
        self.cam = cam
        self.cells = None              # CUDA tensor [nx, ny, nz] u8, the merged world
        self.origin = None
        self.present = None            # numpy bool [ncx, ncy, ncz]

    def build(self, objects, materials):
        cam = self.cam
        cam._push_settings()
        cs = int(cam.settings.chunk_size)
        objects = [o for o in objects if o.visible]
        origin, dims = world_box(objects, cs)
        lib, ctx = cam._lib, cam._ctx
        o3 = (C.c_int32 * 3)(*[int(v) for v in origin])
        d3 = (C.c_int32 * 3)(*[int(v) for v in dims])
        cam._check(lib.vt_set_world(ctx, o3, d3))
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
        cells = torch.empty(tuple(int(v) for v in dims), dtype=torch.uint8, device=cam.device)
        torch.cuda.synchronize(cam.device)
        cam._check(lib.vt_build_world(ctx, arr, len(objects), cells.data_ptr()))
        cam._check(lib.vt_upload_dense(ctx, cells.data_ptr()))
        ncx, ncy, ncz = (int(v) // cs for v in dims)
        self.present = cells.view(ncx, cs, ncy, cs, ncz, cs).amax(dim=(1, 3, 5)).ne(0).cpu().numpy()
        self.cells, self.origin = cells, origin
        cam._dense_mode = True
        cam.chunks = {}
        cam._world_origin = origin
        cam._world_shape = tuple(int(v) for v in dims)
        cam._static_materials = np.ascontiguousarray(materials, dtype=np.float64)
        self.update_lod(visible=None)
        return cells

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
