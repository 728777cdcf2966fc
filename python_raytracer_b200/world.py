"""Host-side world bookkeeping: the camera's chunk dictionary -> bricked device grid.

What the reference keeps per chunk is a ``data.Frame``: a dict of points ``data3`` and a dict of inclusive boxes
``data6`` keyed by *world* cell divided by ``resolution`` (reference data.py:96-145, init.py:417-444).  The device
keeps one global grid of 8^3 bricks in world coordinates plus one byte per chunk slot (0 = no frame, else the
frame's resolution); an LOD-r frame only ever answers for cells whose coordinates are multiples of r, so it is
stored at those cells.  Only chunks whose Frame object changed are re-flattened and only the
bricks they touch are re-uploaded.
"""
from __future__ import annotations

import numpy as np

BRICK = 8


def flatten_frame(frame, post, chunk_size, mat_id):
    """One chunk Frame -> [cs, cs, cs] u8 of material ids at full-resolution cell positions.

    Lookup precedence of Frame.get_voxel (data.py:136-145): a ``data3`` point wins, otherwise the FIRST box of
    ``data6`` in insertion order that contains the key; boxes are therefore painted in reverse order, points last.
    Keys outside the chunk (possible only through direct Frame edits) can never be reached by the trace loop's
    lookups at resolution 1 and are dropped.
    """
    cs, r = int(chunk_size), int(frame.resolution)
    out = np.zeros((cs, cs, cs), np.uint8)
    ox, oy, oz = (int(v) for v in post)

    def axis(lo_key, hi_key, o):
        # keys lo..hi (inclusive) -> cell positions key*r inside [o, o+cs), as a slice of the chunk-local array
        lo = max(lo_key, -(-o // r))                   # ceil(o / r)
        hi = min(hi_key, (o + cs - 1) // r)
        if hi < lo:
            return None
        return slice(lo * r - o, hi * r - o + 1, r)

    for key, mat in reversed(list(frame.data6.items())):
        sx, sy, sz = axis(key[0], key[3], ox), axis(key[1], key[4], oy), axis(key[2], key[5], oz)
        if sx is None or sy is None or sz is None:
            continue
        out[sx, sy, sz] = mat_id(mat)
    for key, mat in frame.data3.items():
        x, y, z = key[0] * r - ox, key[1] * r - oy, key[2] * r - oz
        if 0 <= x < cs and 0 <= y < cs and 0 <= z < cs:
            out[x, y, z] = mat_id(mat)
    return out


class WorldMirror:
    """Dense host mirror of the device grid; turns dirty chunks into the list of bricks to upload."""

    def __init__(self, chunk_size):
        self.cs = int(chunk_size)
        self.origin = None          # np.int64[3]
        self.cells = None           # [nx, ny, nz] u8
        self.chunk_res = None       # [ncx, ncy, ncz] u8

    def box_for(self, posts):
        p = np.asarray(list(posts), dtype=np.int64).reshape(-1, 3)
        lo = p.min(axis=0)
        hi = p.max(axis=0) + self.cs
        return lo, hi - lo

    def covers(self, lo, dims):
        if self.origin is None:
            return False
        return bool(np.all(lo >= self.origin) and np.all(lo + dims <= self.origin + np.array(self.cells.shape)))

    def reset(self, lo, dims):
        self.origin = np.asarray(lo, dtype=np.int64)
        self.cells = np.zeros(tuple(int(d) for d in dims), np.uint8)
        self.chunk_res = np.zeros(tuple(int(d) // self.cs for d in dims), np.uint8)

    def slot(self, post):
        return tuple(int(v) for v in (np.asarray(post, dtype=np.int64) - self.origin) // self.cs)

    def write_chunk(self, post, cells, res):
        """Stores one chunk; returns the set of brick coordinates (relative to origin) that changed."""
        rel = np.asarray(post, dtype=np.int64) - self.origin
        x, y, z = (int(v) for v in rel)
        cs = self.cs
        view = self.cells[x:x + cs, y:y + cs, z:z + cs]
        self.chunk_res[self.slot(post)] = res
        if cells is None:
            cells = np.zeros_like(view)
        if np.array_equal(view, cells):
            return set()
        view[...] = cells
        bx = range(x // BRICK, (x + cs - 1) // BRICK + 1)
        by = range(y // BRICK, (y + cs - 1) // BRICK + 1)
        bz = range(z // BRICK, (z + cs - 1) // BRICK + 1)
        return {(i, j, k) for i in bx for j in by for k in bz}

    def gather_bricks(self, coords):
        """coords: iterable of brick coords -> (int32[m,3], u8[m,512]) in the device's brick layout."""
        coords = sorted(coords)
        m = len(coords)
        out = np.zeros((m, BRICK, BRICK, BRICK), np.uint8)
        nx, ny, nz = self.cells.shape
        for i, (bx, by, bz) in enumerate(coords):
            blk = self.cells[bx * 8:min(bx * 8 + 8, nx), by * 8:min(by * 8 + 8, ny), bz * 8:min(bz * 8 + 8, nz)]
            out[i, :blk.shape[0], :blk.shape[1], :blk.shape[2]] = blk
        return np.asarray(coords, dtype=np.int32).reshape(m, 3), out.reshape(m, 512)


def lod_table(origin, present, cam_pos, chunk_size, chunk_lod, dist_max, visible=None):
    """Per-chunk Frame.resolution as Window.chunk_update picks it (reference init.py:447-450):
        pos = chunk origin + chunk_radius;  lod = min(trunc(dist(pos, cam) / (dist_max / (1 + chunk_lod))), chunk_lod)
    and the camera receives the frame of resolution lod + 1.  ``present`` [ncx,ncy,ncz] marks chunks that hold
    voxels; ``visible`` (optional, same shape) is the culling mask (`post_chunk in traversed`, init.py:447).
    Returns u8 [ncx,ncy,ncz] with 0 for absent / culled chunks."""
    present = np.asarray(present).astype(bool)
    cs = int(chunk_size)
    radius = round(cs / 2)
    ix, iy, iz = np.meshgrid(*(np.arange(n) for n in present.shape), indexing="ij")
    cx = ix * cs + int(origin[0]) + radius
    cy = iy * cs + int(origin[1]) + radius
    cz = iz * cs + int(origin[2]) + radius
    d = np.sqrt((cx - float(cam_pos[0])) ** 2 + (cy - float(cam_pos[1])) ** 2 + (cz - float(cam_pos[2])) ** 2)
    lod = np.minimum(np.trunc(d / (dist_max / (1 + chunk_lod))), chunk_lod).astype(np.uint8)
    res = np.where(present, lod + 1, 0).astype(np.uint8)
    if visible is not None:
        res = np.where(np.asarray(visible).astype(bool), res, 0).astype(np.uint8)
    return res


def chunk_presence(cells, chunk_size):
    """[nx,ny,nz] cells -> bool [ncx,ncy,ncz]: chunk holds at least one voxel (init.py:440-446)."""
    cs = int(chunk_size)
    nx, ny, nz = cells.shape
    v = np.asarray(cells).reshape(nx // cs, cs, ny // cs, cs, nz // cs, cs)
    return v.any(axis=(1, 3, 5))
