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
    """Host mirror of the device grid, kept per chunk (the world box may be much larger than the chunks that hold
    voxels: with culling on it spans the camera's reach, see GpuCamera._sync_world); turns dirty chunks into the list
    of bricks to upload."""

    def __init__(self, chunk_size):
        self.cs = int(chunk_size)
        self.origin = None          # np.int64[3]
        self.dims = None            # np.int64[3], cells per axis
        self.chunk_cells = {}       # slot (cx, cy, cz) -> [cs, cs, cs] u8 (only chunks that hold a voxel)
        self.chunk_res = None       # [ncx, ncy, ncz] u8

    def box_for(self, posts):
        p = np.asarray(list(posts), dtype=np.int64).reshape(-1, 3)
        lo = p.min(axis=0)
        hi = p.max(axis=0) + self.cs
        return lo, hi - lo

    def covers(self, lo, dims):
        if self.origin is None:
            return False
        return bool(np.all(lo >= self.origin) and np.all(lo + dims <= self.origin + self.dims))

    def reset(self, lo, dims):
        self.origin = np.asarray(lo, dtype=np.int64)
        self.dims = np.asarray(dims, dtype=np.int64)
        self.chunk_cells = {}
        self.chunk_res = np.zeros(tuple(int(d) // self.cs for d in dims), np.uint8)

    def slot(self, post):
        return tuple(int(v) for v in (np.asarray(post, dtype=np.int64) - self.origin) // self.cs)

    def write_chunk(self, post, cells, res):
        """Stores one chunk; returns the set of brick coordinates (relative to origin) that changed."""
        slot = self.slot(post)
        x, y, z = (int(v) * self.cs for v in slot)
        cs = self.cs
        self.chunk_res[slot] = res
        old = self.chunk_cells.get(slot)
        if cells is not None and not cells.any():
            cells = None
        if cells is None:
            if old is None:
                return set()
            del self.chunk_cells[slot]
        else:
            if old is not None and np.array_equal(old, cells):
                return set()
            self.chunk_cells[slot] = np.ascontiguousarray(cells, dtype=np.uint8).copy()
        bx = range(x // BRICK, (x + cs - 1) // BRICK + 1)
        by = range(y // BRICK, (y + cs - 1) // BRICK + 1)
        bz = range(z // BRICK, (z + cs - 1) // BRICK + 1)
        return {(i, j, k) for i in bx for j in by for k in bz}

    def gather_bricks(self, coords):
        """coords: iterable of brick coords -> (int32[m,3], u8[m,512]) in the device's brick layout.  A brick may
        straddle chunks when chunk_size is not a multiple of 8."""
        coords = sorted(coords)
        m, cs = len(coords), self.cs
        out = np.zeros((m, BRICK, BRICK, BRICK), np.uint8)
        for i, b in enumerate(coords):
            lo = [b[a] * BRICK for a in range(3)]
            hi = [min(lo[a] + BRICK, int(self.dims[a])) for a in range(3)]
            for cx in range(lo[0] // cs, (hi[0] - 1) // cs + 1):
                for cy in range(lo[1] // cs, (hi[1] - 1) // cs + 1):
                    for cz in range(lo[2] // cs, (hi[2] - 1) // cs + 1):
                        blk = self.chunk_cells.get((cx, cy, cz))
                        if blk is None:
                            continue
                        c0 = (cx * cs, cy * cs, cz * cs)
                        s0 = [max(lo[a], c0[a]) for a in range(3)]
                        s1 = [min(hi[a], c0[a] + cs) for a in range(3)]
                        out[i, s0[0] - lo[0]:s1[0] - lo[0], s0[1] - lo[1]:s1[1] - lo[1], s0[2] - lo[2]:s1[2] - lo[2]] = \
                            blk[s0[0] - c0[0]:s1[0] - c0[0], s0[1] - c0[1]:s1[1] - c0[1], s0[2] - c0[2]:s1[2] - c0[2]]
        return np.asarray(coords, dtype=np.int32).reshape(m, 3), out.reshape(m, 512)

    def dense(self):
        """The whole box as one [nx, ny, nz] u8 array (tests)."""
        out = np.zeros(tuple(int(d) for d in self.dims), np.uint8)
        cs = self.cs
        for (cx, cy, cz), blk in self.chunk_cells.items():
            out[cx * cs:(cx + 1) * cs, cy * cs:(cy + 1) * cs, cz * cs:(cz + 1) * cs] = blk
        return out


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
