"""Goxel text models (reference Sprite.load, data.py:271-289) through the C++ loader in libvoxtrace.so.

Host-only: needs the library but no GPU.  ``load_goxel`` returns the voxel list in file order; ``sprite_grid`` turns
it into the dense [sx, sy, sz] material-id grid the device world builder (``vt_build_world``) consumes -- the cells
``Sprite.get_voxel`` can ever return for positions inside the sprite box (data.py:417-419).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import abi
from .abi import VoxtraceError


def load_goxel(path: str, size_x: int, keys):
    """keys: sequence of colour strings as written in the file ("7f7f7f"); returns (xyz int32 [n,3], key index + 1
    u8 [n]) with positions already remapped to (size_x - x, z, y).  Later rows override earlier ones."""
    lib = abi.load()
    arr = (C.c_char_p * max(1, len(keys)))(*[k.encode() for k in keys])
    xyz, mat, n = C.POINTER(C.c_int32)(), C.POINTER(C.c_uint8)(), C.c_int64()
    err = C.create_string_buffer(512)
    rc = lib.vt_goxel_parse(str(path).encode(), int(size_x), arr, len(keys), C.byref(xyz), C.byref(mat), C.byref(n), err, 512)
    if rc != 0:
        raise VoxtraceError(err.value.decode(errors="replace"))
    try:
        if n.value == 0:
            return np.zeros((0, 3), np.int32), np.zeros((0,), np.uint8)
        a = np.ctypeslib.as_array(xyz, shape=(n.value, 3)).copy()
        m = np.ctypeslib.as_array(mat, shape=(n.value,)).copy()
    finally:
        lib.vt_free(xyz)
        lib.vt_free(mat)
    return a, m


def sprite_grid(xyz, mat, size, ids=None):
    """Voxel list -> dense u8 [sx, sy, sz].  ``ids`` maps key index + 1 to a material id (default: identity).
    Positions outside the sprite box are never drawn by chunk_update (init.py:417-421) and are dropped; of several
    rows at one position the last wins (dict semantics of data.py:286)."""
    sx, sy, sz = (int(v) for v in size)
    grid = np.zeros((sx, sy, sz), np.uint8)
    xyz = np.asarray(xyz, np.int64).reshape(-1, 3)
    m = np.asarray(mat, np.uint8)
    if ids is not None:
        m = np.asarray(ids, np.uint8)[m]
    ok = (xyz[:, 0] >= 0) & (xyz[:, 0] < sx) & (xyz[:, 1] >= 0) & (xyz[:, 1] < sy) & (xyz[:, 2] >= 0) & (xyz[:, 2] < sz)
    xyz, m = xyz[ok], m[ok]
    lin = (xyz[:, 0] * sy + xyz[:, 1]) * sz + xyz[:, 2]
    # last occurrence of every cell: unique() on the reversed list returns first occurrences there
    _, first_rev = np.unique(lin[::-1], return_index=True)
    keep = len(lin) - 1 - first_rev
    grid.reshape(-1)[lin[keep]] = m[keep]
    return grid


def write_goxel(path: str, xyz_file, colours):
    """Writes file-space voxels (x, y, z as Goxel exports them) with a Goxel-style header; used by the synthetic
    large-model workload.  colours: array of 'rrggbb' strings, one per voxel."""
    import gzip
    xyz_file = np.asarray(xyz_file, np.int64).reshape(-1, 3)
    opener = gzip.open if str(path).endswith(".gz") else open
    cols = np.asarray(colours)
    with opener(path, "wt") as f:
        f.write("# Goxel 0.14.0\n# One line per voxel\n# X Y Z RRGGBB\n")
        step = 1 << 20
        for i in range(0, len(xyz_file), step):
            blk = xyz_file[i:i + step]
            c = cols[i:i + step]
            f.write("\n".join("%d %d %d %s" % (a, b, d, e) for (a, b, d), e in zip(blk.tolist(), c.tolist())))
            f.write("\n")
