"""GPU tests of the scene-building kernels (csrc/vt_scene.cuh) on their own: the dense -> brick pack pass against its
inverse, the device-side slot scan, chunk presence, in-place brick updates after a dense upload, and the 8-cells-per-
thread world builder against a per-cell numpy restatement of init.py:413-424 / data.py:338-371 -- at sizes that hit
every edge of the kernels (nz not a multiple of 8, fewer / more than 32 bricks along z, ragged last task, unaligned
grid pointer, empty world)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _bg(ray, settings):
    raise AssertionError("never called")


def _camera(cs):
    from python_raytracer_b200 import GpuCamera, Settings
    return GpuCamera(Settings.from_object(dict(width=8, height=8, chunk_size=cs, threads=1)), background_function=_bg, background=_bg)


def _random_world(dims, seed, fill=0.02, holes=True):
    rng = np.random.default_rng(seed)
    cells = (rng.random(dims) < fill).astype(np.uint8) * rng.integers(1, 200, dims, dtype=np.uint8)
    if holes:                                            # whole empty bricks and slabs
        cells[: dims[0] // 2, :, dims[2] // 3: 2 * dims[2] // 3] = 0
        cells[:, : dims[1] // 2, : dims[2] // 5] = 0
    return cells


def _brick_count(cells):
    nx, ny, nz = cells.shape
    pad = np.zeros((-(-nx // 8) * 8, -(-ny // 8) * 8, -(-nz // 8) * 8), np.uint8)
    pad[:nx, :ny, :nz] = cells
    b = pad.reshape(pad.shape[0] // 8, 8, pad.shape[1] // 8, 8, pad.shape[2] // 8, 8).max(axis=(1, 3, 5))
    return int((b > 0).sum())


MATS = np.zeros((256, 7)); MATS[1:, 4] = 1.0
CASES = [(16, (32, 16, 48)), (8, (8, 16, 264)), (12, (36, 24, 60)), (16, (16, 16, 1024)), (8, (24, 8, 8)), (4, (12, 20, 36)),
         (16, (48, 32, 512))]


@pytest.mark.parametrize("cs,dims", CASES)
def test_pack_roundtrip_count_and_presence(cs, dims):
    cam = _camera(cs)
    cells = _random_world(dims, seed=sum(dims) + cs)
    res = np.ones(tuple(d // cs for d in dims), np.uint8)
    origin = (-cs, 0, 2 * cs)
    cam.set_world_dense(origin, cells, res, MATS)
    assert np.array_equal(cam.download_dense().cpu().numpy(), cells)
    assert cam.world_stats()["bricks"] == _brick_count(cells)
    if cs % 8 == 0:
        exp = cells.reshape(dims[0] // cs, cs, dims[1] // cs, cs, dims[2] // cs, cs).max(axis=(1, 3, 5)) > 0
        assert np.array_equal(cam.chunk_presence(), exp)
    # a second upload into the same context replaces the first (table entries of bricks that became empty are cleared)
    cells2 = _random_world(dims, seed=7, fill=0.001, holes=False)
    cam.set_world_dense(origin, cells2, res, MATS)
    assert np.array_equal(cam.download_dense().cpu().numpy(), cells2)
    assert cam.world_stats()["bricks"] == _brick_count(cells2)
    # a denser world into the existing pool: the pack pass is launched before the count is known and repeated when
    # the pool turns out to be too small (the 48 x 32 x 512 case: 1 536 bricks into a pool of 1 024)
    cells3 = _random_world(dims, seed=9, fill=0.5, holes=False)
    cam.set_world_dense(origin, cells3, res, MATS)
    assert np.array_equal(cam.download_dense().cpu().numpy(), cells3)
    assert cam.world_stats()["bricks"] == _brick_count(cells3)
    cam.close()


def test_pack_empty_and_full_worlds():
    cam = _camera(16)
    res = np.ones((1, 2, 3), np.uint8)
    for cells in (np.zeros((16, 32, 48), np.uint8), np.full((16, 32, 48), 9, np.uint8)):
        cam.set_world_dense((0, 0, 0), cells, res, MATS)
        assert np.array_equal(cam.download_dense().cpu().numpy(), cells)
        assert cam.world_stats()["bricks"] == _brick_count(cells)
    cam.close()


def test_pack_unaligned_grid_pointer():
    """A grid whose device pointer is not 8-byte aligned takes the byte-wise loads and gives the same bricks."""
    cam = _camera(16)
    dims = (16, 32, 64)
    cells = _random_world(dims, seed=3)
    n = int(np.prod(dims))
    big = torch.zeros(n + 16, dtype=torch.uint8, device="cuda")
    view = big[3:3 + n].view(dims)
    view.copy_(torch.from_numpy(cells))
    assert view.data_ptr() % 8 != 0 and view.is_contiguous()
    cam.set_world_dense((0, 0, 0), view, np.ones((1, 2, 4), np.uint8), MATS)
    assert np.array_equal(cam.download_dense().cpu().numpy(), cells)
    cam.close()


def test_brick_updates_after_dense_upload():
    """vt_upload_bricks after vt_upload_dense: the host mirror of the table is fetched from the device first, so slots
    are reused, released and extended consistently."""
    cam = _camera(16)
    dims = (32, 32, 64)
    cells = _random_world(dims, seed=11)
    cam.set_world_dense((0, 0, 0), cells, np.ones((2, 2, 4), np.uint8), MATS)
    n0 = cam.world_stats()["bricks"]
    rng = np.random.default_rng(5)
    occupied = [(x, y, z) for x in range(4) for y in range(4) for z in range(8) if cells[8 * x:8 * x + 8, 8 * y:8 * y + 8, 8 * z:8 * z + 8].any()]
    empty = [(x, y, z) for x in range(4) for y in range(4) for z in range(8) if not cells[8 * x:8 * x + 8, 8 * y:8 * y + 8, 8 * z:8 * z + 8].any()]
    assert len(occupied) >= 3 and len(empty) >= 3
    coords, payload = [], []
    exp = cells.copy()
    for k, (x, y, z) in enumerate(occupied[:2]):         # overwrite one, clear one
        blk = rng.integers(1, 9, (8, 8, 8), dtype=np.uint8) if k == 0 else np.zeros((8, 8, 8), np.uint8)
        coords.append((x, y, z)); payload.append(blk)
        exp[8 * x:8 * x + 8, 8 * y:8 * y + 8, 8 * z:8 * z + 8] = blk
    for (x, y, z) in empty[:2]:                          # two new bricks: one takes the released slot, one extends the pool
        blk = rng.integers(1, 9, (8, 8, 8), dtype=np.uint8)
        coords.append((x, y, z)); payload.append(blk)
        exp[8 * x:8 * x + 8, 8 * y:8 * y + 8, 8 * z:8 * z + 8] = blk
    cc = np.ascontiguousarray(coords, np.int32)
    pp = np.ascontiguousarray(np.stack(payload).reshape(len(payload), 512), np.uint8)
    cam._check(cam._lib.vt_upload_bricks(cam._ctx, cc.ctypes.data, pp.ctypes.data, len(cc)))
    assert np.array_equal(cam.download_dense().cpu().numpy(), exp)
    assert cam.world_stats()["bricks"] == n0 + 1 == _brick_count(exp)
    cam.close()


def _rot(size, angle, x, y, z):
    """Sprite.pos_rotated (data.py:338-371), per cell."""
    ex, ey, ez = size[0] - 1, size[1] - 1, size[2] - 1
    ax, ay, az = angle
    if ax and size[1] == size[2]:
        y, z = {1: (ez - z, y), 2: (ey - y, ez - z), 3: (z, ey - y)}[ax]
    if ay and size[0] == size[2]:
        x, z = {1: (z, ex - x), 2: (ex - x, ez - z), 3: (ez - z, x)}[ay]
    if az and size[0] == size[1]:
        x, y = {1: (ey - y, x), 2: (ex - x, ey - y), 3: (y, ex - x)}[az]
    return x, y, z


@pytest.mark.parametrize("cs", [16, 12])
def test_build_cells_matches_per_cell_restatement(cs):
    """Overlapping objects with rotations and fractional positions: the run-of-8 kernel equals the per-cell rule
    (last object with a voxel wins; cells mins <= p < maxs; rotated lookup), for an aligned and an unaligned nz."""
    from python_raytracer_b200.builder import SceneObject, WorldBuilder
    rng = np.random.default_rng(cs)
    objs = []
    for k, (size, pos, rot) in enumerate([((6, 6, 6), (0.0, 0.0, 0.0), (90.0, 0.0, 0.0)), ((6, 6, 6), (2.5, 1.0, -3.0), (0.0, 180.0, 270.0)),
                                          ((4, 10, 4), (-3.0, 2.0, 5.5), (0.0, 90.0, 0.0)), ((9, 3, 5), (7.0, -4.0, 9.0), (0.0, 0.0, 0.0)),
                                          ((2, 2, 21), (1.0, 1.0, 1.0), (0.0, 0.0, 90.0))]):
        spr = (rng.random(size) < 0.6).astype(np.uint8) * np.uint8(k + 1)
        objs.append(SceneObject(sprite=spr, pos=pos, rot=rot))
    cam = _camera(cs)
    wb = WorldBuilder(cam)
    cells = wb.build(objs, MATS[:8], margin=1).cpu().numpy()
    exp = np.zeros_like(cells)
    for o in objs:
        mins, maxs = o.box()
        ang = o.angles()
        for x in range(mins[0], maxs[0]):
            for y in range(mins[1], maxs[1]):
                for z in range(mins[2], maxs[2]):
                    lx, ly, lz = _rot(o.size, ang, x - mins[0], y - mins[1], z - mins[2])
                    if 0 <= lx < o.size[0] and 0 <= ly < o.size[1] and 0 <= lz < o.size[2] and o.sprite[lx, ly, lz]:
                        exp[x - wb.origin[0], y - wb.origin[1], z - wb.origin[2]] = o.sprite[lx, ly, lz]
    assert np.array_equal(cells, exp)
    assert np.array_equal(cam.download_dense().cpu().numpy(), exp)
    # one object moves: the region update redraws its old and new chunks only and equals a fresh build
    objs[1].pos = (4.5, 1.0, -1.0)
    wb.update(objs, [objs[1]])
    cam2 = _camera(cs)
    wb2 = WorldBuilder(cam2)
    order = [objs[0], objs[2], objs[3], objs[4], objs[1]]      # a redrawn object moves to the end of the override order
    fresh = wb2.build(order, MATS[:8]).cpu().numpy()
    off = [int(a - b) for a, b in zip(wb2.origin, wb.origin)]
    got = wb.cells.cpu().numpy()
    assert min(off) >= 0
    sub = got[off[0]:off[0] + fresh.shape[0], off[1]:off[1] + fresh.shape[1], off[2]:off[2] + fresh.shape[2]]
    assert sub.shape == fresh.shape
    assert np.array_equal(sub, fresh)
    assert int((got > 0).sum()) == int((fresh > 0).sum())
    assert np.array_equal(cam.download_dense().cpu().numpy(), got)
    cam.close(); cam2.close()


@pytest.mark.parametrize("rows", [4, 2, 0])
def test_wide_build_kernel_matches_restatement(rows):
    """build_cells16_kernel (unrotated objects, rows that are multiples of 16 cells): 40 overlapping objects of odd sizes at
    odd and fractional positions -- partial coverage of a thread's 16 cells at every object's z faces, sprite rows that start
    at any byte, sprite tails, rows of a thread that straddle two x, later objects overriding earlier ones -- against the
    per-cell rule painted with numpy; buildrows 4 / 2 (wide) and 0 (build_cells_kernel) give the same cells."""
    from python_raytracer_b200.builder import SceneObject, WorldBuilder
    rng = np.random.default_rng(5)
    objs = []
    for k in range(40):
        size = tuple(int(v) for v in rng.integers(1, 40, 3))
        pos = tuple(float(v) for v in (rng.integers(-60, 60, 3) + rng.choice([0.0, 0.5, 0.25], 3)))
        spr = (rng.random(size) < 0.5).astype(np.uint8) * np.uint8(1 + k % 7)
        objs.append(SceneObject(sprite=spr, pos=pos, rot=(0.0, 0.0, 0.0)))
    objs.append(SceneObject(sprite=np.full((3, 3, 70), 3, np.uint8), pos=(5.0, 5.0, -2.0)))       # long along z: whole chunks
    cam = _camera(16)
    cam._check(cam._lib.vt_set_tuning(cam._ctx, b"buildrows", rows))
    wb = WorldBuilder(cam)
    cells = wb.build(objs, MATS[:8]).cpu().numpy()
    assert cells.shape[2] % 16 == 0
    exp = np.zeros_like(cells)
    for o in objs:
        mins, maxs = o.box()
        n = [maxs[a] - mins[a] for a in range(3)]
        if min(n) <= 0:
            continue
        sub = o.sprite[:n[0], :n[1], :n[2]]
        lo = [mins[a] - int(wb.origin[a]) for a in range(3)]
        view = exp[lo[0]:lo[0] + n[0], lo[1]:lo[1] + n[1], lo[2]:lo[2] + n[2]]
        view[sub > 0] = sub[sub > 0]
    assert np.array_equal(cells, exp)
    # a region update through the same dispatch: one object moves by an odd amount
    objs[7].pos = (objs[7].pos[0] + 3.0, objs[7].pos[1] - 2.5, objs[7].pos[2] + 5.0)
    wb.update(objs, [objs[7]])
    cam2 = _camera(16)
    wb2 = WorldBuilder(cam2)
    order = [o for o in objs if o is not objs[7]] + [objs[7]]
    fresh = wb2.build(order, MATS[:8]).cpu().numpy()
    got = wb.cells.cpu().numpy()
    lo = np.minimum(wb.origin, wb2.origin)
    hi = np.maximum(wb.origin + np.array(got.shape), wb2.origin + np.array(fresh.shape))
    a, b = np.zeros(tuple(hi - lo), np.uint8), np.zeros(tuple(hi - lo), np.uint8)
    o1, o2 = (wb.origin - lo).astype(int), (wb2.origin - lo).astype(int)
    a[o1[0]:o1[0] + got.shape[0], o1[1]:o1[1] + got.shape[1], o1[2]:o1[2] + got.shape[2]] = got
    b[o2[0]:o2[0] + fresh.shape[0], o2[1]:o2[1] + fresh.shape[1], o2[2]:o2[2] + fresh.shape[2]] = fresh
    assert np.array_equal(a, b)
    cam.close(); cam2.close()
