"""GPU test of vt_slice_collide_static, the allocation-free variant of vt_slice_collide (include/voxtrace.h; no counterpart in
the reference).  Its per-cell scratch is a file-scope `static __device__` array
of the library instead of cudaMalloc'ed buffers; the caller passes the material table and the in / out sums as DEVICE memory.
Checked through the C ABI against the existing entry point, bit for bit (the tolerance the physics tests use: none), on two
calls in a row -- a large slice, then a small one that would see the first call's leftovers if the scratch were misused --
and again in the other order.  This file is named to be collected last."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _post6(obj, d):
    mins, maxs = obj.box()
    k = [i for i in range(3) if d[i]][0]
    p = list(mins) + list(maxs)
    if d[k] < 0:
        p[k], p[3 + k] = mins[k] - 1, mins[k]
    else:
        p[k], p[3 + k] = maxs[k], maxs[k] + 1
    return p


def test_static_scratch_slice_test_equals_vt_slice_collide():
    from python_raytracer_b200 import GpuCamera, Settings, abi
    from python_raytracer_b200.physics import PhysObject, PhysicsWorld
    rng = np.random.default_rng(21)
    props = np.array([[0, 0, 0], [1, 0.5, 0.25], [1, 0.125, 0.75], [0, 1.0, 1.0], [1, 0.3, 0.1]], np.float64)   # id 3: solidity 0
    cam = GpuCamera(Settings.from_object(dict(width=8, height=8, chunk_size=16)), background=None)
    world = PhysicsWorld(cam, props, gravity=1.0, friction=0.7, friction_air=0.1, max_velocity=10.0, min_velocity=0.01)
    lib, ctx = cam._lib, cam._ctx

    def grid(shape, fill):
        return ((rng.random(shape) < fill) * rng.integers(1, 5, shape)).astype(np.uint8)

    big_a = PhysObject(sprite=grid((96, 40, 120), 0.5), pos=(1.0, 30.0, -2.0), rot=(0.0, 0.0, 0.0), physics=True)
    big_b = PhysObject(sprite=grid((128, 32, 128), 0.5), pos=(0.0, 0.0, 0.0), rot=(0.0, 180.0, 0.0))
    small_a = PhysObject(sprite=grid((12, 12, 12), 0.6), pos=(0.5, 9.0, 0.0), rot=(0.0, 90.0, 0.0), physics=True)
    small_b = PhysObject(sprite=grid((16, 8, 16), 0.6), pos=(0.0, 0.0, 0.0), rot=(0.0, 0.0, 0.0))
    cases = [(big_a, big_b, (0, -1, 0), 0.25, 1.5), (small_a, small_b, (0, -1, 0), 0.0, 0.0),
             (small_a, small_b, (1, 0, 0), 3.0, 0.5), (big_a, big_b, (0, -1, 0), 0.25, 1.5)]
    assert (96 + 1) * 2 * (120 + 1) <= abi.STATIC_SLICE_MAX_CELLS
    # the caller's inputs and outputs, in device memory (what vt_set_phys_materials / the host copies do in the other entry point)
    mats = torch.zeros((len(props), 4), dtype=torch.float64, device=cam.device)
    mats[:, :3] = torch.from_numpy(props).to(cam.device)
    inout = torch.zeros(3, dtype=torch.float64, device=cam.device)
    blocked_seen = 0
    for obj, other, d, fr0, el0 in cases:                       # calls in a row on the same scratch
        p6 = _post6(obj, d)
        want = world.slice_collide(obj, other, p6, d, fr0, el0)                     # existing entry point (cudaMalloc'ed scratch)
        a, b = world._vt_object(obj), world._vt_object(other)
        inout.copy_(torch.tensor([fr0, el0, -1.0], dtype=torch.float64))
        torch.cuda.synchronize()
        lo, hi = (C.c_int32 * 3)(*p6[:3]), (C.c_int32 * 3)(*p6[3:])
        dd = (C.c_int32 * 3)(*d)
        rc = lib.vt_slice_collide_static(ctx, C.byref(a), C.byref(b), lo, hi, dd, float(world.friction), world.seed,
                                         mats.data_ptr(), inout.data_ptr())
        assert rc == 0, lib.vt_last_error(ctx).decode()
        cam._check(lib.vt_sync(ctx))
        got = inout.cpu().numpy()
        assert (float(got[0]), float(got[1]), bool(got[2] != 0.0)) == (float(want[0]), float(want[1]), bool(want[2])), (d, got, want)
        assert got[2] in (0.0, 1.0)
        blocked_seen += bool(want[2])
    assert blocked_seen >= 2                                    # the scenario has contacts: the sums are not trivially equal
    # a slice beyond the array's size is rejected with a clear error, and nothing is launched
    huge = PhysObject(sprite=grid((200, 8, 200), 0.3), pos=(0.0, 20.0, 0.0), physics=True)
    p6 = _post6(huge, (0, -1, 0))
    a, b = world._vt_object(huge), world._vt_object(big_b)
    before = cam.launch_count
    rc = lib.vt_slice_collide_static(ctx, C.byref(a), C.byref(b), (C.c_int32 * 3)(*p6[:3]), (C.c_int32 * 3)(*p6[3:]),
                                     (C.c_int32 * 3)(0, -1, 0), 1.0, 0, mats.data_ptr(), inout.data_ptr())
    assert rc != 0 and b"VT_STATIC_SLICE_MAX_CELLS" in lib.vt_last_error(ctx) and cam.launch_count == before
    cam.close()
