"""ctypes binding of include/voxtrace.h (libvoxtrace.so).  Fails loudly when the library is missing."""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

SETTINGS_INT = ("width", "height", "samples", "static_noise", "chunk_size", "chunk_radius", "background", "reserved")
SETTINGS_F64 = ("shutter", "falloff", "dof", "dist_min", "dist_max", "max_light", "max_bounces",
                "lod_bounces", "lod_samples", "lod_random", "lod_edge", "proportions")
RAY_RECORD = 16
STATIC_SLICE_MAX_CELLS = 65536
TRACE_RECORD = 20

EXPORTS = ("vt_create", "vt_destroy", "vt_last_error", "vt_abi_version", "vt_set_stream", "vt_sync",
           "vt_set_settings", "vt_set_materials", "vt_set_camera", "vt_set_world", "vt_set_chunk_table",
           "vt_upload_bricks", "vt_upload_dense", "vt_chunk_presence", "vt_download_dense", "vt_world_stats", "vt_render", "vt_region_pixels",
           "vt_launch_count", "vt_build_flags", "vt_build_world", "vt_update_region", "vt_goxel_parse", "vt_free",
           "vt_shared_alloc", "vt_shared_free", "vt_ipc_export", "vt_ipc_open", "vt_ipc_close", "vt_set_frame_target",
           "vt_stream_wait_flags", "vt_stream_signal_flags", "vt_stream_signal_flags_on", "vt_wait_status", "vt_copy_to_host", "vt_fx_compose", "vt_fx_draw",
           "vt_wait_status_peek", "vt_pow_host", "vt_selftest_pow", "vt_set_phys_materials", "vt_slice_collide", "vt_slice_collide_static", "vt_set_tuning", "vt_clear_world", "vt_trace_rays", "vt_set_visited_peer")


class VtSettings(C.Structure):
    _fields_ = [(n, C.c_int32) for n in SETTINGS_INT] + [(n, C.c_double) for n in SETTINGS_F64]


class VtMaterial(C.Structure):
    _fields_ = [("albedo", C.c_double * 3), ("roughness", C.c_double), ("absorption", C.c_double),
                ("ior", C.c_double), ("energy", C.c_double), ("reserved", C.c_double)]


class VtRegion(C.Structure):
    _fields_ = [("x0", C.c_int32), ("y0", C.c_int32), ("x1", C.c_int32), ("y1", C.c_int32),
                ("stripe_rows", C.c_int32), ("stripe_count", C.c_int32), ("stripe_index", C.c_int32),
                ("reserved", C.c_int32), ("dev_pixels", C.c_void_p), ("n_pixels", C.c_int64)]


class VtObject(C.Structure):
    _fields_ = [("dev_sprite", C.c_void_p), ("size", C.c_int32 * 3), ("mins", C.c_int32 * 3), ("maxs", C.c_int32 * 3),
                ("angle", C.c_int32 * 3), ("reserved", C.c_int32)]


class VtPhysObject(C.Structure):
    _fields_ = [("dev_grid", C.c_void_p), ("grid_lo", C.c_int32 * 3), ("grid_dims", C.c_int32 * 3), ("size", C.c_int32 * 3),
                ("mins", C.c_int32 * 3), ("angle", C.c_int32 * 3), ("reserved", C.c_int32)]


class VtFx(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("out_width", C.c_int32), ("out_height", C.c_int32),
                ("iris_on", C.c_int32), ("reserved", C.c_int32), ("spill", C.c_double), ("iris", C.c_double),
                ("bloom", C.c_double), ("bloom_blur", C.c_double), ("subsamples", C.c_double), ("smooth", C.c_double)]


class VoxtraceError(RuntimeError):
    pass


_lib = None


def load(path: str | None = None):
    """Loads libvoxtrace.so (building it with nvcc if the sources are newer).  No fallback of any kind."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or os.environ.get("VT_LIB") or None       # VT_LIB: a tuning / checked build (tools/build_variants.py)
    p = path or _build.LIB_PATH
    if path is None and _build.needs_build():
        _build.build_library()
    if not os.path.exists(p):
        raise VoxtraceError("libvoxtrace.so is missing at %s; build it with python -m python_raytracer_b200.build "
                            "(the GPU path has no CPU fallback)" % p)
    lib = C.CDLL(p)
    vp, i32, i64, u64 = C.c_void_p, C.c_int, C.c_int64, C.c_uint64
    sig = {
        "vt_create": (C.c_int, [i32, C.POINTER(vp)]),
        "vt_destroy": (None, [vp]),
        "vt_last_error": (C.c_char_p, [vp]),
        "vt_abi_version": (C.c_int, []),
        "vt_build_flags": (C.c_int, []),
        "vt_set_stream": (C.c_int, [vp, vp]),
        "vt_sync": (C.c_int, [vp]),
        "vt_set_settings": (C.c_int, [vp, C.POINTER(VtSettings)]),
        "vt_set_materials": (C.c_int, [vp, vp, i32]),
        "vt_set_camera": (C.c_int, [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_double]),
        "vt_set_world": (C.c_int, [vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
        "vt_set_chunk_table": (C.c_int, [vp, vp, i64]),
        "vt_upload_bricks": (C.c_int, [vp, vp, vp, i32]),
        "vt_upload_dense": (C.c_int, [vp, vp]),
        "vt_chunk_presence": (C.c_int, [vp, vp, i64]),
        "vt_download_dense": (C.c_int, [vp, vp]),
        "vt_world_stats": (C.c_int, [vp, C.POINTER(i64), C.POINTER(i64)]),
        "vt_render": (C.c_int, [vp, C.POINTER(VtRegion), vp, vp, vp, vp, vp, u64]),
        "vt_region_pixels": (i64, [C.POINTER(VtRegion)]),
        "vt_launch_count": (i64, [vp]),
        "vt_build_world": (C.c_int, [vp, C.POINTER(VtObject), i32, vp]),
        "vt_update_region": (C.c_int, [vp, C.POINTER(VtObject), i32, vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
        "vt_goxel_parse": (C.c_int, [C.c_char_p, C.c_int32, C.POINTER(C.c_char_p), i32, C.POINTER(C.POINTER(C.c_int32)),
                                     C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(i64), C.c_char_p, i32]),
        "vt_free": (None, [vp]),
        "vt_shared_alloc": (C.c_int, [vp, i64, C.POINTER(vp)]),
        "vt_shared_free": (C.c_int, [vp, vp]),
        "vt_ipc_export": (C.c_int, [vp, vp, C.c_char_p]),
        "vt_ipc_open": (C.c_int, [vp, C.c_char_p, C.POINTER(vp)]),
        "vt_ipc_close": (C.c_int, [vp, vp]),
        "vt_set_frame_target": (C.c_int, [vp, vp, vp, C.c_uint32]),
        "vt_stream_wait_flags": (C.c_int, [vp, vp, i32, C.c_uint32, C.c_double]),
        "vt_stream_signal_flags": (C.c_int, [vp, C.POINTER(vp), i32, C.c_uint32]),
        "vt_stream_signal_flags_on": (C.c_int, [vp, vp, C.POINTER(vp), i32, C.c_uint32]),
        "vt_wait_status": (C.c_int, [vp, C.POINTER(C.c_int)]),
        "vt_wait_status_peek": (C.c_int, [vp, C.POINTER(C.c_int)]),
        "vt_copy_to_host": (C.c_int, [vp, vp, vp, i64]),
        "vt_pow_host": (C.c_double, [C.c_double, C.c_double]),
        "vt_selftest_pow": (C.c_int, [vp, vp, vp, i64, vp]),
        "vt_set_phys_materials": (C.c_int, [vp, vp, i32]),
        "vt_slice_collide": (C.c_int, [vp, C.POINTER(VtPhysObject), C.POINTER(VtPhysObject), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                       C.POINTER(C.c_int32), C.c_double, u64, C.POINTER(C.c_double), C.POINTER(C.c_int32)]),
        "vt_slice_collide_static": (C.c_int, [vp, C.POINTER(VtPhysObject), C.POINTER(VtPhysObject), C.POINTER(C.c_int32),
                                              C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_double, u64, vp, vp]),
        "vt_set_tuning": (C.c_int, [vp, C.c_char_p, i32]),
        "vt_clear_world": (C.c_int, [vp]),
        "vt_trace_rays": (C.c_int, [vp, vp, vp, i64, vp, vp, vp, u64]),
        "vt_set_visited_peer": (C.c_int, [vp, vp, i64]),
        "vt_fx_compose": (C.c_int, [vp, vp, vp, i64]),
        "vt_fx_draw": (C.c_int, [vp, C.POINTER(VtFx), vp, vp, C.POINTER(C.c_int32)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    if lib.vt_abi_version() != 1:
        raise VoxtraceError("libvoxtrace.so ABI version mismatch")
    if path is None:
        _lib = lib
    return lib
