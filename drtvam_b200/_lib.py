"""ctypes binding of ``libdrtvam_b200.so`` (``include/drtvam_b200.h``).

The library is the product: there is no CPU or PyTorch fallback.  Importing
this module raises if the shared object is missing and cannot be built.
"""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

MAX_SURFACES = 8
ABI_VERSION = 13


class DrtvamSurface(C.Structure):
    _fields_ = [
        ("kind", C.c_int32), ("bsdf", C.c_int32),
        ("interior_medium", C.c_int32), ("exterior_medium", C.c_int32),
        ("radius", C.c_float), ("z0", C.c_float), ("z1", C.c_float),
        ("hx", C.c_float), ("hy", C.c_float), ("hz", C.c_float),
        ("eta", C.c_float), ("_pad", C.c_float),
    ]


class DrtvamSceneDesc(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("abi_version", C.c_uint32),
        ("proj_kind", C.c_int32), ("n_patterns", C.c_int32), ("res_x", C.c_int32), ("res_y", C.c_int32),
        ("pixel_x", C.c_float), ("pixel_y", C.c_float),
        ("aperture_radius", C.c_float), ("focus_distance", C.c_float),
        ("motion_analytic", C.c_int32), ("clockwise", C.c_int32),
        ("distance", C.c_float), ("_pad0", C.c_float),
        ("motion_table_host", C.POINTER(C.c_float)),
        ("max_depth", C.c_int32), ("rr_depth", C.c_int32), ("transmission_only", C.c_int32),
        ("regular_sampling", C.c_int32), ("sample_time", C.c_int32),
        ("print_time", C.c_float),
        ("sigma_t", C.c_float), ("albedo", C.c_float),
        ("phase_kind", C.c_int32), ("hg_g", C.c_float),
        ("n_surf", C.c_int32), ("_pad1", C.c_int32),
        ("surf", DrtvamSurface * MAX_SURFACES),
        ("bbox_min", C.c_float * 3), ("bbox_max", C.c_float * 3),
        ("res", C.c_int32 * 3),
        ("surface_aware", C.c_int32), ("sensor_kind", C.c_int32),
        ("majorant", C.c_float),
        ("motion_table_host_f64", C.POINTER(C.c_double)),
    ]


# every symbol include/drtvam_b200.h declares: (restype, argtypes)
_u64, _u32, _i32, _f32, _vp = C.c_uint64, C.c_uint32, C.c_int32, C.c_float, C.c_void_p
SYMBOLS = {
    "drtvam_scene_create": (C.c_int, [C.POINTER(DrtvamSceneDesc), C.c_int, C.POINTER(_vp)]),
    "drtvam_scene_destroy": (None, [_vp]),
    "drtvam_scene_set_depth": (C.c_int, [_vp, _i32, _i32]),
    "drtvam_scene_set_target": (C.c_int, [_vp, _vp, _u32]),
    "drtvam_scene_add_mesh_surface": (C.c_int, [_vp, _vp, _u32, C.c_int32, C.c_float, C.c_int32, C.c_int32]),
    "drtvam_compute_volume": (C.c_int, [_vp, _u32, _vp, _vp]),
    "drtvam_discretize": (C.c_int, [_vp, _vp, _vp]),
    "drtvam_scene_set_inv_volume": (C.c_int, [_vp, _vp]),
    "drtvam_scene_set_z_window": (C.c_int, [_vp, _i32, _i32, C.POINTER(_i32), C.POINTER(_i32)]),
    "drtvam_scene_set_sampler_index": (C.c_int, [_vp, _vp]),
    "drtvam_scene_pin_pixels": (C.c_int, [_vp, _vp, _u64]),
    "drtvam_comm_unique_id": (C.c_int, [_vp]),
    "drtvam_comm_init": (C.c_int, [_vp, _vp, _i32, _i32]),
    "drtvam_comm_destroy": (C.c_int, [_vp]),
    "drtvam_allreduce_dose": (C.c_int, [_vp, _vp, _u64, _vp]),
    "drtvam_allreduce_scalars": (C.c_int, [_vp, _vp, _u64, _vp]),
    "drtvam_scene_plan": (C.c_int, [_vp, _u64, _u32]),
    "drtvam_scene_force_general": (C.c_int, [_vp, _i32]),
    "drtvam_scene_stat": (C.c_int, [_vp, _i32, C.POINTER(_u64)]),
    "drtvam_forward": (C.c_int, [_vp, _vp, _vp, _u64, _u32, _u32, _f32, _vp, _i32, _vp]),
    "drtvam_forward_noalloc": (C.c_int, [_vp, _vp, _vp, _u64, _u32, _u32, _f32, _vp, _i32, _vp]),
    "drtvam_adjoint": (C.c_int, [_vp, _vp, _vp, _u64, _u32, _u32, _f32, _vp, _vp]),
    "drtvam_raygen": (C.c_int, [_vp, _vp, _u64, _u32, _u32, _vp, _vp]),
    "drtvam_radon": (C.c_int, [_vp, _vp, _u64, _u32, _u32, _vp, _vp]),
    "drtvam_corner_filter": (C.c_int, [_vp, _vp, _u64, _u32, _u32, _f32, _f32, _vp, _vp]),
    "drtvam_count_visits": (C.c_int, [_vp, _vp, _u64, _u32, _u32, C.POINTER(_u64)]),
    "drtvam_forward_host": (C.c_int, [_vp, _vp, _vp, _u64, _u32, _u32, _f32, _vp]),
    "drtvam_adjoint_host": (C.c_int, [_vp, _vp, _vp, _u64, _u32, _u32, _f32, _vp]),
    "drtvam_thresholded_loss": (C.c_int, [C.c_int, _vp, _vp, _f32, _vp, _u64, _f32, _f32, _f32, _f32, _f32, _i32, _f32,
                                          _vp, _vp, _vp]),
    "drtvam_thresholded_loss_sa": (C.c_int, [C.c_int, _vp, _vp, _f32, _vp, _u64, _f32, _f32, _f32, _f32, _f32, _i32, _f32,
                                             _vp, _vp, _vp]),
    "drtvam_lbfgs_history": (C.c_int, [C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _u64, _i32, _vp]),
    "drtvam_lbfgs_direction": (C.c_int, [C.c_int, _vp, _i32, _u64, _vp, _i32, _vp, _vp]),
    "drtvam_lbfgs_combine": (C.c_int, [C.c_int, _vp, _i32, _vp, _u64, _vp, _vp]),
    "drtvam_armijo_losses": (C.c_int, [C.c_int, _vp, _vp, _vp, _u64, _f32, _f32, _f32, _f32, _f32, _i32, _i32, _i32, _vp,
                                       _vp, _vp]),
    "drtvam_armijo_select": (C.c_int, [C.c_int, _vp, _i32, _i32, C.c_double, _vp, _vp, _vp, C.c_double, _vp, _vp]),
    "drtvam_lbfgs_update": (C.c_int, [C.c_int, _vp, _vp, _vp, _u64, _vp, _vp]),
    "drtvam_launch_count": (_u64, []),
    "drtvam_last_error": (C.c_char_p, []),
    "drtvam_abi_version": (C.c_int, []),
}

_lib = None


def lib_path() -> str:
    return _build.LIB


def load():
    """Load (building first if the .so is absent and nvcc is present)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        try:
            _build.build()
        except Exception as exc:  # no toolchain: fail loudly, never fall back
            raise ImportError(
                f"drtvam_b200: CUDA library {path} is missing and could not be built ({exc}). "
                "Run `python -m drtvam_b200.build`.") from exc
    lib = C.CDLL(path)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the export is missing
        fn.restype = res
        fn.argtypes = args
    if lib.drtvam_abi_version() != ABI_VERSION:
        raise ImportError(f"drtvam_b200: ABI version mismatch ({lib.drtvam_abi_version()} != {ABI_VERSION}); rebuild")
    _lib = lib
    return lib


class DrtvamError(RuntimeError):
    pass


def check(rc: int):
    if rc != 0:
        msg = load().drtvam_last_error()
        raise DrtvamError((msg or b"unknown error").decode())
