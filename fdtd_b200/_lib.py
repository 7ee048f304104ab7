"""ctypes binding of include/fdtd_b200.h.  Loads the in-tree CUDA library; there is no fallback:
a missing library is an ImportError-grade failure, a missing GPU makes `create` fail (ECUDA)."""
from __future__ import annotations

import ctypes as C
import os

from .build import LIB_PATH

ABI_VERSION = 2
OK, EINVAL, ENOMEM, ECUDA, ESTATE = 0, 1, 2, 3, 4
FIELD_E, FIELD_H = 0, 1
EXACT_ARITHMETIC = 1 << 0
FORCE_GENERIC_KERNELS = 1 << 1
FUSED_STEP = 1 << 2        # ABI 1 name, ignored: the fused step is the default
IN_PLACE_STEP = 1 << 3     # opt out of the fused step (two half-step kernels, fields updated in place)
COLORMAP_CPU, COLORMAP_LINEAR, COLORMAP_APP_Z = 0, 1, 2
(HALO_SEND_EX_TOP, HALO_SEND_EY_TOP, HALO_RECV_EX_LOW, HALO_RECV_EY_LOW,
 HALO_SEND_HX_BOTTOM, HALO_SEND_HY_BOTTOM, HALO_RECV_HX_HIGH, HALO_RECV_HY_HIGH,
 HALO_SEND_EX_BOTTOM, HALO_SEND_EY_BOTTOM, HALO_SEND_EZ_BOTTOM, HALO_RECV_EX_HIGH, HALO_RECV_EY_HIGH, HALO_RECV_EZ_HIGH,
 HALO_SEND_MAT_BOTTOM, HALO_RECV_MAT_HIGH) = range(16)
HALO_PENDING = 0x100
P2P_BLOB_BYTES = 2048
STATIC_SCRATCH_BYTES = 8 << 20   # fdtd_b200_read_field_static: at most STATIC_SCRATCH_BYTES // 12 cells per call


class Material(C.Structure):
    _fields_ = [("relative_permeability", C.c_double), ("magnetic_conductivity", C.c_double),
                ("relative_permittivity", C.c_double), ("eletrical_conductivity", C.c_double)]


class PhysicalConstants(C.Structure):
    _fields_ = [("vacuum_permittivity", C.c_double), ("vacuum_permeability", C.c_double)]


class UpdateCoefficients(C.Structure):
    _fields_ = [("c_a", C.c_double), ("c_b", C.c_double), ("d_a", C.c_double), ("d_b", C.c_double)]


class CoefficientsF32(C.Structure):
    _fields_ = [("c_a", C.c_float), ("c_b", C.c_float), ("d_a", C.c_float), ("d_b", C.c_float)]


class Config(C.Structure):
    _fields_ = [("nx", C.c_uint64), ("ny", C.c_uint64), ("nz", C.c_uint64), ("dx", C.c_double * 3), ("dt", C.c_double),
                ("device", C.c_int32), ("flags", C.c_uint32), ("z_begin", C.c_uint64), ("z_end", C.c_uint64),
                ("probe_ring_steps", C.c_uint32), ("reserved", C.c_uint32)]


class GradedPml(C.Structure):
    _fields_ = [("m", C.c_double), ("m_a", C.c_double), ("sigma_max", C.c_double), ("kappa_max", C.c_double), ("a_max", C.c_double)]


class Source(C.Structure):
    _fields_ = [("cell", C.c_uint64), ("j", C.c_float * 3), ("m", C.c_float * 3)]


class Shape(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("label", C.c_uint32), ("params", C.c_float * 3), ("rot", C.c_float * 9),
                ("translation", C.c_float * 3), ("reserved", C.c_float)]


SHAPE_BALL, SHAPE_CUBOID, SHAPE_CYLINDER = 0, 1, 2

# every symbol include/fdtd_b200.h declares: name -> (restype, argtypes)
_H = C.c_void_p
_u64p = C.POINTER(C.c_uint64)
_u32p = C.POINTER(C.c_uint32)
_fp = C.POINTER(C.c_float)
_dp = C.POINTER(C.c_double)
SYMBOLS = {
    "fdtd_b200_abi_version": (C.c_int, []),
    "fdtd_b200_device_count": (C.c_int, []),
    "fdtd_b200_global_error": (C.c_char_p, []),
    "fdtd_b200_update_coefficients_new": (C.c_int, [C.c_double, C.POINTER(PhysicalConstants), C.POINTER(Material), C.POINTER(UpdateCoefficients)]),
    "fdtd_b200_index_materials": (C.c_int, [C.POINTER(Material), C.c_uint64, _u32p, C.POINTER(Material), C.c_uint32, _u32p]),
    "fdtd_b200_canonicalize_material_index": (C.c_int, [_u32p, C.c_uint64, C.POINTER(Material), C.c_uint32, _u32p, C.POINTER(Material), C.c_uint32, _u32p]),
    "fdtd_b200_memory_required": (C.c_size_t, [C.POINTER(Config), C.c_uint32]),
    "fdtd_b200_choose_step_mode": (C.c_int, [C.POINTER(Config), C.c_uint32, C.c_uint64, _u32p, _u64p]),
    "fdtd_b200_create": (C.c_int, [C.POINTER(Config), C.POINTER(PhysicalConstants), C.POINTER(Material), C.c_uint32, _u32p, C.POINTER(_H)]),
    "fdtd_b200_create_from_table": (C.c_int, [C.POINTER(Config), C.POINTER(CoefficientsF32), C.c_uint32, _u32p, C.POINTER(_H)]),
    "fdtd_b200_destroy": (None, [_H]),
    "fdtd_b200_last_error": (C.c_char_p, [_H]),
    "fdtd_b200_reset": (C.c_int, [_H]),
    "fdtd_b200_set_stream": (C.c_int, [_H, C.c_void_p]),
    "fdtd_b200_reset_stream": (C.c_int, [_H]),
    "fdtd_b200_sync": (C.c_int, [_H]),
    "fdtd_b200_tick": (C.c_uint64, [_H]),
    "fdtd_b200_time": (C.c_double, [_H]),
    "fdtd_b200_set_time": (C.c_int, [_H, C.c_uint64, C.c_double]),
    "fdtd_b200_pml_coefficients": (C.c_int, [C.c_double, C.POINTER(PhysicalConstants), C.c_double, C.c_double, C.c_double, _dp, _dp, _dp]),
    "fdtd_b200_pml_coefficients_graded": (C.c_int, [C.c_double, C.POINTER(PhysicalConstants), C.POINTER(GradedPml), C.c_double, _dp, _dp, _dp]),
    "fdtd_b200_set_pml": (C.c_int, [_H, C.c_int, _dp, _dp, _dp, _dp]),
    "fdtd_b200_has_pml": (C.c_int, [_H]),
    "fdtd_b200_pml_in_kernel": (C.c_int, [_H]),
    "fdtd_b200_pml_state_floats": (C.c_uint64, [_H]),
    "fdtd_b200_read_pml_state": (C.c_int, [_H, _fp]),
    "fdtd_b200_write_pml_state": (C.c_int, [_H, _fp]),
    "fdtd_b200_memory_in_use": (C.c_uint64, [_H]),
    "fdtd_b200_step_path": (C.c_int, [_H, C.c_char_p, C.c_size_t]),
    "fdtd_b200_index_bytes": (C.c_int, [_H]),
    "fdtd_b200_uses_vec4": (C.c_int, [_H]),
    "fdtd_b200_uses_bulk": (C.c_int, [_H]),
    "fdtd_b200_uses_fused": (C.c_int, [_H]),
    "fdtd_b200_kernel_launches": (C.c_uint64, [_H]),
    "fdtd_b200_set_profiling": (C.c_int, [_H, C.c_int]),
    "fdtd_b200_get_profile": (C.c_int, [_H, C.c_int, C.POINTER(C.c_double), _u64p, _u64p]),
    "fdtd_b200_timer_start": (C.c_int, [_H]),
    "fdtd_b200_timer_stop": (C.c_int, [_H, C.POINTER(C.c_double)]),
    "fdtd_b200_read_material_ids": (C.c_int, [_H, _u32p]),
    "fdtd_b200_read_coefficient_table": (C.c_int, [_H, C.POINTER(CoefficientsF32), C.c_uint32, _u32p]),
    "fdtd_b200_set_sources": (C.c_int, [_H, C.c_uint32, C.POINTER(Source)]),
    "fdtd_b200_set_source_table": (C.c_int, [_H, C.c_uint32, _u64p, C.c_uint64, _fp]),
    "fdtd_b200_step": (C.c_int, [_H, C.c_uint64]),
    "fdtd_b200_update_h": (C.c_int, [_H, C.c_uint64, C.c_uint64]),
    "fdtd_b200_update_e": (C.c_int, [_H, C.c_uint64, C.c_uint64]),
    "fdtd_b200_end_step": (C.c_int, [_H]),
    "fdtd_b200_fused_update": (C.c_int, [_H, C.c_uint64, C.c_uint64]),
    "fdtd_b200_fused_commit": (C.c_int, [_H]),
    "fdtd_b200_halo_plane": (C.c_int, [_H, C.c_int, C.POINTER(C.c_void_p), _u64p]),
    "fdtd_b200_halo_copy": (C.c_int, [_H, C.c_int, _H, C.c_int]),
    "fdtd_b200_p2p_export": (C.c_int, [_H, C.c_void_p]),
    "fdtd_b200_p2p_connect": (C.c_int, [_H, C.c_void_p, C.c_void_p]),
    "fdtd_b200_p2p_refresh": (C.c_int, [_H]),
    "fdtd_b200_p2p_step": (C.c_int, [_H, C.c_uint64]),
    "fdtd_b200_p2p_connected": (C.c_int, [_H]),
    "fdtd_b200_rasterize_scene": (C.c_int, [C.c_int, _u64p, C.c_uint64, C.c_uint64, C.POINTER(C.c_double), C.POINTER(Shape), C.c_uint32, C.c_uint32, _u32p]),
    "fdtd_b200_set_materials_from_scene": (C.c_int, [_H, C.POINTER(C.c_double), C.POINTER(Shape), C.c_uint32, C.c_uint32]),
    "fdtd_b200_read_field": (C.c_int, [_H, C.c_int, _u64p, _u64p, _fp]),
    "fdtd_b200_read_field_static": (C.c_int, [_H, C.c_int, _u64p, _u64p, _fp]),
    "fdtd_b200_write_field": (C.c_int, [_H, C.c_int, _u64p, _u64p, _fp]),
    "fdtd_b200_device_field": (C.c_int, [_H, C.c_int, C.POINTER(C.c_void_p)]),
    "fdtd_b200_row_pitch": (C.c_uint64, [_H]),
    "fdtd_b200_register_probes": (C.c_int, [_H, C.c_uint32, _u64p]),
    "fdtd_b200_read_probes": (C.c_int, [_H, C.c_uint64, C.c_uint64, _fp]),
    "fdtd_b200_project": (C.c_int, [_H, C.c_int, _fp, _fp, C.c_int, C.c_uint32, C.c_uint32, C.c_void_p, C.c_int]),
}

_lib = None


class Fdtd_b200Error(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"fdtd_b200 error {code}: {message}")
        self.code = code


def lib() -> C.CDLL:
    """The loaded libfdtd_b200.so.  Raises if it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m fdtd_b200.build` (or __graft_entry__.build()). "
            "fdtd_b200 has no CPU fallback."
        )
    l = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(l, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    if l.fdtd_b200_abi_version() != ABI_VERSION:
        raise ImportError("fdtd_b200 ABI version mismatch")
    _lib = l
    return l


def check(rc: int, handle=None) -> None:
    if rc == OK:
        return
    l = lib()
    msg = l.fdtd_b200_last_error(handle) if handle else l.fdtd_b200_global_error()
    raise Fdtd_b200Error(rc, (msg or b"").decode("utf-8", "replace"))
