"""ctypes binding of libtracer_b200.so (include/tracer_b200.h).

There is no CPU fallback: if the shared library is missing or does not load,
importing the engine raises.  The library is built in-tree by
``tracer_b200/csrc/build.py`` (``__graft_entry__.build()``).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtracer_b200.so")

ABI_VERSION = 15

# enum TrcGeomKind
(G_FLAT_INFINITE, G_RECT_PLATE, G_RECT_EXTRUDED, G_ROUND_PLATE, G_TRIANGLE, G_PARABOLOID,
 G_PARAB_DISH, G_PARAB_HEX, G_PARAB_RECT, G_PARAB_CYL, G_PARAB_TROUGH, G_SPHERE, G_HEMISPHERE,
 G_SPHERE_RECT, G_CONE_INFINITE, G_CONE_FINITE, G_FRUSTUM, G_FRUSTUM_RECT, G_CYL_INFINITE,
 G_CYL_FINITE, G_CYL_RECT, G_QUADRIC_FLAT, G_QUADRIC_RECT, G_SPHERE_CUT, G_ELLIPSOID, G_RECT_PERFORATED,
 G_POLYGON) = range(27)
# enum TrcOpticsKind
(O_TRANSPARENT, O_REFLECTIVE, O_REAL_REFLECTIVE, O_LAMBERTIAN, O_LAMBERTIAN_SPECULAR,
 O_LAMB_DIRECTIONAL, O_REFLECTIVE_SPECTRAL, O_LAMBSPEC_DIRECTIONAL,
 O_LAMB_DIRECTIONAL_SPECTRAL, O_LAMB_PIECEWISE_SPEC, O_LAMBERTIAN_ABSORBANT, O_PERIODIC,
 O_FRESNEL_CONDUCTOR, O_REFRACTIVE) = range(14)
F_ONE_SIDED = 1
F_UNBOUNDED = 2
F_IAM = 4
F_POLY = 8
F_TABULATED = 16
# enum TrcFluxCoord
(FLUX_XY, FLUX_R_ATAN_XY, FLUX_R_ATAN_YX, FLUX_Z_PHI, FLUX_THETA_PHI) = range(5)
BVH_MAX_LEAF = 4

# numpy mirror of struct TrcSurfaceRec (384 bytes)
SURFACE_DTYPE = np.dtype([
    ("frame", "<f8", (12,)), ("inv", "<f8", (12,)), ("gp", "<f8", (12,)), ("op", "<f8", (8,)),
    ("gkind", "<i4"), ("okind", "<i4"), ("flags", "<u4"), ("table_off", "<i4"),
    ("table_len", "<i4"), ("pad_", "<i4", (3,)),
], align=True)
assert SURFACE_DTYPE.itemsize == 384

# numpy mirror of struct TrcBvhNode (64 bytes)
BVH_NODE_DTYPE = np.dtype([
    ("lo", "<f8", (3,)), ("hi", "<f8", (3,)), ("first", "<i4"), ("count", "<i4"), ("pad_", "<i4", (2,)),
], align=True)
assert BVH_NODE_DTYPE.itemsize == 64


HIT_CHUNK = 256
HIT_RECORD_BYTES = 64          # struct TrcHitRecord: v[3], d[3], energy (float64), ray, surf (int32)


class OpticsArgs(C.Structure):
    _fields_ = [
        ("h", C.c_int64),
        ("d_records", C.c_void_p), ("d_slot_rec", C.c_void_p), ("d_wavelengths", C.c_void_p),
        ("d_out_vertices", C.c_void_p), ("d_out_directions", C.c_void_p), ("d_out_energy", C.c_void_p),
        ("d_out_parents", C.c_void_p), ("ld_out", C.c_int64),
        ("d_draws", C.c_void_p * 3),
        ("seed", C.c_uint64), ("bounce", C.c_uint32), ("pad_", C.c_uint32),
        ("min_energy", C.c_double),
        ("d_absorbed", C.c_void_p), ("d_normals", C.c_void_p), ("d_e_in", C.c_void_p),
        ("d_dirs_in", C.c_void_p), ("d_out_wavelengths", C.c_void_p), ("d_cull", C.c_void_p),
        ("d_n_cull", C.c_void_p), ("d_attenuated", C.c_void_p),
        ("d_slot_surf", C.c_void_p), ("d_seg", C.c_void_p), ("d_child_off", C.c_void_p),
        ("d_ref_index", C.c_void_p), ("d_out_ref_index", C.c_void_p), ("d_n_void", C.c_void_p),
        ("d_ray_ids", C.c_void_p), ("ray_id_base", C.c_int64), ("d_out_ray_ids", C.c_void_p),
        ("d_ref_index_im", C.c_void_p), ("d_out_ref_index_im", C.c_void_p),
    ]


class SpectraArgs(C.Structure):
    _fields_ = [
        ("h", C.c_int64), ("n_samples", C.c_int32),
        ("d_slot_surf", C.c_void_p), ("d_parents", C.c_void_p),
        ("d_in_wavelengths", C.c_void_p), ("d_in_spectra", C.c_void_p), ("ld_in", C.c_int64),
        ("d_normals", C.c_void_p), ("d_dirs_in", C.c_void_p), ("ld_cols", C.c_int64),
        ("d_out_wavelengths", C.c_void_p), ("d_out_spectra", C.c_void_p), ("ld_out", C.c_int64),
        ("d_energy", C.c_void_p), ("d_absorbed", C.c_void_p), ("d_lost_spectra", C.c_void_p),
        ("d_cull", C.c_void_p), ("d_n_cull", C.c_void_p), ("min_energy", C.c_double),
    ]


COL_NORMALS, COL_E_IN, COL_DIRS_IN, COL_ATTENUATED = 1, 2, 4, 8


TRACE_MAX_BOUNCES = 8
TRACE_TIME_WALK, TRACE_WALK_DONE, TRACE_SEG, TRACE_ZERO_TALLIES = 1, 2, 4, 8


class TraceArgs(C.Structure):
    _fields_ = [
        ("n", C.c_int64), ("d_vertices", C.c_void_p), ("d_directions", C.c_void_p), ("d_energy", C.c_void_p),
        ("ld", C.c_int64), ("d_wavelengths", C.c_void_p), ("d_ray_ids", C.c_void_p), ("ray_id_base", C.c_int64),
        ("use_bvh", C.c_int32), ("first_bounce", C.c_uint32), ("n_bounces", C.c_int32), ("flags", C.c_int32),
        ("seed", C.c_uint64), ("min_energy", C.c_double), ("columns", C.c_uint32), ("pad_", C.c_uint32),
        ("capacity", C.c_int64 * TRACE_MAX_BOUNCES),
        ("d_keep", C.c_void_p), ("keep_bytes", C.c_int64), ("d_scratch", C.c_void_p), ("scratch_bytes", C.c_int64),
        ("d_absorbed_per_surface", C.c_void_p), ("d_hits_per_surface", C.c_void_p), ("h_counts", C.c_void_p),
    ]


class TraceBounceLayout(C.Structure):
    _fields_ = [(k, C.c_int64) for k in (
        "ld",
        "off_vertices", "off_directions", "off_energy", "off_parents", "off_ray_ids", "off_wavelengths",
        "off_alt_vertices", "off_alt_directions", "off_alt_energy", "off_alt_parents", "off_alt_ray_ids",
        "off_alt_wavelengths",
        "off_absorbed", "off_normals", "off_e_in", "off_dirs_in", "off_attenuated", "off_seg")]


class TracePlan(C.Structure):
    _fields_ = [("keep_bytes", C.c_int64), ("scratch_bytes", C.c_int64), ("counts_words", C.c_int64),
                ("bounce", TraceBounceLayout * TRACE_MAX_BOUNCES),
                ("off_records", C.c_int64), ("off_keys", C.c_int64), ("off_chunk_off", C.c_int64),
                ("off_chunk_cnt", C.c_int64), ("off_total", C.c_int64)]


class BuieTables(C.Structure):
    _fields_ = [
        ("n_seg", C.c_int32), ("has_aureole", C.c_int32),
        ("cdf", C.c_double * 212), ("theta", C.c_double * 212), ("w", C.c_double * 212),
        ("sum_integ_dni", C.c_double), ("gamma", C.c_double), ("kappa", C.c_double),
        ("theta_dni", C.c_double), ("theta_tot", C.c_double),
    ]


# name -> (restype, argtypes); every symbol declared in include/tracer_b200.h
_P = C.c_void_p
_I64 = C.c_int64
_I32 = C.c_int32
_D = C.c_double
_DP = C.POINTER(C.c_double)
SIGNATURES = {
    "trc_abi_version": (C.c_int, []),
    "trc_last_error": (C.c_char_p, []),
    "trc_launch_count": (C.c_uint64, []),
    "trc_scene_create": (C.c_int, [_P, _I32, _P, _I32, C.POINTER(_P)]),
    "trc_scene_update_frames": (C.c_int, [_P, _P, _P, _I32]),
    "trc_bvh_build": (C.c_int, [_P, _P, _I32, _I32, _P, _I32, C.POINTER(_I32), _P]),
    "trc_scene_set_bvh": (C.c_int, [_P, _P, _I32, _P, _I32, _P, _I32]),
    "trc_scene_set_shadow_grid": (C.c_int, [_P, _I32, _P, _P, _P, _P, _P, C.c_double, C.POINTER(_I32)]),
    "trc_spectra": (C.c_int, [_P, _P, _P]),
    "trc_stream_probe": (C.c_int, [_I32, _P, _P, _I64, _I64, C.POINTER(C.c_float)]),
    "trc_scene_destroy": (C.c_int, [_P]),
    "trc_scene_num_surfaces": (_I32, [_P]),
    "trc_nearest_hit": (C.c_int, [_P, _I64, _P, _P, _I64, _I32, _P, _P]),
    "trc_trace_hits": (C.c_int, [_P, _I64, _P, _P, _P, _I64, _I32, _P, _P, _I64, _P, _P, _P, _P]),
    "trc_trace_hits_slice": (C.c_int, [_P, _I64, _P, _P, _P, _I64, _I32, _P, _P, _P, _I64, _P, _P, _P, _I64, _P]),
    "trc_order_work_bytes": (_I64, [_I64, _I64, _I32]),
    "trc_order_hits": (C.c_int, [_I64, _P, _P, _P, _I64, _I32, _P, _P, _P, _P, _P]),
    "trc_sort_work_bytes": (_I64, [_I64, _I32]),
    "trc_sort_pairs": (C.c_int, [_I64, _I32, _P, _P, _P, _P, _P, _P, _P]),
    "trc_sort_pairs_static": (C.c_int, [_I64, _I32, _P, _P, _P, _P, _P, _P]),
    "trc_segment_offsets_dyn": (C.c_int, [_P, _I64, _I32, _P, _P, _P]),
    "trc_optics": (C.c_int, [_P, C.POINTER(OpticsArgs), _P]),
    "trc_gather_rows": (C.c_int, [_I64, _P, _I32, _P, _I64, _P, _I64, _P, _P, _P]),
    "trc_source_pillbox_disk": (C.c_int, [_I64, _I64, C.c_uint64, _DP, _DP, _DP, _D, _D, _D, _D, _D, _D,
                                          _P, _P, _P, _I64, _P]),
    "trc_source_pillbox_rect": (C.c_int, [_I64, _I64, C.c_uint64, _DP, _DP, _DP, _D, _D, _D, _D, _P, _P, _P, _I64, _P]),
    "trc_source_buie_rect": (C.c_int, [_I64, _I64, C.c_uint64, C.POINTER(BuieTables), _DP, _DP, _DP, _D, _D, _D,
                                       _P, _P, _P, _I64, _P]),
    "trc_source_buie_disk": (C.c_int, [_I64, _I64, C.c_uint64, C.POINTER(BuieTables), _DP, _DP, _DP, _D, _D,
                                       _P, _P, _P, _I64, _P]),
    "trc_fluxmap_rect": (C.c_int, [_I64, _P, _I64, _P, _DP, _D, _D, _I32, _I32, _P, _P]),
    "trc_trace_plan": (C.c_int, [_P, C.POINTER(TraceArgs), C.POINTER(TracePlan)]),
    "trc_trace_init": (C.c_int, [_P, C.POINTER(TraceArgs), _P]),
    "trc_trace": (C.c_int, [_P, C.POINTER(TraceArgs), _P]),
    "trc_trace_walk_ms": (C.c_int, [_P, C.POINTER(C.c_float)]),
    "trc_fluxmap_binned": (C.c_int, [_I32, _I64, _P, _I64, _P, _DP, _D, _P, _I32, _P, _I32, _P, _P]),
    "trc_segment_sums": (C.c_int, [_I64, _P, _P, _P, _P]),
    "trc_fma_peak": (C.c_int, [_I32, C.POINTER(C.c_double)]),
}


class TracerB200Error(RuntimeError):
    pass


_lib = None


def load():
    """Load the shared library (once) and bind every declared symbol."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python tracer_b200/csrc/build.py` "
            "(there is no CPU fallback for the bounce loop)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    if lib.trc_abi_version() != ABI_VERSION:
        raise ImportError("libtracer_b200.so ABI version mismatch")
    _lib = lib
    return lib


_STATUS_EXC = {-1: ValueError, -2: NotImplementedError, -3: TracerB200Error, -4: MemoryError}


def check(status):
    """Map a TrcStatus to the Python exception classes the reference raises."""
    if status == 0:
        return
    msg = load().trc_last_error().decode("utf-8", "replace")
    raise _STATUS_EXC.get(status, TracerB200Error)(msg)


def dptr(t):
    """Raw device pointer of a torch tensor (None -> NULL)."""
    if t is None:
        return None
    return C.c_void_p(t.data_ptr())


def host_doubles(a, n):
    arr = np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1))
    if arr.size != n:
        raise ValueError(f"expected {n} doubles, got {arr.size}")
    return arr, arr.ctypes.data_as(_DP)
