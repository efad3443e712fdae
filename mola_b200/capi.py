"""ctypes binding of the C ABI in include/mola_b200/mmg.h (libmola_b200_maps.so).

This is the same binding a maintainer of the reference would write for its FFI
(see INTEGRATION.md); the Python host classes in maps.py sit on top of it. There is
no fallback: if the CUDA library is missing or fails to load, importing raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libmola_b200_maps.so")

MMG_OK = 0
MMG_ERR_INVALID_ARG = -1
MMG_ERR_CUDA = -2
MMG_ERR_NO_DEVICE = -3
MMG_ERR_OUT_OF_MEM = -4
MMG_ERR_UNSUPPORTED = -5
MMG_ERR_NCCL = -6
MMG_ERR_OUT_OF_RANGE = -7
HASHED_VOXEL, SPARSE_VOXEL, SPARSE_TREES = 0, 1, 2
MEM_HOST, MEM_DEVICE = 0, 1


class Params(C.Structure):
    _fields_ = [("voxel_size", C.c_float), ("max_points_per_voxel", C.c_uint32),
                ("remove_voxels_farther_than", C.c_double), ("min_distance_between_points", C.c_float),
                ("minimum_points_clearance", C.c_float), ("sigma_dist", C.c_double),
                ("max_corr_distance", C.c_double), ("decimation", C.c_uint32), ("match_mean", C.c_uint32)]


class TableStats(C.Structure):
    _fields_ = [("capacity", C.c_uint64), ("live_voxels", C.c_uint64), ("points", C.c_uint64),
                ("hbm_bytes", C.c_uint64), ("load_factor", C.c_double), ("mean_probe", C.c_double),
                ("max_probe", C.c_double)]


class MmgError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"mmg error {code}: {msg}")
        self.code = code


# every symbol include/mola_b200/mmg.h declares (tests check that all of them are exported)
EXPORTS = [
    "mmg_default_params", "mmg_create", "mmg_destroy", "mmg_clear", "mmg_clone", "mmg_set_params",
    "mmg_get_params", "mmg_reserve", "mmg_insert_batch", "mmg_insert_scans", "mmg_nn_single_batch", "mmg_nn_multi_batch",
    "mmg_nn_radius_batch", "mmg_nn_radius_fetch", "mmg_match_batch", "mmg_match_planes_batch", "mmg_nn_single_batch_noalloc", "mmg_likelihood", "mmg_bbox", "mmg_is_empty", "mmg_invalidate_bbox",
    "mmg_num_voxels", "mmg_num_points", "mmg_row_capacity", "mmg_export", "mmg_voxel_stats",
    "mmg_get_table_stats", "mmg_mean_stencil_points", "mmg_comm_unique_id", "mmg_comm_init", "mmg_comm_destroy", "mmg_bcast",
    "mmg_kernel_launch_count", "mmg_last_error", "mmg_version", "mmg_set_tuning",
]

_lib = None


def lib():
    """Loads the CUDA library; raises (no CPU fallback) if it was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -m mola_b200.build` "
                "(there is no CPU fallback for the map kernels)")
        _lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        _lib.mmg_last_error.restype = C.c_char_p
        _lib.mmg_version.restype = C.c_char_p
        _lib.mmg_kernel_launch_count.restype = C.c_uint64
    return _lib


def check(rc, allow=()):
    if rc != MMG_OK and rc not in allow:
        raise MmgError(rc, lib().mmg_last_error().decode())
    return rc


def fptr(a):
    """float* from a numpy array (host) or an int device pointer."""
    if a is None:
        return None
    if isinstance(a, int):
        return C.cast(a, C.POINTER(C.c_float))
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_float))


def vptr(a):
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    assert a.flags["C_CONTIGUOUS"]
    return C.c_void_p(a.ctypes.data)


def pose_ptr(pose12):
    if pose12 is None:
        return None, None
    p = np.ascontiguousarray(np.asarray(pose12, dtype=np.float64).reshape(12))
    return p, p.ctypes.data_as(C.POINTER(C.c_double))
