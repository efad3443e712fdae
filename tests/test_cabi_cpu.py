"""CPU-side checks of the drop-in boundary: the C-ABI library loads without a GPU, exports every symbol that
include/mola_b200/mmg.h declares (and the ctypes binding lists exactly those), and refuses to work without a device
instead of falling back to a CPU path."""
import ctypes as C
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "mola_b200", "mmg.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return set(re.findall(r"\b(mmg_[a-z0-9_]+)\s*\(", src))


def test_every_declared_symbol_is_exported(built):
    from mola_b200 import capi
    lib = capi.lib()
    declared = _header_symbols()
    assert declared == set(capi.EXPORTS), (declared ^ set(capi.EXPORTS))
    for name in sorted(declared):
        assert getattr(lib, name) is not None
    assert b"sm_100a" in lib.mmg_version()


def test_default_params_follow_the_reference(built):
    from mola_b200 import capi
    lib = capi.lib()
    p = capi.Params()
    # HashedVoxelPointCloud.h:119, SparseVoxelPointCloud.h:158, SparseTreesPointCloud.h:113,337
    for kind, size in ((capi.HASHED_VOXEL, 1.0), (capi.SPARSE_VOXEL, 0.2), (capi.SPARSE_TREES, 10.0)):
        assert lib.mmg_default_params(kind, C.byref(p)) == capi.MMG_OK
        assert abs(p.voxel_size - size) < 1e-6 and p.max_points_per_voxel == 0 and p.decimation == 10
    assert abs(p.minimum_points_clearance - 0.2) < 1e-6
    assert lib.mmg_default_params(7, C.byref(p)) == capi.MMG_ERR_INVALID_ARG


def test_no_device_means_no_map(built):
    """Without a CUDA device mmg_create must fail loudly (no CPU fallback). Only meaningful on a GPU-less host."""
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present")
    from mola_b200 import capi
    lib = capi.lib()
    p = capi.Params()
    lib.mmg_default_params(capi.HASHED_VOXEL, C.byref(p))
    h = C.c_void_p()
    rc = lib.mmg_create(capi.HASHED_VOXEL, C.byref(p), 0, C.byref(h))
    assert rc == capi.MMG_ERR_NO_DEVICE and not h.value
    assert b"no CPU path" in lib.mmg_last_error()
