import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def built():
    from mola_b200 import build
    return build.build_all()


def compose_host(pose12, xs, ys, zs):
    """CPose3D::composePoint as the oracle defines it (fp64, left-to-right, then float)."""
    if pose12 is None:
        return xs.copy(), ys.copy(), zs.copy()
    T = np.asarray(pose12, np.float64)
    x, y, z = xs.astype(np.float64), ys.astype(np.float64), zs.astype(np.float64)
    g = [(((T[3 * r] * x + T[3 * r + 1] * y) + T[3 * r + 2] * z) + T[9 + r]).astype(np.float32) for r in range(3)]
    return g[0], g[1], g[2]


def assert_same_map(gpu_map, oracle_map):
    """Voxel key set, per-voxel counts and per-voxel ordered point lists, bit-exact."""
    gk, gc, gp, _ = gpu_map.exportVoxels()
    ok, oc, op = oracle_map.export()
    assert gk.shape == ok.shape, f"voxel count differs: gpu {gk.shape[0]} oracle {ok.shape[0]}"
    assert np.array_equal(gk, ok)
    assert np.array_equal(gc, oc)
    assert np.array_equal(gp.view(np.uint32), op.view(np.uint32))
