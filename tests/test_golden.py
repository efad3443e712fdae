"""Golden vectors produced by the reference's own code (tests/golden/make_golden.py) replayed on
(a) the CPU restatement (oracle port)  — runs without a GPU —  and (b) the CUDA path through the C ABI."""
import glob
import os

import numpy as np
import pytest

from conftest import GpuAdapter, check_against_golden

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def _load(path):
    d = dict(np.load(path))
    opts = {str(k): float(v) for k, v in zip(d["opt_names"], d["opt_vals"])}
    for k in ("max_points_per_voxel", "match_mean"):
        if k in opts:
            opts[k] = int(opts[k])
    return d, int(d["kind"]), float(d["size"]), opts


def test_fixtures_exist():
    assert len(GOLDEN) >= 7


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_port_matches_golden(path):
    from oracle import mmo
    d, kind, size, opts = _load(path)
    check_against_golden(mmo.OracleMap(kind, size, which="port", **opts), d)


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_gpu_matches_golden(path, built):
    d, kind, size, opts = _load(path)
    # likelihood: fp64 sum in a different (tree) order -> 1e-12 relative
    check_against_golden(GpuAdapter(kind, size, **opts), d, loglik_rtol=1e-12)
