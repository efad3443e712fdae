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


class GpuAdapter:
    """Gives a mola_b200 map the same call surface as oracle.mmo.OracleMap, so one checker serves both."""

    def __init__(self, kind, size, **opts):
        import mola_b200 as mb
        cls = [mb.HashedVoxelPointCloud, mb.SparseVoxelPointCloud, mb.SparseTreesPointCloud][kind]
        self.kind, self.m = kind, cls(size)
        self.set(**opts)

    def set(self, **opts):
        m = self.m
        for k, v in opts.items():
            if k in ("sigma_dist", "max_corr_distance", "decimation"):
                setattr(m.likelihoodOptions, k, v)
            elif k == "match_mean":
                m.likelihoodOptions.match_mean = bool(v)
            elif k == "remove_voxels_farther_than" and self.kind == 2:
                m.insertionOptions.remove_submaps_farther_than = v
            else:
                setattr(m.insertionOptions, k, v)

    def insert(self, xs, ys, zs, pose=None):
        self.m.insertPointCloud(xs, ys, zs, pose)

    def export(self):
        k, c, p, _ = self.m.exportVoxels()
        return k, c, p

    def bbox(self):
        return self.m.boundingBox()

    def isEmpty(self):
        return self.m.isEmpty()

    def num_points(self):
        return self.m.numPoints()

    def nn_single(self, x, y, z):
        return self.m.nn_single_search_batch(x, y, z)

    def nn_multi(self, x, y, z, k):
        return self.m.nn_multiple_search_batch(x, y, z, k)

    def nn_radius(self, x, y, z, r2, maxp=0):
        return self.m.nn_radius_search_batch(x, y, z, float(r2), int(maxp))

    def likelihood(self, xs, ys, zs, pose=None):
        return self.m.computeObservationLikelihood(xs, ys, zs, pose)

    def voxel_stats(self):
        return self.m.voxelStats()


def bits(a):
    return a.view(np.uint32) if a.dtype == np.float32 else a


def check_against_golden(m, d, loglik_rtol=0.0):
    """Replays a golden case (tests/golden/*.npz, produced by the reference's own code) on map `m`."""
    for k in range(int(d["nscans"])):
        m.insert(d[f"in{k}_xs"], d[f"in{k}_ys"], d[f"in{k}_zs"], d[f"in{k}_pose"])
    keys, counts, pts = m.export()
    assert np.array_equal(keys, d["keys"]) and np.array_equal(counts, d["counts"])
    assert np.array_equal(bits(pts), bits(d["points"]))
    assert np.array_equal(np.concatenate(m.bbox()), d["bbox"])
    q = (d["qx"], d["qy"], d["qz"])
    if "nn1_found" in d:
        f, p, dd, ids = m.nn_single(*q)
        assert np.array_equal(f, d["nn1_found"])
        k = f > 0
        assert np.array_equal(ids[k], d["nn1_ids"][k]) and np.array_equal(bits(dd[k]), bits(d["nn1_d2"][k]))
        assert np.array_equal(bits(p[k]), bits(d["nn1_pts"][k]))
        last = int(d["nscans"]) - 1
        from mola_b200 import synth
        T = synth.perturbed_pose(d[f"in{last}_pose"], 0.12, 0.7)
        ll = m.likelihood(d[f"in{last}_xs"], d[f"in{last}_ys"], d[f"in{last}_zs"], T)
        assert abs(ll - float(d["loglik"])) <= loglik_rtol * abs(float(d["loglik"]))
    if "mean" in d:
        mean, _ = m.voxel_stats()
        assert np.array_equal(bits(mean), bits(d["mean"]))
    for kk in d["ks"]:
        kk = int(kk)
        c, p, dd, ids = m.nn_multi(*q, kk)
        assert np.array_equal(c, d[f"nnk{kk}_count"])
        msk = np.arange(kk)[None, :] < c[:, None]
        assert np.array_equal(ids[msk], d[f"nnk{kk}_ids"][msk]) and np.array_equal(bits(dd[msk]), bits(d[f"nnk{kk}_d2"][msk]))
        assert np.array_equal(bits(p[msk]), bits(d[f"nnk{kk}_pts"][msk]))
    for j, (r, mp) in enumerate(d["radii"]):
        off, p, dd, ids = m.nn_radius(*q, np.float32(r * r), int(mp))
        assert np.array_equal(off, d[f"rad{j}_off"]) and np.array_equal(ids, d[f"rad{j}_ids"])
        assert np.array_equal(bits(dd), bits(d[f"rad{j}_d2"])) and np.array_equal(bits(p), bits(d[f"rad{j}_pts"]))
