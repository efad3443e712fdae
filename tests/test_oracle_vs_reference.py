"""Pins the CPU restatement (oracle/mmo_port.cpp) against the reference's own code
(oracle/_ref/libmmo_ref.so = the reference .cpp files compiled unmodified against shim/mrpt).
Runs wherever the prebuilt _ref library exists (this container, and the GPU box: it ships)."""
import numpy as np
import pytest

from oracle import mmo

pytestmark = pytest.mark.skipif(not mmo.have_reference(), reason="oracle/_ref/libmmo_ref.so not built (needs /root/reference)")


def scans(n=3, sensor="tiny", spacing=2.5):
    from mola_b200 import synth
    return [synth.lidar_scan(sensor, s=spacing * k, scan_id=k) for k in range(n)]


def both(kind, size, **opts):
    return mmo.OracleMap(kind, size, which="port", **opts), mmo.OracleMap(kind, size, which="reference", **opts)


def same_export(a, b):
    ka, ca, pa = a.export()
    kb, cb, pb = b.export()
    assert np.array_equal(ka, kb) and np.array_equal(ca, cb)
    assert np.array_equal(pa.view(np.uint32), pb.view(np.uint32))
    assert a.num_points() == b.num_points() and a.isEmpty() == b.isEmpty()


def queries(sc, shift=(0.07, -0.03, 0.02)):
    xs, ys, zs, pose = sc
    T = np.asarray(pose)
    x, y, z = xs.astype(np.float64), ys.astype(np.float64), zs.astype(np.float64)
    g = [(((T[3 * r] * x + T[3 * r + 1] * y) + T[3 * r + 2] * z) + T[9 + r] + shift[r]).astype(np.float32) for r in range(3)]
    return g


def same_hits(ra, rb):
    for x, y in zip(ra, rb):
        assert x.shape == y.shape
        assert np.array_equal(x.view(np.uint32) if x.dtype == np.float32 else x, y.view(np.uint32) if y.dtype == np.float32 else y)


HASHED_OPTS = [dict(), dict(max_points_per_voxel=4), dict(min_distance_between_points=0.05),
               dict(min_distance_between_points=0.08, max_points_per_voxel=6), dict(remove_voxels_farther_than=15.0)]


@pytest.mark.parametrize("opts", HASHED_OPTS)
def test_hashed(opts):
    a, b = both(0, 0.5, **opts)
    sc = scans(4, "tiny", 6.0)
    for s in sc:
        a.insert(*s)
        b.insert(*s)
        same_export(a, b)
    assert np.array_equal(np.concatenate(a.bbox()), np.concatenate(b.bbox()))
    q = queries(sc[-1])
    fa, fb = a.nn_single(*q), b.nn_single(*q)
    assert np.array_equal(fa[0], fb[0])
    k = fa[0] > 0
    same_hits([r[k] for r in fa[1:]], [r[k] for r in fb[1:]])
    for kk in (1, 2, 3, 4):
        ma, mb = a.nn_multi(*q, kk), b.nn_multi(*q, kk)
        assert np.array_equal(ma[0], mb[0])
        m = np.arange(kk)[None, :] < ma[0][:, None]
        same_hits([r[m] for r in ma[1:]], [r[m] for r in mb[1:]])
    for r, mp in ((0.3, 0), (0.9, 0), (0.9, 2), (1.6, 5)):
        same_hits(a.nn_radius(*q, np.float32(r * r), mp), b.nn_radius(*q, np.float32(r * r), mp))
    la, lb = a.likelihood(*sc[-1][:3], sc[-1][3]), b.likelihood(*sc[-1][:3], sc[-1][3])
    assert la == lb


def test_hashed_full_scan_and_no_pose():
    a, b = both(0, 0.5)
    sc = scans(2, "hdl64", 4.0)
    for s in sc:
        a.insert(*s)
        b.insert(*s)
    a.insert(*sc[0][:3])  # no robot pose
    b.insert(*sc[0][:3])
    same_export(a, b)
    q = queries(sc[1])
    q = [c[::11] for c in q]
    same_hits(a.nn_single(*q), b.nn_single(*q))
    assert abs(a.mean_stencil_points(*q) - b.mean_stencil_points(*q)) < 1e-9


def test_hashed_uncached_bbox():
    a, b = both(0, 0.5)
    s = scans(1)[0]
    assert np.array_equal(np.concatenate(a.bbox()), np.concatenate(b.bbox()))  # empty: zeros, and cached
    a.insert(*s)
    b.insert(*s)
    assert np.array_equal(np.concatenate(a.bbox()), np.concatenate(b.bbox()))


@pytest.mark.parametrize("opts", [dict(), dict(max_points_per_voxel=3), dict(max_points_per_voxel=16), dict(match_mean=1),
                                  dict(match_mean=1, max_points_per_voxel=20)])
def test_sparse_voxel(opts):
    a, b = both(1, 0.2, **opts)
    sc = scans(3, "tiny", 1.0)
    for s in sc:
        a.insert(*s)
        b.insert(*s)
        same_export(a, b)
    ma, mb = a.voxel_stats(), b.voxel_stats()
    assert np.array_equal(ma[0].view(np.uint32), mb[0].view(np.uint32))  # the reference's running mean
    assert np.array_equal(np.concatenate(a.bbox()), np.concatenate(b.bbox()))
    q = queries(sc[-1])
    for r, mp in ((0.15, 0), (0.45, 0), (0.45, 3), (0.8, 16), (0.3, 1)):
        same_hits(a.nn_radius(*q, np.float32(r * r), mp), b.nn_radius(*q, np.float32(r * r), mp))
    for m in (a, b):  # nn_single / nn_multiple throw in the reference
        with pytest.raises(RuntimeError):
            m.nn_single(*q)
        with pytest.raises(RuntimeError):
            m.nn_multi(*q, 2)


def test_sparse_voxel_uncached_bbox_and_ties():
    a, b = both(1, 0.2)
    # exact .5 voxel boundaries (ties-to-even rounding) and duplicated points (equal d2 collapse, N2)
    x = np.array([0.1, 0.3, 0.5, -0.1, -0.3, 0.7, 0.7, 0.7, 0.25, 0.25], np.float32)
    y = np.array([0.1, 0.1, 0.1, 0.1, 0.1, 0.2, 0.2, 0.2, 0.25, 0.25], np.float32)
    z = np.zeros_like(x)
    for m in (a, b):
        m.insert(x, y, z)
    same_export(a, b)
    q = [np.array([0.6, 0.2], np.float32), np.array([0.2, 0.2], np.float32), np.zeros(2, np.float32)]
    for mp in (0, 1, 2, 8):
        same_hits(a.nn_radius(*q, np.float32(1.0), mp), b.nn_radius(*q, np.float32(1.0), mp))
    a2, b2 = both(1, 0.2)
    a2.insert(x, y, z), b2.insert(x, y, z)
    assert np.array_equal(np.concatenate(a2.bbox()), np.concatenate(b2.bbox()))


@pytest.mark.parametrize("opts", [dict(), dict(minimum_points_clearance=0.5), dict(minimum_points_clearance=0.05, remove_voxels_farther_than=25.0)])
def test_sparse_trees(opts):
    a, b = both(2, 10.0, **opts)
    sc = scans(4, "tiny", 9.0)
    for s in sc:
        a.insert(*s)
        b.insert(*s)
        same_export(a, b)
    assert np.array_equal(np.concatenate(a.bbox()), np.concatenate(b.bbox()))
    q = queries(sc[-1])
    same_hits(a.nn_single(*q), b.nn_single(*q))
    for kk in (1, 3, 8):
        ma, mb = a.nn_multi(*q, kk), b.nn_multi(*q, kk)
        assert np.array_equal(ma[0], mb[0])
        m = np.arange(kk)[None, :] < ma[0][:, None]
        same_hits([r[m] for r in ma[1:]], [r[m] for r in mb[1:]])
    for r, mp in ((0.2, 0), (1.0, 0), (2.0, 0), (2.0, 4), (6.0, 2)):
        same_hits(a.nn_radius(*q, np.float32(r * r), mp), b.nn_radius(*q, np.float32(r * r), mp))
    assert a.likelihood(*sc[-1][:3], sc[-1][3]) == b.likelihood(*sc[-1][:3], sc[-1][3])
