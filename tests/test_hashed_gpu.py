"""GPU parity of the HashedVoxelPointCloud path against the CPU oracle, through the C ABI.
Bit-exact: voxel key set, per-voxel ordered point lists, found flags, ids, points, d2."""
import numpy as np
import pytest

from conftest import assert_same_map, compose_host

pytestmark = pytest.mark.gpu


def _maps(voxel=0.5, **opts):
    from mola_b200 import HashedVoxelPointCloud
    from oracle import mmo
    g = HashedVoxelPointCloud(voxel)
    for k, v in opts.items():
        setattr(g.insertionOptions, k, v)
    o = mmo.OracleMap(0, voxel, **opts)
    return g, o


def _scan(sensor="hdl64", s=0.0, scan_id=0):
    from mola_b200 import synth
    return synth.lidar_scan(sensor, s=s, scan_id=scan_id)


def _check_nn_single(g, o, qx, qy, qz):
    f, p, d, i = g.nn_single_search_batch(qx, qy, qz)
    of, op, od, oi = o.nn_single(qx, qy, qz)
    assert np.array_equal(f, of)
    k = of > 0
    assert np.array_equal(i[k], oi[k])
    assert np.array_equal(p[k].view(np.uint32), op[k].view(np.uint32))
    assert np.array_equal(d[k].view(np.uint32), od[k].view(np.uint32))
    return k.mean()


def test_c1_insert_and_nn_single(built):
    """BASELINE config 1: one 120k-pt HDL-64 scan, voxel 0.5 m, insert + nn_single of the re-posed scan."""
    from mola_b200 import synth
    g, o = _maps(0.5)
    xs, ys, zs, pose = _scan()
    assert g.isEmpty()
    g.insertPointCloud(xs, ys, zs, pose)
    o.insert(xs, ys, zs, pose)
    assert not g.isEmpty()
    assert_same_map(g, o)
    assert g.numPoints() == o.num_points()
    qx, qy, qz = compose_host(synth.perturbed_pose(pose), xs, ys, zs)
    frac = _check_nn_single(g, o, qx, qy, qz)
    assert frac > 0.9
    # same queries composed on the device from the sensor-frame scan
    f, p, d, i = g.nn_single_search_batch(xs, ys, zs, pose=synth.perturbed_pose(pose))
    of, op, od, oi = o.nn_single(qx, qy, qz)
    assert np.array_equal(f, of) and np.array_equal(i[of > 0], oi[of > 0])


@pytest.mark.parametrize("opts", [dict(max_points_per_voxel=5), dict(max_points_per_voxel=32), dict(max_points_per_voxel=40),
                                  dict(min_distance_between_points=0.05), dict(min_distance_between_points=0.10, max_points_per_voxel=7),
                                  dict(min_distance_between_points=0.02)])
def test_decimation_options_multi_scan(built, opts):
    g, o = _maps(0.5, **opts)
    for k in range(4):
        xs, ys, zs, pose = _scan(s=3.0 * k, scan_id=k)
        g.insertPointCloud(xs, ys, zs, pose)
        o.insert(xs, ys, zs, pose)
        assert_same_map(g, o)
    gmin, gmax = g.boundingBox()
    omin, omax = o.bbox()
    assert np.array_equal(gmin, omin) and np.array_equal(gmax, omax)
    xs, ys, zs, pose = _scan(s=5.0, scan_id=77)
    qx, qy, qz = compose_host(pose, xs, ys, zs)
    _check_nn_single(g, o, qx[:30000], qy[:30000], qz[:30000])


def test_point_by_point_equals_batch(built):
    """Property: one batch == the same points inserted in many small batches (sequential semantics)."""
    g, o = _maps(0.4, min_distance_between_points=0.03, max_points_per_voxel=20)
    from mola_b200 import HashedVoxelPointCloud
    g2 = HashedVoxelPointCloud(0.4)
    g2.insertionOptions.min_distance_between_points = 0.03
    g2.insertionOptions.max_points_per_voxel = 20
    xs, ys, zs, pose = _scan("tiny")
    g.insertPointCloud(xs, ys, zs, pose)
    o.insert(xs, ys, zs, pose)
    for a in range(0, xs.size, 37):
        g2.insertPointCloud(xs[a:a + 37], ys[a:a + 37], zs[a:a + 37], pose)
    assert_same_map(g, o)
    assert_same_map(g2, o)


def test_eviction(built):
    g, o = _maps(0.5, remove_voxels_farther_than=20.0)
    for k in range(6):
        xs, ys, zs, pose = _scan(s=10.0 * k, scan_id=k)
        g.insertPointCloud(xs, ys, zs, pose)
        o.insert(xs, ys, zs, pose)
        assert_same_map(g, o)
    # no robot pose given: no eviction, identity compose
    xs, ys, zs, _ = _scan(s=0.0, scan_id=9)
    g.insertPointCloud(xs, ys, zs)
    o.insert(xs, ys, zs)
    assert_same_map(g, o)


@pytest.mark.parametrize("k", [1, 2, 3, 5])
def test_nn_multiple_search(built, k):
    """k<=3 follows the reference's 3-slot buffer, including its forward-copy shift."""
    g, o = _maps(0.5)
    xs, ys, zs, pose = _scan()
    g.insertPointCloud(xs, ys, zs, pose)
    o.insert(xs, ys, zs, pose)
    from mola_b200 import synth
    qx, qy, qz = compose_host(synth.perturbed_pose(pose), xs[::7], ys[::7], zs[::7])
    c, p, d, i = g.nn_multiple_search_batch(qx, qy, qz, k)
    oc, op, od, oi = o.nn_multi(qx, qy, qz, k)
    assert np.array_equal(c, oc)
    m = np.arange(k)[None, :] < oc[:, None]
    assert np.array_equal(i[m], oi[m])
    assert np.array_equal(d[m].view(np.uint32), od[m].view(np.uint32))
    assert np.array_equal(p[m].view(np.uint32), op[m].view(np.uint32))


@pytest.mark.parametrize("r,maxp", [(0.2, 0), (0.7, 0), (1.3, 0), (0.7, 2), (1.0, 10), (2.0, 1)])
def test_nn_radius_search(built, r, maxp):
    g, o = _maps(0.5)
    xs, ys, zs, pose = _scan("tiny")
    for s in (0.0, 2.0):
        xs, ys, zs, pose = _scan("tiny", s=s, scan_id=int(s))
        g.insertPointCloud(xs, ys, zs, pose)
        o.insert(xs, ys, zs, pose)
    qx, qy, qz = compose_host(pose, xs, ys, zs)
    off, p, d, i = g.nn_radius_search_batch(qx, qy, qz, r * r, maxp)
    ooff, op, od, oi = o.nn_radius(qx, qy, qz, np.float32(r * r), maxp)
    assert np.array_equal(off, ooff)
    assert np.array_equal(i, oi)
    assert np.array_equal(d.view(np.uint32), od.view(np.uint32))
    assert np.array_equal(p.view(np.uint32), op.view(np.uint32))
    assert off[-1] > 0


def test_likelihood_and_bbox(built):
    g, o = _maps(0.5)
    # bounding box before any point was cached: from voxel corners (reference :872-902)
    xs, ys, zs, pose = _scan("tiny")
    assert np.array_equal(np.concatenate(g.boundingBox()), np.zeros(6, np.float32))
    assert g.computeObservationLikelihood(xs, ys, zs, pose) == 0.0
    g.insertPointCloud(xs, ys, zs, pose)
    o.bbox()
    o.insert(xs, ys, zs, pose)
    assert np.array_equal(np.concatenate(g.boundingBox()), np.concatenate(o.bbox()))
    from mola_b200 import synth
    q = synth.perturbed_pose(pose, 0.3, 2.0)
    a, b = g.computeObservationLikelihood(xs, ys, zs, q), o.likelihood(xs, ys, zs, q)
    assert a < 0 and abs(a - b) <= 1e-9 * abs(b)
    g.likelihoodOptions.decimation = 3
    g.likelihoodOptions.max_corr_distance = 0.05
    o.set(decimation=3, max_corr_distance=0.05)
    a, b = g.computeObservationLikelihood(xs, ys, zs, q), o.likelihood(xs, ys, zs, q)
    assert abs(a - b) <= 1e-9 * abs(b)


def test_uncached_bbox_from_voxels(built):
    g, o = _maps(0.5)
    xs, ys, zs, pose = _scan("tiny")
    # a clone made through export/import would lose the cache; here: fresh maps, insert with the
    # cache pre-seeded to zeros by an early boundingBox() call on the empty map (reference quirk)
    g.boundingBox(), o.bbox()
    g.insertPointCloud(xs, ys, zs, pose)
    o.insert(xs, ys, zs, pose)
    assert np.array_equal(np.concatenate(g.boundingBox()), np.concatenate(o.bbox()))


def test_edge_cases(built):
    from mola_b200 import MapError
    g, o = _maps(0.5)
    e = np.zeros(0, np.float32)
    g.insertPointCloud(e, e, e)
    assert g.isEmpty() and g.numVoxels() == 0
    f, p, d, i = g.nn_single_search_batch(np.ones(5, np.float32), np.ones(5, np.float32), np.ones(5, np.float32))
    assert not f.any()
    with pytest.raises(MapError):
        g.nn_multiple_search_batch(e, e, e, 0)
    with pytest.raises(MapError):
        g.nn_single_search_2d(0, 0)
    g.insertPoint((1.0, 2.0, 3.0))
    assert not g.isEmpty()
    ok, pt, d2, idx = g.nn_single_search((1.1, 2.0, 3.0))
    assert ok and np.allclose(pt, [1, 2, 3]) and abs(d2 - 0.01) < 1e-6
    # points around the origin: truncation toward zero makes index 0 span (-v, v)
    x = np.array([-0.4, 0.4, -0.6, 0.6, -0.0, 1e-30], np.float32)
    g2, o2 = _maps(0.5)
    g2.insertPointCloud(x, x, x)
    o2.insert(x, x, x)
    assert_same_map(g2, o2)
    # clear keeps working, clone is deep
    c = g2.clone()
    g2.clear()
    assert g2.isEmpty() and not c.isEmpty()
    assert_same_map(c, o2)
    # non-finite / out-of-range coordinates are reported, valid ones still inserted
    y = np.array([1.0, np.nan, 3e6, 2.0], np.float32)
    g3, o3 = _maps(0.5)
    assert g3.insertPointCloud(y, y, y) is True
    o3.insert(y[[0, 3]], y[[0, 3]], y[[0, 3]])
    assert_same_map(g3, o3)


def _scan_batch(n_scans, sensor="tiny", step=2.0):
    scans = [_scan(sensor, s=step * k, scan_id=k) for k in range(n_scans)]
    xs, ys, zs = (np.concatenate([s[a] for s in scans]) for a in range(3))
    off = np.concatenate([[0], np.cumsum([s[0].size for s in scans])]).astype(np.uint64)
    poses = np.stack([s[3] for s in scans])
    return scans, xs, ys, zs, off, poses


@pytest.mark.parametrize("opts", [dict(), dict(max_points_per_voxel=5), dict(max_points_per_voxel=40),
                                  dict(min_distance_between_points=0.05), dict(remove_voxels_farther_than=15.0)])
def test_insert_scans_equals_sequential(built, opts):
    """mmg_insert_scans (one fused device pass when the options allow it, scan by scan otherwise) ==
    insertObservation once per scan in order (HashedVoxelPointCloud.cpp:289-331)."""
    g, o = _maps(0.5, **opts)
    scans, xs, ys, zs, off, poses = _scan_batch(5, "tiny", step=4.0)
    # a map that already holds something, so n0 > 0 rows are exercised too
    g.insertPointCloud(*scans[2])
    o.insert(*scans[2])
    g.insertPointClouds(xs, ys, zs, off, poses)
    for s in scans:
        o.insert(*s)
    assert_same_map(g, o)
    gmin, gmax = g.boundingBox()
    omin, omax = o.bbox()
    assert np.array_equal(gmin, omin) and np.array_equal(gmax, omax)


def test_insert_scans_hdl64_and_device_buffers(built):
    """Four 120k-pt scans in one call, device-resident inputs, then nn_single parity on the result."""
    import torch
    g, o = _maps(0.5)
    scans, xs, ys, zs, off, poses = _scan_batch(4, "hdl64", step=3.0)
    d = [torch.from_numpy(a).cuda() for a in (xs, ys, zs)]
    g.insertPointClouds(d[0], d[1], d[2], off, poses)
    for s in scans:
        o.insert(*s)
    assert_same_map(g, o)
    # the ranking scratch (float4.w of free slots) must be clean again: a second batch still matches
    g.insertPointClouds(d[0], d[1], d[2], off, poses)
    for s in scans:
        o.insert(*s)
    assert_same_map(g, o)
    qx, qy, qz = compose_host(poses[1], scans[1][0] + np.float32(0.07), scans[1][1], scans[1][2])
    _check_nn_single(g, o, qx[:50000], qy[:50000], qz[:50000])


def test_insert_adversarial_order(built):
    """Ranking stress: many points of few voxels in shuffled order, duplicates, a voxel filled across
    several batches and partially filled rows."""
    rng = np.random.default_rng(5)
    g, o = _maps(1.0)
    for it in range(4):
        n = 40000
        cells = rng.integers(0, 6, size=(n, 3)).astype(np.float32)
        p = cells + rng.random((n, 3), dtype=np.float32) * np.float32(0.9) + np.float32(0.05)
        p[::7] = p[3]  # exact duplicates
        if it == 0:
            p = p[:50]  # leaves partially filled rows for the next batches
        x, y, z = (np.ascontiguousarray(p[:, a]) for a in range(3))
        g.insertPointCloud(x, y, z)
        o.insert(x, y, z)
        assert_same_map(g, o)


@pytest.mark.parametrize("thr", [0.05, 0.2, 0.75, float("inf"), 0.0])
def test_match_points_batch(built, thr):
    """Matcher = nn_single_search per point + 'd2 < thr^2' filter + order-preserving compaction
    (oracle: the reference's nn_single_search restated, filtered in numpy)."""
    from mola_b200 import synth
    g, o = _maps(0.5)
    xs, ys, zs, pose = _scan()
    g.insertPointCloud(xs, ys, zs, pose)
    o.insert(xs, ys, zs, pose)
    qpose = synth.perturbed_pose(pose, dx=0.15, dyaw_deg=0.8)
    qx, qy, qz = compose_host(qpose, xs, ys, zs)
    of, op, od, oi = o.nn_single(qx, qy, qz)
    t2 = np.float32(thr) * np.float32(thr)
    keep = (of > 0) & (od < t2)
    idx, pts, d2, ids = g.match_points_batch(xs, ys, zs, pose=qpose, max_dist_sqr=float(t2))
    assert np.array_equal(idx, np.nonzero(keep)[0].astype(np.uint32))
    assert np.array_equal(ids, oi[keep])
    assert np.array_equal(d2.view(np.uint32), od[keep].view(np.uint32))
    assert np.array_equal(pts.view(np.uint32), op[keep].view(np.uint32))
    if 0 < thr < 1:
        assert 0 < idx.size < xs.size
    # device buffers give the same pairings
    import torch
    d = [torch.from_numpy(a).cuda() for a in (xs, ys, zs)]
    idx2, pts2, d22, ids2 = g.match_points_batch(d[0], d[1], d[2], pose=qpose, max_dist_sqr=float(t2))
    assert np.array_equal(idx2.cpu().numpy().view(np.uint32), idx) and np.array_equal(ids2.cpu().numpy().view(np.uint64), ids)
