"""BASELINE config 2 at full size (10M-point HashedVoxelPointCloud, 1M queries) on the GPU: size-independent
properties plus an oracle check on a sample. The oracle twin is built from the same scans (CPU, ~2 s)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAP_POINTS = 10_000_000


@pytest.fixture(scope="module")
def c2(built):
    import bench
    from mola_b200 import HashedVoxelPointCloud
    from oracle import mmo
    g = HashedVoxelPointCloud(bench.VOXEL)
    o = mmo.OracleMap(0, bench.VOXEL)

    def ins(xs, ys, zs, pose):
        g.insertPointCloud(xs, ys, zs, pose)
        o.insert(xs, ys, zs, pose)
    nscans, n_in = bench.map_scans(MAP_POINTS, g.numPoints, ins)
    q = bench.query_batch(nscans)
    return g, o, q, nscans


def test_c2_map_matches_oracle(c2):
    g, o, q, nscans = c2
    assert g.numPoints() == o.num_points() >= MAP_POINTS
    assert g.numVoxels() == o.num_voxels()
    gk, gc, gp, _ = g.exportVoxels()
    ok, oc, op = o.export()
    assert np.array_equal(gk, ok) and np.array_equal(gc, oc)
    assert int(gc.max()) <= 32 and int(gc.sum()) == g.numPoints()
    assert np.array_equal(gp.view(np.uint32), op.view(np.uint32))
    assert np.array_equal(np.concatenate(g.boundingBox()), np.concatenate(o.bbox()))


def test_c2_queries_sample_vs_oracle_and_exhaustive_kernel(c2):
    g, o, q, _ = c2
    f, p, d, i = g.nn_single_search_batch(*q)
    assert f.mean() > 0.99
    # (a) the pruned, grouped kernel == the exhaustive one-thread-per-query walk (k = 1), all 1M queries
    c1, p1, d1, i1 = g.nn_multiple_search_batch(*q, 1)
    assert np.array_equal(c1, f.astype(np.uint32))
    k = f > 0
    assert np.array_equal(i1[k, 0], i[k]) and np.array_equal(d1[k, 0].view(np.uint32), d[k].view(np.uint32))
    # (b) oracle on a strided 50k sample
    s = slice(0, None, 20)
    of, op, od, oi = o.nn_single(q[0][s], q[1][s], q[2][s], nthreads=8)
    assert np.array_equal(f[s], of)
    ks = of > 0
    assert np.array_equal(i[s][ks], oi[ks]) and np.array_equal(d[s][ks].view(np.uint32), od[ks].view(np.uint32))
    assert np.array_equal(p[s][ks].view(np.uint32), op[ks].view(np.uint32))
    # (c) d2 really is the squared distance to the returned point
    dd = (p[k] - np.stack(q, 1)[k]).astype(np.float32)
    ref = (dd[:, 0] * dd[:, 0] + dd[:, 1] * dd[:, 1]).astype(np.float32) + dd[:, 2] * dd[:, 2]
    assert np.array_equal(ref.astype(np.float32).view(np.uint32), d[k].view(np.uint32))


def test_c2_stored_points_find_themselves(c2):
    """Idempotence: every stored point queried against the map comes back with d2 == 0."""
    g, o, q, _ = c2
    gk, gc, gp, _ = g.exportVoxels()
    sel = np.flatnonzero(gc > 0)[::4]
    pts = gp[sel, 0, :]  # slot 0 of every 4th voxel
    f, p, d, i = g.nn_single_search_batch(np.ascontiguousarray(pts[:, 0]), np.ascontiguousarray(pts[:, 1]), np.ascontiguousarray(pts[:, 2]))
    assert f.all() and not d.any()
    # the id names the voxel of the point: 5 slot bits + 3 x 20 index bits
    cx = ((i >> np.uint64(5)) & np.uint64(0xFFFFF)).astype(np.int64)
    cx = np.where(cx >= 1 << 19, cx - (1 << 20), cx)
    assert np.array_equal(cx, gk[sel, 0].astype(np.int64))


def test_c5_fused_insert_of_64_scans_matches_sequential_oracle(built):
    """BASELINE config 5 at full size: 64 OS1-128 scans (16.8M points) inserted by ONE mmg_insert_scans call ==
    the oracle inserting them one after the other (voxel set, counts, ordered point lists, bounding box), and a
    second identical batch changes nothing but the already-full rows' neighbours (still equal to the oracle)."""
    import torch
    from mola_b200 import HashedVoxelPointCloud, synth
    from oracle import mmo
    ns = 64
    scans = [synth.lidar_scan("os1_128", s=2.0 * k, seed=5, scan_id=k) for k in range(ns)]
    off = np.concatenate([[0], np.cumsum([s[0].size for s in scans])]).astype(np.uint64)
    poses = np.stack([s[3] for s in scans])
    d = [torch.from_numpy(np.concatenate([s[a] for s in scans])).cuda() for a in range(3)]
    g = HashedVoxelPointCloud(0.5)
    o = mmo.OracleMap(0, 0.5)
    for rep in range(2):
        g.insertPointClouds(d[0], d[1], d[2], off, poses)
        for s in scans:
            o.insert(*s)
        assert g.numPoints() == o.num_points() and g.numVoxels() == o.num_voxels()
        gk, gc, gp, _ = g.exportVoxels()
        ok, oc, op = o.export()
        assert np.array_equal(gk, ok) and np.array_equal(gc, oc)
        assert np.array_equal(gp.view(np.uint32), op.view(np.uint32))
    assert np.array_equal(np.concatenate(g.boundingBox()), np.concatenate(o.bbox()))
    # every scan queried against the finished map, one scan checked against the oracle
    s = scans[31]
    qpose = synth.perturbed_pose(s[3])
    f, p, dd, ii = g.nn_single_search_batch(s[0], s[1], s[2], pose=qpose)
    from conftest import compose_host
    of, op_, od, oi = o.nn_single(*compose_host(qpose, s[0], s[1], s[2]), nthreads=8)
    k = of > 0
    assert np.array_equal(f, of) and np.array_equal(ii[k], oi[k]) and np.array_equal(dd[k].view(np.uint32), od[k].view(np.uint32))
