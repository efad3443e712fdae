"""GPU parity of the SparseVoxelPointCloud and SparseTreesPointCloud paths against the CPU oracle,
through the C ABI. Bit-exact: block / voxel sets, ordered point lists, ids, points, d2, CSR offsets."""
import numpy as np
import pytest

from conftest import GpuAdapter, bits, compose_host

pytestmark = pytest.mark.gpu


def _pair(kind, size, **opts):
    from oracle import mmo
    return GpuAdapter(kind, size, **opts), mmo.OracleMap(kind, size, **opts)


def _scan(sensor="tiny", s=0.0, scan_id=0):
    from mola_b200 import synth
    return synth.lidar_scan(sensor, s=s, scan_id=scan_id)


def _same_export(g, o):
    gk, gc, gp = g.export()
    ok, oc, op = o.export()
    assert np.array_equal(gk, ok) and np.array_equal(gc, oc)
    assert gp.shape == op.shape and np.array_equal(bits(gp), bits(op))


def _same_radius(g, o, q, r, mp):
    a = g.nn_radius(*q, np.float32(r * r), mp)
    b = o.nn_radius(*q, np.float32(r * r), mp)
    assert np.array_equal(a[0], b[0]), f"offsets differ r={r} maxp={mp}"
    assert np.array_equal(a[3], b[3])
    assert np.array_equal(bits(a[2]), bits(b[2])) and np.array_equal(bits(a[1]), bits(b[1]))
    return int(b[0][-1])


@pytest.mark.parametrize("opts", [dict(), dict(max_points_per_voxel=3), dict(match_mean=1), dict(match_mean=1, max_points_per_voxel=20)])
def test_sparse_voxel_insert_radius(built, opts):
    g, o = _pair(1, 0.2, **opts)
    for k in range(3):
        sc = _scan("tiny", s=1.0 * k, scan_id=k)
        g.insert(*sc)
        o.insert(*sc)
        _same_export(g, o)
    assert np.array_equal(np.concatenate(g.bbox()), np.concatenate(o.bbox()))
    gm, gcov = g.voxel_stats()
    om, ocov = o.voxel_stats()
    assert np.array_equal(bits(gm), bits(om))
    assert np.allclose(gcov, ocov, rtol=1e-5, atol=1e-9)
    q = compose_host(sc[3], sc[0] + np.float32(0.03), sc[1], sc[2])
    hits = 0
    for r, mp in ((0.15, 0), (0.45, 0), (0.45, 3), (0.8, 16), (0.3, 1)):
        hits += _same_radius(g, o, q, r, mp)
    assert hits > 0


def test_sparse_voxel_c3_knn_extension(built):
    """BASELINE config 3 (extension): k=8 stencil k-NN == brute force over the 27-cell stencil."""
    from mola_b200 import MapError, SparseVoxelPointCloud
    from oracle import mmo
    m = SparseVoxelPointCloud(0.2)
    o = mmo.OracleMap(1, 0.2, which="port")  # the k-NN extension exists in the port only (the reference throws)
    xs, ys, zs, pose = _scan("hdl64")
    m.insertPointCloud(xs, ys, zs, pose)
    o.insert(xs, ys, zs, pose)
    q = [c[::13].copy() for c in compose_host(pose, xs + np.float32(0.02), ys, zs)]
    with pytest.raises(MapError):
        m.nn_multiple_search_batch(*q, 8)
    m.enable_knn_extension = True
    cnt, pts, d2, ids = m.nn_multiple_search_batch(*q, 8)
    # brute force over the exported map, restricted to the 27-cell stencil of round-based indices
    keys, counts, P = o.export()
    lut = {tuple(k): i for i, k in enumerate(keys)}
    inv = np.float32(1.0) / np.float32(0.2)
    for j in range(0, q[0].size, 97):
        c = [int(np.rint(np.float32(q[a][j]) * inv)) for a in range(3)]
        cand = []
        for dx in (-1, 0, 1):
            for dy in (-1, 0, 1):
                for dz in (-1, 0, 1):
                    i = lut.get((c[0] + dx, c[1] + dy, c[2] + dz))
                    if i is None:
                        continue
                    for s in range(counts[i]):
                        p = P[i, s]
                        dd = [np.float32(p[a]) - np.float32(q[a][j]) for a in range(3)]
                        cand.append((np.float32(np.float32(dd[0] * dd[0] + dd[1] * dd[1]) + dd[2] * dd[2]), len(cand)))
        cand.sort()
        k = min(8, len(cand))
        assert cnt[j] == k
        assert np.array_equal(d2[j, :k], np.array([c_[0] for c_ in cand[:k]], np.float32))
    # the C++ port's extension mode (what bench_configs.py times) agrees on every query, ids included
    o.enable_knn_extension()
    oc, op, od, oi = o.nn_multi(*q, 8)
    mk = np.arange(8)[None, :] < oc[:, None]
    assert np.array_equal(cnt, oc) and np.array_equal(ids[mk], oi[mk]) and np.array_equal(bits(d2[mk]), bits(od[mk]))
    assert np.array_equal(bits(pts[mk]), bits(op[mk]))


@pytest.mark.parametrize("opts", [dict(), dict(minimum_points_clearance=0.5), dict(minimum_points_clearance=0.05, remove_voxels_farther_than=25.0)])
def test_sparse_trees(built, opts):
    g, o = _pair(2, 10.0, **opts)
    for k in range(4):
        sc = _scan("tiny", s=9.0 * k, scan_id=k)
        g.insert(*sc)
        o.insert(*sc)
        _same_export(g, o)
        assert g.num_points() == o.num_points()
    assert np.array_equal(np.concatenate(g.bbox()), np.concatenate(o.bbox()))
    q = compose_host(sc[3], sc[0] + np.float32(0.07), sc[1] - np.float32(0.03), sc[2])
    a, b = g.nn_single(*q), o.nn_single(*q)
    assert np.array_equal(a[0], b[0])
    k = b[0] > 0
    assert k.any()
    assert np.array_equal(a[3][k], b[3][k]) and np.array_equal(bits(a[2][k]), bits(b[2][k])) and np.array_equal(bits(a[1][k]), bits(b[1][k]))
    for kk in (1, 3, 8, 32):
        ma, mb = g.nn_multi(*q, kk), o.nn_multi(*q, kk)
        assert np.array_equal(ma[0], mb[0])
        msk = np.arange(kk)[None, :] < mb[0][:, None]
        assert np.array_equal(ma[3][msk], mb[3][msk]) and np.array_equal(bits(ma[2][msk]), bits(mb[2][msk]))
        assert np.array_equal(bits(ma[1][msk]), bits(mb[1][msk]))
    for r, mp in ((0.2, 0), (1.0, 0), (2.0, 0), (2.0, 4), (6.0, 2)):
        _same_radius(g, o, q, r, mp)
    la, lb = g.likelihood(*sc[:3], sc[3]), o.likelihood(*sc[:3], sc[3])
    assert abs(la - lb) <= 1e-12 * abs(lb)


def test_sparse_trees_full_scan_far_queries(built):
    """Exactness of the branch-and-bound search: 120k-pt scans, queries far from any stored point,
    negative coordinates and the double-width block 0."""
    g, o = _pair(2, 10.0, minimum_points_clearance=0.1)
    for k in range(2):
        sc = _scan("hdl64", s=7.0 * k, scan_id=k)
        g.insert(*sc)
        o.insert(*sc)
    _same_export(g, o)
    rng = np.random.default_rng(5)
    n = 4000
    q = [rng.uniform(-60, 60, n).astype(np.float32), rng.uniform(-25, 25, n).astype(np.float32), rng.uniform(-4, 12, n).astype(np.float32)]
    a, b = g.nn_single(*q), o.nn_single(*q)
    assert np.array_equal(a[0], b[0])
    k = b[0] > 0
    assert k.sum() > n // 4
    assert np.array_equal(a[3][k], b[3][k]) and np.array_equal(bits(a[2][k]), bits(b[2][k]))
    ma, mb = g.nn_multi(*q, 5), o.nn_multi(*q, 5)
    assert np.array_equal(ma[0], mb[0])
    msk = np.arange(5)[None, :] < mb[0][:, None]
    assert np.array_equal(ma[3][msk], mb[3][msk]) and np.array_equal(bits(ma[2][msk]), bits(mb[2][msk]))
    qs = [c[:300].copy() for c in q]
    _same_radius(g, o, qs, 1.5, 0)
    _same_radius(g, o, qs, 3.0, 7)


def test_sparse_trees_clone_clear_empty(built):
    from mola_b200 import SparseTreesPointCloud
    m = SparseTreesPointCloud(10.0)
    e = np.zeros(0, np.float32)
    m.insertPointCloud(e, e, e)
    assert m.isEmpty() and m.numVoxels() == 0 and m.numPoints() == 0
    f, p, d, i = m.nn_single_search_batch(np.ones(3, np.float32), np.ones(3, np.float32), np.ones(3, np.float32))
    assert not f.any()
    off, p, d, i = m.nn_radius_search_batch(np.ones(3, np.float32), np.ones(3, np.float32), np.ones(3, np.float32), 4.0)
    assert off[-1] == 0
    xs, ys, zs, pose = _scan("tiny")
    m.insertPointCloud(xs, ys, zs, pose)
    c = m.clone()
    n0 = m.numPoints()
    m.clear()
    assert m.isEmpty() and c.numPoints() == n0 and n0 > 0
    k1, c1, p1, _ = c.exportVoxels()
    assert int(c1.sum()) == n0


@pytest.mark.parametrize("opts", [dict(), dict(max_points_per_voxel=3)])
def test_sparse_voxel_insert_scans(built, opts):
    """Fused multi-scan insert for SparseVoxel: ordered rows and running means equal scan-by-scan insertion."""
    g, o = _pair(1, 0.2, **opts)
    scans = [_scan("tiny", s=1.5 * k, scan_id=k) for k in range(4)]
    xs, ys, zs = (np.concatenate([s[a] for s in scans]) for a in range(3))
    off = np.concatenate([[0], np.cumsum([s[0].size for s in scans])]).astype(np.uint64)
    g.insert(*scans[1])
    o.insert(*scans[1])
    g.m.insertPointClouds(xs, ys, zs, off, np.stack([s[3] for s in scans]))
    for s in scans:
        o.insert(*s)
    _same_export(g, o)
    gm, _ = g.voxel_stats()
    om, _ = o.voxel_stats()
    assert np.array_equal(bits(gm), bits(om))


def test_sparse_trees_insert_scans_falls_back(built):
    g, o = _pair(2, 10.0)
    scans = [_scan("tiny", s=1.5 * k, scan_id=k) for k in range(3)]
    xs, ys, zs = (np.concatenate([s[a] for s in scans]) for a in range(3))
    off = np.concatenate([[0], np.cumsum([s[0].size for s in scans])]).astype(np.uint64)
    g.m.insertPointClouds(xs, ys, zs, off, np.stack([s[3] for s in scans]))
    for s in scans:
        o.insert(*s)
    _same_export(g, o)


def test_sparse_trees_match_points(built):
    g, o = _pair(2, 10.0, minimum_points_clearance=0.05)
    sc = _scan("tiny")
    g.insert(*sc)
    o.insert(*sc)
    q = compose_host(sc[3], sc[0] + np.float32(0.04), sc[1], sc[2])
    of, op, od, oi = o.nn_single(*q)
    t2 = np.float32(np.median(od[of > 0]))  # a threshold that some pairs reach exactly: '<' is strict
    keep = (of > 0) & (od < t2)
    idx, pts, d2, ids = g.m.match_points_batch(*q, max_dist_sqr=float(t2))
    assert 0 < idx.size < q[0].size
    assert np.array_equal(idx, np.nonzero(keep)[0].astype(np.uint32)) and np.array_equal(ids, oi[keep])
    assert np.array_equal(bits(d2), bits(od[keep])) and np.array_equal(bits(pts), bits(op[keep]))


@pytest.mark.parametrize("kind,size,cap", [(0, 1.0, 0), (0, 1.0, 5), (1, 1.0, 0), (1, 1.0, 7)])
def test_insert_hot_voxels_out_of_order(built, kind, size, cap):
    """The free-slot set of k_insert_rank under the heaviest contention: 1.5 M points of 27 voxels in one batch, so every
    voxel is hit by thousands of warps at once and the earliest points of a voxel routinely arrive after later ones have
    taken its slots (replace-the-minimum path, published bounds, retries). Three batches: the first leaves rows partially
    filled, the second fills them, the third must be rejected by the probe alone. Hashed (cap 32) and SparseVoxel
    (cap 16, running means), hard cap and max_points_per_voxel."""
    rng = np.random.default_rng(11 + kind + cap)
    opts = dict(max_points_per_voxel=cap) if cap else dict()
    g, o = _pair(kind, size, **opts)
    for n in (60, 1_500_000, 300_000):
        cells = rng.integers(0, 3, size=(n, 3)).astype(np.float32)
        p = cells + rng.random((n, 3), dtype=np.float32) * np.float32(0.8) + np.float32(0.1)
        if kind == 1:
            p -= np.float32(0.5)  # SparseVoxel cells are centred on k * size
        x, y, z = (np.ascontiguousarray(p[:, a]) for a in range(3))
        g.insert(x, y, z)
        o.insert(x, y, z)
        _same_export(g, o)
    if kind == 1:
        gm, _ = g.voxel_stats()
        om, _ = o.voxel_stats()
        assert np.array_equal(bits(gm), bits(om))
