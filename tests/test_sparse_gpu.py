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


def test_sparse_trees_radius_duplicates_ties_and_chunks(built):
    """std::map<float> semantics under stress (SparseTreesPointCloud.cpp:663-707): thousands of duplicated
    points (equal d2 inside a block: the highest block-local index survives), mirror images across a block boundary
    (equal d2 in two blocks: the later block survives), queries with more hits than the sort kernel's shared memory holds
    (global-memory path), and a batch cut into many chunks of queries by a small hit limit (the round-1 library refused
    batches of 2^31 hits instead of chunking)."""
    import ctypes as C
    from mola_b200 import capi
    g, o = _pair(2, 10.0, minimum_points_clearance=0.0)
    rng = np.random.default_rng(11)
    base = np.stack([rng.uniform(2.0, 8.0, 350), rng.uniform(-3.0, 3.0, 350), rng.uniform(0.0, 2.0, 350)], 1).astype(np.float32)
    base = np.round(base * 64) / 64                                    # exactly representable offsets
    mirror = base.copy()
    mirror[:, 0] = np.float32(10.0) - base[:, 0]                        # x = 5 is the boundary between blocks 0 and 1
    pts = np.concatenate([np.repeat(base, 60, axis=0), np.repeat(mirror, 60, axis=0)]).astype(np.float32)
    pts = pts[rng.permutation(pts.shape[0])]
    xs, ys, zs = (np.ascontiguousarray(pts[:, a]) for a in range(3))
    g.insert(xs, ys, zs, None)                                          # one batch: no intra-batch clearance test
    o.insert(xs, ys, zs, None)
    assert g.num_points() == pts.shape[0] == o.num_points()
    _same_export(g, o)
    nq = 240
    q = [np.full(nq, 5.0, np.float32), rng.uniform(-2, 2, nq).astype(np.float32), rng.uniform(0, 2, nq).astype(np.float32)]
    q[0][nq // 2:] = rng.uniform(3, 7, nq - nq // 2).astype(np.float32)   # half on the boundary plane, half around it
    lib = capi.lib()
    lib.mmg_set_tuning.argtypes = [C.c_char_p, C.c_int]
    try:
        for limit in (1 << 26, 300_000):
            capi.check(lib.mmg_set_tuning(b"radius_chunk_hits", limit))
            for r, mp in ((6.0, 0), (6.0, 16), (6.0, 100), (1.0, 0), (1.0, 3)):
                hits = _same_radius(g, o, q, r, mp)
                assert hits > 0
        # r = 6 m: every query sees tens of thousands of raw hits (> 16384: global-memory sort) that collapse to <= 700
        raw = g.nn_radius(*q, np.float32(36.0), 0)
        assert int(np.diff(raw[0]).max()) <= 700
        # small maxPoints take the per-block bounded k-nearest path (k_trees_radius_topk): same answers as the reference and
        # as the hit-list path, with 60-fold duplicates at every distance and mirror ties across the block boundary
        # (r = 2 m: 8 blocks per query, 8 x 32 = 256 candidates = the path's limit; r = 6 m, maxPoints 9: 27 x 9 = 243)
        for r, mp in ((2.0, 1), (2.0, 7), (2.0, 32), (3.5, 16), (6.0, 9)):
            capi.check(lib.mmg_set_tuning(b"radius_topk", 1))
            assert _same_radius(g, o, q, r, mp) > 0
            a = g.nn_radius(*q, np.float32(r * r), mp)
            capi.check(lib.mmg_set_tuning(b"radius_topk", 0))
            b = g.nn_radius(*q, np.float32(r * r), mp)
            assert all(np.array_equal(np.asarray(x).view(np.uint8), np.asarray(y).view(np.uint8)) for x, y in zip(a, b))
    finally:
        capi.check(lib.mmg_set_tuning(b"radius_chunk_hits", 1 << 26))
        capi.check(lib.mmg_set_tuning(b"radius_topk", 1))


def test_sparse_voxel_radius_split_batches(built):
    """A SparseVoxel radius batch whose hit list outgrows one pass is answered in pieces of queries (the round-1 library
    refused it): forced here with a small per-pass limit, results identical to the single pass and to the reference."""
    import ctypes as C
    from mola_b200 import capi
    g, o = _pair(1, 0.2)
    for k in range(2):
        sc = _scan("hdl64", s=1.0 * k, scan_id=k)
        g.insert(*sc)
        o.insert(*sc)
    q = [c[::40].copy() for c in compose_host(sc[3], sc[0] + np.float32(0.03), sc[1], sc[2])]
    lib = capi.lib()
    lib.mmg_set_tuning.argtypes = [C.c_char_p, C.c_int]
    try:
        for limit in (0x7FFFFFFF, 20_000):
            capi.check(lib.mmg_set_tuning(b"sv_split_hits", limit))
            for r, mp in ((0.5, 0), (0.5, 8), (0.3, 2)):
                assert _same_radius(g, o, q, r, mp) > (20_000 if mp == 0 else 0)
    finally:
        capi.check(lib.mmg_set_tuning(b"sv_split_hits", 0x7FFFFFFF))


def test_point_to_plane_pairing(built):
    """Extension: point-to-plane pairing against the per-voxel statistics of a SparseVoxel
    map, pinned by a CPU restatement: own voxel by round-to-nearest index, >= min_points stored points, distance to the
    voxel's running mean below the threshold, plane normal = eigenvector of the smallest eigenvalue of the stored points'
    covariance (numpy eigh, fp64), 1e-5 on the normal (up to the sign convention), the plane distance and the eigenvalues."""
    from mola_b200 import MapError, HashedVoxelPointCloud, SparseVoxelPointCloud, synth
    g = SparseVoxelPointCloud(0.4)
    scans = [_scan("hdl64", s=0.7 * k, scan_id=k) for k in range(3)]
    for sc in scans:
        g.insertPointCloud(*sc)
    keys, counts, pts, cap = g.exportVoxels()
    mean, _ = g.voxelStats()
    table = {tuple(k): i for i, k in enumerate(keys.tolist())}
    xs, ys, zs, pose = scans[1]
    qpose = synth.perturbed_pose(pose, dx=0.05, dyaw_deg=0.2)
    sel = slice(0, None, 7)
    lx, ly, lz = xs[sel].copy(), ys[sel].copy(), zs[sel].copy()
    q = np.stack(compose_host(qpose, lx, ly, lz), 1)
    thr2, minp = np.float32(0.09), 6
    idx, gm, gn, gd, ge = g.match_planes_batch(lx, ly, lz, pose=qpose, max_dist_sqr=float(thr2), min_points=minp)
    inv = np.float32(1.0) / np.float32(0.4)
    cell = np.rint(q * inv).astype(np.int64)
    exp_idx, checked = [], 0
    j = 0
    for i in range(q.shape[0]):
        v = table.get(tuple(cell[i].tolist()))
        if v is None or counts[v] < minp:
            continue
        d = (q[i] - mean[v]).astype(np.float32)
        if not (np.float32(np.float32(d[0] * d[0] + d[1] * d[1]) + d[2] * d[2]) < thr2):
            continue
        P = pts[v, :counts[v]].astype(np.float64)
        c = P.mean(0)
        cov = (P - c).T @ (P - c) / (counts[v] - 1)
        w, V = np.linalg.eigh(cov)
        if not (w[2] > 0):
            continue
        exp_idx.append(i)
        assert j < idx.size and idx[j] == i, (i, j)
        assert np.array_equal(bits(gm[j]), bits(mean[v]))
        assert abs(ge[j, 0] - w[0]) <= 1e-5 * w[2] + 1e-12 and abs(ge[j, 1] - w[1]) <= 1e-5 * w[2] + 1e-12
        if w[1] - w[0] > 1e-3 * w[2]:  # a well-defined plane: the normal is unique up to sign
            n = V[:, 0]
            a = int(np.argmax(np.abs(n)))
            n = n if n[a] > 0 else -n
            if abs(abs(n[a]) - np.sort(np.abs(n))[1]) > 1e-3:  # sign convention unambiguous
                assert np.allclose(gn[j], n, atol=1e-5), (gn[j], n)
                assert abs(gd[j] - float(n @ (q[i].astype(np.float64) - c))) <= 1e-5
                checked += 1
        j += 1
    assert j == idx.size and len(exp_idx) > 2000 and checked > 1000
    # device buffers: the same pairings
    import torch
    d = [torch.from_numpy(a).cuda() for a in (lx, ly, lz)]
    r2 = g.match_planes_batch(d[0], d[1], d[2], pose=qpose, max_dist_sqr=float(thr2), min_points=minp)
    assert np.array_equal(r2[0].cpu().numpy().view(np.uint32), idx) and np.array_equal(bits(r2[2].cpu().numpy()), bits(gn))
    with pytest.raises(MapError):
        HashedVoxelPointCloud(0.5).match_planes_batch(lx, ly, lz)
