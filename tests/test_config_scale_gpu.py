"""BASELINE configs 4 and 3 at config scale, checked against the reference library (oracle/_ref) itself.

C4 (SparseTreesPointCloud, grid 10 m, clearance 0.05): the GPU map holds > 5 M points in > 300 blocks (multi-row cell
chains, thousands of hits per query at r = 2 m). The reference's CPU build of such a map is impractical (its per-block
nearest-neighbour test is a linear pass here: MRPT's kd-tree is not on the image), but SparseTreesPointCloud decomposes
by block: the clearance test of a point looks at the PRE-batch points of its OWN block only
(SparseTreesPointCloud.cpp:916-959 with :613-628), and a radius search reads only the blocks its index box names
(:659-700). So the reference map is built from exactly those input points that fall into a block subset S (same scans,
same order, composed with the same arithmetic), and
  * the blocks of S must come out bit-identical in both maps (block-local point order included), and
  * queries whose search box lies inside S must get bit-identical answers from the full GPU map and the S-only
    reference map, for the whole sweep r in {0.2, 0.5, 1, 1.5, 2} x maxPoints in {0, 16}.
C3 (SparseVoxelPointCloud 0.2 m, 40 HDL-64 scans as in tests/bench_configs.py): whole-map export, running means and
radius searches against the reference.
"""
import threading

import numpy as np
import pytest

from conftest import bits, compose_host

pytestmark = pytest.mark.gpu

G = np.float32(10.0)
G_HALF = np.float32(0.5) * G            # grid_size_half_
G_INV = np.float32(1.0) / G             # grid_size_inv_


def outer_idx(v):
    """coordToOuterIdx (SparseTreesPointCloud.h:70-77): (int32)((x + g/2) * (1/g)), fp32, truncation."""
    return ((v.astype(np.float32) + G_HALF) * G_INV).astype(np.int32)


def in_blocks(x, y, z, S):
    k = np.stack([outer_idx(x), outer_idx(y), outer_idx(z)], 1)
    m = np.zeros(x.size, bool)
    for b in S:
        m |= (k[:, 0] == b[0]) & (k[:, 1] == b[1]) & (k[:, 2] == b[2])
    return m


def _same_radius(a, b):
    assert np.array_equal(a[0], b[0]), "CSR offsets differ"
    assert np.array_equal(a[3], b[3]), "ids differ"
    assert np.array_equal(bits(a[2]), bits(b[2])) and np.array_equal(bits(a[1]), bits(b[1]))
    return int(b[0][-1])


def test_c4_sparse_trees_config_scale(built):
    from mola_b200 import SparseTreesPointCloud, synth
    from oracle import mmo
    which = mmo.checker()
    NSCANS, SPACING = 420, 3.8
    g = SparseTreesPointCloud(10.0)
    g.insertionOptions.minimum_points_clearance = 0.05
    # block subset S: 2 x 2 x 2 blocks around the middle of the trajectory (queries near their common corner touch all 8)
    mid = synth.lidar_scan("hdl64", s=SPACING * (NSCANS // 2), seed=4, scan_id=NSCANS // 2)
    c = [int(outer_idx(np.array([mid[3][9 + a]], np.float32))[0]) for a in range(3)]
    S = [(c[0] + i, c[1] + j, k) for i in (0, 1) for j in (0, 1) for k in (0, 1)]
    # a second, scattered subset: insert parity only, one reference map per block, in parallel threads
    rng = np.random.default_rng(44)
    o = mmo.OracleMap(2, 10.0, which=which, minimum_points_clearance=0.05)
    sub_inputs = []
    for k in range(NSCANS):
        xs, ys, zs, pose = synth.lidar_scan("hdl64", s=SPACING * k, seed=4, scan_id=k)
        g.insertPointCloud(xs, ys, zs, pose)
        gx, gy, gz = compose_host(pose, xs, ys, zs)
        if abs(k - NSCANS // 2) * SPACING < 160.0:  # scans that can reach S at all
            m = in_blocks(gx, gy, gz, S)
            if m.any():
                sub_inputs.append((gx[m].copy(), gy[m].copy(), gz[m].copy()))
    keys, counts, pts, cap = g.exportVoxels()
    stored, nblocks = int(counts.sum()), int(keys.shape[0])
    assert stored == g.numPoints()
    assert stored >= 5_000_000 and nblocks >= 300, (stored, nblocks)
    # the reference twin of S: the same points in the same order (already composed: identity pose)
    for gx, gy, gz in sub_inputs:
        o.insert(gx, gy, gz, None)
    ok_, oc_, op_ = o.export()
    assert ok_.shape[0] >= 6
    row = {tuple(k): i for i, k in enumerate(keys.tolist())}
    checked = 0
    for i, k in enumerate(ok_.tolist()):
        j = row[tuple(k)]
        n = int(oc_[i])
        assert int(counts[j]) == n, f"block {k}: {counts[j]} vs {n} points"
        assert np.array_equal(bits(pts[j, :n]), bits(op_[i, :n])), f"block {k}: stored points differ"
        checked += n
    assert checked >= 100_000
    assert int(counts.max()) > 32 * 32  # blocks far beyond one row of one cell
    # queries: points of re-posed scans around the middle, kept when their search box stays inside S
    qs = [[], [], []]
    for j in range(6):
        s = synth.lidar_scan("hdl64", s=SPACING * (NSCANS // 2) + 0.9 + 1.7 * j, seed=4, scan_id=50000 + j)
        q = compose_host(synth.perturbed_pose(s[3]), *s[:3])
        r = np.float32(2.0)
        m = in_blocks(q[0] - r, q[1] - r, q[2] - r, S) & in_blocks(q[0] + r, q[1] + r, q[2] + r, S)
        for a in range(3):
            qs[a].append(q[a][m])
    q = [np.ascontiguousarray(np.concatenate(x)[:3000]) for x in qs]
    assert q[0].size >= 2000, q[0].size
    threads = max(1, len(__import__("os").sched_getaffinity(0)))
    total = 0
    for r in (0.2, 0.5, 1.0, 1.5, 2.0):
        for mp in (0, 16):
            r2 = np.float32(r) * np.float32(r)
            ro = o.nn_radius(*q, r2, mp, nthreads=threads)
            rg = g.nn_radius_search_batch(*q, float(r2), mp)
            total += _same_radius(rg, ro)
    assert total > 5_000_000  # millions of hits compared
    # nn_single / nn_multiple see the query's own block only: same twin
    a, b = g.nn_single_search_batch(*q), o.nn_single(*q)
    assert np.array_equal(a[0], b[0])
    k = b[0] > 0
    assert k.mean() > 0.9
    assert np.array_equal(a[3][k], b[3][k]) and np.array_equal(bits(a[2][k]), bits(b[2][k])) and np.array_equal(bits(a[1][k]), bits(b[1][k]))


def test_c4_insert_parity_scattered_blocks(built):
    """Insert parity on blocks spread over the whole config-scale map: one reference map per sampled block (the
    clearance test never looks outside a point's own block), built in parallel host threads."""
    from mola_b200 import SparseTreesPointCloud, synth
    from oracle import mmo
    which = mmo.checker()
    NSCANS, SPACING = 420, 3.8
    g = SparseTreesPointCloud(10.0)
    g.insertionOptions.minimum_points_clearance = 0.05
    scans = []
    for k in range(NSCANS):
        xs, ys, zs, pose = synth.lidar_scan("hdl64", s=SPACING * k, seed=4, scan_id=k)
        g.insertPointCloud(xs, ys, zs, pose)
        if k % 3 == 0:
            scans.append((k, compose_host(pose, xs, ys, zs)))
    # blocks are sampled from a map built from every third scan only (cheaper on the CPU side): build that map too
    g3 = SparseTreesPointCloud(10.0)
    g3.insertionOptions.minimum_points_clearance = 0.05
    for k, (gx, gy, gz) in scans:
        g3.insertPointCloud(gx, gy, gz, None)
    keys, counts, pts, cap = g3.exportVoxels()
    assert keys.shape[0] >= 300
    rng = np.random.default_rng(7)
    pick = rng.choice(keys.shape[0], 24, replace=False)
    results = {}

    def work(j):
        S = [tuple(int(v) for v in keys[j])]
        o = mmo.OracleMap(2, 10.0, which=which, minimum_points_clearance=0.05)
        for k, (gx, gy, gz) in scans:
            m = in_blocks(gx, gy, gz, S)
            if m.any():
                o.insert(gx[m].copy(), gy[m].copy(), gz[m].copy(), None)
        results[j] = o.export()
    th = [threading.Thread(target=work, args=(int(j),)) for j in pick]
    for t in th:
        t.start()
    for t in th:
        t.join()
    npts = 0
    for j in pick:
        ok_, oc_, op_ = results[int(j)]
        assert ok_.shape[0] == 1 and tuple(ok_[0]) == tuple(keys[j])
        n = int(oc_[0])
        assert int(counts[j]) == n
        assert np.array_equal(bits(pts[j, :n]), bits(op_[0, :n]))
        npts += n
    assert npts > 50_000
    assert g.numPoints() >= 5_000_000


def test_c3_sparse_voxel_bench_size(built):
    """BASELINE config 3 at the size tests/bench_configs.py runs it (40 HDL-64 scans, voxel 0.2 m): whole-map export
    (per-voxel ordered points), running means and radius searches (plain and match_mean, with the std::map semantics of
    maxPoints != 0) against the reference; the k-NN / covariance extension against the port (the reference throws)."""
    from mola_b200 import SparseVoxelPointCloud, synth
    from oracle import mmo
    which = mmo.checker()
    g = SparseVoxelPointCloud(0.2)
    o = mmo.OracleMap(1, 0.2, which=which)
    for k in range(40):
        s = synth.lidar_scan("hdl64", s=2.0 * k, seed=3, scan_id=k)
        g.insertPointCloud(*s)
        o.insert(*s)
    gk, gc, gp, _ = g.exportVoxels()
    ok_, oc_, op_ = o.export()
    assert gk.shape[0] > 100_000
    assert np.array_equal(gk, ok_) and np.array_equal(gc, oc_) and np.array_equal(bits(gp), bits(op_))
    gm, _ = g.voxelStats()
    om, _ = o.voxel_stats()
    assert np.array_equal(bits(gm), bits(om))
    s = synth.lidar_scan("hdl64", s=2.0 * 20 + 0.7, seed=3, scan_id=90000)
    q = [np.ascontiguousarray(c[::24]) for c in compose_host(synth.perturbed_pose(s[3]), *s[:3])]
    assert q[0].size >= 5000
    threads = max(1, len(__import__("os").sched_getaffinity(0)))
    hits = 0
    for mean in (0, 1):
        g.likelihoodOptions.match_mean = bool(mean)
        o.set(match_mean=mean)
        for r, mp in ((0.2, 0), (0.5, 0), (0.5, 8), (1.0, 16), (1.0, 0)):
            r2 = np.float32(r) * np.float32(r)
            # SparseVoxel queries of the reference are not thread-safe (block cache): one thread
            ro = o.nn_radius(*q, r2, mp, nthreads=1)
            rg = g.nn_radius_search_batch(*q, float(r2), mp)
            hits += _same_radius(rg, ro)
    assert hits > 1_000_000
