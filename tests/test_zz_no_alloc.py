"""mmg_nn_single_batch_noalloc (static-scratch nn_single_search, include/mola_b200/mmg.h) against mmg_nn_single_batch,
through the C-ABI, bit for bit, on two calls in a row (the static work area keeps the previous call's contents), plus the
size limit. Kept in a file pytest collects last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_Q = 65536  # MMG_NOALLOC_MAX_QUERIES


def _nn(lib, capi, h, dq, n, noalloc, stream, pose=None):
    import torch
    dev = dq[0].device
    out = [torch.full((n,), 7, dtype=torch.uint8, device=dev), torch.full((n, 3), -5.0, dtype=torch.float32, device=dev),
           torch.full((n,), -5.0, dtype=torch.float32, device=dev), torch.full((n,), 99, dtype=torch.int64, device=dev)]
    fp = lambda t: C.cast(t.data_ptr(), C.POINTER(C.c_float))
    vp = lambda t: C.c_void_p(t.data_ptr())
    pk, pp = capi.pose_ptr(pose)
    if noalloc:
        rc = lib.mmg_nn_single_batch_noalloc(h, fp(dq[0]), fp(dq[1]), fp(dq[2]), C.c_size_t(n), pp, vp(out[0]), vp(out[1]), vp(out[2]), vp(out[3]), stream)
    else:
        rc = lib.mmg_nn_single_batch(h, fp(dq[0]), fp(dq[1]), fp(dq[2]), C.c_size_t(n), pp, vp(out[0]), vp(out[1]), vp(out[2]), vp(out[3]),
                                     capi.MEM_DEVICE, stream)
    capi.check(rc)
    torch.cuda.synchronize()
    return [o.cpu().numpy() for o in out]


@pytest.mark.parametrize("kind", ["hashed", "sparse_voxel"])
def test_noalloc_equals_nn_single_twice(built, kind):
    import torch
    from mola_b200 import HashedVoxelPointCloud, SparseVoxelPointCloud, capi, synth
    lib = capi.lib()
    g = HashedVoxelPointCloud(0.5) if kind == "hashed" else SparseVoxelPointCloud(0.2)
    for k in range(6):
        g.insertPointCloud(*synth.lidar_scan("hdl64", s=1.5 * k, seed=11, scan_id=k))
    assert g.numPoints() > 50_000
    dev = torch.device("cuda", 0)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    # two different batches, the second smaller than the first: stale queue entries of call 1 lie beyond call 2's tails
    batches = []
    for j, n in enumerate((MAX_Q, 20_011)):
        s = synth.lidar_scan("hdl64", s=3.1 + 2.2 * j, seed=11, scan_id=700 + j)
        q = [np.ascontiguousarray(np.resize(c, n)) for c in s[:3]]
        batches.append(([torch.from_numpy(a).to(dev) for a in q], n, synth.perturbed_pose(s[3])))
    for dq, n, pose in batches:
        ref = _nn(lib, capi, g._h, dq, n, False, stream, pose)
        got = _nn(lib, capi, g._h, dq, n, True, stream, pose)
        assert ref[0].mean() > 0.3  # a real workload: many queries find a neighbour, some do not
        for a, b in zip(ref, got):  # bit for bit (the tolerance of every nn_single parity test here: none)
            assert np.array_equal(a.view(np.uint8), b.view(np.uint8))
    # and the first batch again after the smaller one
    dq, n, pose = batches[0]
    ref, got = _nn(lib, capi, g._h, dq, n, False, stream, pose), _nn(lib, capi, g._h, dq, n, True, stream, pose)
    for a, b in zip(ref, got):
        assert np.array_equal(a.view(np.uint8), b.view(np.uint8))


def test_noalloc_rejects_sizes_beyond_the_static_area(built):
    import torch
    from mola_b200 import HashedVoxelPointCloud, SparseTreesPointCloud, capi, synth
    lib = capi.lib()
    g = HashedVoxelPointCloud(0.5)
    g.insertPointCloud(*synth.lidar_scan("hdl64", s=0.0, seed=11, scan_id=0))
    dev = torch.device("cuda", 0)
    n = MAX_Q + 1
    dq = [torch.zeros(n, dtype=torch.float32, device=dev) for _ in range(3)]
    with pytest.raises(capi.MmgError) as e:
        _nn(lib, capi, g._h, dq, n, True, None)
    assert e.value.code == capi.MMG_ERR_INVALID_ARG and "MMG_NOALLOC_MAX_QUERIES" in str(e.value)
    # n = 0 is a no-op; SparseTrees maps are refused (their nn_single has no scratch of this kind)
    assert lib.mmg_nn_single_batch_noalloc(g._h, None, None, None, C.c_size_t(0), None, None, None, None, None, None) == capi.MMG_OK
    t = SparseTreesPointCloud(10.0)
    with pytest.raises(capi.MmgError) as e:
        _nn(lib, capi, t._h, dq, 16, True, None)
    assert e.value.code == capi.MMG_ERR_UNSUPPORTED
