"""Host-side mirror of the reference's map classes over the C ABI.

Same class names, option names, method names and error behaviour as
mola_metric_maps (HashedVoxelPointCloud.h:285-364, SparseVoxelPointCloud.h,
SparseTreesPointCloud.h); every query additionally has a `*_batch` form, which is
what the GPU is for (the single-query virtuals are 1-element batches).

Inputs may be numpy float32 arrays (host memory: the library stages H2D/D2H
itself) or torch CUDA tensors (device memory: results come back as torch CUDA
tensors, enqueued on torch's current stream).
"""
import ctypes as C
from types import SimpleNamespace

import numpy as np

from . import capi


class MapError(RuntimeError):
    """Raised where the reference throws (THROW_EXCEPTION / ASSERT_)."""


def pose_from_xyz_ypr(x, y, z, yaw=0.0, pitch=0.0, roll=0.0):
    """Row-major R (3x3) followed by t, as 12 doubles (mrpt::poses::CPose3D::FromXYZYawPitchRoll)."""
    cy, sy, cp, sp, cr, sr = np.cos(yaw), np.sin(yaw), np.cos(pitch), np.sin(pitch), np.cos(roll), np.sin(roll)
    R = np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                  [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                  [-sp, cp * sr, cp * cr]], dtype=np.float64)
    return np.concatenate([R.reshape(9), np.array([x, y, z], dtype=np.float64)])


def _is_torch(a):
    return hasattr(a, "data_ptr") and hasattr(a, "is_cuda")


class _Map:
    KIND = None
    NAME = ""

    def __init__(self, size, device=0):
        self._lib = capi.lib()
        self._p = capi.Params()
        capi.check(self._lib.mmg_default_params(self.KIND, C.byref(self._p)))
        self._p.voxel_size = float(size)
        self._h = C.c_void_p()
        capi.check(self._lib.mmg_create(self.KIND, C.byref(self._p), int(device), C.byref(self._h)))
        self.device = int(device)
        self.insertionOptions = SimpleNamespace()
        self.likelihoodOptions = SimpleNamespace(sigma_dist=0.5, max_corr_distance=1.0, decimation=10)
        self._init_options()

    # -- lifetime ---------------------------------------------------------
    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            self._lib.mmg_destroy(h)
            self._h = C.c_void_p()

    def clone(self):
        """Deep copy (the reference classes are copy-constructible)."""
        self._sync_params()
        o = object.__new__(type(self))
        o._lib, o._p, o.device = self._lib, capi.Params.from_buffer_copy(self._p), self.device
        o._h = C.c_void_p()
        capi.check(self._lib.mmg_clone(self._h, C.byref(o._h)))
        o.insertionOptions = SimpleNamespace(**vars(self.insertionOptions))
        o.likelihoodOptions = SimpleNamespace(**vars(self.likelihoodOptions))
        return o

    # -- options ----------------------------------------------------------
    def _init_options(self):
        raise NotImplementedError

    def _fill_params(self, p):
        raise NotImplementedError

    def _sync_params(self):
        # fast path (every insert / query call comes through here): nothing to do while the option values are the same objects
        key = (*vars(self.insertionOptions).values(), *vars(self.likelihoodOptions).values())
        if key == getattr(self, "_opt_key", None):
            return
        self._opt_key = key
        p = capi.Params.from_buffer_copy(self._p)
        lo = self.likelihoodOptions
        p.sigma_dist, p.max_corr_distance, p.decimation = float(lo.sigma_dist), float(lo.max_corr_distance), int(lo.decimation)
        self._fill_params(p)
        if bytes(p) != bytes(self._p):
            capi.check(self._lib.mmg_set_params(self._h, C.byref(p)))
            self._p = p

    # -- helpers ----------------------------------------------------------
    def _stream(self):
        import torch
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _xyz(self, xs, ys, zs):
        """Returns (mem, stream, keepalive, px, py, pz, n, torch_or_None)."""
        if _is_torch(xs):
            import torch
            assert xs.is_cuda and xs.dtype == torch.float32
            xs, ys, zs = xs.contiguous(), ys.contiguous(), zs.contiguous()
            return capi.MEM_DEVICE, self._stream(), (xs, ys, zs), capi.fptr(xs.data_ptr()), capi.fptr(ys.data_ptr()), \
                capi.fptr(zs.data_ptr()), xs.numel(), torch
        xs = np.ascontiguousarray(xs, dtype=np.float32)
        ys = np.ascontiguousarray(ys, dtype=np.float32)
        zs = np.ascontiguousarray(zs, dtype=np.float32)
        assert xs.shape == ys.shape == zs.shape and xs.ndim == 1
        return capi.MEM_HOST, None, (xs, ys, zs), capi.fptr(xs), capi.fptr(ys), capi.fptr(zs), xs.size, None

    def _alloc(self, th, shape, np_dtype, torch_dtype_name):
        if th is not None:
            return th.empty(shape, dtype=getattr(th, torch_dtype_name), device=f"cuda:{self.device}")
        return np.empty(shape, dtype=np_dtype)

    @staticmethod
    def _ptr(a):
        return C.c_void_p(a.data_ptr()) if _is_torch(a) else C.c_void_p(a.ctypes.data)

    # -- CMetricMap -------------------------------------------------------
    def clear(self):
        capi.check(self._lib.mmg_clear(self._h))

    def isEmpty(self):
        e = C.c_int()
        capi.check(self._lib.mmg_is_empty(self._h, C.byref(e)))
        return bool(e.value)

    def insertPointCloud(self, xs, ys, zs, robotPose=None):
        """insertObservation(CObservationPointCloud, robotPose): points are composed with
        robotPose (identity if None) and inserted in input order; a given robotPose also
        triggers the far-voxel eviction (HashedVoxelPointCloud.cpp:289-331,410-425).
        Returns the number of points whose voxel index was out of the supported range."""
        self._sync_params()
        mem, st, keep, px, py, pz, n, _ = self._xyz(xs, ys, zs)
        pk, pp = capi.pose_ptr(robotPose)
        rc = self._lib.mmg_insert_batch(self._h, px, py, pz, C.c_size_t(n), pp, 1 if robotPose is not None else 0, mem, st)
        capi.check(rc, allow=(capi.MMG_ERR_OUT_OF_RANGE,))
        return rc == capi.MMG_ERR_OUT_OF_RANGE

    def reserve(self, voxels):
        """voxels_.reserve(): pre-sizes the table and the row store (HashedVoxelPointCloud.cpp:1070)."""
        capi.check(self._lib.mmg_reserve(self._h, C.c_uint64(int(voxels))))

    def insertPointClouds(self, xs, ys, zs, scan_offsets, robotPoses, with_eviction=True):
        """A sequence of insertObservation calls in one batch: scan s is points
        [scan_offsets[s], scan_offsets[s+1]) with pose robotPoses[s] (12 doubles each). Same result as
        calling insertPointCloud once per scan in order; fused into one device pass when the options
        allow it (mmg_insert_scans)."""
        self._sync_params()
        mem, st, keep, px, py, pz, n, _ = self._xyz(xs, ys, zs)
        off = np.ascontiguousarray(scan_offsets, dtype=np.uint64)
        poses = np.ascontiguousarray(np.asarray(robotPoses, dtype=np.float64).reshape(-1, 12))
        assert off.ndim == 1 and off.size == poses.shape[0] + 1 and int(off[-1]) - int(off[0]) <= n
        rc = self._lib.mmg_insert_scans(self._h, px, py, pz, C.c_size_t(poses.shape[0]), off.ctypes.data_as(C.POINTER(C.c_uint64)),
                                        poses.ctypes.data_as(C.POINTER(C.c_double)), 1 if with_eviction else 0, mem, st)
        capi.check(rc, allow=(capi.MMG_ERR_OUT_OF_RANGE,))
        return rc == capi.MMG_ERR_OUT_OF_RANGE

    def internal_insertPointCloud3D(self, pc_in_map, xs, ys, zs):
        """The reference's private helper (:1065-1082): compose + insert, no eviction."""
        self._sync_params()
        mem, st, keep, px, py, pz, n, _ = self._xyz(xs, ys, zs)
        pk, pp = capi.pose_ptr(pc_in_map)
        capi.check(self._lib.mmg_insert_batch(self._h, px, py, pz, C.c_size_t(n), pp, 0, mem, st), allow=(capi.MMG_ERR_OUT_OF_RANGE,))

    def insertPoint(self, pt):
        x, y, z = pt
        self.insertPointCloud(np.array([x], np.float32), np.array([y], np.float32), np.array([z], np.float32))

    def boundingBox(self):
        b = (C.c_float * 6)()
        capi.check(self._lib.mmg_bbox(self._h, b))
        return np.array(b[:3], np.float32), np.array(b[3:], np.float32)

    def computeObservationLikelihood(self, xs, ys, zs, takenFrom=None):
        self._sync_params()
        mem, st, keep, px, py, pz, n, _ = self._xyz(xs, ys, zs)
        pk, pp = capi.pose_ptr(takenFrom)
        out = C.c_double()
        rc = self._lib.mmg_likelihood(self._h, px, py, pz, C.c_size_t(n), pp, C.byref(out), mem, st)
        if rc in (capi.MMG_ERR_UNSUPPORTED, capi.MMG_ERR_INVALID_ARG):
            raise MapError(self._lib.mmg_last_error().decode())
        capi.check(rc)
        return out.value

    # -- contents ---------------------------------------------------------
    def numVoxels(self):
        v = C.c_uint64()
        capi.check(self._lib.mmg_num_voxels(self._h, C.byref(v)))
        return v.value

    def numPoints(self):
        v = C.c_uint64()
        capi.check(self._lib.mmg_num_points(self._h, C.byref(v)))
        return v.value

    def rowCapacity(self):
        v = C.c_uint32()
        capi.check(self._lib.mmg_row_capacity(self._h, C.byref(v)))
        return v.value

    def exportVoxels(self):
        """(keys[V,3] int32, counts[V] uint32, points[V,cap,3] float32), sorted by (cx,cy,cz)."""
        V, cap = self.numVoxels(), self.rowCapacity()
        keys = np.zeros((V, 3), np.int32)
        counts = np.zeros(V, np.uint32)
        pts = np.zeros((V, cap, 3), np.float32)
        if V:
            capi.check(self._lib.mmg_export(self._h, capi.vptr(keys), capi.vptr(counts), capi.vptr(pts), capi.MEM_HOST, None))
        order = np.lexsort((keys[:, 2], keys[:, 1], keys[:, 0])) if V else np.zeros(0, np.int64)
        return keys[order], counts[order], pts[order], order

    def visitAllPoints(self, f):
        keys, counts, pts, _ = self.exportVoxels()
        for v in range(len(counts)):
            for j in range(counts[v]):
                f(pts[v, j])

    def tableStats(self):
        s = capi.TableStats()
        capi.check(self._lib.mmg_get_table_stats(self._h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in capi.TableStats._fields_}

    # -- NearestNeighborsCapable -----------------------------------------
    def nn_has_indices_or_ids(self):
        return False

    def nn_index_count(self):
        return 0

    def nn_prepare_for_3d_queries(self):
        pass

    def nn_single_search_batch(self, xs, ys, zs, pose=None):
        """Batched nn_single_search. Returns (found[n] uint8, pts[n,3], d2[n], ids[n] uint64)."""
        self._sync_params()
        mem, st, keep, px, py, pz, n, th = self._xyz(xs, ys, zs)
        found = self._alloc(th, (n,), np.uint8, "uint8")
        pts = self._alloc(th, (n, 3), np.float32, "float32")
        d2 = self._alloc(th, (n,), np.float32, "float32")
        ids = self._alloc(th, (n,), np.uint64, "int64")
        pk, pp = capi.pose_ptr(pose)
        rc = self._lib.mmg_nn_single_batch(self._h, px, py, pz, C.c_size_t(n), pp, self._ptr(found), self._ptr(pts), self._ptr(d2),
                                           self._ptr(ids), mem, st)
        self._raise_like_reference(rc)
        return found, pts, d2, ids

    def match_points_batch(self, xs, ys, zs, pose=None, max_dist_sqr=float("inf")):
        """Batched distance-threshold matcher (what mp2p_icp's Matcher_Points_DistanceThreshold does with one
        nn_single_search per point): returns (local_idx[m] uint32, map_pts[m,3], d2[m], ids[m]) for the points
        whose nearest map point lies at d2 < max_dist_sqr, in local-point order."""
        self._sync_params()
        mem, st, keep, px, py, pz, n, th = self._xyz(xs, ys, zs)
        idx = self._alloc(th, (n,), np.uint32, "int32")
        pts = self._alloc(th, (n, 3), np.float32, "float32")
        d2 = self._alloc(th, (n,), np.float32, "float32")
        ids = self._alloc(th, (n,), np.uint64, "int64")
        cnt = C.c_uint64()
        pk, pp = capi.pose_ptr(pose)
        rc = self._lib.mmg_match_batch(self._h, px, py, pz, C.c_size_t(n), pp, C.c_float(max_dist_sqr), C.byref(cnt), self._ptr(idx),
                                       self._ptr(pts), self._ptr(d2), self._ptr(ids), mem, st)
        self._raise_like_reference(rc)
        m = int(cnt.value)
        return idx[:m], pts[:m], d2[:m], ids[:m]

    def match_planes_batch(self, xs, ys, zs, pose=None, max_dist_sqr=float("inf"), min_points=5):
        """Point-to-plane pairing (SparseVoxelPointCloud only; extension): returns (local_idx[m], mean[m,3], normal[m,3],
        plane_dist[m], eig[m,2]) for the points whose own voxel stores >= min_points points and whose distance to the
        voxel's running mean is below sqrt(max_dist_sqr), in local-point order."""
        self._sync_params()
        mem, st, keep, px, py, pz, n, th = self._xyz(xs, ys, zs)
        idx = self._alloc(th, (n,), np.uint32, "int32")
        mean = self._alloc(th, (n, 3), np.float32, "float32")
        nrm = self._alloc(th, (n, 3), np.float32, "float32")
        dist = self._alloc(th, (n,), np.float32, "float32")
        eig = self._alloc(th, (n, 2), np.float32, "float32")
        cnt = C.c_uint64()
        pk, pp = capi.pose_ptr(pose)
        rc = self._lib.mmg_match_planes_batch(self._h, px, py, pz, C.c_size_t(n), pp, C.c_float(max_dist_sqr), C.c_uint32(int(min_points)),
                                              C.byref(cnt), self._ptr(idx), self._ptr(mean), self._ptr(nrm), self._ptr(dist), self._ptr(eig), mem, st)
        self._raise_like_reference(rc)
        m = int(cnt.value)
        return idx[:m], mean[:m], nrm[:m], dist[:m], eig[:m]

    def nn_multiple_search_batch(self, xs, ys, zs, N, pose=None):
        """Batched nn_multiple_search. Returns (count[n], pts[n,N,3], d2[n,N], ids[n,N])."""
        if N <= 0:
            raise MapError("ASSERT_(N > 0)")  # HashedVoxelPointCloud.cpp:702
        self._sync_params()
        mem, st, keep, px, py, pz, n, th = self._xyz(xs, ys, zs)
        cnt = self._alloc(th, (n,), np.uint32, "int32")
        pts = self._alloc(th, (n, N, 3), np.float32, "float32")
        d2 = self._alloc(th, (n, N), np.float32, "float32")
        ids = self._alloc(th, (n, N), np.uint64, "int64")
        pk, pp = capi.pose_ptr(pose)
        rc = self._lib.mmg_nn_multi_batch(self._h, px, py, pz, C.c_size_t(n), pp, C.c_size_t(N), self._ptr(cnt), self._ptr(pts),
                                          self._ptr(d2), self._ptr(ids), mem, st)
        self._raise_like_reference(rc)
        return cnt, pts, d2, ids

    def nn_radius_search_batch(self, xs, ys, zs, search_radius_sqr, maxPoints=0, pose=None):
        """Batched nn_radius_search, CSR: (offsets[n+1] uint64, pts[T,3], d2[T], ids[T])."""
        self._sync_params()
        mem, st, keep, px, py, pz, n, th = self._xyz(xs, ys, zs)
        offs = self._alloc(th, (n + 1,), np.uint64, "int64")
        total = C.c_uint64()
        pk, pp = capi.pose_ptr(pose)
        rc = self._lib.mmg_nn_radius_batch(self._h, px, py, pz, C.c_size_t(n), pp, C.c_float(search_radius_sqr), C.c_size_t(maxPoints),
                                           self._ptr(offs), C.byref(total), mem, st)
        self._raise_like_reference(rc)
        T = total.value
        pts = self._alloc(th, (T, 3), np.float32, "float32")
        d2 = self._alloc(th, (T,), np.float32, "float32")
        ids = self._alloc(th, (T,), np.uint64, "int64")
        if T:
            capi.check(self._lib.mmg_nn_radius_fetch(self._h, self._ptr(pts), self._ptr(d2), self._ptr(ids), mem, st))
        return offs, pts, d2, ids

    def _raise_like_reference(self, rc):
        if rc == capi.MMG_ERR_UNSUPPORTED:
            raise MapError(self._lib.mmg_last_error().decode())
        capi.check(rc)

    # single-query forms of the virtual interface (1-element batches)
    def nn_single_search(self, query):
        q = np.asarray(query, np.float32).reshape(3)
        f, p, d, i = self.nn_single_search_batch(q[0:1].copy(), q[1:2].copy(), q[2:3].copy())
        if not f[0]:
            return False, None, None, None
        return True, p[0], float(d[0]), int(i[0])

    def nn_multiple_search(self, query, N):
        q = np.asarray(query, np.float32).reshape(3)
        c, p, d, i = self.nn_multiple_search_batch(q[0:1].copy(), q[1:2].copy(), q[2:3].copy(), N)
        k = int(c[0])
        return p[0, :k], d[0, :k], i[0, :k]

    def nn_radius_search(self, query, search_radius_sqr, maxPoints=0):
        q = np.asarray(query, np.float32).reshape(3)
        o, p, d, i = self.nn_radius_search_batch(q[0:1].copy(), q[1:2].copy(), q[2:3].copy(), search_radius_sqr, maxPoints)
        return p, d, i

    def nn_single_search_2d(self, *a):
        raise MapError(f"Cannot run a 2D search on a {self.NAME}")  # HashedVoxelPointCloud.cpp:638-664

    nn_multiple_search_2d = nn_radius_search_2d = nn_single_search_2d

    # -- multi-GPU --------------------------------------------------------
    def broadcast(self, comm, root=0):
        """Replicates the map held by rank `root` onto every rank of `comm` (NCCL)."""
        self._sync_params()
        capi.check(self._lib.mmg_bcast(self._h, comm._h, int(root), None))
        capi.check(self._lib.mmg_get_params(self._h, C.byref(self._p)))


class HashedVoxelPointCloud(_Map):
    """mola::HashedVoxelPointCloud (include/mola_metric_maps/HashedVoxelPointCloud.h)."""
    KIND, NAME = capi.HASHED_VOXEL, "HashedVoxelPointCloud"
    SSO_MAX_POINTS_PER_VOXEL = 32
    HARD_MAX_MATCHES = 3

    def __init__(self, voxel_size=1.00, device=0):
        super().__init__(voxel_size, device)

    def _init_options(self):
        self.insertionOptions = SimpleNamespace(max_points_per_voxel=0, remove_voxels_farther_than=0.0,
                                                min_distance_between_points=0.0)

    def _fill_params(self, p):
        io = self.insertionOptions
        p.max_points_per_voxel = int(io.max_points_per_voxel) & 0xFFFFFFFF  # read as int into a uint32 (.cpp:1041)
        p.remove_voxels_farther_than = float(io.remove_voxels_farther_than)
        p.min_distance_between_points = float(io.min_distance_between_points)

    def setVoxelProperties(self, voxel_size):
        self._p.voxel_size = -1.0  # force the clear even for an equal size (.cpp:189-200)
        p = capi.Params.from_buffer_copy(self._p)
        p.voxel_size = float(voxel_size)
        capi.check(self._lib.mmg_set_params(self._h, C.byref(p)))
        self._p = p


class SparseVoxelPointCloud(_Map):
    """mola::SparseVoxelPointCloud. nn_single/multiple_search throw like the reference unless
    `enable_knn_extension` is set (k<=8 stencil k-NN + per-voxel mean/covariance)."""
    KIND, NAME = capi.SPARSE_VOXEL, "SparseVoxelPointCloud"
    HARDLIMIT_MAX_POINTS_PER_VOXEL = 16

    def __init__(self, voxel_size=0.20, device=0):
        self.enable_knn_extension = False
        super().__init__(voxel_size, device)

    def _init_options(self):
        self.insertionOptions = SimpleNamespace(max_points_per_voxel=0)
        self.likelihoodOptions.match_mean = False

    def _fill_params(self, p):
        p.max_points_per_voxel = int(self.insertionOptions.max_points_per_voxel) & 0xFFFFFFFF
        p.match_mean = 1 if self.likelihoodOptions.match_mean else 0

    def nn_single_search_batch(self, *a, **k):
        if not self.enable_knn_extension:
            raise MapError("n-search not available in this map type")  # SparseVoxelPointCloud.cpp:695-719
        return super().nn_single_search_batch(*a, **k)

    def nn_multiple_search_batch(self, *a, **k):
        if not self.enable_knn_extension:
            raise MapError("n-search not available in this map type")
        return super().nn_multiple_search_batch(*a, **k)

    def voxelStats(self):
        """(mean[V,3], cov[V,6]) in the order of exportVoxels()."""
        V = self.numVoxels()
        mean, cov = np.zeros((V, 3), np.float32), np.zeros((V, 6), np.float32)
        if V:
            capi.check(self._lib.mmg_voxel_stats(self._h, capi.vptr(mean), capi.vptr(cov), capi.MEM_HOST, None))
        _, _, _, order = self.exportVoxels()
        return mean[order], cov[order]


class SparseTreesPointCloud(_Map):
    """mola::SparseTreesPointCloud (grid of coarse blocks, exact NN inside the query's block)."""
    KIND, NAME = capi.SPARSE_TREES, "SparseTreesPointCloud"

    def __init__(self, grid_size=10.0, device=0):
        super().__init__(grid_size, device)

    def _init_options(self):
        self.insertionOptions = SimpleNamespace(minimum_points_clearance=0.20, remove_submaps_farther_than=0.0)

    def _fill_params(self, p):
        io = self.insertionOptions
        p.minimum_points_clearance = float(io.minimum_points_clearance)
        # read as uint64_t into a float member: fractions are truncated (SparseTreesPointCloud.cpp:893)
        p.remove_voxels_farther_than = float(np.float32(int(io.remove_submaps_farther_than)))


class Communicator:
    """NCCL communicator owned by the C library; the unique id is exchanged by the caller
    (e.g. torch.distributed.broadcast_object_list)."""

    def __init__(self, unique_id, rank, nranks, device):
        self._lib = capi.lib()
        self._h = C.c_void_p()
        buf = (C.c_uint8 * 128).from_buffer_copy(bytes(unique_id))
        capi.check(self._lib.mmg_comm_init(buf, int(rank), int(nranks), int(device), C.byref(self._h)))

    @staticmethod
    def unique_id():
        buf = (C.c_uint8 * 128)()
        capi.check(capi.lib().mmg_comm_unique_id(buf))
        return bytes(buf)

    def __del__(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.mmg_comm_destroy(self._h)
            self._h = C.c_void_p()
