"""ctypes binding of oracle/mmo.h for the two CPU checkers. TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs, never by
mola_b200/ (the product)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PORT_PATH = os.path.join(_HERE, "_build", "libmmo_port.so")
REF_PATH = os.path.join(_HERE, "_ref", "libmmo_ref.so")


class Params(C.Structure):
    _fields_ = [("voxel_size", C.c_float), ("max_points_per_voxel", C.c_uint32),
                ("remove_voxels_farther_than", C.c_double), ("min_distance_between_points", C.c_float),
                ("minimum_points_clearance", C.c_float), ("sigma_dist", C.c_double),
                ("max_corr_distance", C.c_double), ("decimation", C.c_uint32), ("match_mean", C.c_uint32)]


_libs = {}


def load(which="port"):
    """which: 'port' (our restatement, built on demand) or 'reference' (oracle/_ref, prebuilt)."""
    if which not in _libs:
        path = PORT_PATH if which == "port" else REF_PATH
        if which == "port" and not os.path.exists(path):
            import subprocess
            os.makedirs(os.path.dirname(path), exist_ok=True)
            subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-fPIC", "-shared", "-o", path,
                                   os.path.join(_HERE, "mmo_port.cpp"), "-lpthread"])
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        lib = C.CDLL(path)
        lib.mmo_create.restype = C.c_void_p
        lib.mmo_impl_name.restype = C.c_char_p
        lib.mmo_num_voxels.restype = C.c_uint64
        lib.mmo_num_points.restype = C.c_uint64
        lib.mmo_row_capacity.restype = C.c_uint32
        lib.mmo_mean_stencil_points.restype = C.c_double
        _libs[which] = lib
    return _libs[which]


def have_reference():
    return os.path.exists(REF_PATH)


def checker():
    """The CPU twin the parity tests compare against: the reference's own code (oracle/_ref, reference sources
    compiled against the MRPT stand-in) wherever it was built, the port otherwise. MMO_CHECKER=port|reference overrides."""
    w = os.environ.get("MMO_CHECKER")
    if w in ("port", "reference"):
        return w
    return "reference" if have_reference() else "port"


def _f(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _pose(p):
    if p is None:
        return None, None
    k = np.ascontiguousarray(np.asarray(p, np.float64).reshape(12))
    return k, k.ctypes.data_as(C.POINTER(C.c_double))


class OracleMap:
    """Same surface as mola_b200.maps._Map, on the CPU."""

    def __init__(self, kind, size, which=None, **opts):
        self.which = which or checker()
        self.lib = load(self.which)
        self.kind = kind
        self.p = Params(voxel_size=size, max_points_per_voxel=0, remove_voxels_farther_than=0.0,
                        min_distance_between_points=0.0, minimum_points_clearance=0.20, sigma_dist=0.5,
                        max_corr_distance=1.0, decimation=10, match_mean=0)
        for k, v in opts.items():
            setattr(self.p, k, v)
        self.h = C.c_void_p(self.lib.mmo_create(kind, C.byref(self.p)))
        assert self.h.value

    def __del__(self):
        if getattr(self, "h", None) is not None and self.h.value:
            self.lib.mmo_destroy(self.h)
            self.h = C.c_void_p()

    def set(self, **opts):
        for k, v in opts.items():
            setattr(self.p, k, v)
        self.lib.mmo_set_params(self.h, C.byref(self.p))

    def clear(self):
        self.lib.mmo_clear(self.h)

    def enable_knn_extension(self, on=True):
        """Port only: SparseVoxel k-NN (extension N4) instead of the reference's throw."""
        if self.lib.mmo_enable_knn_extension(self.h, 1 if on else 0):
            raise RuntimeError("the k-NN extension exists for the SparseVoxel port only")

    def isEmpty(self):
        return bool(self.lib.mmo_is_empty(self.h))

    @staticmethod
    def _xyz(xs, ys, zs):
        return (np.ascontiguousarray(xs, np.float32), np.ascontiguousarray(ys, np.float32),
                np.ascontiguousarray(zs, np.float32))

    def insert(self, xs, ys, zs, pose=None, has_robot_pose=None):
        xs, ys, zs = self._xyz(xs, ys, zs)
        k, pp = _pose(pose)
        hp = (pose is not None) if has_robot_pose is None else has_robot_pose
        self.lib.mmo_insert(self.h, _f(xs), _f(ys), _f(zs), C.c_size_t(xs.size), pp, 1 if hp else 0)

    def nn_single(self, xs, ys, zs, nthreads=1):
        xs, ys, zs = self._xyz(xs, ys, zs)
        n = xs.size
        found, pts = np.zeros(n, np.uint8), np.zeros((n, 3), np.float32)
        d2, ids = np.zeros(n, np.float32), np.zeros(n, np.uint64)
        rc = self.lib.mmo_nn_single(self.h, _f(xs), _f(ys), _f(zs), C.c_size_t(n), C.c_void_p(found.ctypes.data), _f(pts),
                                    _f(d2), C.c_void_p(ids.ctypes.data), nthreads)
        if rc:
            raise RuntimeError(f"oracle rc={rc}")
        return found, pts, d2, ids

    def nn_multi(self, xs, ys, zs, k, nthreads=1):
        xs, ys, zs = self._xyz(xs, ys, zs)
        n = xs.size
        cnt, pts = np.zeros(n, np.uint32), np.zeros((n, k, 3), np.float32)
        d2, ids = np.zeros((n, k), np.float32), np.zeros((n, k), np.uint64)
        rc = self.lib.mmo_nn_multi(self.h, _f(xs), _f(ys), _f(zs), C.c_size_t(n), C.c_size_t(k), C.c_void_p(cnt.ctypes.data),
                                   _f(pts), _f(d2), C.c_void_p(ids.ctypes.data), nthreads)
        if rc:
            raise RuntimeError(f"oracle rc={rc}")
        return cnt, pts, d2, ids

    def nn_radius(self, xs, ys, zs, r2, max_points=0, nthreads=1):
        xs, ys, zs = self._xyz(xs, ys, zs)
        n = xs.size
        offs = np.zeros(n + 1, np.uint64)
        rc = self.lib.mmo_nn_radius(self.h, _f(xs), _f(ys), _f(zs), C.c_size_t(n), C.c_float(r2), C.c_size_t(max_points),
                                    C.c_void_p(offs.ctypes.data), nthreads)
        if rc:
            raise RuntimeError(f"oracle rc={rc}")
        T = int(offs[-1])
        pts, d2, ids = np.zeros((T, 3), np.float32), np.zeros(T, np.float32), np.zeros(T, np.uint64)
        if T:
            self.lib.mmo_nn_radius_fetch(self.h, _f(pts), _f(d2), C.c_void_p(ids.ctypes.data))
        return offs, pts, d2, ids

    def likelihood(self, xs, ys, zs, pose=None):
        xs, ys, zs = self._xyz(xs, ys, zs)
        k, pp = _pose(pose)
        out = C.c_double()
        rc = self.lib.mmo_likelihood(self.h, _f(xs), _f(ys), _f(zs), C.c_size_t(xs.size), pp, C.byref(out))
        if rc:
            raise RuntimeError(f"oracle rc={rc}")
        return out.value

    def bbox(self):
        b = (C.c_float * 6)()
        self.lib.mmo_bbox(self.h, b)
        return np.array(b[:3], np.float32), np.array(b[3:], np.float32)

    def num_voxels(self):
        return self.lib.mmo_num_voxels(self.h)

    def num_points(self):
        return self.lib.mmo_num_points(self.h)

    def export(self):
        V, cap = self.num_voxels(), self.lib.mmo_row_capacity(self.h)
        keys, counts = np.zeros((V, 3), np.int32), np.zeros(V, np.uint32)
        pts = np.zeros((V, cap, 3), np.float32)
        if V:
            self.lib.mmo_export(self.h, C.c_void_p(keys.ctypes.data), C.c_void_p(counts.ctypes.data), _f(pts))
        return keys, counts, pts

    def voxel_stats(self):
        V = self.num_voxels()
        mean, cov = np.zeros((V, 3), np.float32), np.zeros((V, 6), np.float32)
        if V:
            self.lib.mmo_voxel_stats(self.h, _f(mean), _f(cov))
        return mean, cov

    def mean_stencil_points(self, xs, ys, zs):
        xs, ys, zs = self._xyz(xs, ys, zs)
        return self.lib.mmo_mean_stencil_points(self.h, _f(xs), _f(ys), _f(zs), C.c_size_t(xs.size))
