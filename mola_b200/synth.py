"""Deterministic synthetic LiDAR scans (ctypes binding of csrc/synth.cpp; host only).

Shapes follow SURVEY.md §8(d): HDL-64 = 64 rings x 1875 azimuth steps = 120,000 points,
OS1-128 = 128 x 2048 = 262,144 points, cast into a procedural street scene.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

SENSORS = {
    "hdl64": dict(rings=64, az=1875, top=2.0, bot=-24.8, n=120000),
    "os1_128": dict(rings=128, az=2048, top=22.5, bot=-22.5, n=262144),
    "tiny": dict(rings=16, az=128, top=10.0, bot=-25.0, n=2048),
}


def _lib():
    global _LIB
    if _LIB is None:
        p = os.path.join(_HERE, "_lib", "libmola_b200_synth.so")
        if not os.path.exists(p):
            from . import build
            build.build_synth()
        _LIB = C.CDLL(p)
        _LIB.mbs_lidar_scan.restype = C.c_size_t
    return _LIB


def trajectory_pose(s):
    px, py, yaw = C.c_double(), C.c_double(), C.c_double()
    _lib().mbs_trajectory_pose(C.c_double(s), C.byref(px), C.byref(py), C.byref(yaw))
    return px.value, py.value, yaw.value


def lidar_scan(sensor="hdl64", s=0.0, seed=1, scan_id=0, noise=0.02, max_range=120.0, n=None):
    """One scan taken at arc-length s of the procedural trajectory.
    Returns (xs, ys, zs) in the sensor frame and the sensor pose (12 doubles: R row-major, t)."""
    cfg = SENSORS[sensor]
    n = cfg["n"] if n is None else int(n)
    xs, ys, zs = (np.empty(n, np.float32) for _ in range(3))
    pose = np.empty(12, np.float64)
    px, py, yaw = trajectory_pose(s)
    fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    got = _lib().mbs_lidar_scan(cfg["rings"], cfg["az"], C.c_float(cfg["top"]), C.c_float(cfg["bot"]), C.c_double(px),
                                C.c_double(py), C.c_double(yaw), C.c_uint64(seed), C.c_uint64(scan_id), C.c_float(noise),
                                C.c_float(max_range), C.c_size_t(n), fp(xs), fp(ys), fp(zs),
                                pose.ctypes.data_as(C.POINTER(C.c_double)))
    assert got == n, f"scene produced only {got} of {n} points"
    return xs, ys, zs, pose


def perturbed_pose(pose12, dx=0.10, dyaw_deg=0.5, dy=0.0):
    """pose12 composed with a small SE(3) step: mimics the pose guess of one ICP iteration."""
    R = pose12[:9].reshape(3, 3)
    t = pose12[9:]
    a = np.deg2rad(dyaw_deg)
    dR = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
    return np.concatenate([(R @ dR).reshape(9), t + R @ np.array([dx, dy, 0.0])])
