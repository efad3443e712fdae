"""Where does a per-scan insert call spend its time? Host time per mmg_insert_batch call for several batch sizes."""
import ctypes as C
import sys
import time
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from mola_b200 import HashedVoxelPointCloud, capi, synth

lib = capi.lib()
g = HashedVoxelPointCloud(0.5)
scans = [synth.lidar_scan("hdl64", s=5.0 * k, seed=2, scan_id=k) for k in range(40)]
dev = [[torch.from_numpy(a).cuda() for a in s[:3]] for s in scans]
for d, s in zip(dev[:8], scans[:8]):
    g.insertPointCloud(d[0], d[1], d[2], s[3])
torch.cuda.synchronize()
fp = lambda t: C.cast(t.data_ptr(), C.POINTER(C.c_float))
for n in (1000, 20000, 120000):
    t = []
    for k in range(8, 40):
        d, pose = dev[k], np.ascontiguousarray(scans[k][3])
        pp = pose.ctypes.data_as(C.POINTER(C.c_double))
        t0 = time.perf_counter()
        lib.mmg_insert_batch(g._h, fp(d[0]), fp(d[1]), fp(d[2]), C.c_size_t(n), pp, 1, 1, None)
        t.append(time.perf_counter() - t0)
    print(f"n={n}: C call median {1e6*np.median(t):.1f} us  min {1e6*min(t):.1f} us")
t = []
for k in range(8, 40):
    d = dev[k]
    t0 = time.perf_counter()
    g.insertPointCloud(d[0], d[1], d[2], scans[k][3])
    t.append(time.perf_counter() - t0)
print(f"python insertPointCloud 120k: median {1e6*np.median(t):.1f} us")

# the bench scenario: a clone of a large map, new territory with every scan
import bench
big = HashedVoxelPointCloud(0.5)
ns, _ = bench.map_scans(3_000_000, big.numPoints, lambda xs, ys, zs, pose: big.insertPointCloud(
    torch.from_numpy(xs).cuda(), torch.from_numpy(ys).cuda(), torch.from_numpy(zs).cuda(), pose))
more = [synth.lidar_scan("hdl64", s=5.0 * (ns + j), seed=2, scan_id=ns + j) for j in range(24)]
dm = [[torch.from_numpy(a).cuda() for a in s[:3]] for s in more]
m = big.clone()
torch.cuda.synchronize()
t = []
for d, s in zip(dm, more):
    t0 = time.perf_counter()
    m.insertPointCloud(d[0], d[1], d[2], s[3])
    t.append(1e6 * (time.perf_counter() - t0))
print("clone of 3M-pt map, new territory, us per call:", " ".join(f"{x:.0f}" for x in t))
t = []
for d, s in zip(dm, more):
    t0 = time.perf_counter()
    big.insertPointCloud(d[0], d[1], d[2], s[3])
    t.append(1e6 * (time.perf_counter() - t0))
print("original map, same scans, us per call:", " ".join(f"{x:.0f}" for x in t))
