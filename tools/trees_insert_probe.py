"""Host-side cost of SparseTreesPointCloud inserts: call overhead, an all-rejected scan, and new scans along a trajectory
(time until the call returns / until the device is done). A profiling aid: `python tools/trees_insert_probe.py [scans]`."""
import time, torch, numpy as np, sys
sys.path.insert(0, "/root/repo")
from mola_b200 import SparseTreesPointCloud, synth
g = SparseTreesPointCloud(10.0)
g.insertionOptions.minimum_points_clearance = 0.05
dev = torch.device("cuda", 0)
xs, ys, zs, pose = synth.lidar_scan("hdl64", s=0.0, seed=4, scan_id=0)
d = [torch.from_numpy(a).to(dev) for a in (xs, ys, zs)]
e = [torch.empty(0, dtype=torch.float32, device=dev) for _ in range(3)]
g.insertPointCloud(*d, pose)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(2000):
    g.insertPointCloud(*e, pose)
t1 = time.perf_counter()
print("empty insert call overhead us:", (t1 - t0) / 2000 * 1e6)
# same scan again and again: everything rejected (phase 1 + sync only)
for k in range(3):
    g.insertPointCloud(*d, pose)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(200):
    g.insertPointCloud(*d, pose)
torch.cuda.synchronize()
t1 = time.perf_counter()
print("repeat-scan insert (all rejected) us:", (t1 - t0) / 200 * 1e6)
st = torch.cuda.current_stream()
ms = []
NS = int(sys.argv[1]) if len(sys.argv) > 1 else 60
for k in range(1, NS):
    xs, ys, zs, pose = synth.lidar_scan("hdl64", s=2.0 * k, seed=4, scan_id=k)
    d2 = [torch.from_numpy(a).to(dev) for a in (xs, ys, zs)]
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    g.insertPointCloud(*d2, pose)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    ms.append(((t1 - t0) * 1e6, (t2 - t0) * 1e6))
a = np.array(ms[10:])
print("new scans: call returns after us (median)", np.median(a[:, 0]), " done after us", np.median(a[:, 1]))
for i in range(0, len(ms), 50):
    b = np.array(ms[i:i + 50])
    print(f"scans {i+1}..{i+len(b)}: done after us median {np.median(b[:,1]):.0f} mean {np.mean(b[:,1]):.0f} max {np.max(b[:,1]):.0f}")
print("points", g.numPoints())
