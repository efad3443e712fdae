"""Per-call trace of bench.py's map-build leg: which mmg_insert_batch calls are slow, and why (table / row-store growth).
Prints wall and CUDA-event time of every call above 0.3 ms plus the totals. A profiling aid."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench
from mola_b200 import HashedVoxelPointCloud, synth

dev = torch.device("cuda", 0)
g = HashedVoxelPointCloud(bench.VOXEL)
stream = torch.cuda.current_stream(dev)
rows = []
t_all = time.perf_counter()
for k in range(int(os.environ.get("TRACE_SCANS", "200"))):
    xs, ys, zs, pose = synth.lidar_scan("hdl64", s=bench.SPACING * k, seed=2, scan_id=k)
    d = [torch.from_numpy(a).to(dev) for a in (xs, ys, zs)]
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record(stream)
    g.insertPointCloud(d[0], d[1], d[2], pose)
    e1.record(stream)
    t1 = time.perf_counter()
    e1.synchronize()
    rows.append((k, 1e3 * (t1 - t0), e0.elapsed_time(e1)))
st = g.tableStats()
print(f"{len(rows)} scans, wall {time.perf_counter() - t_all:.2f} s; call wall sum {sum(r[1] for r in rows):.1f} ms, event sum {sum(r[2] for r in rows):.1f} ms; "
      f"voxels {st['live_voxels']}, capacity {st['capacity']}")
print("median call wall %.3f ms, event %.3f ms" % (np.median([r[1] for r in rows]), np.median([r[2] for r in rows])))
for k, w, e in rows:
    if w > 0.3 or e > 0.3:
        print(f"  scan {k}: wall {w:.3f} ms, event {e:.3f} ms")
