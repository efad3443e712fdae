"""Small insert + query workload for `compute-sanitizer --tool racecheck` (shared-memory hazards of the insert
and query kernels): three scans into a Hashed and a SparseVoxel map (cap-only path), one with min-distance (sorted path),
then one batched query each. Sizes are small: racecheck slows kernels down by two orders of magnitude."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from mola_b200 import HashedVoxelPointCloud, SparseVoxelPointCloud, synth

for cls, size, opts in ((HashedVoxelPointCloud, 0.5, {}), (SparseVoxelPointCloud, 0.2, {}), (HashedVoxelPointCloud, 0.5, {"min_distance_between_points": 0.05})):
    m = cls(size)
    for k, v in opts.items():
        setattr(m.insertionOptions, k, v)
    for k in range(3):
        xs, ys, zs, pose = synth.lidar_scan("hdl64", s=2.0 * k, scan_id=k)
        xs, ys, zs = (np.ascontiguousarray(a[::8]) for a in (xs, ys, zs))
        m.insertPointCloud(xs, ys, zs, pose)
    if cls is HashedVoxelPointCloud:
        f, p, d, i = m.nn_single_search_batch(xs, ys, zs, pose=synth.perturbed_pose(pose))
        print(cls.__name__, opts, m.numPoints(), "points,", int(f.sum()), "of", f.size, "queries matched")
    else:
        off, p, d, i = m.nn_radius_search_batch(xs[:2000], ys[:2000], zs[:2000], 0.16, 4, pose=pose)
        print(cls.__name__, m.numPoints(), "points,", int(off[-1]), "radius hits")
print("racecheck workload done")
