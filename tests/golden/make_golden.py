"""Generates the committed golden vectors under tests/golden/ by running the REFERENCE's own code
(oracle/_ref/libmmo_ref.so: reference sources compiled unmodified against shim/mrpt) on small seeded
inputs. Run here (where /root/reference exists):  python tests/golden/make_golden.py
The inputs are stored too, so the fixtures do not depend on the synthetic generator."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from mola_b200 import synth  # noqa: E402
from oracle import mmo  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))

CASES = {
    # name: (kind, size, options, radius list [(r, maxPoints)], k list)
    "hashed_default": (0, 0.5, dict(), [(0.4, 0), (1.1, 0), (1.1, 2)], [1, 2, 3]),
    "hashed_decim": (0, 0.5, dict(max_points_per_voxel=6, min_distance_between_points=0.06), [(0.8, 0), (0.8, 3)], [3]),
    "hashed_evict": (0, 0.4, dict(remove_voxels_farther_than=12.0), [(0.5, 0)], [2]),
    "sparsevoxel_default": (1, 0.2, dict(), [(0.15, 0), (0.5, 0), (0.5, 4)], []),
    "sparsevoxel_mean": (1, 0.25, dict(match_mean=1, max_points_per_voxel=5), [(0.6, 0), (0.6, 3)], []),
    "sparsetrees_default": (2, 10.0, dict(minimum_points_clearance=0.2), [(0.5, 0), (2.0, 0), (2.0, 3)], [1, 4]),
    "sparsetrees_evict": (2, 8.0, dict(minimum_points_clearance=0.05, remove_voxels_farther_than=20.0), [(1.5, 0)], [2]),
}


def main():
    for name, (kind, size, opts, radii, ks) in CASES.items():
        m = mmo.OracleMap(kind, size, which="reference", **opts)
        d = {"kind": kind, "size": np.float32(size), "opt_names": np.array(list(opts.keys())), "opt_vals": np.array([float(v) for v in opts.values()])}
        nsc = 4
        for k in range(nsc):
            xs, ys, zs, pose = synth.lidar_scan("tiny", s=5.0 * k, seed=11, scan_id=k)
            m.insert(xs, ys, zs, pose)
            d[f"in{k}_xs"], d[f"in{k}_ys"], d[f"in{k}_zs"], d[f"in{k}_pose"] = xs, ys, zs, pose
        d["nscans"] = nsc
        d["keys"], d["counts"], d["points"] = m.export()
        d["bbox"] = np.concatenate(m.bbox())
        T = synth.perturbed_pose(pose, 0.12, 0.7)
        x, y, z = xs.astype(np.float64), ys.astype(np.float64), zs.astype(np.float64)
        q = [(((T[3 * r] * x + T[3 * r + 1] * y) + T[3 * r + 2] * z) + T[9 + r]).astype(np.float32)[::2] for r in range(3)]
        d["qx"], d["qy"], d["qz"] = q
        if kind != 1:
            d["nn1_found"], d["nn1_pts"], d["nn1_d2"], d["nn1_ids"] = m.nn_single(*q)
            d["loglik"] = np.float64(m.likelihood(xs, ys, zs, T))
        else:
            d["mean"], _ = m.voxel_stats()
        for kk in ks:
            d[f"nnk{kk}_count"], d[f"nnk{kk}_pts"], d[f"nnk{kk}_d2"], d[f"nnk{kk}_ids"] = m.nn_multi(*q, kk)
        d["ks"] = np.array(ks, np.int64)
        d["radii"] = np.array(radii, np.float64).reshape(-1, 2)
        for j, (r, mp) in enumerate(radii):
            d[f"rad{j}_off"], d[f"rad{j}_pts"], d[f"rad{j}_d2"], d[f"rad{j}_ids"] = m.nn_radius(*q, np.float32(r * r), int(mp))
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
        print(name, "voxels", len(d["counts"]), "points", int(d["counts"].sum()))


if __name__ == "__main__":
    main()
