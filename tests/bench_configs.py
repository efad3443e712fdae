#!/usr/bin/env python
"""tests/bench_configs.py — the other BASELINE.json configurations (bench.py is config 2, the headline).
Lives under tests/ because every leg checks the CUDA path against the CPU oracle (test infrastructure) while timing it.
Run: python tests/bench_configs.py [--configs c1,c3,c4,c5]   (or under torchrun for the sharded legs)

  c1  one 120k-pt HDL-64 scan: HashedVoxelPointCloud (0.5 m) insert + nn_single_search, GPU vs the CPU reference, full parity
  c3  SparseVoxelPointCloud (0.2 m): nn_multiple_search k=8 (extension N4) + per-voxel mean/covariance
  c4  SparseTreesPointCloud (grid 10 m, clearance 0.05): nn_radius_search sweep r = 0.2 .. 2 m, maxPoints 0 / 16, queries sharded
  c5  64 OS1-128 scans (262,144 pts each) inserted in one fused call, then all of them queried against the replicated map

One JSON line per config on rank 0. Under torchrun the map is built on rank 0, broadcast over NCCL, and the
queries (c4: points, c5: whole scans) are sharded across ranks; times are CUDA-event times, max over ranks.
The CPU legs (oracle/_ref when built, else the port) run on rank 0 on bounded samples and are parity-checked.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (compose_host, cpu_threads, cpu_lib_kind)


def ev_time(stream, fn, reps=3, warm=1):
    """best-of-`reps` CUDA-event time (ms) of fn() on `stream`."""
    import torch
    for _ in range(warm):
        fn()
    best = None
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(stream)
        fn()
        e1.record(stream)
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    return best


def max_over_ranks(x, world, dev):
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def to_dev(arrs, dev):
    import torch
    return [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in arrs]


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


# ------------------------------------------------------------------------------------------ c1
def run_c1(ctx):
    import torch
    from mola_b200 import HashedVoxelPointCloud, synth
    from oracle import mmo
    if ctx.rank != 0:
        return None
    xs, ys, zs, pose = synth.lidar_scan("hdl64")
    qpose = synth.perturbed_pose(pose)
    kind = bench.cpu_lib_kind()
    g = HashedVoxelPointCloud(0.5, device=ctx.local_rank)
    g.insertPointCloud(xs, ys, zs, pose)      # allocations, first-launch costs
    g.nn_single_search_batch(xs, ys, zs, pose=qpose)
    # host buffers through the public API: this is what a single-scan caller pays
    t = []
    for _ in range(5):
        g.clear()
        t0 = time.perf_counter()
        g.insertPointCloud(xs, ys, zs, pose)
        t1 = time.perf_counter()
        f, p, d, i = g.nn_single_search_batch(xs, ys, zs, pose=qpose)
        t2 = time.perf_counter()
        t.append((t1 - t0, t2 - t1))
    ins_s, q_s = min(a for a, _ in t), min(b for _, b in t)
    # device-resident inputs, CUDA events
    dx = to_dev((xs, ys, zs), ctx.dev)

    def ins_dev():
        g.clear()
        g.insertPointCloud(dx[0], dx[1], dx[2], pose)
    ins_ms = ev_time(ctx.stream, ins_dev)
    q_ms = ev_time(ctx.stream, lambda: g.nn_single_search_batch(dx[0], dx[1], dx[2], pose=qpose))
    # CPU reference, 1 thread (the reference classes are single-threaded) and all threads for the const queries
    o = mmo.OracleMap(0, 0.5, which=kind)
    t0 = time.perf_counter()
    o.insert(xs, ys, zs, pose)
    cpu_ins = time.perf_counter() - t0
    qx, qy, qz = bench.compose_host(qpose, xs, ys, zs)
    t0 = time.perf_counter()
    of, op, od, oi = o.nn_single(qx, qy, qz)
    cpu_q1 = time.perf_counter() - t0
    thr = bench.cpu_threads()
    t0 = time.perf_counter()
    o.nn_single(qx, qy, qz, nthreads=thr)
    cpu_qn = time.perf_counter() - t0
    k = of > 0
    gk, gc, gp, _ = g.exportVoxels()
    ok_, oc, op_ = o.export()
    parity = bool(np.array_equal(gk, ok_) and np.array_equal(gc, oc) and np.array_equal(bits(gp), bits(op_)) and np.array_equal(f, of)
                  and np.array_equal(i[k], oi[k]) and np.array_equal(bits(d[k]), bits(od[k])) and np.array_equal(bits(p[k]), bits(op[k])))
    n = xs.size
    return {"config": "c1", "workload": "HashedVoxelPointCloud 0.5 m: insert + nn_single_search of one 120k-pt HDL-64 scan", "points": n,
            "voxels": int(g.numVoxels()), "found_frac": float(k.mean()), "parity_bit_exact": parity,
            "gpu": {"insert_ms_device_inputs": ins_ms, "nn_single_ms_device_inputs": q_ms, "insert_ms_host_api": 1e3 * ins_s,
                    "nn_single_ms_host_api": 1e3 * q_s, "insert_mpts_s": n / ins_ms / 1e3, "nn_single_qps": n / q_ms * 1e3,
                    "nn_single_qps_host_api": n / q_s},
            "cpu": {"kind": kind, "insert_ms_1thread": 1e3 * cpu_ins, "nn_single_ms_1thread": 1e3 * cpu_q1,
                    "nn_single_ms_all_threads": 1e3 * cpu_qn, "threads": thr, "nn_single_qps_1thread": n / cpu_q1,
                    "nn_single_qps_all_threads": n / cpu_qn}}


# ------------------------------------------------------------------------------------------ c3
def run_c3(ctx):
    from mola_b200 import SparseVoxelPointCloud, synth
    from oracle import mmo
    if ctx.rank != 0:
        return None
    ns = ctx.args.c3_scans
    scans = [synth.lidar_scan("hdl64", s=2.0 * k, seed=3, scan_id=k) for k in range(ns)]
    off = np.concatenate([[0], np.cumsum([s[0].size for s in scans])]).astype(np.uint64)
    poses = np.stack([s[3] for s in scans])
    d = to_dev([np.concatenate([s[a] for s in scans]) for a in range(3)], ctx.dev)
    g = SparseVoxelPointCloud(0.2, device=ctx.local_rank)
    g.enable_knn_extension = True

    def build():
        g.clear()
        g.insertPointClouds(d[0], d[1], d[2], off, poses)
    ins_ms = ev_time(ctx.stream, build, reps=2)
    qs = synth.lidar_scan("hdl64", s=2.0 * (ns // 2) + 0.7, seed=3, scan_id=9000)
    qpose = synth.perturbed_pose(qs[3])
    q = bench.compose_host(qpose, *qs[:3])
    dq = to_dev(q, ctx.dev)
    knn_ms = ev_time(ctx.stream, lambda: g.nn_multiple_search_batch(dq[0], dq[1], dq[2], 8))
    V = g.numVoxels()
    t0 = time.perf_counter()
    gm, gcov = g.voxelStats()
    stats_s = time.perf_counter() - t0
    # CPU twin (port: the k-NN / covariance extension has no reference behaviour)
    o = mmo.OracleMap(1, 0.2, which="port")  # k-NN extension: port only (the reference throws)
    o.enable_knn_extension()
    t0 = time.perf_counter()
    for s in scans:
        o.insert(*s)
    cpu_ins = time.perf_counter() - t0
    sample = min(q[0].size, 20000)
    t0 = time.perf_counter()
    oc, op, od, oi = o.nn_multi(q[0][:sample], q[1][:sample], q[2][:sample], 8)
    cpu_knn = time.perf_counter() - t0
    gc_, gp_, gd_, gi_ = (a.cpu().numpy() for a in g.nn_multiple_search_batch(dq[0], dq[1], dq[2], 8))
    gc_, gi_ = gc_.view(np.uint32)[:sample], gi_.view(np.uint64)[:sample]
    m = np.arange(8)[None, :] < oc[:, None]
    om, ocov = o.voxel_stats()
    parity = bool(np.array_equal(gc_, oc) and np.array_equal(gi_[m], oi[m]) and np.array_equal(bits(gd_[:sample][m]), bits(od[m]))
                  and np.array_equal(bits(gm), bits(om)) and np.allclose(gcov, ocov, rtol=1e-5, atol=1e-9))
    n_in, nq = int(off[-1]), q[0].size
    return {"config": "c3", "workload": f"SparseVoxelPointCloud 0.2 m, {ns} HDL-64 scans; nn_multiple_search k=8 (extension N4) of one 120k-pt scan; "
            "per-voxel mean + covariance", "points_in": n_in, "stored_points": int(g.numPoints()), "voxels": int(V),
            "parity": parity, "parity_note": "k-NN ids/d2 bit-exact on the CPU sample, means bit-exact, covariance rtol 1e-5",
            "gpu": {"insert_ms": ins_ms, "insert_mpts_s": n_in / ins_ms / 1e3, "knn8_ms": knn_ms, "knn8_qps": nq / knn_ms * 1e3,
                    "voxel_stats_ms_host_api": 1e3 * stats_s},
            "cpu": {"kind": "port", "threads": 1, "insert_mpts_s": n_in / cpu_ins / 1e6, "knn8_qps": sample / cpu_knn, "sample": sample}}


# ------------------------------------------------------------------------------------------ c4
def run_c4(ctx):
    import torch
    import torch.distributed as dist
    from mola_b200 import SparseTreesPointCloud, synth
    from mola_b200.maps import Communicator
    from oracle import mmo
    a = ctx.args
    g = SparseTreesPointCloud(10.0, device=ctx.local_rank)
    g.insertionOptions.minimum_points_clearance = 0.05
    build_s, nscans, n_in = 0.0, 0, 0
    per_scan = []
    if ctx.rank == 0:
        # warm-up: a throw-away map of 160 scans takes every kernel and growth step of the build once (first use of a kernel or
        # of a driver memset variant loads its module; on a fresh box that is a read from a cold page cache: 0.1 - 0.4 s each)
        w = SparseTreesPointCloud(10.0, device=ctx.local_rank)
        w.insertionOptions.minimum_points_clearance = 0.05
        for k in range(160):
            xs, ys, zs, pose = synth.lidar_scan("hdl64", s=a.c4_spacing * k, seed=5, scan_id=k)
            w.insertPointCloud(*to_dev((xs, ys, zs), ctx.dev), pose)
        w.nn_radius_search_batch(*to_dev((xs[:1000], ys[:1000], zs[:1000]), ctx.dev), 1.0, 0)
        del w
        torch.cuda.synchronize()
        while nscans < a.c4_max_scans:  # the insert alone is timed (CUDA events); scan synthesis and the H2D copy are not
            xs, ys, zs, pose = synth.lidar_scan("hdl64", s=a.c4_spacing * nscans, seed=4, scan_id=nscans)
            d = to_dev((xs, ys, zs), ctx.dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(ctx.stream)
            g.insertPointCloud(d[0], d[1], d[2], pose)
            e1.record(ctx.stream)
            e1.synchronize()
            per_scan.append(e0.elapsed_time(e1))
            build_s += 1e-3 * per_scan[-1]
            n_in += xs.size
            nscans += 1
            if nscans % 32 == 0 and g.numPoints() >= a.c4_points:
                break
    bcast_s = 0.0
    if ctx.world > 1:
        obj = [Communicator.unique_id() if ctx.rank == 0 else None, nscans]
        dist.broadcast_object_list(obj, src=0)
        comm = Communicator(obj[0], ctx.rank, ctx.world, ctx.local_rank)
        nscans = obj[1]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        g.broadcast(comm, 0)
        torch.cuda.synchronize()
        bcast_s = time.perf_counter() - t0
    # queries: scans re-posed mid-trajectory; rank r takes its contiguous shard of the batch (mola_b200.sharding)
    from mola_b200 import sharding
    nq_total = a.c4_queries
    qs = []
    j = 0
    while sum(x[0].size for x in qs) < nq_total:
        s = synth.lidar_scan("hdl64", s=a.c4_spacing * (nscans // 2) + 0.9 + 1.3 * j, seed=4, scan_id=50000 + j)
        qs.append(bench.compose_host(synth.perturbed_pose(s[3]), *s[:3]))
        j += 1
    q = [np.concatenate([x[k] for x in qs])[:nq_total] for k in range(3)]
    lo, hi = sharding.shard_range(nq_total, ctx.rank, ctx.world)
    dq = to_dev([x[lo:hi] for x in q], ctx.dev)
    sweep = []
    for r in (0.2, 0.5, 1.0, 1.5, 2.0):
        for mp in (0, 16):
            r2 = float(np.float32(r) * np.float32(r))
            holder = {}

            def run():
                holder["o"] = g.nn_radius_search_batch(dq[0], dq[1], dq[2], r2, mp)
            ms = max_over_ranks(ev_time(ctx.stream, run, reps=2), ctx.world, ctx.dev)
            hits = torch.tensor([int(holder["o"][0][-1])], dtype=torch.float64, device=ctx.dev)
            if ctx.world > 1:
                dist.all_reduce(hits)
            sweep.append({"r": r, "maxPoints": mp, "ms": ms, "qps": nq_total / ms * 1e3, "hits": int(hits[0])})
    nn_ms = max_over_ranks(ev_time(ctx.stream, lambda: g.nn_single_search_batch(dq[0], dq[1], dq[2]), reps=3), ctx.world, ctx.dev)
    if ctx.rank != 0:
        return None
    out = {"config": "c4", "workload": f"SparseTreesPointCloud grid 10 m, clearance 0.05: nn_radius_search sweep, {nq_total} queries sharded x{ctx.world}",
           "n_gpus": ctx.world, "stored_points": int(g.numPoints()), "blocks": int(g.numVoxels()), "scans": nscans, "points_in": n_in,
           "insert_s": build_s, "insert_mpts_s": n_in / build_s / 1e6 if build_s else None,
           "insert_ms_per_scan": {"median": float(np.median(per_scan)), "p90": float(np.percentile(per_scan, 90)), "max": float(np.max(per_scan)),
                                  "top10_sum": float(np.sort(per_scan)[-10:].sum())} if per_scan else None, "bcast_s": bcast_s, "nn_single_ms": nn_ms,
           "nn_single_qps": nq_total / nn_ms * 1e3, "sweep": sweep}
    # CPU parity + baseline on a reduced twin (the reference's per-block kd-tree rebuild makes a 50M-pt CPU build impractical here)
    if a.c4_cpu_scans > 0:
        kind = bench.cpu_lib_kind()
        g2 = SparseTreesPointCloud(10.0, device=ctx.local_rank)
        g2.insertionOptions.minimum_points_clearance = 0.05
        o = mmo.OracleMap(2, 10.0, which=kind, minimum_points_clearance=0.05)
        for k in range(a.c4_cpu_scans):
            s = synth.lidar_scan("hdl64", s=a.c4_spacing * k, seed=4, scan_id=k)
            g2.insertPointCloud(*s)
            o.insert(*s)
        s = synth.lidar_scan("hdl64", s=a.c4_spacing * (a.c4_cpu_scans // 2) + 0.9, seed=4, scan_id=50000)
        cq = [x[:a.c4_cpu_queries] for x in bench.compose_host(synth.perturbed_pose(s[3]), *s[:3])]
        cpu, ok = [], True
        for r, mp in ((0.2, 0), (1.0, 16), (2.0, 0)):
            r2 = np.float32(r) * np.float32(r)
            t0 = time.perf_counter()
            ro = o.nn_radius(*cq, r2, mp, nthreads=bench.cpu_threads())
            dt = time.perf_counter() - t0
            rg = g2.nn_radius_search_batch(*cq, float(r2), mp)
            ok = ok and np.array_equal(rg[0], ro[0]) and np.array_equal(rg[3], ro[3]) and np.array_equal(bits(rg[2]), bits(ro[2]))
            cpu.append({"r": r, "maxPoints": mp, "qps": cq[0].size / dt, "hits": int(ro[0][-1])})
        out["cpu"] = {"kind": kind, "threads": bench.cpu_threads(), "map_points": int(o.num_points()), "queries": int(cq[0].size), "sweep": cpu,
                      "parity_bit_exact_on_reduced_map": bool(ok)}
    return out


# ------------------------------------------------------------------------------------------ c5
def run_c5(ctx):
    import torch
    import torch.distributed as dist
    from mola_b200 import HashedVoxelPointCloud, synth
    from mola_b200.maps import Communicator
    from oracle import mmo
    ns = ctx.args.c5_scans
    scans = [synth.lidar_scan("os1_128", s=2.0 * k, seed=5, scan_id=k) for k in range(ns)]
    off = np.concatenate([[0], np.cumsum([s[0].size for s in scans])]).astype(np.uint64)
    poses = np.stack([s[3] for s in scans])
    g = HashedVoxelPointCloud(0.5, device=ctx.local_rank)
    ins_ms = None
    if ctx.rank == 0:
        d = to_dev([np.concatenate([s[a] for s in scans]) for a in range(3)], ctx.dev)

        def build():
            g.clear()
            g.insertPointClouds(d[0], d[1], d[2], off, poses)
        ins_ms = ev_time(ctx.stream, build, reps=3)
        del d
    bcast_ms = 0.0
    if ctx.world > 1:
        obj = [Communicator.unique_id() if ctx.rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        comm = Communicator(obj[0], ctx.rank, ctx.world, ctx.local_rank)
        g.broadcast(comm, 0)  # warm-up (NCCL channel set-up), then timed
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        g.broadcast(comm, 0)
        torch.cuda.synchronize()
        bcast_ms = max_over_ranks(1e3 * (time.perf_counter() - t0), ctx.world, ctx.dev)
    # queries: whole scans are dealt to the ranks (scan s -> rank s % world), each re-posed by one ICP-style step
    mine = [s for s in range(ns) if s % ctx.world == ctx.rank]
    dq = [to_dev(scans[s][:3], ctx.dev) for s in mine]
    qposes = [synth.perturbed_pose(scans[s][3]) for s in mine]
    res = {}

    def query():
        for k, s in enumerate(mine):
            res[s] = g.nn_single_search_batch(dq[k][0], dq[k][1], dq[k][2], pose=qposes[k])
    q_ms = max_over_ranks(ev_time(ctx.stream, query, reps=3), ctx.world, ctx.dev)
    # parity on one scan per rank against a CPU twin of the whole map (port; ~1 s per 4M inserted points)
    ok = True
    if ctx.args.c5_check:
        o = mmo.OracleMap(0, 0.5)
        for s in scans:
            o.insert(*s)
        s = mine[len(mine) // 2]
        qx, qy, qz = bench.compose_host(qposes[len(mine) // 2], *scans[s][:3])
        of, op, od, oi = o.nn_single(qx, qy, qz, nthreads=max(1, bench.cpu_threads() // ctx.world))
        f, p, dd, ii = (x.cpu().numpy() for x in res[s])
        k = of > 0
        ok = bool(np.array_equal(f, of) and np.array_equal(ii.view(np.uint64)[k], oi[k]) and np.array_equal(bits(dd[k]), bits(od[k])))
        if ctx.rank == 0:
            ok = ok and g.numPoints() == o.num_points() and g.numVoxels() == o.num_voxels()
    t = torch.tensor([1.0 if ok else 0.0], dtype=torch.float64, device=ctx.dev)
    if ctx.world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if ctx.rank != 0:
        return None
    n = int(off[-1])
    return {"config": "c5", "workload": f"{ns} OS1-128 scans (262,144 pts each): one fused insert into HashedVoxelPointCloud 0.5 m, NCCL replicate, "
            f"every scan queried (nn_single_search) on rank scan % {ctx.world}", "n_gpus": ctx.world, "points": n, "stored_points": int(g.numPoints()),
            "voxels": int(g.numVoxels()), "insert_ms": ins_ms, "insert_mpts_s": n / ins_ms / 1e3, "bcast_ms": bcast_ms, "query_ms": q_ms,
            "query_qps": n / q_ms * 1e3, "parity_bit_exact_sample": bool(t[0] > 0.5) if ctx.args.c5_check else None}


class Ctx:
    pass


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c1,c3,c4,c5")
    ap.add_argument("--c3-scans", type=int, default=40)
    ap.add_argument("--c4-points", type=int, default=50_000_000)
    ap.add_argument("--c4-max-scans", type=int, default=6000)
    ap.add_argument("--c4-spacing", type=float, default=2.0)
    ap.add_argument("--c4-queries", type=int, default=200_000)
    ap.add_argument("--c4-cpu-scans", type=int, default=6)
    ap.add_argument("--c4-cpu-queries", type=int, default=5000)
    ap.add_argument("--c5-scans", type=int, default=64)
    ap.add_argument("--c5-check", type=int, default=1)
    args = ap.parse_args()
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench_configs.py needs a CUDA device: the map kernels have no CPU fallback")
    ctx = Ctx()
    ctx.args = args
    ctx.rank, ctx.world, ctx.local_rank = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(ctx.local_rank)
    ctx.dev = torch.device("cuda", ctx.local_rank)
    ctx.stream = torch.cuda.current_stream(ctx.dev)
    if ctx.world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=ctx.dev)
    try:
        for c in args.configs.split(","):
            t0 = time.perf_counter()
            line = {"c1": run_c1, "c3": run_c3, "c4": run_c4, "c5": run_c5}[c.strip()](ctx)
            if line is not None:
                line["wall_s"] = time.perf_counter() - t0
                print(json.dumps(line), flush=True)
    finally:
        if ctx.world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
