#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (BASELINE.json config 2).

Workload: a ~10M-point HashedVoxelPointCloud (voxel 0.5 m, cap 32/voxel) built by inserting
synthetic HDL-64 scans along a procedural street, then batched nn_single_search of 1M query
points (9 consecutive re-posed scans, truncated) per step. One step = one pass of 1M queries.

  value  : queries/s with queries resident in HBM (device pointers through the C ABI)
  e2e    : queries/s through the C ABI with HOST (pinned) buffers, H2D + D2H inside the timed region
  roofline / cpu_baseline / clocks: see DESIGN.md "Measurement"

`--impl reference` times the reference's own CPU code (oracle/_ref when it was built, else the
oracle port) on the same map and a bounded sample of the same queries.
Under torchrun (N>1): rank 0 builds the map, broadcasts it with NCCL, every rank queries its own
1M-point batch (weak scaling); time = max over ranks.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = "C2: 10M-pt HashedVoxelPointCloud (voxel 0.5 m), batched nn_single_search of 1M points per step per GPU"
VOXEL = 0.5
MAP_POINTS = 10_000_000
N_QUERIES = 1_000_000
SPACING = 5.0          # metres between map scans
MAX_SCANS = 900


def make_config(map_points, map_voxels, map_scans, queries):
    """The `config` object, identical in both arms (ours and --impl reference) for the same workload."""
    return {"workload": WORKLOAD, "map_points": map_points, "map_voxels": map_voxels, "map_scans": map_scans,
            "queries_per_step_per_gpu": queries,
            "l2": "GPU arm: L2 flushed between steps (256 MiB write); CPU arm: the 10M-point map (>1 GB of buckets) exceeds the LLC"}


def source_stamp():
    """Hash of the CUDA sources and the C-ABI header: profiles/traffic.json carries the stamp of the tree its ncu capture
    ran on (tools/ncu_summary.py), and its counters are quoted only when that stamp is this tree's."""
    import glob
    import hashlib
    h = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(ROOT, "mola_b200", "csrc", "*")) + glob.glob(os.path.join(ROOT, "include", "mola_b200", "*.h"))):
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


def compose_host(pose12, xs, ys, zs):
    T = np.asarray(pose12, np.float64)
    x, y, z = xs.astype(np.float64), ys.astype(np.float64), zs.astype(np.float64)
    return [(((T[3 * r] * x + T[3 * r + 1] * y) + T[3 * r + 2] * z) + T[9 + r]).astype(np.float32) for r in range(3)]


def map_scans(target_points, count_fn, insert_fn, log=None):
    """Feeds scans to insert_fn until count_fn() >= target_points. Returns (#scans, #input points)."""
    from mola_b200 import synth
    n_in, k = 0, 0
    while k < MAX_SCANS:
        xs, ys, zs, pose = synth.lidar_scan("hdl64", s=SPACING * k, seed=2, scan_id=k)
        insert_fn(xs, ys, zs, pose)
        n_in += xs.size
        k += 1
        if k % 16 == 0 or k < 4:
            c = count_fn()
            if log:
                log(f"  scan {k}: {c} stored points")
            if c >= target_points:
                break
    return k, n_in


def query_batch(n_scans_in_map, rank=0, n=N_QUERIES, world=1):
    """n query points in the map frame: consecutive scans taken mid-trajectory, each re-posed by a small SE(3) step (one
    ICP iteration's pose guess), truncated to n. Several ranks shard "scan points x ICP iterations": every rank holds the
    same scans under its own pose guess (a 0.10 m step in its own direction, +-0.5 deg of yaw; rank 0 = the N = 1 batch),
    so all ranks work on the same stretch of road. (A stretch of road per rank, 50 m apart, made the slowest stretch set
    the step, 0.172-0.193 ms at N = 8; dealing one run of scans round-robin spread each rank's queries over 8 x the map
    area, 0.187-0.199 ms: profiles/r1i_bench_n8.json, r1j_bench_n8_round_robin_queries.json.)"""
    from mola_b200 import synth
    out = [[], [], []]
    have, j = 0, 0
    s0 = SPACING * (n_scans_in_map // 2) + 1.7
    phi = 2.0 * np.pi * rank / max(world, 1)
    while have < n:
        xs, ys, zs, pose = synth.lidar_scan("hdl64", s=s0 + 1.0 * j, seed=2, scan_id=100000 + j)
        guess = synth.perturbed_pose(pose, dx=0.10 * np.cos(phi), dy=0.10 * np.sin(phi), dyaw_deg=0.5 if rank % 2 == 0 else -0.5)
        g = compose_host(guess, xs, ys, zs)
        for a in range(3):
            out[a].append(g[a])
        have += xs.size
        j += 1
    return [np.ascontiguousarray(np.concatenate(o)[:n]) for o in out]


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def bind_to_gpu_numa_node(local_rank):
    """Pins this process to the CPUs next to its GPU (sysfs local_cpulist of the GPU's PCI function) before any pinned
    buffer is allocated, so the host side of the e2e copies stays on the GPU's NUMA node. torchrun does not bind ranks.
    Returns the cpulist string, or None when the topology cannot be read (nothing is changed then)."""
    try:
        out = subprocess.run(["nvidia-smi", f"--id={local_rank}", "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                             capture_output=True, text=True, timeout=20).stdout.strip().lower()
        bus = out[-12:] if len(out) >= 12 else out          # 00000000:1B:00.0 -> 0000:1b:00.0
        path = f"/sys/bus/pci/devices/{bus}/local_cpulist"
        cl = open(path).read().strip()
        cpus = set()
        for part in cl.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= set(os.sched_getaffinity(0))
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cl
    except Exception:
        return None


def cpu_threads():
    return max(1, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))


def cpu_lib_kind():
    from oracle import mmo
    return "reference" if mmo.have_reference() else "port"


def time_cpu_queries(omap, q, sample, threads, repeats=3):
    """queries/s of the CPU checker on the first `sample` queries, best of `repeats`."""
    best = None
    for _ in range(repeats):
        t = time.perf_counter()
        omap.nn_single(q[0][:sample], q[1][:sample], q[2][:sample], nthreads=threads)
        dt = time.perf_counter() - t
        best = dt if best is None else min(best, dt)
    return sample / best, best


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the same path on the host cores."""
    if rank != 0:
        return
    from oracle import mmo
    kind = cpu_lib_kind()
    log = (lambda s: print(s, file=sys.stderr, flush=True)) if args.verbose else None
    om = mmo.OracleMap(0, VOXEL, which=kind)
    t0 = time.perf_counter()
    ins_t = [0.0]

    def ins(xs, ys, zs, pose):
        t = time.perf_counter()
        om.insert(xs, ys, zs, pose)
        ins_t[0] += time.perf_counter() - t
    nscans, n_in = map_scans(args.map_points, om.num_points, ins, log)
    q = query_batch(nscans)
    threads = cpu_threads()
    sample = N_QUERIES if args.cpu_sample <= 0 else min(N_QUERIES, args.cpu_sample)  # default: the full 1M-query step
    for _ in range(args.warmup):
        om.nn_single(q[0][:sample], q[1][:sample], q[2][:sample], nthreads=threads)
    t = time.perf_counter()
    for _ in range(args.steps):
        om.nn_single(q[0][:sample], q[1][:sample], q[2][:sample], nthreads=threads)
    dt = time.perf_counter() - t
    qps = sample * args.steps / dt
    qps1, _ = time_cpu_queries(om, q, min(sample, 100_000), 1, repeats=2)
    line = {"impl": "reference", "metric": "nn_single_search queries/s", "value": qps, "unit": "queries/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": make_config(int(om.num_points()), int(om.num_voxels()), nscans, sample),
            "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": threads, "kind": kind,
                             "sample": f"all {sample} queries of the step, {threads} threads (std::thread shards of the const batch); 1 thread on the first 100k: {qps1:.0f} q/s",
                             "value_1thread": qps1, "insert_mpts_s_1thread": n_in / ins_t[0] / 1e6},
            "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "setup_s": time.perf_counter() - t0}
    emit(json.dumps(line))


def bench_batch_insert(args, dev, local_rank, stream):
    """Insert throughput on a batch large enough to be bandwidth- rather than launch-bound: BASELINE config 5's shape,
    `--insert-scans` OS1-128 scans (262,144 points each, 2 m apart) inserted by ONE mmg_insert_scans call into an empty
    HashedVoxelPointCloud, inputs resident in HBM, CUDA-event time. Algorithmic bytes/point = 36 + 12*a."""
    import torch
    from mola_b200 import HashedVoxelPointCloud, synth
    ns = args.insert_scans
    if ns <= 0:
        return None
    scans = [synth.lidar_scan("os1_128", s=2.0 * k, seed=5, scan_id=k) for k in range(ns)]
    off = np.concatenate([[0], np.cumsum([s[0].size for s in scans])]).astype(np.uint64)
    poses = np.stack([s[3] for s in scans])
    d = [torch.from_numpy(np.concatenate([s[a] for s in scans])).to(dev) for a in range(3)]
    best, cold, stored = None, None, 0
    m = HashedVoxelPointCloud(VOXEL, device=local_rank)
    for rep in range(4):  # rep 0 = cold (table / row-store growth inside the timed region), then clear() + re-insert
        m.clear()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        m.insertPointClouds(d[0], d[1], d[2], off, poses)
        e1.record(stream)
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        if rep == 0:
            cold = ms
        else:
            best = ms if best is None else min(best, ms)
        stored = m.numPoints()
    # the same 64 scans offered again to the map that now holds them: every voxel they hit is full or already has these
    # points' earlier copies, so nothing is accepted (a = 0) and the pass is the probe path alone: compose, voxel index,
    # one 16-byte table load per voxel and warp -- the steady state of a vehicle re-observing mapped ground
    warm = None
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        m.insertPointClouds(d[0], d[1], d[2], off, poses)
        e1.record(stream)
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        warm = ms if warm is None else min(warm, ms)
    stored_after = m.numPoints()
    del m
    n = int(off[-1])
    a = stored / n
    bpp = 36.0 + 12.0 * a
    return {"mpts_s": n / (best * 1e-3) / 1e6, "points_in": n, "scans": ns, "stored_points": int(stored), "accepted_frac": a,
            "ms": best, "ms_cold_map": cold, "algorithmic_bytes_per_point": bpp, "achieved_gbs": bpp * n / (best * 1e-3) / 1e9,
            "reobserve": {"ms": warm, "mpts_s": n / (warm * 1e-3) / 1e6, "accepted_points": int(stored_after - stored),
                          "algorithmic_bytes_per_point": 36.0 + 12.0 * (stored_after - stored) / n,
                          "achieved_gbs": (36.0 + 12.0 * (stored_after - stored) / n) * n / (warm * 1e-3) / 1e9,
                          "note": "the same scans inserted again into the map that holds them (mostly rejected by the probe alone)"},
            "note": "one mmg_insert_scans call (k_insert_rank + k_insert_commit per chunk), cleared map with its row store kept (ms_cold_map: first insert into a new map, allocation included), best of 3"}


def bench_scan_insert(gmap, nscans, dev, stream, k=32):
    """Steady-state mapping: k further HDL-64 scans (120k points each, continuing the trajectory) inserted one call per scan
    into a clone of the built 10M-point map, inputs resident in HBM, one CUDA-event bracket around the whole loop (the
    build loop's own timing is interleaved with scan synthesis and, at N=1, the CPU twin, which leaves the GPU idle between calls)."""
    import torch
    from mola_b200 import synth
    scans = []
    for j in range(k):
        xs, ys, zs, pose = synth.lidar_scan("hdl64", s=SPACING * (nscans + j), seed=2, scan_id=nscans + j)
        scans.append(([torch.from_numpy(a).to(dev) for a in (xs, ys, zs)], pose))
    best = None
    for rep in range(3):
        m = gmap.clone()
        m.reserve(int(m.numVoxels() * 1.5) + 16384 * k)  # voxels_.reserve(): table / row-store growth stays out of the timed loop
        m.insertPointCloud(*scans[0][0], scans[0][1])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for d, pose in scans[1:]:
            m.insertPointCloud(d[0], d[1], d[2], pose)
        e1.record(stream)
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
        del m
    n = sum(int(d[0].numel()) for d, _ in scans[1:])
    return {"mpts_s": n / (best * 1e-3) / 1e6, "points_in": n, "scans": k - 1, "ms_per_scan": best / (k - 1),
            "note": "one mmg_insert_batch call per 120k-point scan into a clone of the 10M-point map, device-resident inputs, best of 3"}


def run_gpu(args, rank, world, local_rank):
    import ctypes as C
    import torch
    import torch.distributed as dist
    from mola_b200 import HashedVoxelPointCloud, capi
    from mola_b200.maps import Communicator
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 else None
    log = (lambda s: print(s, file=sys.stderr, flush=True)) if (args.verbose and rank == 0) else None
    lib = capi.lib()
    t_setup = time.perf_counter()
    # process warm-up, untimed: the first launch of every kernel loads its module and the first pool allocation creates the
    # pool (130 ms on a fresh process, tools/build_trace.py); these are start-up costs of the process, not of a map
    from mola_b200 import synth as _synth
    _w = HashedVoxelPointCloud(VOXEL, device=local_rank)
    _s = _synth.lidar_scan("tiny")
    _w.insertPointCloud(_s[0], _s[1], _s[2], _s[3])
    _w.nn_single_search_batch(_s[0], _s[1], _s[2])
    del _w
    gmap = HashedVoxelPointCloud(VOXEL, device=local_rank)
    stream = torch.cuda.current_stream(dev)
    sp = C.c_void_p(stream.cuda_stream)

    if args.only == "insert_batched":  # profiling aid: just the batched-insert leg
        emit(json.dumps({"insert_batched": bench_batch_insert(args, dev, local_rank, stream)}))
        return
    # ---- map build (rank 0), CPU twin for the cpu_baseline leg (rank 0, N == 1) ----------------
    want_cpu = (rank == 0 and world == 1 and not args.no_cpu)
    om = None
    if want_cpu:
        from oracle import mmo
        cpu_kind = cpu_lib_kind()
        om = mmo.OracleMap(0, VOXEL, which=cpu_kind)
    ins_ms, cpu_ins_s, nscans, n_in = [0.0], [0.0], 0, 0
    if rank == 0:
        def ins(xs, ys, zs, pose):
            dx, dy, dz = (torch.from_numpy(a).to(dev) for a in (xs, ys, zs))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            gmap.insertPointCloud(dx, dy, dz, pose)
            e1.record(stream)
            e1.synchronize()
            ins_ms[0] += e0.elapsed_time(e1)
            if om is not None:
                t = time.perf_counter()
                om.insert(xs, ys, zs, pose)
                cpu_ins_s[0] += time.perf_counter() - t
        nscans, n_in = map_scans(args.map_points, gmap.numPoints, ins, log)
    if world > 1:
        obj = [Communicator.unique_id() if rank == 0 else None, nscans]
        dist.broadcast_object_list(obj, src=0)
        comm = Communicator(obj[0], rank, world, local_rank)
        nscans = obj[1]
        # the first collective on a new communicator pays NCCL's lazy channel set-up (hundreds of ms); the second one is
        # the transfer itself (the receivers already hold the map then, the bytes still cross NVLink)
        torch.cuda.synchronize()
        dist.barrier()
        tb = time.perf_counter()
        gmap.broadcast(comm, 0)
        torch.cuda.synchronize()
        comm_init_s = time.perf_counter() - tb
        dist.barrier()
        tb = time.perf_counter()
        gmap.broadcast(comm, 0)
        torch.cuda.synchronize()
        bcast_s = time.perf_counter() - tb
    else:
        bcast_s, comm_init_s = 0.0, 0.0
    stats = gmap.tableStats()
    batch_insert = bench_batch_insert(args, dev, local_rank, stream) if rank == 0 else None
    scan_insert = bench_scan_insert(gmap, nscans, dev, stream) if rank == 0 else None

    # ---- queries ---------------------------------------------------------------------------
    q = query_batch(nscans, rank, world=world)
    n = q[0].size
    hq = [torch.from_numpy(a).pin_memory() for a in q]                       # pinned host inputs
    dq = [a.to(dev, non_blocking=True) for a in hq]                          # HBM-resident inputs
    d_found = torch.empty(n, dtype=torch.uint8, device=dev)
    d_xyz = torch.empty((n, 3), dtype=torch.float32, device=dev)
    d_d2 = torch.empty(n, dtype=torch.float32, device=dev)
    d_id = torch.empty(n, dtype=torch.int64, device=dev)
    h_found = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_xyz = torch.empty((n, 3), dtype=torch.float32).pin_memory()
    h_d2 = torch.empty(n, dtype=torch.float32).pin_memory()
    h_id = torch.empty(n, dtype=torch.int64).pin_memory()
    fp = lambda t: C.cast(t.data_ptr(), C.POINTER(C.c_float))
    vp = lambda t: C.c_void_p(t.data_ptr())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)            # > 126 MB L2

    def step_device():
        capi.check(lib.mmg_nn_single_batch(gmap._h, fp(dq[0]), fp(dq[1]), fp(dq[2]), C.c_size_t(n), None, vp(d_found), vp(d_xyz), vp(d_d2),
                                           vp(d_id), capi.MEM_DEVICE, sp))

    def step_host():
        capi.check(lib.mmg_nn_single_batch(gmap._h, fp(hq[0]), fp(hq[1]), fp(hq[2]), C.c_size_t(n), None, vp(h_found), vp(h_xyz), vp(h_d2),
                                           vp(h_id), capi.MEM_HOST, sp))

    pbar = C.c_double()
    capi.check(lib.mmg_mean_stencil_points(gmap._h, fp(dq[0]), fp(dq[1]), fp(dq[2]), C.c_size_t(n), capi.MEM_DEVICE, sp, C.byref(pbar)))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        step_device()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = lib.mmg_kernel_launch_count()
    evs = []
    t_wall = time.perf_counter()
    for _ in range(args.steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        step_device()
        e1.record(stream)
        evs.append((e0, e1))
    barrier()
    wall_s = time.perf_counter() - t_wall
    launches = lib.mmg_kernel_launch_count() - launches0
    dev_ms = sum(a.elapsed_time(b) for a, b in evs)
    # ---- end to end: host buffers through the C ABI -------------------------------------------
    for _ in range(2):
        step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_host()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    found_frac = float(d_found.float().mean().item())

    # ---- multi-GPU extras: identical shards (every rank runs rank 0's batch) and strong scaling (ONE ICP iteration: the
    # N = 1 batch of 1M queries cut into `world` contiguous shards, results gathered on every rank over NCCL) ------------
    extra = None
    if world > 1:
        q0 = query_batch(nscans, 0, world=world)
        dq0 = [torch.from_numpy(a).to(dev) for a in q0]

        def step_ident():
            capi.check(lib.mmg_nn_single_batch(gmap._h, fp(dq0[0]), fp(dq0[1]), fp(dq0[2]), C.c_size_t(n), None, vp(d_found), vp(d_xyz), vp(d_d2),
                                               vp(d_id), capi.MEM_DEVICE, sp))
        a, b = n * rank // world, n * (rank + 1) // world
        sh = [t[a:b].contiguous() for t in dq0]
        ns = b - a
        res = torch.empty((ns, 8), dtype=torch.float32, device=dev)       # 32 B/query: xyz, d2, id (2 words), found, pad
        gat = torch.empty((world * ((n + world - 1) // world), 8), dtype=torch.float32, device=dev)
        s_found = torch.empty(ns, dtype=torch.uint8, device=dev)
        s_xyz = torch.empty((ns, 3), dtype=torch.float32, device=dev)
        s_d2 = torch.empty(ns, dtype=torch.float32, device=dev)
        s_id = torch.empty(ns, dtype=torch.int64, device=dev)
        pad = (n + world - 1) // world

        def step_strong():
            capi.check(lib.mmg_nn_single_batch(gmap._h, fp(sh[0]), fp(sh[1]), fp(sh[2]), C.c_size_t(ns), None, vp(s_found), vp(s_xyz), vp(s_d2),
                                               vp(s_id), capi.MEM_DEVICE, sp))
            res[:, 0:3] = s_xyz
            res[:, 3] = s_d2
            res[:, 4:6] = s_id.view(torch.float32).view(ns, 2)
            res[:, 6] = s_found.to(torch.float32)
            buf = res if ns == pad else torch.cat([res, res.new_zeros((pad - ns, 8))])
            dist.all_gather_into_tensor(gat, buf)

        def timed(fn):
            for _ in range(3):
                flush.zero_()
                fn()
            barrier()
            ev = []
            for _ in range(args.steps):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                fn()
                e1.record(stream)
                ev.append((e0, e1))
            barrier()
            t = torch.tensor([sum(x.elapsed_time(y) for x, y in ev) / args.steps], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t[0])
        ident_ms = timed(step_ident)
        strong_ms = timed(step_strong)
        extra = {"weak_identical_shards": {"ms_per_step": ident_ms, "value": n * world / (ident_ms * 1e-3), "unit": "queries/s",
                                           "note": "every rank runs rank 0's 1M-query batch on its replica: isolates contention from workload differences"},
                 "strong": {"ms_per_iteration": strong_ms, "queries": n, "value": n / (strong_ms * 1e-3), "unit": "queries/s",
                            "gathered_bytes_per_rank": int(gat.numel() * 4),
                            "note": "one ICP iteration: the N=1 batch split into contiguous shards, 32 B/query of results all-gathered over NCCL; "
                                    "time = max over ranks, CUDA events around kernel + pack + all_gather"}}

    t = torch.tensor([dev_ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    per_rank = [dev_ms / args.steps]
    if world > 1:
        every = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(every, t)
        per_rank = [float(x[0]) / args.steps for x in every]
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = float(t[0]), float(t[1])
    if rank != 0:
        return
    total_q = n * world
    qps = total_q * args.steps / (dev_ms * 1e-3)
    e2e_qps = total_q * args.steps / (e2e_ms * 1e-3)
    # roofline: algorithmic bytes per query = 361 + 12 * P-bar; one launch = one step
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    bytes_per_q = 361.0 + 12.0 * pbar.value
    if batch_insert:
        batch_insert["frac_of_hbm_peak"] = batch_insert["achieved_gbs"] / peak
        batch_insert["reobserve"]["frac_of_hbm_peak"] = batch_insert["reobserve"]["achieved_gbs"] / peak
    step_s = dev_ms / args.steps * 1e-3
    # What bounds the step. The SURVEY 8(d) figure (361 + 12 P-bar bytes per query) counts every stencil probe and every
    # candidate point as memory traffic; exact pruning and L1/L2 reuse remove nearly all of it, so it is kept as a labelled
    # extra and NOT used as the roofline. `achieved` = the bytes a step cannot avoid moving (12 B query in + 25 B result out
    # per query) over the step time; `traffic` = DRAM bytes per step measured by ncu on this very source tree (else null).
    # The real limiter is instruction issue: `issue` gives warp instructions per step (ncu) over the issue slots of the step.
    compulsory = 37.0 * n
    achieved = compulsory / step_s / 1e9
    roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
            "kernel": "one step = k_nn_own + k_nn_tasks + k_nn_finish (mmg_nn_single_batch, device pointers)",
            "algorithmic_bytes_per_query": 37.0, "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
            "limiter": "instruction issue + L2 latency, not DRAM bandwidth (see `issue`)",
            "survey_8d_formula": {"bytes_per_query": bytes_per_q, "mean_stencil_points": pbar.value, "gbs": bytes_per_q * n / step_s / 1e9,
                                  "x_of_hbm_peak": bytes_per_q * n / step_s / 1e9 / peak,
                                  "note": "361 + 12*P-bar; > 1 x peak means these bytes never reach DRAM (pruning + L1/L2), it is not a roofline"}}
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        lab = tj.get("labels", {}).get("nn_single_step")
        if tj.get("source_stamp") != source_stamp():
            roof["ncu"] = f"profiles/traffic.json was captured on another source tree (stamp {tj.get('source_stamp')}): not quoted"
        elif lab:
            tot = lab["total"]
            traffic = tot["dram_bytes_read"] + tot["dram_bytes_write"]
            sm_hz = 1e6 * float((clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz") or 1965.0)
            slots = step_s * 148 * 4 * sm_hz
            roof["traffic"] = traffic
            roof["dram_gbs_actual"] = traffic / step_s / 1e9
            roof["dram_frac_of_peak"] = traffic / step_s / 1e9 / peak
            roof["issue"] = {"warp_instructions_per_step": tot["warp_instructions"], "issue_slots_per_step": slots,
                             "frac_of_issue_peak": tot["warp_instructions"] / slots, "sm_mhz": sm_hz / 1e6,
                             "note": "ncu smsp__inst_executed.sum of the step's kernels / (step time x 148 SMs x 4 schedulers x SM clock)"}
            roof["ncu"] = {"source": lab.get("source"), "kernels": lab["kernels"]}
    except Exception as e:  # no capture for this tree yet
        roof["ncu"] = f"no usable profiles/traffic.json ({type(e).__name__})"
    line = {"metric": "nn_single_search queries/s", "value": qps, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": make_config(int(stats["points"]), int(stats["live_voxels"]), nscans, n),
            "sharding": {"parallelism": f"map replicated by NCCL broadcast, queries sharded x{world} (scan points x ICP pose guesses: one guess per rank)",
                         "rank0_cpu_binding": numa, "map_hbm_bytes": int(stats["hbm_bytes"])},
            "e2e": {"value": e2e_qps, "unit": "queries/s", "h2d_bytes_per_step": 12 * n, "d2h_bytes_per_step": 25 * n},
            "gpu_launches": int(launches), "ms_per_step_by_rank": per_rank,
            "roofline": roof,
            "clocks": clocks, "found_frac": found_frac, "wall_ms_per_step_incl_flush": 1e3 * wall_s / args.steps,
            "insert": {"mpts_s": n_in / (ins_ms[0] * 1e-3) / 1e6 if ins_ms[0] else None, "points_in": n_in, "scans": nscans,
                       "note": "the map build itself: per-scan calls interleaved with scan synthesis (and the CPU twin at N=1), summed CUDA-event time; "
                               "see insert_per_scan for back-to-back calls"},
            "insert_per_scan": scan_insert, "insert_batched": batch_insert,
            "table": {k: stats[k] for k in ("capacity", "load_factor", "mean_probe", "max_probe")},
            "bcast_s": bcast_s, "comm_init_s": comm_init_s, "multi_gpu": extra, "setup_s": time.perf_counter() - t_setup}
    if om is not None:
        threads = cpu_threads()
        sample = min(n, args.cpu_sample if args.cpu_sample > 0 else 200_000)
        qps_mt, _ = time_cpu_queries(om, q, sample, threads, repeats=2)
        qps_1, _ = time_cpu_queries(om, q, min(sample, 100_000), 1, repeats=2)
        line["cpu_baseline"] = {"value": qps_mt, "unit": "queries/s", "cores": threads, "kind": cpu_kind,
                                "sample": f"first {sample} of the 1M queries on the same 10M-pt map, {threads} threads (1 thread: {qps_1:.0f} q/s)",
                                "value_1thread": qps_1, "insert_mpts_s_1thread": n_in / cpu_ins_s[0] / 1e6 if cpu_ins_s[0] else None}
    emit(json.dumps(line))


def run_config(args, rank, world, local_rank):
    """--config c4 | c5: the multi-GPU BASELINE configurations as bench lines (the legs themselves live in
    tests/bench_configs.py because each also checks the CUDA path against the CPU reference while timing it).
      c4: 50M-pt SparseTreesPointCloud, nn_radius_search sweep r = 0.2 .. 2 m x maxPoints 0 / 16, queries sharded over the ranks
      c5: 64 OS1-128 scans inserted in one call, map replicated over NCCL, every scan queried on rank scan % N"""
    import argparse as ap
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import bench_configs as bc
    ctx = bc.Ctx()
    ctx.args = ap.Namespace(c3_scans=40, c4_points=args.c4_points, c4_max_scans=6000, c4_spacing=2.0, c4_queries=200_000,
                            c4_cpu_scans=0 if args.no_cpu else 6, c4_cpu_queries=5000, c5_scans=args.insert_scans or 64,
                            c5_check=0 if args.no_cpu else 1)
    ctx.rank, ctx.world, ctx.local_rank = rank, world, local_rank
    torch.cuda.set_device(local_rank)
    ctx.dev = torch.device("cuda", local_rank)
    ctx.stream = torch.cuda.current_stream(ctx.dev)
    t0 = time.perf_counter()
    out = (bc.run_c4 if args.config == "c4" else bc.run_c5)(ctx)
    if rank != 0 or out is None:
        return
    base = {"n_gpus": world, "steps": 2, "warmup": 1, "higher_is_better": True, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "gpu_launches": int(__import__("mola_b200").capi.lib().mmg_kernel_launch_count()), "wall_s": time.perf_counter() - t0}
    if args.config == "c4":
        head = [x for x in out["sweep"] if x["r"] == 2.0 and x["maxPoints"] == 0][0]
        line = dict(base, metric="nn_radius_search queries/s (r = 2 m, maxPoints = 0)", value=head["qps"], unit="queries/s", ms_per_step=head["ms"],
                    scaling="strong", config={"workload": "C4: " + out["workload"], "stored_points": out["stored_points"], "blocks": out["blocks"],
                                              "map_scans": out["scans"]},
                    result_stream_gbs=24.0 * head["hits"] / (head["ms"] * 1e-3) / 1e9, sweep=out["sweep"],
                    insert={"mpts_s": out["insert_mpts_s"], "points_in": out["points_in"], "s": out["insert_s"], "ms_per_scan": out.get("insert_ms_per_scan")},
                    bcast_s=out["bcast_s"],
                    nn_single_qps=out["nn_single_qps"], cpu_baseline=out.get("cpu"))
    else:
        line = dict(base, metric="nn_single_search queries/s (64 OS1-128 scans against the replicated map)", value=out["query_qps"], unit="queries/s",
                    ms_per_step=out["query_ms"], scaling="strong", config={"workload": "C5: " + out["workload"], "points": out["points"],
                                                                              "stored_points": out["stored_points"], "voxels": out["voxels"]},
                    insert={"mpts_s": out["insert_mpts_s"], "ms": out["insert_ms"]}, bcast_ms=out["bcast_ms"], parity_bit_exact_sample=out["parity_bit_exact_sample"])
    emit(json.dumps(line))


_REAL_STDOUT = None


def quiet_stdout():
    """Points fd 1 at stderr for the run: NCCL's version banner and any other library chatter written to the C-level
    stdout must not share it with the one JSON line the driver parses. emit() writes to the saved descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        print(line, flush=True)
    else:
        os.write(_REAL_STDOUT, (line + "\n").encode())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--map-points", type=int, default=MAP_POINTS)
    ap.add_argument("--cpu-sample", type=int, default=0, help="--impl reference: queries per step (0 = the full 1M-query step); "
                    "ours: size of the cpu_baseline sample (0 = 200k)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--insert-scans", type=int, default=64, help="OS1-128 scans in the batched-insert measurement (0 = skip)")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--only", default="", choices=["", "insert_batched"], help="profiling aid: run one leg only")
    ap.add_argument("--config", default="c2", choices=["c2", "c4", "c5"], help="BASELINE.json configuration: c2 = the headline (default); "
                    "c4 = 50M-pt SparseTrees radius sweep, queries sharded; c5 = 64 OS1-128 scans inserted, replicated, queried")
    ap.add_argument("--c4-points", type=int, default=50_000_000)
    args = ap.parse_args()
    quiet_stdout()
    rank, world, local_rank = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the map kernels have no CPU fallback")
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # keeps NCCL's version banner out of stdout: the JSON line stays alone there
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        if args.config != "c2":
            run_config(args, rank, world, local_rank)
        else:
            run_gpu(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
