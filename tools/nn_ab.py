#!/usr/bin/env python
"""A/B timing of the nn_single_search kernel organisations on the bench workload (C2 map, 1M queries).

Builds the 10M-point map like bench.py, then for every setting of mmg_set_tuning() given on the command line
(e.g. `nn_kernel=1` `nn_kernel=0,nn_gw=8,nn_minb=10`) runs the query batch, compares all outputs bit for bit with
the first setting's, and prints CUDA-event times (L2 flushed between steps). A profiling aid, not a bench line."""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    from mola_b200 import HashedVoxelPointCloud, capi
    settings = [a for a in sys.argv[1:] if "=" in a] or ["nn_kernel=1", "nn_kernel=0"]
    steps = int(os.environ.get("AB_STEPS", "20"))
    map_points = int(os.environ.get("AB_MAP_POINTS", str(bench.MAP_POINTS)))
    dev = torch.device("cuda", 0)
    lib = capi.lib()
    lib.mmg_set_tuning.argtypes = [C.c_char_p, C.c_int]
    gmap = HashedVoxelPointCloud(bench.VOXEL, device=0)
    stream = torch.cuda.current_stream(dev)
    sp = C.c_void_p(stream.cuda_stream)

    def ins(xs, ys, zs, pose):
        gmap.insertPointCloud(*(torch.from_numpy(a).to(dev) for a in (xs, ys, zs)), pose)
    nscans, _ = bench.map_scans(map_points, gmap.numPoints, ins)
    q = bench.query_batch(nscans)
    n = q[0].size
    dq = [torch.from_numpy(a).to(dev) for a in q]
    fp = lambda t: C.cast(t.data_ptr(), C.POINTER(C.c_float))
    vp = lambda t: C.c_void_p(t.data_ptr())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ref = None
    for s in settings:
        for kv in s.split(","):
            k, v = kv.split("=")
            capi.check(lib.mmg_set_tuning(k.encode(), int(v)))
        out = [torch.zeros(n, dtype=torch.uint8, device=dev), torch.zeros((n, 3), dtype=torch.float32, device=dev),
               torch.zeros(n, dtype=torch.float32, device=dev), torch.zeros(n, dtype=torch.int64, device=dev)]

        def step():
            capi.check(lib.mmg_nn_single_batch(gmap._h, fp(dq[0]), fp(dq[1]), fp(dq[2]), C.c_size_t(n), None, vp(out[0]), vp(out[1]),
                                               vp(out[2]), vp(out[3]), capi.MEM_DEVICE, sp))
        for _ in range(3):
            flush.zero_()
            step()
        torch.cuda.synchronize()
        ms = []
        for _ in range(steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            step()
            e1.record(stream)
            e1.synchronize()
            ms.append(e0.elapsed_time(e1))
        st = None
        if os.environ.get("AB_E2E"):
            import time
            hq = [torch.from_numpy(a).pin_memory() for a in q]
            ho = [torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty((n, 3), dtype=torch.float32).pin_memory(),
                  torch.empty(n, dtype=torch.float32).pin_memory(), torch.empty(n, dtype=torch.int64).pin_memory()]

            def step_host():
                capi.check(lib.mmg_nn_single_batch(gmap._h, fp(hq[0]), fp(hq[1]), fp(hq[2]), C.c_size_t(n), None, vp(ho[0]), vp(ho[1]),
                                                   vp(ho[2]), vp(ho[3]), capi.MEM_HOST, sp))
            for _ in range(3):
                step_host()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(steps):
                step_host()
            torch.cuda.synchronize()
            st = {"e2e_ms": 1e3 * (time.perf_counter() - t0) / steps, "e2e_same": bool(all(
                np.array_equal(a.numpy().view(np.uint8), b.cpu().numpy().view(np.uint8)) for a, b in zip(ho, out)))}
        if os.environ.get("AB_MATCH"):
            # mmg_match_batch (threshold 0.5 m) on the same batch, device buffers: rate relative to nn_single, pairings checked
            thr2 = 0.25
            midx = torch.zeros(n, dtype=torch.int32, device=dev)
            mxyz = torch.zeros((n, 3), dtype=torch.float32, device=dev)
            md2 = torch.zeros(n, dtype=torch.float32, device=dev)
            mid = torch.zeros(n, dtype=torch.int64, device=dev)
            cnt = C.c_uint64()

            def step_match():
                capi.check(lib.mmg_match_batch(gmap._h, fp(dq[0]), fp(dq[1]), fp(dq[2]), C.c_size_t(n), None, C.c_float(thr2), C.byref(cnt),
                                               vp(midx), vp(mxyz), vp(md2), vp(mid), capi.MEM_DEVICE, sp))
            for _ in range(3):
                flush.zero_()
                step_match()
            mm = []
            for _ in range(steps):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                step_match()
                e1.record(stream)
                e1.synchronize()
                mm.append(e0.elapsed_time(e1))
            k = int(cnt.value)
            keep = (out[0] > 0) & (out[2] < thr2)
            ok = bool(k == int(keep.sum().item()) and torch.equal(midx[:k].long(), torch.nonzero(keep).flatten()) and
                      torch.equal(mid[:k], out[3][keep]) and torch.equal(md2[:k], out[2][keep]))
            st = dict(st or {}, match_ms=float(np.mean(mm)), match_pairings=k, match_equals_filtered_nn_single=ok,
                      match_rate_vs_nn_single=float(np.mean(ms) / np.mean(mm)))
        res = [o.cpu().numpy() for o in out]
        same = None
        if ref is None:
            ref = res
        else:
            same = all(np.array_equal(a.view(np.uint8), b.view(np.uint8)) for a, b in zip(ref, res))
        print(json.dumps({"setting": s, "ms_mean": float(np.mean(ms)), "ms_min": float(np.min(ms)), "gqps": n / np.mean(ms) / 1e6,
                          "bit_identical_to_first": same, "found": float(res[0].mean()), "stats": st}), flush=True)


if __name__ == "__main__":
    main()
