"""PCIe / host-memory ceiling for the e2e leg of bench.py: pinned H2D 12 MB, D2H 25 MB, and both at once on two streams.

Single process: `python tools/pcie_probe.py`. Under torchrun (`--nproc-per-node N`) every rank drives its own GPU at the
same time (ranks bound to their GPU's NUMA node like bench.py does), and rank 0 prints the per-rank and aggregate figures:
the host-side ceiling the N-GPU e2e number of bench.py runs into."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
numa = None
if world > 1:
    import torch.distributed as dist
    import bench
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    numa = bench.bind_to_gpu_numa_node(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
h_in = torch.empty(12_000_000, dtype=torch.uint8).pin_memory()
h_out = torch.empty(25_000_000, dtype=torch.uint8).pin_memory()
d_in = torch.empty_like(h_in, device=dev)
d_out = torch.empty_like(h_out, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def t(fn, n=20):
    fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n


def h2d():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


def both():
    h2d()
    d2h()


a, b, c = t(h2d), t(d2h), t(both)
mine = {"rank": rank, "cpus": numa, "h2d_gbs": 12e-3 / a, "d2h_gbs": 25e-3 / b, "both_ms": c * 1e3, "both_gbs": 37e-3 / c, "gqps_ceiling": 1e6 / c / 1e9}
if world > 1:
    every = [None] * world
    dist.all_gather_object(every, mine)
    if rank == 0:
        print(json.dumps({"n_gpus": world, "aggregate_both_gbs": sum(e["both_gbs"] for e in every),
                          "aggregate_gqps_ceiling": 1e6 * world / max(e["both_ms"] * 1e-3 for e in every) / 1e9, "ranks": every}))
    dist.destroy_process_group()
else:
    print(f"H2D 12MB {a*1e3:.3f} ms {12e-3/a:.1f} GB/s | D2H 25MB {b*1e3:.3f} ms {25e-3/b:.1f} GB/s | both {c*1e3:.3f} ms -> {1e6/c/1e9:.2f} G queries/s ceiling")
