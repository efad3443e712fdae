#!/usr/bin/env python
"""Turns `ncu -i <rep> --page raw --csv` output into profiles/traffic.json, the per-kernel counters bench.py quotes.

  python tools/ncu_summary.py <label>=<raw.csv> [<label>=<raw.csv> ...] > profiles/traffic.json

Every kernel launch found in the csv files is summarised (DRAM bytes, duration, warp instructions, active lanes per
instruction, SM-active share, occupancy, cache hit rates); launches of the same kernel inside one label are summed.
The file is stamped with a hash of the kernel sources (`source_stamp`, bench.source_stamp()): bench.py prints these
numbers only when the stamp equals the one of the tree it runs from, so a stale capture can never be quoted.
"""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

WANT = {
    "gpu__time_duration.sum": "duration_ns",
    "dram__bytes_read.sum": "dram_bytes_read",
    "dram__bytes_write.sum": "dram_bytes_write",
    "smsp__inst_executed.sum": "warp_instructions",
    "smsp__thread_inst_executed_per_inst_executed.ratio": "active_lanes_per_instruction",
    "sm__cycles_active.avg": "sm_cycles_active_avg",
    "sm__cycles_elapsed.max": "sm_cycles_elapsed_max",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct",
    "l1tex__t_sector_hit_rate.pct": "l1_hit_pct",
    "launch__registers_per_thread": "registers",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
}
UNIT_SCALE = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6, "ns": 1.0,
              "nsecond": 1.0, "s": 1e9, "second": 1e9}


def num(x):
    try:
        return float(str(x).replace(",", ""))
    except ValueError:
        return None


def parse(path):
    rows = list(csv.reader(open(path, newline="")))
    # skip ncu's "==PROF==" chatter: the header row is the one that has "Kernel Name"
    h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    head, units = rows[h], rows[h + 1]
    out = []
    for r in rows[h + 2:]:
        if len(r) != len(head):
            continue
        d = dict(zip(head, r))
        k = {"kernel": re.sub(r"\(.*$", "", d["Kernel Name"]).strip()}
        for col, name in WANT.items():
            if col in d:
                v = num(d[col])
                u = units[head.index(col)]
                if v is not None and (name.startswith("dram_bytes") or name == "duration_ns"):
                    v *= UNIT_SCALE.get(u, 1.0)
                k[name] = v
        out.append(k)
    return out


def main():
    import bench
    res = {"source_stamp": bench.source_stamp(), "how": "ncu --set full --clock-control none, one capture per label; `ncu -i rep --page raw --csv` "
           "summarised by tools/ncu_summary.py", "labels": {}}
    for arg in sys.argv[1:]:
        label, path = arg.split("=", 1)
        ks = {}
        for k in parse(path):
            a = ks.setdefault(k["kernel"], {"launches": 0})
            a["launches"] += 1
            for name, v in k.items():
                if name == "kernel" or v is None:
                    continue
                if name in ("duration_ns", "dram_bytes_read", "dram_bytes_write", "warp_instructions", "sm_cycles_active_avg", "sm_cycles_elapsed_max"):
                    a[name] = a.get(name, 0.0) + v
                else:  # ratios: launch-duration weighted mean would need two passes; keep the value of the longest launch
                    if k.get("duration_ns", 0) >= a.get("_longest", -1):
                        a[name] = v
            a["_longest"] = max(a.get("_longest", -1), k.get("duration_ns", 0))
        for a in ks.values():
            a.pop("_longest", None)
            if "sm_cycles_active_avg" in a and a.get("sm_cycles_elapsed_max"):
                a["sm_active_frac_of_elapsed"] = a["sm_cycles_active_avg"] / a["sm_cycles_elapsed_max"]
        tot = {n: sum(a.get(n, 0.0) for a in ks.values()) for n in ("duration_ns", "dram_bytes_read", "dram_bytes_write", "warp_instructions")}
        res["labels"][label] = {"source": os.path.relpath(path, ROOT) if os.path.isabs(path) else path, "kernels": ks, "total": tot}
    json.dump(res, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
