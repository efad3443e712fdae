"""bench.py's contract with the driver, checked without a GPU through the reference arm: stdout carries exactly one JSON
line, with the keys the contract names; the arm times the reference's own CPU code (oracle/_ref) or the port."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line(built):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--map-points", "200000", "--cpu-sample", "20000"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout[:500]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "nn_single_search queries/s" and d["unit"] == "queries/s"
    for k in ("value", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config",
              "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["higher_is_better"] is True and d["vs_baseline"] is None and d["value"] > 0
    assert "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly(built):
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_config_object_is_the_same_in_both_arms_and_stamp_is_stable(built):
    """`config` comes from one function for both arms (the driver compares them), and the source stamp that guards
    profiles/traffic.json is a pure function of the kernel sources."""
    sys.path.insert(0, ROOT)
    import bench
    a = bench.make_config(10165305, 389152, 496, 1_000_000)
    b = bench.make_config(10165305, 389152, 496, 1_000_000)
    assert a == b and set(a) == {"workload", "map_points", "map_voxels", "map_scans", "queries_per_step_per_gpu", "l2"}
    s1, s2 = bench.source_stamp(), bench.source_stamp()
    assert s1 == s2 and len(s1) == 16
    tj = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tj):
        t = json.load(open(tj))
        if "labels" in t:  # a capture of another tree is allowed to exist, bench.py then refuses to quote it
            assert "source_stamp" in t and "nn_single_step" in t["labels"]
