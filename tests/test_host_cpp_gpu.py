"""The C++ drop-in classes (mola_b200/host): build + symbol checks on CPU, and the C++ test program
(tests/cpp/test_host_classes.cpp: reference virtual interface vs the oracle) on the GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_host_library_builds_and_exports_reference_class_names(built):
    from mola_b200 import build
    lib = build.build_host()
    syms = subprocess.run(["nm", "-DC", "--defined-only", lib], capture_output=True, text=True, check=True).stdout
    for cls in ("HashedVoxelPointCloud", "SparseVoxelPointCloud", "SparseTreesPointCloud"):
        assert f"mola::{cls}::internal_CreateFromMapDefinition" in syms
        assert f"mola::{cls}::serializeTo" in syms
        assert f"mola::{cls}::runtimeClassId()" in syms
    assert "mola::GpuPointMapBase::nn_single_search_batch" in syms
    assert "mola::registerGpuMetricMaps()" in syms
    # only the C ABI crosses into the CUDA library: every undefined mmg_* symbol is declared in mmg.h
    und = subprocess.run(["nm", "-D", "--undefined-only", lib], capture_output=True, text=True, check=True).stdout
    from mola_b200 import capi
    used = {l.split()[-1] for l in und.splitlines() if " mmg_" in l}
    assert used and used <= set(capi.EXPORTS)
    assert os.path.exists(build.build_host_test())


@pytest.mark.gpu
def test_host_classes_against_oracle(built):
    from mola_b200 import build
    from oracle import mmo
    exe = build.build_host_test()
    ora = mmo.REF_PATH if mmo.have_reference() else mmo.PORT_PATH
    mmo.load("port")
    r = subprocess.run([exe, ora], capture_output=True, text=True, timeout=600)
    print(r.stdout[-3000:], r.stderr[-2000:])
    assert r.returncode == 0 and "ALL OK" in r.stdout
    if mmo.have_reference():  # serializeTo/From of all three maps against the reference classes' own archives
        assert "archives cross-checked with the reference classes: 6 cases" in r.stdout
