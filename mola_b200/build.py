"""In-tree build of the native libraries (no JIT cache: the .so files travel with the repo).

  mola_b200/_lib/libmola_b200_maps.so   CUDA kernels + C ABI (include/mola_b200/mmg.h), sm_100a
  mola_b200/_lib/libmola_b200_synth.so  host-only synthetic scan generator
  oracle/_build/libmmo_port.so          CPU restatement (test infrastructure)
  oracle/_ref/libmmo_ref.so             reference .cpp files + MRPT shim (only where /root/reference exists)
"""
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "mola_b200", "csrc")
LIBDIR = os.path.join(ROOT, "mola_b200", "_lib")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--extended-lambda", "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-shared", "-cudart", "shared"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd, verbose):
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("build failed: " + " ".join(cmd))
    if verbose and (r.stdout or r.stderr):
        print(r.stdout + r.stderr)


def build_maps(force=False, verbose=False, ptxas_v=False):
    os.makedirs(LIBDIR, exist_ok=True)
    out = os.path.join(LIBDIR, "libmola_b200_maps.so")
    cus = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    deps = cus + glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) + \
        glob.glob(os.path.join(ROOT, "include", "mola_b200", "*.h"))
    if force or _newer(out, deps):
        cmd = [NVCC] + NVCC_FLAGS + (["-Xptxas", "-v"] if ptxas_v else []) + ["-o", out] + cus + ["-ldl"]
        _run(cmd, verbose or ptxas_v)
    return out


def build_synth(force=False, verbose=False):
    os.makedirs(LIBDIR, exist_ok=True)
    out = os.path.join(LIBDIR, "libmola_b200_synth.so")
    src = os.path.join(CSRC, "synth.cpp")
    if force or _newer(out, [src]):
        _run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", out, src, "-lpthread"], verbose)
    return out


def build_oracle(force=False, verbose=False):
    odir = os.path.join(ROOT, "oracle")
    os.makedirs(os.path.join(odir, "_build"), exist_ok=True)
    out = os.path.join(odir, "_build", "libmmo_port.so")
    srcs = [os.path.join(odir, "mmo_port.cpp"), os.path.join(odir, "mmo.h")]
    if force or _newer(out, srcs):
        _run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-fPIC", "-shared", "-o", out, srcs[0], "-lpthread"], verbose)
    return out


def build_reference(force=False, verbose=False):
    """Compiles the reference's own sources where they lie (needs /root/reference). Returns the
    path of oracle/_ref/libmmo_ref.so, or None when neither the sources nor a prebuilt file exist."""
    odir = os.path.join(ROOT, "oracle")
    out = os.path.join(odir, "_ref", "libmmo_ref.so")
    script = os.path.join(odir, "build_ref.sh")
    if os.path.isdir("/root/reference/mola_metric_maps") and os.path.exists(script):
        deps = [script, os.path.join(odir, "ref_bridge.cpp"), os.path.join(odir, "mmo.h")] + \
            glob.glob(os.path.join(ROOT, "shim", "mrpt", "**", "*.h"), recursive=True)
        if force or _newer(out, deps):
            _run(["bash", script], verbose)
    return out if os.path.exists(out) else None


def build_host(force=False, verbose=False, namespace="mola"):
    """The C++ drop-in classes (mola_b200/host) against the stand-in MRPT headers of shim/mrpt:
    mola_b200/_lib/libmola_b200_host.so. In a MOLA workspace the same sources build against MRPT."""
    os.makedirs(LIBDIR, exist_ok=True)
    out = os.path.join(LIBDIR, "libmola_b200_host.so" if namespace == "mola" else f"libmola_b200_host_{namespace}.so")
    hdir = os.path.join(ROOT, "mola_b200", "host")
    srcs = [os.path.join(hdir, "GpuMetricMaps.cpp"), os.path.join(hdir, "register.cpp")]
    deps = srcs + [os.path.join(hdir, "GpuMetricMaps.h")] + glob.glob(os.path.join(ROOT, "shim", "mrpt", "**", "*.h"), recursive=True)
    maps = build_maps(force, verbose)
    if force or _newer(out, deps + [maps]):
        _run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DMOLA_B200_MRPT_SHIM", f"-DMOLA_B200_NAMESPACE={namespace}",
              "-I" + os.path.join(ROOT, "shim"), "-I" + os.path.join(ROOT, "include"), "-o", out] + srcs +
             ["-L" + LIBDIR, "-lmola_b200_maps", "-Wl,-rpath,$ORIGIN"], verbose)
    return out


def build_host_test(force=False, verbose=False):
    """tests/cpp/test_host_classes.cpp -> tests/cpp/_build/test_host_classes (links the host classes and the C ABI)."""
    host = build_host(force, verbose, namespace="mola_gpu")
    tdir = os.path.join(ROOT, "tests", "cpp")
    os.makedirs(os.path.join(tdir, "_build"), exist_ok=True)
    out = os.path.join(tdir, "_build", "test_host_classes")
    src = os.path.join(tdir, "test_host_classes.cpp")
    if force or _newer(out, [src, host, os.path.join(ROOT, "oracle", "mmo.h")]):
        _run(["g++", "-O1", "-std=c++17", "-DMOLA_B200_MRPT_SHIM", "-DMOLA_B200_NAMESPACE=mola_gpu", "-I" + os.path.join(ROOT, "shim"),
              "-I" + os.path.join(ROOT, "include"), "-o", out, src, "-L" + LIBDIR, "-lmola_b200_host_mola_gpu", "-lmola_b200_maps", "-ldl",
              "-Wl,-rpath," + LIBDIR], verbose)
    return out


def build_all(force=False, verbose=False):
    return {"maps": build_maps(force, verbose), "synth": build_synth(force, verbose),
            "oracle": build_oracle(force, verbose), "reference": build_reference(force, verbose),
            "host": build_host(force, verbose), "host_test": build_host_test(force, verbose)}


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv, verbose=True))
