"""Builds gpu-raytracer_b200/libokrt.so (the C-ABI library of include/okrt.h) in-tree.

Host builders are compiled with g++ (-ffp-contract=off), the device side with nvcc for sm_100a only
(-gencode arch=compute_100a,code=sm_100a -lineinfo, -fmad=false: see csrc/okrt_kernels.cuh for why).
nvcc cross-compiles without a GPU; the resulting .so travels to the GPU box with the repo snapshot.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libokrt.so")
OBJ = os.path.join(HERE, "_obj")

HOST_SOURCES = ["okrt_host.cpp", "okrt_accel.cpp", "okrt_obj.cpp", "okrt_image.cpp"]
DEVICE_SOURCES = ["okrt_device.cu", "okrt_build_device.cu", "okrt_layout_device.cu"]
HEADERS = ["okrt_internal.h", "okrt_accel.h", "okrt_kernels.cuh", "okrt_wavefront.cuh", "okrt_math.cuh", "okrt_wide.cuh", "okrt_layout_device.h"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _host_cxx():
    # the image exports CXX=/opt/gcc/bin/g++ (a wrapper without libgomp specs); the system g++ is what nvcc uses
    for cand in ("/usr/bin/g++", shutil.which("g++")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("g++ not found")


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


OUT_STATS = os.path.join(HERE, "libokrt_stats.so")
OUT_CHECKS = os.path.join(HERE, "libokrt_checks.so")


def build_stats_library(force=False, verbose=False):
    """The same library with -DOKRT_TRACE_STATS: the traversal kernel counts its node steps, triangle tests and busy lanes
    (okrt_get_trace_stats).  bench.py renders one step through it to quote the kernel's executed work; the timed arms use
    libokrt.so, which carries no counters."""
    return build_library(force=force, verbose=verbose, extra_nvcc=("-DOKRT_TRACE_STATS",), out=OUT_STATS, obj_dir=OBJ + "_stats")


def build_checks_library(force=False, verbose=False):
    """The same library with -DOKRT_DEBUG_CHECKS: device-side asserts on every traversal-stack push / pop and every node,
    triangle and instance index the kernels form (OKRT_CHECK).  compute-sanitizer is closed on the GPU pool, so the GPU suite is
    run once per round against this build instead (OKRT_LIB=.../libokrt_checks.so python -m pytest tests -m gpu)."""
    return build_library(force=force, verbose=verbose, extra_nvcc=("-DOKRT_DEBUG_CHECKS",), out=OUT_CHECKS, obj_dir=OBJ + "_checks")


def build_library(force=False, verbose=False, extra_nvcc=(), out=None, obj_dir=None):
    global OUT, OBJ
    saved = (OUT, OBJ)
    try:
        if out:
            OUT = out
        if obj_dir:
            OBJ = obj_dir
        return _build_library(force, verbose, extra_nvcc)
    finally:
        OUT, OBJ = saved


def _build_library(force, verbose, extra_nvcc):
    os.makedirs(OBJ, exist_ok=True)
    inc = ["-I" + os.path.join(ROOT, "include"), "-I" + CSRC]
    common_deps = [os.path.join(ROOT, "include", "okrt.h")] + [os.path.join(CSRC, h) for h in HEADERS]
    objs = []
    for src in HOST_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src + ".o")
        objs.append(o)
        if force or _stale(o, [s] + common_deps):
            cmd = [_host_cxx(), "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-pthread",
                   "-fvisibility=hidden", "-Wall", "-c", s, "-o", o] + inc + [f for f in extra_nvcc if f.startswith("-D")]
            if verbose:
                print(" ".join(cmd))
            subprocess.run(cmd, check=True)
    for src in DEVICE_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src + ".o")
        objs.append(o)
        if force or _stale(o, [s] + common_deps):
            cmd = [_nvcc(), "-ccbin", _host_cxx(), "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
                   "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
                   "-Xcompiler", "-fPIC,-ffp-contract=off,-fvisibility=hidden", "-Xptxas", "-v",
                   "-c", s, "-o", o] + inc + list(extra_nvcc)
            if verbose:
                print(" ".join(cmd))
            res = subprocess.run(cmd, check=True, capture_output=True, text=True)
            with open(os.path.join(OBJ, src + ".ptxas.log"), "w") as f:
                f.write(res.stderr)
            if verbose:
                print(res.stderr)
    if force or _stale(OUT, objs):
        cmd = [_nvcc(), "-ccbin", _host_cxx(), "-shared", "-o", OUT] + objs + ["-lpthread"]
        if verbose:
            print(" ".join(cmd))
        subprocess.run(cmd, check=True)
    return OUT


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose=True))
    print(build_stats_library(force="--force" in sys.argv, verbose=False))
