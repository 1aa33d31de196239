"""The C++ host mirror (host/RayTracer.hpp) compiles against include/okrt.h, links libokrt.so, and fails loudly
without a GPU (exit code 3 = OKRT_ERR_NO_DEVICE surfaced as okrt::Error)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_mirror_builds_and_runs(okrt, tmp_path):
    exe = str(tmp_path / "okrt_example")
    pkg = os.path.join(ROOT, "gpu-raytracer_b200")
    subprocess.run(["/usr/bin/g++", "-std=c++17", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(pkg, "host"),
                    os.path.join(pkg, "host", "example_main.cpp"), "-L" + pkg, "-lokrt", "-Wl,-rpath," + pkg, "-o", exe], check=True)
    r = subprocess.run([exe], capture_output=True, text=True)
    if okrt.abi.lib().okrt_device_count() > 0:
        assert r.returncode == 0 and "frames 1" in r.stdout, r.stdout
    else:
        assert r.returncode == 3 and "no CUDA device" in r.stdout, r.stdout


def test_header_is_plain_c(tmp_path):
    src = tmp_path / "c_only.c"
    src.write_text('#include "okrt.h"\nint main(void){ okrt_options o; okrt_options_default(&o); return (int)sizeof(okrt_render_data) - 240; }\n')
    subprocess.run(["/usr/bin/gcc", "-std=c99", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), "-c", str(src), "-o", str(tmp_path / "c_only.o")], check=True)


def test_nccl_sample_sharded_example_builds(okrt, tmp_path):
    """host/sample_sharded_nccl.cpp (the sample-sharded exchange from a C++ host: ncclAllReduce inside the
    okrt_reduce_accumulation callback) compiles and links against include/okrt.h, libokrt.so and the system NCCL; without
    a GPU it exits 3 (no CPU path)."""
    import pytest
    if not os.path.exists("/usr/include/nccl.h"):
        pytest.skip("no system NCCL headers")
    exe = str(tmp_path / "okrt_nccl_example")
    pkg = os.path.join(ROOT, "gpu-raytracer_b200")
    subprocess.run(["/usr/bin/g++", "-std=c++17", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(pkg, "host"),
                    "-I/usr/local/cuda/include", os.path.join(pkg, "host", "sample_sharded_nccl.cpp"), "-L" + pkg, "-lokrt", "-lnccl",
                    "-L/usr/local/cuda/lib64", "-lcudart", "-Wl,-rpath," + pkg, "-o", exe], check=True)
    if okrt.abi.lib().okrt_device_count() <= 0:
        r = subprocess.run([exe, str(tmp_path / "id")], capture_output=True, text=True)
        assert r.returncode == 3 and "no CUDA device" in r.stdout, r.stdout
