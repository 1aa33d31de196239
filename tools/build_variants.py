"""Builds tuning variants of libokrt.so into gpurun_out-independent files: tools/_variants/<name>.so"""
import importlib.util, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("okrt_build", os.path.join(ROOT, "gpu-raytracer_b200", "build.py"))
b = importlib.util.module_from_spec(spec); spec.loader.exec_module(b)
out_dir = os.path.join(ROOT, "tools", "_variants"); os.makedirs(out_dir, exist_ok=True)
for arg in sys.argv[1:]:
    name, flags = arg.split("=", 1)
    b.build_library(force=True, extra_nvcc=flags.split(), out=os.path.join(out_dir, name + ".so"))
    log = os.path.join(ROOT, "gpu-raytracer_b200", "_obj", "okrt_device.cu.ptxas.log")
    lines = open(log).read().splitlines()
    for i, l in enumerate(lines):
        if "Compiling entry function" in l and ("traceFastKernel" in l or "traceWideKernel" in l or "shadeKernelILb0" in l):
            print("  ", l.split("'")[1][:40], "|", lines[i + 2].strip(), "|", lines[i + 3].strip()[:60])
    print("built", name)
b.build_library(force=True, out=os.path.join(ROOT, "gpu-raytracer_b200", "libokrt.so"))
