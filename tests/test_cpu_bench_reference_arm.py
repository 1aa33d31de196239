"""bench.py --impl reference (the CPU arm the driver runs beside the CUDA arm) works without a GPU and prints the
contract's JSON line."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         cwd=ROOT, capture_output=True, text=True, timeout=600, check=True).stdout
    line = json.loads(out.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "Mrays/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["steps"] == 1 and line["n_gpus"] == 1
    assert line["cpu_baseline"]["kind"] in ("port", "reference") and line["cpu_baseline"]["cores"] >= 1
    assert line["cpu_baseline"]["value"] == line["value"] == line["e2e"]["value"]
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in line["config"] and line["gpu_launches"] == 0


def test_cuda_arm_fails_loudly_without_a_device():
    """No CPU fallback on the product path: without a GPU the CUDA arm must exit non-zero, not print a number."""
    import ctypes
    try:
        ctypes.CDLL("libcuda.so.1")
        has_driver = True
    except OSError:
        has_driver = False
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "3", "--no-cpu-baseline"],
                         cwd=ROOT, capture_output=True, text=True, timeout=600)
    if res.returncode == 0:
        # a GPU is present (this test then just checks the line); on the CPU-only container the branch below runs
        assert has_driver and json.loads(res.stdout.strip().splitlines()[-1])["gpu_launches"] > 0
    else:
        assert "value" not in res.stdout
