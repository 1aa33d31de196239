#!/bin/bash
# the GPU suite against the -DOKRT_DEBUG_CHECKS build (device asserts on stack and index bounds)
OKRT_LIB=$PWD/gpu-raytracer_b200/libokrt_checks.so timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2p_checks_pytest.log; cat gpurun_out/r2p_checks_pytest.log
OKRT_LIB=$PWD/gpu-raytracer_b200/libokrt_checks.so python tools/sanitizer_case.py
python tools/sanitizer_case.py
