#!/bin/bash
# GPU run r2b: full GPU suite on the wide kernel, then bench A/B of the variants, then step statistics
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2b_pytest.log
tail -3 gpurun_out/r2b_pytest.log
tools/run_variants.sh --steps 10 > gpurun_out/r2b_variants.log 2>&1
cat gpurun_out/r2b_variants.log
OKRT_LIB=$PWD/tools/_variants/stats.so OKRT_DEBUG_COUNTS=1 python bench.py --no-cpu-baseline --steps 1 --warmup 3 2>&1 | grep -E "bounce|Mrays" | tail -12 > gpurun_out/r2b_stats.log
cat gpurun_out/r2b_stats.log
