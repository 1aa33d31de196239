#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2c_pytest.log
tail -3 gpurun_out/r2c_pytest.log
python bench.py --no-cpu-baseline --steps 10 2>&1 | tail -1 | cut -c1-200
python bench.py --no-cpu-baseline --steps 10 2>&1 | tail -1 | cut -c1-200
