#!/bin/bash
# racecheck + memcheck of a small render, then the launch list of the bench's render kernels
python tools/sanitizer_case.py > gpurun_out/r2o_plain.txt 2>&1; cat gpurun_out/r2o_plain.txt
timeout 900 compute-sanitizer --tool racecheck --racecheck-report all python tools/sanitizer_case.py > gpurun_out/r2o_racecheck.txt 2>&1; tail -5 gpurun_out/r2o_racecheck.txt
timeout 900 compute-sanitizer --tool memcheck python tools/sanitizer_case.py > gpurun_out/r2o_memcheck.txt 2>&1; tail -4 gpurun_out/r2o_memcheck.txt
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs"
$B > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'generateKernel|trace|shadeKernel|accumulateKernel' -c 700 --csv --log-file gpurun_out/r2o_launches.csv $B > gpurun_out/ncu_l.log 2>&1
python tools/launch_summary.py gpurun_out/r2o_launches.csv | cut -c1-120
python -c "
import json; d=json.loads(open('gpurun_out/plain.log').read().strip().splitlines()[-1]); print(d['unit_peaks'])"
