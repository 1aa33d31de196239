#!/bin/bash
# usage (on the GPU box, via gpurun): tools/gpu_profile.sh <tag> [kernel-regex]   -> gpurun_out/<tag>_{bench.json,launches.csv,prof.ncu-rep}
tag=$1; k=${2:-traceFast}
python bench.py --no-cpu-baseline --steps 10 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || { tail -5 gpurun_out/${tag}_bench.err; exit 1; }
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:$k -s 12 -c 1 -o gpurun_out/${tag}_prof -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_f.log 2>&1
cut -c80-140 gpurun_out/${tag}_bench.json
