#!/bin/bash
# evidence set r2_s: launch list of the render kernels (one batch in flight so that the serialised list matches the events),
# ncu --set full of the traversal, shade and generate kernels with per-source-line tables made on the box (matching library)
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs"
$B > gpurun_out/r2s_plain.json 2> gpurun_out/r2s_plain.err || { tail -5 gpurun_out/r2s_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'generateKernel|trace|shadeKernel|accumulateKernel' -c 700 --csv --log-file gpurun_out/r2s_launches.csv $B > gpurun_out/ncu_l.log 2>&1
python tools/launch_summary.py gpurun_out/r2s_launches.csv | cut -c1-120 | tee gpurun_out/r2s_launches_summary.txt
for k in traceFast shadeKernel generateKernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 12 -c 1 -o gpurun_out/r2s_prof_$k -f $B > gpurun_out/ncu_$k.log 2>&1
  python tools/ncu_summary.py gpurun_out/r2s_prof_$k.ncu-rep > gpurun_out/r2s_ncu_full_${k}_summary.txt
done
python tools/sass_lines.py gpurun_out/r2s_prof_traceFast.ncu-rep gpu-raytracer_b200/libokrt.so traceFastKernel 45 > gpurun_out/r2s_ncu_source_traceFastKernel_lines.txt 2>&1
python tools/sass_lines.py gpurun_out/r2s_prof_shadeKernel.ncu-rep gpu-raytracer_b200/libokrt.so shadeKernelILb0 45 > gpurun_out/r2s_ncu_source_shadeKernel_lines.txt 2>&1
grep -E "gpu__time_duration|dram__bytes" gpurun_out/r2s_ncu_full_*_summary.txt
