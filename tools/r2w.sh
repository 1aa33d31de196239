#!/bin/bash
# final evidence set r2_w: GPU suite, smoke, full default bench line + reference arm, launch list, ncu full of the three main kernels
t0=$(date +%s)
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r2w_pytest.log; cat gpurun_out/r2w_pytest.log
OKRT_LIB=$PWD/gpu-raytracer_b200/libokrt_checks.so python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > gpurun_out/r2w_checks_pytest.log; cat gpurun_out/r2w_checks_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/r2w_bench.json 2> gpurun_out/r2w_bench.err; echo "bench rc=$? $(( $(date +%s) - t0 )) s"; tail -3 gpurun_out/r2w_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2w_bench_ref.json 2>/dev/null
# ncu serialises the kernels: one batch in flight (OKRT_LANES=1) gives every kernel the grid it has when it runs alone
export OKRT_LANES_FOR_NCU=1
B="env OKRT_LANES=1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs"
$B > gpurun_out/r2w_plain.json 2> gpurun_out/r2w_plain.err || { tail -5 gpurun_out/r2w_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'generateKernel|trace|shadeKernel|accumulateKernel' -c 700 --csv --log-file gpurun_out/r2w_launches.csv $B > gpurun_out/ncu_l.log 2>&1
python tools/launch_summary.py gpurun_out/r2w_launches.csv | cut -c1-120 | tee gpurun_out/r2w_launches_summary.txt
for k in traceFast shadeKernel generateKernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 12 -c 1 -o gpurun_out/r2w_prof_$k -f $B > gpurun_out/ncu_$k.log 2>&1
  python tools/ncu_summary.py gpurun_out/r2w_prof_$k.ncu-rep > gpurun_out/r2w_ncu_full_${k}_summary.txt
done
python tools/sass_lines.py gpurun_out/r2w_prof_traceFast.ncu-rep gpu-raytracer_b200/libokrt.so traceFastKernel 45 > gpurun_out/r2w_ncu_source_traceFastKernel_lines.txt 2>&1
python tools/sass_lines.py gpurun_out/r2w_prof_shadeKernel.ncu-rep gpu-raytracer_b200/libokrt.so shadeKernelILb0 45 > gpurun_out/r2w_ncu_source_shadeKernel_lines.txt 2>&1
grep -E "gpu__time_duration|dram__bytes" gpurun_out/r2w_ncu_full_*_summary.txt
echo "total $(( $(date +%s) - t0 )) s"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2w_bench.json'))
print({k:d[k] for k in ('value','ms_per_step','first_call_ms','fallback_segments')}, d['e2e']['value'], d.get('parity',{}).get('bit_exact'), d.get('cpu_baseline',{}).get('value'), d.get('one_batch_in_flight',{}).get('value'))
r=d.get('roofline',{}); print({k:r.get(k) for k in ('achieved','peak','frac','frac_of_l2_sector_rate','kernel_share_of_step','kernel_ms_per_launch')})
for k,v in d.get('configs',{}).items(): print(k, round(v['value'],1), v.get('ms_per_frame'), (v.get('cpu_baseline') or {}).get('value'), (v.get('parity') or {}).get('bit_exact'))
print(json.load(open('gpurun_out/r2w_bench_ref.json'))['value'])
PY
