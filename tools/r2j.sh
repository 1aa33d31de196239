#!/bin/bash
# GPU run r2j: GPU suite, the full default bench line, the reference arm, the ncu launch list and two full captures
t0=$(date +%s)
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2j_pytest.log; cat gpurun_out/r2j_pytest.log
echo "pytest $(( $(date +%s) - t0 )) s"
for i in 1 2; do python bench.py --no-cpu-baseline --no-configs --steps 20 2>&1 | tail -1 | tee gpurun_out/r2j_short_$i.json | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('value %9.1f  e2e %9.1f  ms/step %6.3f launches %d fallback %d' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['gpu_launches'], d['fallback_segments']))"; done
python bench.py > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err; echo "bench rc=$? $(( $(date +%s) - t0 )) s"; tail -3 gpurun_out/r2j_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2j_bench_ref.json 2>/dev/null
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs"
$B > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2j_launches.csv $B > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:traceFast -s 12 -c 1 -o gpurun_out/r2j_prof_trace -f $B > gpurun_out/ncu_f.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:shadeKernel -s 12 -c 1 -o gpurun_out/r2j_prof_shade -f $B > gpurun_out/ncu_s.log 2>&1
echo "total $(( $(date +%s) - t0 )) s"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2j_bench.json'))
print({k:d[k] for k in ('value','ms_per_step','first_call_ms','fallback_segments')}, d['e2e']['value'], d.get('parity'), d.get('cpu_baseline',{}).get('value'))
r=d.get('roofline',{}); print({k:r.get(k) for k in ('achieved','peak','frac','frac_of_l2_sector_rate','per_ray','simt_efficiency','kernel_share_of_step')})
print(d.get('unit_peaks'))
for k,v in d.get('configs',{}).items(): print(k, round(v['value'],1), v.get('ms_per_frame'), v.get('first_call_ms'), v.get('fallback_segments'), v.get('wall_ms'))
PY
