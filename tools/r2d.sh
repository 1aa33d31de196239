#!/bin/bash
# GPU run r2d: the full default bench line, the reference arm, the ncu launch list and two full captures
t0=$(date +%s)
python bench.py > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc=$? $(( $(date +%s) - t0 )) s"; tail -3 gpurun_out/r2d_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2d_bench_ref.json 2>/dev/null
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs"
$B > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2d_launches.csv $B > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:traceFast -s 12 -c 1 -o gpurun_out/r2d_prof_trace -f $B > gpurun_out/ncu_f.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:shadeKernel -s 12 -c 1 -o gpurun_out/r2d_prof_shade -f $B > gpurun_out/ncu_s.log 2>&1
echo "total $(( $(date +%s) - t0 )) s"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2d_bench.json'))
print({k:d[k] for k in ('value','ms_per_step','first_call_ms','fallback_segments')}, d['e2e']['value'], d.get('parity'), d.get('cpu_baseline',{}).get('value'))
r=d.get('roofline',{}); print({k:r.get(k) for k in ('achieved','peak','frac','frac_of_l2_sector_rate','per_ray','simt_efficiency','kernel_share_of_step')})
print(d.get('unit_peaks'))
for k,v in d.get('configs',{}).items(): print(k, round(v['value'],1), v.get('ms_per_frame'), v.get('first_call_ms'), v.get('fallback_segments'), v.get('wall_ms'))
PY
