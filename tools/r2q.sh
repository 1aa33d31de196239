#!/bin/bash
t0=$(date +%s)
python bench.py > gpurun_out/r2q_bench.json 2> gpurun_out/r2q_bench.err; echo "bench rc=$? $(( $(date +%s) - t0 )) s"; tail -3 gpurun_out/r2q_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2q_bench_ref.json 2>/dev/null; echo "ref arm $(( $(date +%s) - t0 )) s"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2q_bench.json'))
print({k:d[k] for k in ('value','ms_per_step','first_call_ms','fallback_segments')}, d['e2e']['value'], d.get('parity',{}).get('bit_exact'), d.get('cpu_baseline',{}).get('value'), d.get('one_batch_in_flight'))
r=d.get('roofline',{}); print({k:r.get(k) for k in ('achieved','peak','frac','frac_of_l2_sector_rate','per_ray','simt_efficiency','kernel_share_of_step','wasted_local_sectors_per_global_sector','traffic')})
print(d.get('unit_peaks'))
for k,v in d.get('configs',{}).items(): print(k, round(v['value'],1), v.get('ms_per_frame'), v.get('first_call_ms'), v.get('fallback_segments'), v.get('cpu_baseline'), v.get('parity'))
print(open('gpurun_out/r2q_bench_ref.json').read()[:600])
PY
