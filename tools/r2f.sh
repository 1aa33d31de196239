#!/bin/bash
python -m pytest tests -m gpu -x -q -s 2>&1 | grep -E "passed|failed|FAILED|Error|scene edit|dragon:|assert" | tail -12 > gpurun_out/r2f_pytest.log
cat gpurun_out/r2f_pytest.log
t0=$(date +%s)
python bench.py > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo "bench rc=$? $(( $(date +%s) - t0 )) s"; tail -3 gpurun_out/r2f_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2f_bench.json'))
print({k:d[k] for k in ('value','ms_per_step','first_call_ms','fallback_segments')}, 'e2e', d['e2e']['value'], d.get('parity',{}).get('bit_exact'), d.get('cpu_baseline',{}).get('value'))
print(d.get('layout_build'))
r=d.get('roofline',{}); print({k:r.get(k) for k in ('achieved','peak','frac','frac_of_l2_sector_rate','per_ray','simt_efficiency','kernel_share_of_step')})
for k,v in d.get('configs',{}).items(): print(k, round(v['value'],1), v.get('ms_per_frame'), v.get('first_call_ms'), v.get('layout'), v.get('fallback_segments'), v.get('wall_ms'))
PY
