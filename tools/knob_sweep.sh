#!/bin/bash
# usage (on the GPU box): tools/knob_sweep.sh <out.log> "VAR=val VAR2=val" "..." ; one bench.py run per environment setting
out=$1; shift
for cfg in "$@"; do
  echo -n "$cfg : " >> $out
  env $cfg python bench.py --no-cpu-baseline --steps 10 2>&1 | tail -1 | python -c "
import sys,json
t=sys.stdin.read()
try:
    d=json.loads(t); print('%.1f Mrays/s  %.3f ms/step  e2e %.1f  fallback %d  trace share %.3f' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['fallback_segments'], d.get('trace_kernel_ms', 0) / (d['ms_per_step'] * d['steps'])))
except Exception: print('FAILED', t[-300:])
" >> $out
done
cat $out
