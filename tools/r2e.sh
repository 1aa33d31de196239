#!/bin/bash
python -m pytest tests -m gpu -x -q -s 2>&1 | grep -E "passed|failed|FAILED|Error|scene edit|dragon:|assert" | tail -15 > gpurun_out/r2e_pytest.log
cat gpurun_out/r2e_pytest.log
run() { python bench.py --no-cpu-baseline --no-configs --steps 10 2>&1 | tail -1 | python -c "
import sys,json
t=sys.stdin.read()
try:
    d=json.loads(t); print('%-28s %8.1f Mrays/s  %7.3f ms/step  e2e %8.1f  first call %6.1f ms  fallback %d' % (sys.argv[1], d['value'], d['ms_per_step'], d['e2e']['value'], d['first_call_ms'], d['fallback_segments']))
except Exception: print(sys.argv[1], 'FAILED', t[-400:])
" "$1"; }
run device-layout
OKRT_LAYOUT_HOST=1 run host-layout
for f in tools/_variants/*.so; do OKRT_LIB=$PWD/$f run $f; done
