#!/bin/bash
# usage: tools/run_variants.sh [bench args]; runs bench.py for the default build and every tools/_variants/*.so
run() { python bench.py --no-cpu-baseline "$@" 2>&1 | tail -1 | python -c "
import sys,json
t=sys.stdin.read()
try:
    d=json.loads(t); print('%-28s %8.1f Mrays/s  %7.3f ms/step  fallback %d' % (sys.argv[1], d['value'], d['ms_per_step'], d['fallback_segments']))
except Exception: print(sys.argv[1], 'FAILED', t[-300:])
" "$NAME"; }
NAME=default run "$@"
for f in tools/_variants/*.so; do NAME=$f; OKRT_LIB=$PWD/$f run "$@"; done
