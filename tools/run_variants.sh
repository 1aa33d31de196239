#!/bin/bash
# usage: tools/run_variants.sh [bench args]; runs bench.py for the default build and every tools/_variants/*.so
python bench.py --no-cpu-baseline "$@" 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('default', round(d['value'],1), d['ms_per_step'], d['fallback_segments'])"
for f in tools/_variants/*.so; do
  OKRT_LIB=$PWD/$f python bench.py --no-cpu-baseline "$@" 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$f', round(d['value'],1), d['ms_per_step'], d['fallback_segments'])"
done
