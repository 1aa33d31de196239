#!/bin/bash
N=${1:-8}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 --no-configs 2>gpurun_out/r2h_err.log | tail -1 | python -c "
import sys,json
t=sys.stdin.read()
try:
    d=json.loads(t); print('%-14s value %9.1f  e2e %9.1f  ms/step %6.3f rank_ms %s %s' % (sys.argv[1], d['value'], d['e2e']['value'], d['ms_per_step'], d.get('rank_ms'), d.get('band_balance')))
except Exception as e: print(sys.argv[1], 'FAILED', t[-300:]); print(open('gpurun_out/r2h_err.log').read()[-1500:])
" "$1"; }
run feedback
run feedback-again
