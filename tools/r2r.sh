#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2r_pytest.log; cat gpurun_out/r2r_pytest.log
for i in 1 2; do python bench.py --no-cpu-baseline --no-configs --steps 20 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('value %9.1f  e2e %9.1f  ms/step %6.3f launches %d fallback %d one-batch %.1f' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['gpu_launches'], d['fallback_segments'], d['one_batch_in_flight']['value']))"; done
python tools/small_batch.py 8 2>&1 | tail -7
