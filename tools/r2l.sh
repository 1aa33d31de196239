#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2l_pytest.log; cat gpurun_out/r2l_pytest.log
for l in 1 2 3 4; do echo "--- lanes $l, rank 0 of 8"; OKRT_LANES=$l python tools/small_batch.py 8 2>&1 | tail -7; done
for l in 1 3; do echo "--- lanes $l, unsharded"; OKRT_LANES=$l python tools/small_batch.py 1 2>&1 | tail -7 | head -6; done
python bench.py --no-cpu-baseline --no-configs --steps 20 2>&1 | tail -1 > gpurun_out/r2l_short.json; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2l_short.json'))
print('value %9.1f  e2e %9.1f  ms/step %6.3f launches %d fallback %d' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['gpu_launches'], d['fallback_segments']), d.get('one_batch_in_flight'))
r=d.get('roofline',{}); print({k:r.get(k) for k in ('achieved','peak','frac','kernel_share_of_step','kernel_ms_per_launch')})
PY
