#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2m_pytest.log; cat gpurun_out/r2m_pytest.log
for l in 3 4; do echo "--- lanes $l, rank 0 of 8"; OKRT_LANES=$l python tools/small_batch.py 8 2>&1 | tail -7; done
python bench.py --no-cpu-baseline --no-configs --steps 20 2>&1 | tail -1 > gpurun_out/r2m_short.json; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2m_short.json'))
print('value %9.1f  e2e %9.1f  ms/step %6.3f launches %d fallback %d' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['gpu_launches'], d['fallback_segments']), d.get('one_batch_in_flight'))
PY
