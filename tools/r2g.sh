#!/bin/bash
# run on N GPUs (gpurun --gpus N): the bench under torchrun + the C++ NCCL example
N=${1:-2}
t0=$(date +%s)
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r2g_bench_${N}gpu.json 2> gpurun_out/r2g_bench_${N}gpu.err
echo "bench N=$N rc=$? $(( $(date +%s) - t0 )) s"; tail -2 gpurun_out/r2g_bench_${N}gpu.err | cut -c1-300
python - $N <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.loads(open('gpurun_out/r2g_bench_%sgpu.json'%n).read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step','n_gpus')}, 'e2e', d['e2e']['value'], 'rank_ms', d.get('rank_ms'))
    print(d.get('configs'))
except Exception as e: print('parse failed', e)
PY
pkg=gpu-raytracer_b200
/usr/bin/g++ -std=c++17 -O1 -Iinclude -I$pkg/host -I/usr/local/cuda/include $pkg/host/sample_sharded_nccl.cpp -L$pkg -lokrt -lnccl -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,$PWD/$pkg -o /tmp/okrt_nccl || echo "nccl example: compile failed"
rm -f /tmp/okrt_nccl_id
for r in $(seq 1 $((N-1))); do RANK=$r WORLD_SIZE=$N timeout 120 /tmp/okrt_nccl /tmp/okrt_nccl_id 8 > gpurun_out/r2g_nccl_rank$r.log 2>&1 & done
RANK=0 WORLD_SIZE=$N timeout 120 /tmp/okrt_nccl /tmp/okrt_nccl_id 8 2>&1 | tail -2
wait
cat gpurun_out/r2g_nccl_rank1.log | tail -1
