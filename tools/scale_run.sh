#!/bin/bash
# 1/2/4/8-GPU bench series on one box (strong scaling of the 1080p image); one JSON line per N in gpurun_out/scale_<tag>.jsonl
# usage: scale_run.sh <tag> [extra bench.py flags, e.g. --no-configs]
tag=${1:-run}; shift
out=gpurun_out/scale_${tag}.jsonl
: > $out
for n in 1 2 4 8; do
  if [ $n -eq 1 ]; then
    timeout 400 python bench.py --gpus 1 --no-cpu-baseline "$@" >> $out 2> gpurun_out/scale_${tag}_n$n.err
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n "$@" >> $out 2> gpurun_out/scale_${tag}_n$n.err
  fi
done
python - <<PY
import json
rows=[json.loads(l) for l in open("$out") if l.startswith("{")]
base=rows[0]["value"] if rows else 1
ebase=rows[0]["e2e"]["value"] if rows else 1
for r in rows:
    print(r["n_gpus"], round(r["value"],1), "Mrays/s  eff", round(r["value"]/base/r["n_gpus"],3), "e2e", round(r["e2e"]["value"],1), "eff", round(r["e2e"]["value"]/ebase/r["n_gpus"],3), "sm_mhz", r["clocks"].get("sm_mhz"), r.get("rank_ms"))
    c5 = (r.get("configs") or {}).get("C5")
    if c5: print("   C5", {k: c5.get(k) for k in ("value", "wall_ms", "all_reduce_ms", "all_reduce_gbs", "render_ms")})
PY
