#!/bin/bash
# f-1 / f-2 through the STOCK reference engine (baseline/_ref), 8B shapes, 1 GPU:
#   config 4 (256 requests sharing a 2048-token prefix) with admission-time prefix reuse on both arms and without it,
#   and the bench workload with the delta wire format bound to Executor / Worker
mkdir -p gpurun_out
O=gpurun_out
T=${1:-r2f}
run() { # name, args...
  n=$1; shift
  timeout 420 python benchmarks/ref_engine.py --model 8b --runs 2 --out $O/${T}_$n.json "$@" > $O/${T}_$n.log 2>&1
  echo "$n rc=$? $(python - <<PY
import json
try:
    d=json.load(open('$O/${T}_$n.json')); print([(round(r['wall_s'],2), round(r['tok_s_generated'])) for r in d['runs']], {k:d.get(k) for k in ('prefix_stats','wire_stats') if k in d})
except Exception as e: print('no json', e)
PY
)"
}
run sp_stock_pc   --backend stock --workload shared-prefix --prefix-cache --nccl-port 29701
run sp_b200_pc    --backend b200  --workload shared-prefix --prefix-cache --nccl-port 29702
run sp_b200_nopc  --backend b200  --workload shared-prefix --nccl-port 29703
run bench_b200_delta --backend b200 --wire delta --nccl-port 29704
tail -3 $O/${T}_sp_b200_pc.log | cut -c1-300
