#!/bin/bash
# multi-GPU bench line without the extra keys (decode value, roofline, e2e only):  r2_multi_slim.sh N [tag]
N=${1:-8}
T=${2:-r2slim}
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus $N --steps 100 --warmup 5 --no-engine --no-prefill --no-gpu-reference --no-cpu-baseline > $O/${T}_bench_n$N.json 2> $O/${T}_bench_n$N.err
echo "bench n=$N rc=$?"
grep "^{" $O/${T}_bench_n$N.json | cut -c1-3000
tail -3 $O/${T}_bench_n$N.err | cut -c1-300
