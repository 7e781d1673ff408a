#!/bin/bash
# multi-GPU bench line (strong scaling over KV heads, engine arms at tp = N):  r2_multi.sh N [tag]
N=${1:-2}
T=${2:-r2multi}
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 100 --warmup 5 > $O/${T}_bench_n$N.json 2> $O/${T}_bench_n$N.err
echo "bench n=$N rc=$?"
cut -c1-4000 $O/${T}_bench_n$N.json
tail -5 $O/${T}_bench_n$N.err | cut -c1-300
