# bench.py with the CUDA-graph timed region vs eager, at N=1 and under torchrun at N=$1
N=${1:-2}
mkdir -p gpurun_out
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/s28_bench_graph_n1.json 2> gpurun_out/s28_bench_graph_n1.err; echo "n1 graph rc=$?"
timeout 300 python bench.py --no-cpu-baseline --no-graph > gpurun_out/s28_bench_eager_n1.json 2> gpurun_out/s28_bench_eager_n1.err; echo "n1 eager rc=$?"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29521 bench.py --gpus $N > gpurun_out/s28_bench_graph_n$N.json 2> gpurun_out/s28_bench_graph_n$N.err; echo "n$N graph rc=$?"
timeout 300 $TR --master-port 29522 bench.py --gpus $N --no-graph > gpurun_out/s28_bench_eager_n$N.json 2> gpurun_out/s28_bench_eager_n$N.err; echo "n$N eager rc=$?"
for f in gpurun_out/s28_bench_graph_n1.json gpurun_out/s28_bench_eager_n1.json gpurun_out/s28_bench_graph_n$N.json gpurun_out/s28_bench_eager_n$N.json; do
python - $f <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split('/')[-1], 'value',round(d['value']),'ms',round(d['ms_per_step'],4),'eager_ms',round(d['timing']['eager_ms_per_step'],4),'frac',round(d['roofline']['frac'],3),'e2e',round(d['e2e']['value']),'launches',d['gpu_launches'],'err',d['self_check_max_abs_err'])
except Exception as e:
    print(sys.argv[1], 'FAILED', e); print(open(sys.argv[1].replace('.json','.err')).read()[-1500:])
P
done
