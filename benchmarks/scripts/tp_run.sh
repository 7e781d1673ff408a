N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29511 benchmarks/engine_bench.py --model 8b --backend b200 --fused-qk > gpurun_out/s14_eng_8b_tp${N}_b200.json 2> gpurun_out/s14_eng_8b_tp${N}_b200.err; echo "b200 rc=$?"; tail -c 700 gpurun_out/s14_eng_8b_tp${N}_b200.json
timeout 400 $TR --master-port 29512 benchmarks/engine_bench.py --model 8b --backend flashinfer > gpurun_out/s14_eng_8b_tp${N}_fi.json 2> gpurun_out/s14_eng_8b_tp${N}_fi.err; echo "fi rc=$?"; tail -c 700 gpurun_out/s14_eng_8b_tp${N}_fi.json
timeout 200 $TR --master-port 29513 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/s14_bench_tp${N}.json 2> gpurun_out/s14_bench_tp${N}.err; echo "bench rc=$?"; tail -c 400 gpurun_out/s14_bench_tp${N}.json
nproc
