# 0.6B engine bench, one GPU: both arms with the sampler each really runs, plus the cheaper torch sampler on both
E="python benchmarks/engine_bench.py --model 0.6b"
timeout 300 $E --backend b200 --fused-qk --sampler b200 > gpurun_out/s14_eng06_b200_sb200.json 2> gpurun_out/s14_eng06_b200_sb200.err; tail -c 500 gpurun_out/s14_eng06_b200_sb200.json
timeout 300 $E --backend b200 --fused-qk --sampler torch > gpurun_out/s14_eng06_b200_storch.json 2> gpurun_out/s14_eng06_b200_storch.err; tail -c 300 gpurun_out/s14_eng06_b200_storch.json
timeout 300 $E --backend flashinfer --sampler reference > gpurun_out/s14_eng06_fi_sref.json 2> gpurun_out/s14_eng06_fi_sref.err; tail -c 500 gpurun_out/s14_eng06_fi_sref.json
timeout 300 $E --backend flashinfer --sampler torch > gpurun_out/s14_eng06_fi_storch.json 2> gpurun_out/s14_eng06_fi_storch.err; tail -c 300 gpurun_out/s14_eng06_fi_storch.json
