# Round verification on one B200: gpu tests, smoke, bench line (+ reference arm), launch list, prefill sweep.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/s38_bench.json 2> gpurun_out/s38_bench.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/s38_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/s38_bench_ref.json 2> gpurun_out/s38_bench_ref.err; echo "ref rc=$?"; tail -c 600 gpurun_out/s38_bench_ref.json
# (prefill sweep: run benchmarks/sweep.py prefill separately)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/s38_bench_launches.csv python bench.py --steps 3 --warmup 3 > gpurun_out/s38_ncu_bench.log 2>&1; echo "ncu rc=$?"
