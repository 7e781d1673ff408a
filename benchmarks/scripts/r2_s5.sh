#!/bin/bash
# round 2, session 5: prefill stage 2 (early QK, P through smem, TMA Q, hybrid K/V producers): parity, sweep, ncu, bench
mkdir -p gpurun_out
O=gpurun_out
rm -f $O/r2s5_status.txt
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_vs_flashinfer.py -x -q -k "prefill or module or smoke" > $O/r2s5_pytest_prefill.log 2>&1
echo "pytest prefill rc=$?" >> $O/r2s5_status.txt
tail -15 $O/r2s5_pytest_prefill.log | cut -c1-300
timeout 900 python benchmarks/sweep.py prefill --no-flashinfer > $O/r2s5_sweep_prefill.jsonl 2> $O/r2s5_sweep_prefill.err
echo "sweep prefill rc=$?" >> $O/r2s5_status.txt
cat $O/r2s5_sweep_prefill.jsonl | cut -c1-260
tail -3 $O/r2s5_sweep_prefill.err
for w in prefill prefill_short; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:prefill_sm100_persist -c 1 -s 2 -o $O/r2s5_${w} -f python benchmarks/prof_one.py $w 3 > $O/r2s5_ncu_${w}.log 2>&1
  echo "ncu $w rc=$?" >> $O/r2s5_status.txt
  ncu -i $O/r2s5_${w}.ncu-rep --page raw --csv > $O/r2s5_${w}_raw.csv 2>/dev/null
done
timeout 1500 python bench.py --steps 20 --warmup 5 --no-engine > $O/r2s5_bench.json 2> $O/r2s5_bench.err
echo "bench rc=$?" >> $O/r2s5_status.txt
cat $O/r2s5_status.txt
cut -c1-2500 $O/r2s5_bench.json
tail -3 $O/r2s5_bench.err
