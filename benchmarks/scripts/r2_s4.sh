#!/bin/bash
# round 2, session 4: TMA-fed prefill kernel (parity, sweep, ncu), fixed bench.py
mkdir -p gpurun_out
O=gpurun_out
rm -f $O/r2s4_status.txt
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_vs_flashinfer.py -x -q -k "prefill or smoke or module" > $O/r2s4_pytest_prefill.log 2>&1
echo "pytest prefill rc=$?" >> $O/r2s4_status.txt
tail -4 $O/r2s4_pytest_prefill.log
timeout 900 python benchmarks/sweep.py prefill --no-flashinfer > $O/r2s4_sweep_prefill.jsonl 2> $O/r2s4_sweep_prefill.err
echo "sweep prefill rc=$?" >> $O/r2s4_status.txt
cat $O/r2s4_sweep_prefill.jsonl | cut -c1-260
timeout 1500 python bench.py --steps 20 --warmup 5 > $O/r2s4_bench.json 2> $O/r2s4_bench.err
echo "bench rc=$?" >> $O/r2s4_status.txt
for w in prefill prefill_short; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:prefill_sm100_persist -c 1 -s 2 -o $O/r2s4_${w} -f python benchmarks/prof_one.py $w 3 > $O/r2s4_ncu_${w}.log 2>&1
  echo "ncu $w rc=$?" >> $O/r2s4_status.txt
  ncu -i $O/r2s4_${w}.ncu-rep --page raw --csv > $O/r2s4_${w}_raw.csv 2>/dev/null
done
cat $O/r2s4_status.txt
cut -c1-4000 $O/r2s4_bench.json
tail -3 $O/r2s4_bench.err
