#!/bin/bash
# round 2, session 7: state after the decode rewrite (flat warp streams): full gpu suite, smoke, bench line (+ reference
# arm), launch list, decode + prefill sweeps, ncu full of the prefill kernel (long + short) with the warp-role split and of
# the decode kernel at the TP8 and the short-context shapes
mkdir -p gpurun_out
O=gpurun_out
T=${1:-r2s7}
rm -f $O/${T}_status.txt
timeout 2400 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1
echo "pytest rc=$?" >> $O/${T}_status.txt
tail -4 $O/${T}_pytest.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1
echo "smoke rc=$?" >> $O/${T}_status.txt
timeout 1500 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err
echo "bench rc=$?" >> $O/${T}_status.txt
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err
echo "bench ref rc=$?" >> $O/${T}_status.txt
timeout 900 python benchmarks/sweep.py decode > $O/${T}_sweep_decode.jsonl 2> $O/${T}_sweep_decode.err
echo "sweep decode rc=$?" >> $O/${T}_status.txt
timeout 900 python benchmarks/sweep.py prefill --no-flashinfer > $O/${T}_sweep_prefill.jsonl 2> $O/${T}_sweep_prefill.err
echo "sweep prefill rc=$?" >> $O/${T}_status.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_bench_launches.csv python bench.py --steps 3 --warmup 3 --no-engine --no-cpu-baseline > $O/${T}_ncu_bench.log 2>&1
echo "ncu launches rc=$?" >> $O/${T}_status.txt
for w in prefill prefill_short; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:prefill_sm100_persist -c 1 -s 2 -o $O/${T}_${w} -f python benchmarks/prof_one.py $w 3 > $O/${T}_ncu_${w}.log 2>&1
  echo "ncu $w rc=$?" >> $O/${T}_status.txt
  ncu -i $O/${T}_${w}.ncu-rep --page raw --csv > $O/${T}_${w}_raw.csv 2>/dev/null
  ncu -i $O/${T}_${w}.ncu-rep --page source --csv --print-source sass > $O/${T}_${w}_src.csv 2>/dev/null
  python benchmarks/ncu_roles.py $O/${T}_${w}_src.csv 6 > $O/${T}_${w}_roles.txt 2>&1
done
for w in decode_tp8 decode_short_tp8; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -c 1 -s 2 -o $O/${T}_${w} -f python benchmarks/prof_one.py $w 3 > $O/${T}_ncu_${w}.log 2>&1
  echo "ncu $w rc=$?" >> $O/${T}_status.txt
  ncu -i $O/${T}_${w}.ncu-rep --page raw --csv > $O/${T}_${w}_raw.csv 2>/dev/null
done
rm -f $O/*.ncu-rep
cat $O/${T}_status.txt
cut -c1-3500 $O/${T}_bench.json
tail -3 $O/${T}_bench.err
