#!/bin/bash
# round 2, session 9: (1) decode with guided dynamic chunk claiming at Hkv = 1: parity + sweep (static-split numbers: r2s7);
# (2) prefill design A (S/P aliased in TMEM, TS PV) with the hybrid TMA / cp.async producers against design B (emu0)
mkdir -p gpurun_out
O=gpurun_out
T=r2s9
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_vs_flashinfer.py -x -q -k "decode or module or config5" > $O/${T}_pytest_decode.log 2>&1
echo "pytest decode rc=$? $(tail -1 $O/${T}_pytest_decode.log | cut -c1-100)"
timeout 900 python benchmarks/sweep.py decode --no-flashinfer > $O/${T}_sweep_decode.jsonl 2> $O/${T}_sweep_decode.err
echo "sweep decode rc=$?"
python - <<PY
import json
for l in open('$O/${T}_sweep_decode.jsonl'):
    r=json.loads(l); print(r['case'][:30].ljust(30), r['B'],r['L'],r['Hq'],r['Hkv'],r['ps'], 'ms %.4f'%r['ms'], 'frac %.3f'%r['frac_measured'], 'graph_ms %s'%(('%.4f'%r['graph_ms']) if 'graph_ms' in r else '-'))
PY
for w in decode_tp8 decode_short_tp8; do
  timeout 600 ncu --set full --clock-control none -k regex:decode_kernel -c 1 -s 2 -o $O/${T}_${w} -f python benchmarks/prof_one.py $w 3 > $O/${T}_ncu_${w}.log 2>&1
  ncu -i $O/${T}_${w}.ncu-rep --page raw --csv > $O/${T}_${w}_raw.csv 2>/dev/null
  rm -f $O/${T}_${w}.ncu-rep
  python - <<PY
import csv
r=list(csv.reader(open('$O/${T}_${w}_raw.csv'))); h=r[0]
print('$w', {k:r[2][h.index(k)] for k in ('gpu__time_duration.sum','sm__cycles_active.min','sm__cycles_active.avg','sm__cycles_active.max','sm__cycles_elapsed.max','dram__bytes_read.sum') if k in h})
PY
done
bash benchmarks/scripts/r2_prefill_ab.sh $T "designA emu0" "designA"
