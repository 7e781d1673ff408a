#!/bin/bash
# decode iteration: parity subset, sweep, per-kernel ncu times at the TP8 shapes
mkdir -p gpurun_out
O=gpurun_out
T=${1:-r2s12}
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_vs_flashinfer.py -x -q -k "decode or module or config5" > $O/${T}_pytest_decode.log 2>&1
echo "pytest decode rc=$? $(tail -1 $O/${T}_pytest_decode.log | cut -c1-100)"
timeout 900 python benchmarks/sweep.py decode --no-flashinfer > $O/${T}_sweep_decode.jsonl 2> $O/${T}_sweep_decode.err
echo "sweep decode rc=$?"
python - <<PY
import json
for l in open('$O/${T}_sweep_decode.jsonl'):
    r=json.loads(l); print(r['case'][:30].ljust(30), r['B'],r['L'],r['Hq'],r['Hkv'],r['ps'], 'ms %.4f'%r['ms'], 'frac %.3f'%r['frac_measured'], 'graph_ms %s'%(('%.4f'%r['graph_ms']) if 'graph_ms' in r else '-'))
PY
for w in decode_tp8 decode_tp8_p1 decode_short_tp8; do
  timeout 600 ncu --metrics gpu__time_duration.sum,sm__cycles_active.min,sm__cycles_active.avg,sm__cycles_active.max,sm__cycles_elapsed.max --clock-control none -k regex:decode -c 2 -s 6 --csv --log-file $O/${T}_${w}_launches.csv python benchmarks/prof_one.py $w 4 > $O/${T}_ncu_${w}.log 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open('$O/${T}_${w}_launches.csv')) if len(r)>10 and 'decode' in r[4]]
d={}
for r in rows: d.setdefault(r[4][:22],{})[r[-3]]=r[-1]
for k,v in d.items(): print('$w',k,v)
PY
done
