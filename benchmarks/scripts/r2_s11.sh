#!/bin/bash
# round 2, session 11: full gpu suite at the new head (prefill design A + packed softmax, decode with control block /
# self-resetting counter / parallel merge), decode sweep, ncu of the merge + decode kernels at the TP8 shapes
mkdir -p gpurun_out
O=gpurun_out
T=${1:-r2s11}
timeout 2400 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1
echo "pytest rc=$? $(tail -1 $O/${T}_pytest.log | cut -c1-120)"
grep -E "FAILED|Error|error" $O/${T}_pytest.log | head -10 | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1
echo "smoke rc=$?"
timeout 900 python benchmarks/sweep.py decode --no-flashinfer > $O/${T}_sweep_decode.jsonl 2> $O/${T}_sweep_decode.err
echo "sweep decode rc=$?"
python - <<PY
import json
for l in open('$O/${T}_sweep_decode.jsonl'):
    r=json.loads(l); print(r['case'][:30].ljust(30), r['B'],r['L'],r['Hq'],r['Hkv'],r['ps'], 'ms %.4f'%r['ms'], 'frac %.3f'%r['frac_measured'], 'graph_ms %s'%(('%.4f'%r['graph_ms']) if 'graph_ms' in r else '-'))
PY
for w in decode_tp8 decode_short_tp8; do
  timeout 600 ncu --metrics gpu__time_duration.sum,sm__cycles_active.min,sm__cycles_active.avg,sm__cycles_active.max,sm__cycles_elapsed.max --clock-control none -k regex:decode -c 4 -s 4 --csv --log-file $O/${T}_${w}_launches.csv python benchmarks/prof_one.py $w 4 > $O/${T}_ncu_${w}.log 2>&1
  echo "ncu $w rc=$?"; grep -E "decode" $O/${T}_${w}_launches.csv | cut -d, -f5,13- | cut -c1-200 | tail -20
done
