#!/bin/bash
# prefill kernel iteration: parity, sweep, ncu (tag = $1)
T=${1:-iter}
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_vs_flashinfer.py -x -q -k "prefill or module or smoke" > $O/${T}_pytest_prefill.log 2>&1
echo "pytest prefill rc=$?"
tail -3 $O/${T}_pytest_prefill.log | cut -c1-300
timeout 900 python benchmarks/sweep.py prefill --no-flashinfer > $O/${T}_sweep_prefill.jsonl 2> $O/${T}_sweep_prefill.err
echo "sweep prefill rc=$?"
python - <<PY
import json
for l in open('$O/${T}_sweep_prefill.jsonl'):
    d=json.loads(l); print(d['case'][:40].ljust(40), d['Hq'],d['Hkv'], 'ms %.3f'%d['ms'], 'TF %.0f'%d['TFLOPs'], 'frac %.3f'%d['frac_measured_bf16'])
PY
tail -3 $O/${T}_sweep_prefill.err
for w in prefill prefill_short; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:prefill_sm100_persist -c 1 -s 2 -o $O/${T}_${w} -f python benchmarks/prof_one.py $w 3 > $O/${T}_ncu_${w}.log 2>&1
  echo "ncu $w rc=$?"
  ncu -i $O/${T}_${w}.ncu-rep --page raw --csv > $O/${T}_${w}_raw.csv 2>/dev/null
  ncu -i $O/${T}_${w}.ncu-rep --page source --csv --print-source sass > $O/${T}_${w}_src.csv 2>/dev/null
  python benchmarks/ncu_roles.py $O/${T}_${w}_src.csv 4 | cut -c1-180
done
