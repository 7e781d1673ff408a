#!/bin/bash
# Same-box A/B of prefill variants (benchmarks/_ab/*.so built by build_ab.sh): parity tests + sweep for each, ncu for
# the ones listed in $NCU.   usage: r2_prefill_ab.sh TAG "variants" "ncu variants"
T=${1:-ab}
VARS=${2:-"base emu0 emu1 emu2"}
NCU=${3:-""}
mkdir -p gpurun_out
O=gpurun_out
L=nano-vllm-x_b200/libb200attn.so
cp $L /tmp/orig.so
for v in $VARS; do
  cp benchmarks/_ab/$v.so $L
  timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_vs_flashinfer.py -x -q -k "prefill or module" > $O/${T}_${v}_pytest.log 2>&1
  echo "$v pytest rc=$? $(tail -1 $O/${T}_${v}_pytest.log | cut -c1-100)"
  for run in 1 2; do
    timeout 600 python benchmarks/sweep.py prefill --no-flashinfer > $O/${T}_${v}_sweep$run.jsonl 2> $O/${T}_${v}_sweep$run.err
    echo "$v run$run rc=$? $(grep -o '"frac_measured_bf16": [0-9.]*' $O/${T}_${v}_sweep$run.jsonl | cut -c24-28 | paste -sd' ')"
  done
done
for v in $NCU; do
  cp benchmarks/_ab/$v.so $L
  for w in prefill prefill_short; do
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:prefill_sm100_persist -c 1 -s 2 -o $O/${T}_${v}_${w} -f python benchmarks/prof_one.py $w 3 > $O/${T}_${v}_ncu_${w}.log 2>&1
    echo "ncu $v $w rc=$?"
    ncu -i $O/${T}_${v}_${w}.ncu-rep --page raw --csv > $O/${T}_${v}_${w}_raw.csv 2>/dev/null
    ncu -i $O/${T}_${v}_${w}.ncu-rep --page source --csv --print-source sass > $O/${T}_${v}_${w}_src.csv 2>/dev/null
    python benchmarks/ncu_roles.py $O/${T}_${v}_${w}_src.csv 6 > $O/${T}_${v}_${w}_roles.txt 2>&1
    python benchmarks/ncu_roles.py $O/${T}_${v}_${w}_src.csv 0 reasons 2>&1 | grep "region @" >> $O/${T}_${v}_${w}_roles.txt
    grep -E "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed|gpu__time_duration.sum|sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed" $O/${T}_${v}_${w}_raw.csv | head -0
    python - <<PY
import csv
r=list(csv.reader(open('$O/${T}_${v}_${w}_raw.csv')))
h=r[0]
for k in ('gpu__time_duration.sum','sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed','sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed','l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed','sm__issue_active.avg.pct_of_peak_sustained_elapsed'):
    if k in h: print('  $v $w', k, r[2][h.index(k)])
PY
    rm -f $O/${T}_${v}_${w}.ncu-rep
  done
done
cp /tmp/orig.so $L
