#!/bin/bash
# flake hunt: the full-bench-step prefill property test (+ prefix split + oracle cases) N times per library variant
# usage: r2_stress.sh TAG "variants ('' = the library in the tree)" N
T=${1:-stress}; VARS=${2:-"tree"}; N=${3:-20}
mkdir -p gpurun_out; O=gpurun_out
L=nano-vllm-x_b200/libb200attn.so
cp $L /tmp/orig.so
for v in $VARS; do
  [ "$v" != "tree" ] && cp benchmarks/_ab/$v.so $L
  fails=0
  for i in $(seq 1 $N); do
    timeout 600 python -m pytest tests/test_gpu_attention.py -x -q -k "full_bench_step or prefix_split or prefill_vs_oracle or repeatable" > $O/${T}_${v}_run.log 2>&1
    rc=$?
    if [ $rc -ne 0 ]; then fails=$((fails+1)); cp $O/${T}_${v}_run.log $O/${T}_${v}_fail$fails.log; grep -E "^E |assert|FAILED" $O/${T}_${v}_run.log | head -6 | cut -c1-220; fi
  done
  echo "$v: $fails failures in $N runs"
  cp /tmp/orig.so $L
done
