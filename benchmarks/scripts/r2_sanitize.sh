#!/bin/bash
# compute-sanitizer memcheck + racecheck over the small oracle-parity cases of decode / prefill / store / meta
# (the library build without the mbarrier watchdog: build_sanitize.sh)
mkdir -p gpurun_out
O=gpurun_out
T=${1:-r2san}
L=nano-vllm-x_b200/libb200attn.so
cp $L /tmp/orig.so; cp benchmarks/_ab/sanitize.so $L
SEL='test_prefill_vs_oracle or test_decode_vs_oracle or test_store or test_build_metadata'
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --error-exitcode 77 --target-processes all python -m pytest tests/test_gpu_attention.py tests/test_gpu_store_meta.py -x -q -k "$SEL" -p no:cacheprovider > $O/${T}_$tool.log 2>&1
  echo "$tool rc=$? $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $O/${T}_$tool.log | tail -1) | $(grep -E 'passed|failed' $O/${T}_$tool.log | tail -1 | cut -c1-80)"
done
cp /tmp/orig.so $L
