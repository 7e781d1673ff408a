#!/bin/bash
# benchmarks/_ab/sanitize.so: the whole library with the mbarrier watchdog compiled out, for compute-sanitizer runs
set -e
cd "$(dirname "$0")/../.."
C=nano-vllm-x_b200/csrc
mkdir -p benchmarks/_ab $C/build/san
FL="-O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -DB200_MBAR_TIMEOUT_NS=0"
for f in api store_kv meta decode qk_norm_rope sample prefill_mma prefill_sm100; do nvcc $FL -c $C/$f.cu -o $C/build/san/$f.o & done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o benchmarks/_ab/sanitize.so $C/build/san/*.o -lcudart
ls -la benchmarks/_ab/sanitize.so
