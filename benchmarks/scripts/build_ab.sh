#!/bin/bash
# Builds same-box A/B variants of libb200attn.so that differ only in prefill_sm100.cu:
#   benchmarks/_ab/emu{0,1,2}.so   working tree with -DB200_PREFILL_EMU_PAIRS=N
#   benchmarks/_ab/base.so         prefill_sm100.cu of a reference commit (default 6ef47b9)
# (*.so is git-ignored; the files travel to the GPU box with the gpurun snapshot.)
set -e
cd "$(dirname "$0")/../.."
C=nano-vllm-x_b200/csrc
BASE=${1:-6ef47b9}
make -C $C -j8 > /dev/null
mkdir -p benchmarks/_ab $C/build/ab
FL="-O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC"
OTHERS="$C/build/api.o $C/build/store_kv.o $C/build/meta.o $C/build/decode.o $C/build/qk_norm_rope.o $C/build/sample.o $C/build/prefill_mma.o"
for n in 0 1 2; do
  nvcc $FL -DB200_PREFILL_EMU_PAIRS=$n -c $C/prefill_sm100.cu -o $C/build/ab/prefill_emu$n.o &
done
git show $BASE:$C/prefill_sm100.cu > $C/prefill_sm100_base_tmp.cu
nvcc $FL -c $C/prefill_sm100_base_tmp.cu -o $C/build/ab/prefill_base.o &
wait
rm -f $C/prefill_sm100_base_tmp.cu
for v in emu0 emu1 emu2 base; do
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o benchmarks/_ab/$v.so $OTHERS $C/build/ab/prefill_$v.o -lcudart
done
ls -la benchmarks/_ab/
