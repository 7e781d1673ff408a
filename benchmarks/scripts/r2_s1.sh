#!/bin/bash
# round 2, session 1: baseline GPU tests + stock reference engine with / without the drop-in
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > $O/r2s1_smi.txt 2>&1
nproc >> $O/r2s1_smi.txt
timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_zz_gpu_ref_engine.py > $O/r2s1_pytest_base.log 2>&1
echo "pytest_base rc=$?" >> $O/r2s1_status.txt
for arm in stock b200; do
  timeout 900 python benchmarks/ref_engine.py --model 0.6b --backend $arm --greedy --runs 2 --stats --out $O/r2s1_eng06_${arm}_greedy.json > $O/r2s1_eng06_${arm}_greedy.log 2>&1
  echo "eng06 $arm rc=$?" >> $O/r2s1_status.txt
done
timeout 900 python benchmarks/token_identity.py --stock $O/r2s1_eng06_stock_greedy.json --b200 $O/r2s1_eng06_b200_greedy.json --out $O/r2s1_token_identity_06b.json > $O/r2s1_token_identity.log 2>&1
echo "token_identity rc=$?" >> $O/r2s1_status.txt
for arm in stock b200; do
  timeout 900 python benchmarks/ref_engine.py --model 8b --backend $arm --runs 2 --stats --out $O/r2s1_eng8b_${arm}.json > $O/r2s1_eng8b_${arm}.log 2>&1
  echo "eng8b $arm rc=$?" >> $O/r2s1_status.txt
done
timeout 900 python benchmarks/ref_engine.py --model 0.6b --backend b200 --fused-qk --b200-sampler --runs 2 --stats --out $O/r2s1_eng06_b200_all.json > $O/r2s1_eng06_b200_all.log 2>&1
echo "eng06 b200 all rc=$?" >> $O/r2s1_status.txt
cat $O/r2s1_status.txt
grep -h REF_ENGINE_RESULT $O/r2s1_eng*.log | cut -c1-1500
grep -h TOKEN_IDENTITY $O/r2s1_token_identity.log
