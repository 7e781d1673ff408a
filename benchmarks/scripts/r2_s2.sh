#!/bin/bash
# round 2, session 2: stock reference engine with / without the drop-in (sampler eager in both arms)
mkdir -p gpurun_out
O=gpurun_out
rm -f $O/r2s2_status.txt
for arm in stock b200; do
  timeout 900 python benchmarks/ref_engine.py --model 0.6b --backend $arm --greedy --runs 2 --stats --out $O/r2s2_eng06_${arm}_greedy.json > $O/r2s2_eng06_${arm}_greedy.log 2>&1
  echo "eng06 $arm rc=$?" >> $O/r2s2_status.txt
done
timeout 900 python benchmarks/token_identity.py --stock $O/r2s2_eng06_stock_greedy.json --b200 $O/r2s2_eng06_b200_greedy.json --out $O/r2s2_token_identity_06b.json > $O/r2s2_token_identity.log 2>&1
echo "token_identity rc=$?" >> $O/r2s2_status.txt
for arm in stock b200; do
  timeout 900 python benchmarks/ref_engine.py --model 8b --backend $arm --runs 2 --stats --out $O/r2s2_eng8b_${arm}.json > $O/r2s2_eng8b_${arm}.log 2>&1
  echo "eng8b $arm rc=$?" >> $O/r2s2_status.txt
done
timeout 900 python benchmarks/ref_engine.py --model 0.6b --backend stock --runs 2 --stats --out $O/r2s2_eng06_stock.json > $O/r2s2_eng06_stock.log 2>&1
echo "eng06 stock t0.6 rc=$?" >> $O/r2s2_status.txt
timeout 900 python benchmarks/ref_engine.py --model 0.6b --backend b200 --runs 2 --stats --out $O/r2s2_eng06_b200.json > $O/r2s2_eng06_b200.log 2>&1
echo "eng06 b200 t0.6 rc=$?" >> $O/r2s2_status.txt
cat $O/r2s2_status.txt
grep -h REF_ENGINE_RESULT $O/r2s2_eng*.log | cut -c1-1800
grep -h TOKEN_IDENTITY $O/r2s2_token_identity.log
tail -5 $O/r2s2_token_identity.log
