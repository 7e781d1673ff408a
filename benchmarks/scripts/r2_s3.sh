#!/bin/bash
# round 2, session 3: full GPU suite (new parity tests), token-identity arbitration on the full bench workload, bench.py
mkdir -p gpurun_out
O=gpurun_out
rm -f $O/r2s3_status.txt
timeout 2400 python -m pytest tests -m gpu -x -q > $O/r2s3_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2s3_status.txt
tail -5 $O/r2s3_pytest.log
for arm in stock b200; do
  timeout 900 python benchmarks/ref_engine.py --model 0.6b --backend $arm --argmax --runs 2 --out $O/r2s3_eng06_${arm}_argmax.json > $O/r2s3_eng06_${arm}_argmax.log 2>&1
  echo "eng06 argmax $arm rc=$?" >> $O/r2s3_status.txt
done
timeout 1200 python benchmarks/token_identity.py --stock $O/r2s3_eng06_stock_argmax.json --b200 $O/r2s3_eng06_b200_argmax.json --out $O/r2s3_token_identity_06b.json > $O/r2s3_token_identity.log 2>&1
echo "token_identity rc=$?" >> $O/r2s3_status.txt
timeout 1500 python bench.py --steps 20 --warmup 5 > $O/r2s3_bench.json 2> $O/r2s3_bench.err
echo "bench rc=$?" >> $O/r2s3_status.txt
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/r2s3_bench_ref.json 2> $O/r2s3_bench_ref.err
echo "bench ref rc=$?" >> $O/r2s3_status.txt
cat $O/r2s3_status.txt
grep -h TOKEN_IDENTITY $O/r2s3_token_identity.log | cut -c1-1500
tail -3 $O/r2s3_token_identity.log | cut -c1-300
cut -c1-3000 $O/r2s3_bench.json
tail -3 $O/r2s3_bench.err
