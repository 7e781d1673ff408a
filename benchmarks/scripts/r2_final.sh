#!/bin/bash
# round-2 closing verification: full gpu suite, smoke, bench line + reference arm, launch list, ncu full of the decode
# kernel at the config-5 TP1 shape (DRAM traffic for profiles/decode_traffic.json)
mkdir -p gpurun_out
O=gpurun_out
T=${1:-r2final}
timeout 2400 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1
echo "pytest rc=$? $(tail -1 $O/${T}_pytest.log | cut -c1-120)"
grep -E "^FAILED|^ERROR" $O/${T}_pytest.log | head -5 | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1
echo "smoke rc=$?"
timeout 1500 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err
echo "bench rc=$?"; grep "^{" $O/${T}_bench.json | cut -c1-1200
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err
echo "bench ref rc=$?"; grep "^{" $O/${T}_bench_ref.json | cut -c1-600
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_bench_launches.csv python bench.py --steps 3 --warmup 3 --no-engine --no-cpu-baseline > $O/${T}_ncu_bench.log 2>&1
echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -c 1 -s 2 -o $O/${T}_decode_cfg5 -f python benchmarks/prof_one.py decode_cfg5 3 > $O/${T}_ncu_decode.log 2>&1
echo "ncu decode rc=$?"
ncu -i $O/${T}_decode_cfg5.ncu-rep --page raw --csv > $O/${T}_decode_cfg5_raw.csv 2>/dev/null
rm -f $O/${T}_decode_cfg5.ncu-rep
python - <<PY
import csv
r=list(csv.reader(open('$O/${T}_decode_cfg5_raw.csv'))); h=r[0]
print({k:r[2][h.index(k)] for k in ('gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','sm__cycles_active.min','sm__cycles_active.avg','sm__cycles_active.max','sm__cycles_elapsed.max') if k in h})
PY
