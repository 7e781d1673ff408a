#!/bin/bash
# round 3 closing verification of the committed tree: -m gpu suite, smoke, default bench.py, reference arm, ncu launch list.
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r3final_gpu_all.log 2>&1; echo "gpu tests rc=$?"; tail -2 gpurun_out/r3final_gpu_all.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3final_smoke.log 2>&1; echo "smoke rc=$?"
timeout 1500 python bench.py > gpurun_out/r3final_bench.json 2> gpurun_out/r3final_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r3final_bench_ref.json 2> gpurun_out/r3final_bench_ref.err; echo "ref rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r3final_bench_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-engine --no-cpu-baseline --no-gpu-reference > gpurun_out/r3final_ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r3final_bench.json").read().strip().splitlines()[-1])
print("value", d["value"], "frac", d["roofline"]["frac"], "e2e", d["e2e"]["value"], "e2e_sync", d["e2e_sync"]["value"])
print("prefill", [(c["workload"][:22], round(c["frac_burst"], 3)) for c in d["prefill"]["cases"]])
print("page1", d["decode_page1"]["frac"], "store", [round(x["frac"], 3) for x in d["store_kv"]], "engine x", d["engine"].get("b200_over_flashinfer"))
PY
