#!/bin/bash
# round 3, step 2: whole -m gpu suite (new engine test included), then the default bench.py run as the driver launches it,
# the reference arm, and a launch list of a short bench run.
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r3s2_gpu_all.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/r3s2_gpu_all.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3s2_smoke.log 2>&1; echo "smoke rc=$?"
timeout 1500 python bench.py > gpurun_out/r3s2_bench.json 2> gpurun_out/r3s2_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r3s2_bench_ref.json 2> gpurun_out/r3s2_bench_ref.err; echo "ref rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r3s2_bench.json").read().strip().splitlines()[-1])
for k in ("value", "ms_per_step"):
    print(k, d[k])
print("roofline.frac", d["roofline"]["frac"], "traffic", d["roofline"]["traffic"])
print("e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "e2e_sync", d["e2e_sync"] and (d["e2e_sync"]["value"], d["e2e_sync"]["ms_per_step"]))
print("decode_page1", d["decode_page1"])
print("store_kv", d["store_kv"])
print("meta", d["meta"])
print("prefill", [(c["workload"][:24], round(c["tflops"], 1), round(c["frac_burst"], 3)) for c in d["prefill"].get("cases", [])] if d["prefill"] else None, (d["prefill"] or {}).get("error"))
print("engine", json.dumps(d["engine"])[:1200])
PY
