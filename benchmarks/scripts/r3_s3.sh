#!/bin/bash
# round 3, step 3: same-box A/B of the prefill kernel before / after its dead-code trim (the in-tree .so is the build of the
# untrimmed source; `make` on the box rebuilds from the trimmed one), then the whole -m gpu suite, smoke, the default
# bench.py run as the driver launches it, and the reference arm -- all on the trimmed build.
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
AB="--no-engine --no-cpu-baseline --no-m3 --no-e2e --no-gpu-reference --steps 20"
timeout 600 python bench.py $AB > gpurun_out/r3s3_prefill_before.json 2> gpurun_out/r3s3_prefill_before.err; echo "before rc=$?"
touch nano-vllm-x_b200/csrc/prefill_sm100.cu
make -C nano-vllm-x_b200/csrc -j8 > gpurun_out/r3s3_make.log 2>&1; echo "make rc=$?"
timeout 600 python bench.py $AB > gpurun_out/r3s3_prefill_after.json 2> gpurun_out/r3s3_prefill_after.err; echo "after rc=$?"
timeout 600 python bench.py $AB > gpurun_out/r3s3_prefill_after2.json 2> gpurun_out/r3s3_prefill_after2.err; echo "after2 rc=$?"
python - <<'PY'
import json
for tag in ("before", "after", "after2"):
    try:
        d = json.loads(open(f"gpurun_out/r3s3_prefill_{tag}.json").read().strip().splitlines()[-1])
        print(tag, [(c["workload"][:22], round(c["kernel_ms"], 4), round(c["frac_burst"], 4), c["self_check_max_abs_err"]) for c in d["prefill"]["cases"]])
    except Exception as e:
        print(tag, "unreadable:", e)
PY
python -m pytest tests -x -q -m gpu > gpurun_out/r3s3_gpu_all.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/r3s3_gpu_all.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3s3_smoke.log 2>&1; echo "smoke rc=$?"
timeout 1500 python bench.py > gpurun_out/r3s3_bench.json 2> gpurun_out/r3s3_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r3s3_bench_ref.json 2> gpurun_out/r3s3_bench_ref.err; echo "ref rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r3s3_bench.json").read().strip().splitlines()[-1])
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
