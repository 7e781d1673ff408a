#!/bin/bash
# host-path check: module-level GPU tests (eager + graph replay drive pack_step), host-bound e2e (ctx 2048) and the
# full-size line without the extra keys
mkdir -p gpurun_out
O=gpurun_out
T=${1:-r2s29}
timeout 600 python -m pytest tests/test_gpu_attention.py tests/test_gpu_store_meta.py -x -q -k "module or readmission or page_size_16 or meta or graph" > $O/${T}_pytest.log 2>&1
echo "pytest rc=$? $(tail -1 $O/${T}_pytest.log | cut -c1-100)"
for extra in "--ctx 2048 --steps 200" ""; do
timeout 600 python bench.py $extra --no-engine --no-cpu-baseline --no-prefill --no-gpu-reference > $O/${T}_bench.json 2> $O/${T}_bench.err
echo "bench [$extra] rc=$?"; python - <<PY
import json
d=json.loads([l for l in open('$O/${T}_bench.json') if l.startswith('{')][-1])
print('step ms',d['ms_per_step'],'value',d['value'],'e2e ms/step',d['e2e']['ms_per_step'],'e2e',d['e2e']['value'],d['e2e']['self_check_max_abs_err'],d['e2e']['full_resends_in_timed_steps'])
PY
done
