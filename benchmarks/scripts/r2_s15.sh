#!/bin/bash
# full gpu suite + bench line + compute-sanitizer pass at the current head
mkdir -p gpurun_out
O=gpurun_out
T=${1:-r2s15}
timeout 2400 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1
echo "pytest rc=$? $(tail -1 $O/${T}_pytest.log | cut -c1-120)"
grep -E "^FAILED|^ERROR" $O/${T}_pytest.log | head -5 | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1
echo "smoke rc=$?"
timeout 1500 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err
echo "bench rc=$?"; cut -c1-2600 $O/${T}_bench.json
