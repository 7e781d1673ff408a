#!/bin/bash
# round 3, step 1: new every-row FlashInfer tests (config 4 shared prefix, full bench prefill step) + an experiment:
# is the drop-in arm run-to-run token-identical under argmax, and does the delta wire keep it so?
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
python -m pytest tests/test_vs_flashinfer.py -x -q -m gpu -k "config4 or full_bench" > gpurun_out/r3s1_fi.log 2>&1; echo "fi rc=$?"; tail -3 gpurun_out/r3s1_fi.log
for arm in pickleA pickleB delta deltafull; do
  extra=""; port=29551
  case $arm in pickleB) port=29553;; delta) extra="--wire delta"; port=29555;; deltafull) extra="--wire delta --fused-qk --b200-sampler"; port=29557;; esac
  python benchmarks/ref_engine.py --model 0.6b --backend b200 --num-seqs 48 --max-bs 32 --argmax --runs 1 --nccl-port $port $extra --out gpurun_out/r3s1_eng_$arm.json > gpurun_out/r3s1_eng_$arm.log 2>&1; echo "$arm rc=$?"
done
python - <<'PY'
import json
t = {a: json.load(open(f"gpurun_out/r3s1_eng_{a}.json")) for a in ("pickleA", "pickleB", "delta", "deltafull")}
base = t["pickleA"]["tokens"]
for a in ("pickleB", "delta", "deltafull"):
    same = sum(x == y for x, y in zip(base, t[a]["tokens"]))
    print(a, "identical sequences vs pickleA:", same, "/", len(base), "wall", t[a]["runs"][0]["wall_s"])
print("pickleA wall", t["pickleA"]["runs"][0]["wall_s"])
PY
