"""CPU: the harness that drops the module into the STOCK reference engine (benchmarks/ref_engine.py) drives the real
LLM -> EngineClient -> Engine -> Scheduler -> Executor -> Worker -> ModelRunner stack of baseline/_ref.  Here with the
oracle attention arm (gloo, device cpu, eager): it checks the plumbing only (spawned processes see the patches, the
reference bench.py request generator, token ids coming back through the stub tokenizer); the GPU arms are compared in
tests/test_gpu_ref_engine.py."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(900)
def test_stock_engine_runs_under_the_harness_on_cpu(tmp_path):
    if not os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "nanovllmx", "core")):
        pytest.skip("baseline/_ref not installed (python baseline/install_ref.py; needs /root/reference)")
    out = tmp_path / "eng.json"
    r = subprocess.run([sys.executable, "benchmarks/ref_engine.py", "--model", "tiny", "--backend", "oracle", "--num-seqs", "3",
                        "--max-bs", "2", "--context-len", "2048", "--argmax", "--runs", "1", "--nccl-port", "29577", "--out", str(out)],
                       cwd=ROOT, capture_output=True, text=True, timeout=850)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    d = json.load(open(out))
    from benchmarks.ref_engine import bench_workload
    prompts, max_tokens, _ = bench_workload(3)
    # core/engine.py:149: max_tokens caps prompt + generated; a request whose prompt is already longer stops after 1 token
    expect = [max(1, m - len(p)) for p, m in zip(prompts, max_tokens)]
    assert [len(t) for t in d["tokens"]] == expect
    assert d["runs"][0]["generated"] == sum(expect) and d["runs"][0]["reference_formula_tokens"] == sum(max_tokens)
    assert all(0 <= tok < 12000 for t in d["tokens"] for tok in t)
