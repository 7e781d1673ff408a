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


def _engine(out, port, *extra):
    if not os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "nanovllmx", "core")):
        pytest.skip("baseline/_ref not installed (python baseline/install_ref.py; needs /root/reference)")
    r = subprocess.run([sys.executable, "benchmarks/ref_engine.py", "--model", "tiny", "--backend", "oracle", "--num-seqs", "3",
                        "--max-bs", "2", "--context-len", "2048", "--argmax", "--nccl-port", str(port), "--out", str(out), *extra],
                       cwd=ROOT, capture_output=True, text=True, timeout=850)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return json.load(open(out))


@pytest.fixture(scope="module")
def plain_run(tmp_path_factory):
    """Pickle transport, no prefix cache: the reference as shipped (core/executor.py:86-97)."""
    return _engine(tmp_path_factory.mktemp("eng") / "plain.json", 29577, "--runs", "1")


@pytest.mark.timeout(900)
def test_stock_engine_runs_under_the_harness_on_cpu(plain_run):
    d = plain_run
    from benchmarks.ref_engine import bench_workload
    prompts, max_tokens, _ = bench_workload(3)
    # core/engine.py:149: max_tokens caps prompt + generated; a request whose prompt is already longer stops after 1 token
    expect = [max(1, m - len(p)) for p, m in zip(prompts, max_tokens)]
    assert [len(t) for t in d["tokens"]] == expect
    assert d["runs"][0]["generated"] == sum(expect) and d["runs"][0]["reference_formula_tokens"] == sum(max_tokens)
    assert all(0 <= tok < 12000 for t in d["tokens"] for tok in t)


@pytest.mark.timeout(900)
def test_delta_wire_and_prefix_reuse_in_the_real_engine_change_no_token(plain_run, tmp_path):
    """f-1 + f-2 through the stock engine, not only through their unit tests: ``wire_binding.install()`` (per-step delta
    messages over the reference's own queues, decoder mirrors in the worker, finish / preempt hooks in the scheduler) and
    ``prefix_reuse.install()`` (admission-time radix reuse) are switched on together and the workload is served twice.
    The second pass is admitted with every prompt already in the radix tree, so its prefills start from cached slots
    and its sequence ids meet workers that have seen them before.  With a plain argmax sampler both passes must emit
    exactly the tokens of the pickle / no-cache arm."""
    d = _engine(tmp_path / "delta.json", 29579, "--runs", "2", "--wire", "delta", "--prefix-cache")
    assert d["cfg"]["wire"] == "delta" and d["cfg"]["prefix_cache"] is True
    assert len(d["tokens_runs"]) == 2
    assert d["tokens_runs"][0] == plain_run["tokens"], "first pass (cold tree) differs from the pickle / no-cache arm"
    assert d["tokens_runs"][1] == plain_run["tokens"], "second pass (prefix hits, reused workers) differs"
    assert d["runs"][1]["generated"] == plain_run["runs"][0]["generated"]
