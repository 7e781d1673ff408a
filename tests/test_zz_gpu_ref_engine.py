"""GPU: the drop-in, dropped in.  The STOCK reference engine (baseline/_ref: LLM -> Engine -> Scheduler/RadixTree ->
Executor -> Worker -> ModelRunner with CUDA-graph capture at start-up and graph replay for every decode step,
core/model_runner.py:133-157, 276-330, 332-417) runs the reference's bench.py request generator twice: as shipped
(FlashInfer + Triton) and with INTEGRATION.md section 1's monkey-patch.  Greedy tokens (plain argmax in both arms: the reference's top_k=1 form
breaks exact bf16 ties with torch.multinomial, which makes even two runs of ONE arm differ) are compared; every
divergence is arbitrated against an fp32-attention forward of the same model on the same prefix
(benchmarks/token_identity.py)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
NUM_SEQS = 48


def _run(args, timeout):
    r = subprocess.run([sys.executable] + args, cwd=ROOT, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, f"{' '.join(args)} failed:\n{r.stdout[-3000:]}\n{r.stderr[-3000:]}"
    return r.stdout


@pytest.fixture(scope="module")
def engine_runs():
    if not os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "nanovllmx", "core")):
        pytest.fail("baseline/_ref is missing: run `python baseline/install_ref.py` before shipping the tree to the GPU box")
    os.makedirs(OUT, exist_ok=True)
    paths = {}
    for arm, port in (("stock", 29531), ("b200", 29541)):
        paths[arm] = os.path.join(OUT, f"test_ref_engine_{arm}.json")
        _run(["benchmarks/ref_engine.py", "--model", "0.6b", "--backend", arm, "--num-seqs", str(NUM_SEQS), "--max-bs", "32",
              "--argmax", "--runs", "1", "--nccl-port", str(port), "--out", paths[arm]], timeout=900)
    return paths


@pytest.mark.timeout(2400)
def test_real_model_runner_with_dropin_completes_and_matches(engine_runs):
    """>= 32 requests through the real ModelRunner (max_bs 32 < 48 requests, so the scheduler queues, the graphs for
    bs 1..32 are captured and padded replays happen as sequences finish)."""
    a = json.load(open(engine_runs["stock"]))
    b = json.load(open(engine_runs["b200"]))
    assert a["cfg"]["backend"] == "stock" and b["cfg"]["backend"] == "b200" and not b["cfg"]["enforce_eager"]
    ta, tb = a["tokens"], b["tokens"]
    assert len(ta) == len(tb) == NUM_SEQS
    # ignore_eos + max_tokens: the number of generated tokens per request does not depend on the values
    assert [len(x) for x in ta] == [len(x) for x in tb]
    assert a["runs"][0]["generated"] == b["runs"][0]["generated"] > NUM_SEQS
    # first generated token of every request comes from the prefill step alone
    first_equal = sum(x[0] == y[0] for x, y in zip(ta, tb))
    assert first_equal >= int(0.9 * NUM_SEQS), f"only {first_equal}/{NUM_SEQS} first tokens agree"


@pytest.mark.timeout(2400)
def test_every_token_divergence_is_a_bf16_tie(engine_runs):
    rep_path = os.path.join(OUT, "test_ref_engine_token_identity.json")
    _run(["benchmarks/token_identity.py", "--stock", engine_runs["stock"], "--b200", engine_runs["b200"], "--out", rep_path],
         timeout=900)
    rep = json.load(open(rep_path))
    assert rep["arbitrated"] == rep["divergences"]
    for c in rep["cases"]:
        # the flip is between the fp32 model's own top candidates ...
        assert c["rank_of_stock_pick"] <= 3 and c["rank_of_b200_pick"] <= 3, c
        # ... and is decided by a margin below the stated bf16 bound
        assert c["margin_ulps"] <= rep["margin_ulps_bound"], c
    if rep["cases"]:
        # b200's logits are not further from the fp32-attention logits than FlashInfer's (25 % noise allowance)
        assert rep["mean_rms_b200"] <= 1.25 * rep["mean_rms_fi"] + 1e-6, (rep["mean_rms_b200"], rep["mean_rms_fi"])


@pytest.mark.timeout(2400)
def test_delta_wire_and_shared_prefix_reuse_keep_every_token(engine_runs):
    """f-2 and f-1 inside the real engine on the GPU, drop-in arm, plain argmax.

    (1) ``--wire delta`` changes how a step reaches the worker (core/executor.py:86-97, core/worker_client.py:87-94), not
        what the worker computes: its tokens must be IDENTICAL to the pickle run of the same arm, sequence by sequence.
    (2) Config 4 through the stock scheduler: a shared-prefix workload served with admission-time prefix reuse
        (core/scheduler.py:67-74, core/kv_cache.py:159-161) + the delta wire, twice.  The second pass is admitted
        entirely from the radix tree.  Every request must still produce its full quota of tokens, the two passes must
        agree on at least the first generated token of (nearly) every request (the prefix rows are re-read from other
        slots through another code path of the prefill kernel, so later tokens may flip on bf16 ties), and nothing may
        be left in a half-admitted state (the run ends, i.e. no sequence is lost)."""
    delta = os.path.join(OUT, "test_ref_engine_b200_delta.json")
    _run(["benchmarks/ref_engine.py", "--model", "0.6b", "--backend", "b200", "--num-seqs", str(NUM_SEQS), "--max-bs", "32",
          "--argmax", "--runs", "1", "--wire", "delta", "--nccl-port", "29545", "--out", delta], timeout=900)
    a = json.load(open(engine_runs["b200"]))["tokens"]
    b = json.load(open(delta))["tokens"]
    same = sum(x == y for x, y in zip(a, b))
    assert len(a) == len(b) == NUM_SEQS and same == NUM_SEQS, f"delta wire changed tokens: {same}/{NUM_SEQS} sequences identical"

    sp = os.path.join(OUT, "test_ref_engine_b200_shared_prefix.json")
    _run(["benchmarks/ref_engine.py", "--model", "0.6b", "--backend", "b200", "--num-seqs", str(NUM_SEQS), "--max-bs", "32",
          "--argmax", "--runs", "2", "--workload", "shared-prefix", "--prefix-len", "512", "--gen", "32", "--prefix-cache",
          "--wire", "delta", "--nccl-port", "29547", "--out", sp], timeout=900)
    d = json.load(open(sp))
    assert d["cfg"]["prefix_cache"] is True and d["cfg"]["workload"] == "shared-prefix"
    r0, r1 = d["tokens_runs"]
    assert [len(t) for t in r0] == [32] * NUM_SEQS and [len(t) for t in r1] == [32] * NUM_SEQS
    first_equal = sum(x[0] == y[0] for x, y in zip(r0, r1))
    assert first_equal >= int(0.9 * NUM_SEQS), f"only {first_equal}/{NUM_SEQS} first tokens agree between the cold and the cached pass"
