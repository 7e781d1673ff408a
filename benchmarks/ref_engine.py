#!/usr/bin/env python
"""Run the STOCK reference engine (baseline/_ref = unmodified izmttk/nano-vllm-x: LLM -> EngineClient -> Engine ->
Scheduler/RadixTree -> Executor -> WorkerClient -> Worker -> ModelRunner with its CUDA-graph capture/replay) on the
reference's own bench.py workload, once per attention arm:

  --backend stock   the reference as shipped: FlashInfer wrappers + Triton slot write (layers/attention.py, core/kv_cache.py)
  --backend b200    the same engine with INTEGRATION.md section 1's monkey-patch: nanovllmx_b200.{Attention,
                    AttentionBackend, attention_kv_cache, KVCachePool, store_kvcache} swapped in, nothing else
  --backend oracle  CPU debugging arm (no GPU): oracle attention, gloo; exercises the harness plumbing only

What the harness adds around the unmodified tree (the same for every arm, none of it inside baseline/_ref):
  * the import fix the reference's bench.py needs (bench.py:4-5 import ``core.llm``; the package is ``nanovllmx``);
  * a stub tokenizer (nanovllmx/llm.py:8-14 wants a tokenizer directory; prompts here are token ids, bench.py:30);
  * a synthetic model directory (config.json of the named Qwen3 shape) and seeded random-init weights in place of the
    safetensors loader (model_loader.py:42-48) -- there are no checkpoints on the box.
The reference spawns its engine and worker processes with the ``spawn`` start method (core/engine_client.py:35,
core/worker_client.py:38), which re-imports this file in every child: the patches below are applied at import time
from the REF_ENGINE_CFG environment variable, so every process of one run sees the same arm.

Optional extensions on top of the b200 arm: --fused-qk (f-3), --b200-sampler (f-4),
--prefix-cache (f-1, both arms), --wire delta (f-2, both arms).

  python benchmarks/ref_engine.py --model 0.6b --backend b200 --greedy --out gpurun_out/eng.json
  python benchmarks/ref_engine.py --model 8b --backend stock --tp 2
"""
from __future__ import annotations

import argparse
import asyncio
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "baseline", "_ref")
CFG_ENV = "REF_ENGINE_CFG"
for p in (ROOT, REF_DIR):
    if p not in sys.path:
        sys.path.insert(0, p)

MODELS = {
    "0.6b": dict(num_hidden_layers=28, hidden_size=1024, intermediate_size=3072, num_attention_heads=16,
                 num_key_value_heads=8, head_dim=128, vocab_size=151936),
    "8b": dict(num_hidden_layers=36, hidden_size=4096, intermediate_size=12288, num_attention_heads=32,
               num_key_value_heads=8, head_dim=128, vocab_size=151936),
    "tiny": dict(num_hidden_layers=2, hidden_size=256, intermediate_size=512, num_attention_heads=8,
                 num_key_value_heads=4, head_dim=128, vocab_size=12000),
}


def write_model_dir(name: str, base: str = "/tmp/ref_engine_models") -> str:
    """config.json of a Qwen3 dense model of the named shape (what AutoConfig.from_pretrained reads,
    core/model_runner.py:81-96); no weights, no tokenizer files."""
    d = os.path.join(base, f"qwen3-{name}")
    os.makedirs(d, exist_ok=True)
    cfg = dict(architectures=["Qwen3ForCausalLM"], model_type="qwen3", hidden_act="silu", rms_norm_eps=1e-6,
               max_position_embeddings=40960, rope_theta=1000000, attention_bias=False, tie_word_embeddings=False,
               torch_dtype="bfloat16", dtype="bfloat16", bos_token_id=151643, eos_token_id=151645, **MODELS[name])
    with open(os.path.join(d, "config.json"), "w") as f:
        json.dump(cfg, f)
    return d


class StubTokenizer:
    """Stands in for AutoTokenizer (llm.py:8-14): text -> byte ids, ids -> their decimal text (so the token ids
    the engine produced come back through LLM.generate's ``token_str`` unchanged)."""
    eos_token_id = 151645
    pad_token = eos_token = unk_token = "<eos>"

    def __call__(self, texts):
        return {"input_ids": [[b + 1 for b in t.encode()] for t in texts]}

    def batch_decode(self, ids, skip_special_tokens=True):
        return [" ".join(str(int(i)) for i in row) for row in ids]


# =============================================================================================================
# patches (applied at import time in every process of a run)
# =============================================================================================================
def _seeded_init(cfg):
    import torch

    def load_model(model, hf_folder):
        """Stands in for model_loader.load_model (:42-48): normal(0, init_std) matrices, unit norm weights, from a
        seeded generator on the parameter's own device (every TP rank draws its shard from the same stream)."""
        dev = next(model.parameters()).device
        g = torch.Generator(device=dev)
        g.manual_seed(int(cfg.get("seed", 0)))
        with torch.no_grad():
            for name, p in model.named_parameters():
                if p.dim() >= 2:
                    p.normal_(0.0, float(cfg.get("init_std", 0.02)), generator=g)
                elif "norm" in name:
                    p.fill_(1.0)
                else:
                    p.zero_()
        return model.eval()

    return load_model


def _patch_b200(cfg):
    """INTEGRATION.md section 1, second form, verbatim."""
    import nanovllmx_b200 as b200
    import nanovllmx.layers.attention as A
    import nanovllmx.core.kv_cache as K
    import nanovllmx.core.model_runner as R
    import nanovllmx.models.qwen3 as M
    A.Attention = M.Attention = b200.Attention
    A.AttentionBackend = R.AttentionBackend = b200.AttentionBackend
    A.attention_kv_cache = R.attention_kv_cache = b200.attention_kv_cache
    K.KVCachePool = R.KVCachePool = b200.KVCachePool
    K.store_kvcache = b200.store_kvcache
    if cfg.get("fused_qk"):
        def forward(self, positions, hidden_states):
            """models/qwen3.py:117-136 with lines 123-134 replaced by the f-3 kernel."""
            qkv, _ = self.qkv_proj(hidden_states)
            o = self.attn.forward_fused_qkv(qkv, positions, self.q_norm.weight, self.k_norm.weight,
                                            self.q_norm.variance_epsilon, self.rotary_emb.cos_sin_cache)
            out, _ = self.o_proj(o)
            return out
        M.Qwen3Attention.forward = forward
    if cfg.get("b200_sampler"):
        R.Sampler = b200.Sampler


def _patch_cpu_debug(cfg):
    """No-GPU plumbing arm: gloo, device cpu, eager, oracle attention (test infrastructure)."""
    import torch
    import nanovllmx.core.worker as W
    import nanovllmx.core.model_runner as R
    import nanovllmx.layers.attention as A
    import nanovllmx.core.kv_cache as K
    import nanovllmx.models.qwen3 as M
    from benchmarks import _cpu_attention as C

    def init_environment(self):
        W.init_distributed_environment(word_size=self.world_size, rank=self.rank, backend="gloo",
                                       init_method=f"tcp://127.0.0.1:{self.nccl_port}")
        W.initialize_model_parallel(self.tp_size, self.pp_size, self.tp_rank, self.pp_rank)
        self.device = torch.device("cpu")
        self.model_runner = R.ModelRunner(model=self.model, max_bs=self.max_bs, rank=self.rank, device=self.device,
                                          enforce_eager=True, context_len=self.context_len)

    def destroy_environment(self):
        W.destroy_model_parallel()
        W.destroy_distributed_environment()

    W.Worker.init_environment = init_environment
    W.Worker.destroy_environment = destroy_environment
    R.set_cuda_arch = lambda: None
    R.ModelRunner.profile_kv_cache_size = lambda self, util=0.9: int(cfg.get("cpu_kv_slots", 8192))
    A.Attention = M.Attention = C.Attention
    A.AttentionBackend = R.AttentionBackend = C.AttentionBackend
    A.attention_kv_cache = R.attention_kv_cache = C.attention_kv_cache
    K.KVCachePool = R.KVCachePool = C.KVCachePool


def _patch_stats(cfg):
    """Per-step wall times of Engine.step (core/engine.py:81-100), dumped at shutdown: harness instrumentation."""
    import nanovllmx.core.engine as E
    import nanovllmx.core.scheduler as S
    steps = []
    orig_step, orig_shutdown, orig_sched = E.Engine.step, E.Engine.shutdown, S.Scheduler.schedule
    last = {}

    def schedule(self):
        t0 = time.perf_counter()
        b = orig_sched(self)
        last["sched_s"] = time.perf_counter() - t0
        last["mode"] = None if b is None else b.forward_mode.name
        last["bs"] = 0 if b is None else b.num_seqs
        last["new_tokens"] = 0 if b is None else sum(len(s.token_ids) - s.cached_kv_len for s in b.seqs)
        return b

    def step(self):
        t0 = time.perf_counter()
        out = orig_step(self)
        if last.get("mode"):
            steps.append((last["mode"], last["bs"], last["new_tokens"], time.perf_counter() - t0, last["sched_s"]))
        return out

    def shutdown(self):
        try:
            with open(cfg["stats_path"], "w") as f:
                json.dump({"steps": steps}, f)
        finally:
            orig_shutdown(self)

    S.Scheduler.schedule = schedule
    E.Engine.step = step
    E.Engine.shutdown = shutdown


def apply_patches(cfg: dict) -> None:
    import torch  # noqa: F401
    import nanovllmx.llm as L
    import nanovllmx.core.model_runner as R
    import nanovllmx.core.engine as E
    L.init_tokenizer = lambda model: StubTokenizer()
    R.load_model = _seeded_init(cfg)
    if not cfg.get("compiled_sampler"):
        # torch 2.11's Inductor cannot compile the reference sampler once its batch dimension turns symbolic
        # (layers/sampler.py:19 @torch.compile; the top-p cumsum over 151 936 columns becomes a SplitScan and
        # codegen dies with "list indices must be integers or slices, not NoneType": gpurun_out/r2s1_eng06_stock_greedy.log).
        # An environment bug in a component outside the path: Sampler.forward runs eagerly, in EVERY arm.
        import nanovllmx.layers.sampler as SM
        SM.Sampler.forward = getattr(SM.Sampler.forward, "_torchdynamo_orig_callable", SM.Sampler.forward)
    if cfg.get("argmax"):
        # Token-identity experiments only: plain argmax in EVERY arm.  The reference's deterministic form (top_k=1,
        # layers/sampler.py:68-74) keeps all logits tied with the maximum and lets torch.multinomial pick among them,
        # so exact bf16 ties (frequent on random-init weights) make even two runs of one arm differ.
        import nanovllmx.layers.sampler as SM
        SM.Sampler.forward = lambda self, logits, *a, **k: logits.argmax(dim=-1)
    if not cfg.get("verbose"):
        E.print = lambda *a, **k: None          # "Added sequence <id>" x 256 (core/engine.py:70)
    if cfg["backend"] == "b200":
        _patch_b200(cfg)
    elif cfg["backend"] == "oracle":
        _patch_cpu_debug(cfg)
    if cfg.get("prefix_cache"):          # SURVEY.md 8 f-1 (either arm): admission-time radix-tree reuse + cache_sequence fixes
        from nanovllmx_b200.core import prefix_reuse
        prefix_reuse.install()
    if cfg.get("wire") == "delta":       # SURVEY.md 8 f-2 (either arm): per-step delta messages over the reference's own queues
        from nanovllmx_b200.core import wire_binding
        wire_binding.install()
    if cfg.get("stats_path"):
        _patch_stats(cfg)


if os.environ.get(CFG_ENV):          # spawned engine / worker process (or the parent after main() set it)
    apply_patches(json.loads(os.environ[CFG_ENV]))


# =============================================================================================================
# workloads
# =============================================================================================================
def bench_workload(num_seqs=256, max_input_len=1024, max_output_len=1024):
    """bench.py:13-29 verbatim (python ``random``, seed 0; max_tokens drawn after all prompts)."""
    from random import randint, seed
    seed(0)
    prompts = [[randint(0, 10000) for _ in range(randint(100, max_input_len))] for _ in range(num_seqs)]
    max_tokens = [randint(100, max_output_len) for _ in range(num_seqs)]
    return prompts, max_tokens, None


def shared_prefix_workload(num_seqs=256, prefix_len=2048, gen=128):
    """BASELINE.json configs[3] / SURVEY.md 8d config 4: one 2048-token system prompt (seed 1) + per-request suffix of
    randint(16,256) tokens, 128 new tokens each."""
    import random
    rng = random.Random(1)
    prefix = [rng.randint(0, 10000) for _ in range(prefix_len)]
    prompts = [prefix + [rng.randint(0, 10000) for _ in range(rng.randint(16, 256))] for _ in range(num_seqs)]
    return prompts, None, gen


async def _consume(gen, sink):
    async for out in gen:
        if out.token_str:
            sink.append(int(out.token_str))


async def run(cfg: dict) -> dict:
    from nanovllmx.llm import LLM
    from nanovllmx.core.common import SamplingParams
    if cfg["workload"] == "bench":
        prompts, max_tokens, max_new = bench_workload(cfg["num_seqs"])
    else:
        prompts, max_tokens, max_new = shared_prefix_workload(cfg["num_seqs"], cfg.get("prefix_len", 2048), cfg.get("gen", 128))
    tp = cfg["tp"]
    t_init = time.time()
    llm = LLM(model=cfg["model_dir"], gpu_memory_utilization=0.8, max_bs=cfg["max_bs"], tp_size=tp, pp_size=1,
              nccl_port=cfg["nccl_port"], device_ids=None if cfg["backend"] == "oracle" else (cfg.get("device_ids") or list(range(tp))),
              enforce_eager=cfg["enforce_eager"], context_len=cfg["context_len"])

    def params(i):
        kw = dict(ignore_eos=True)
        if max_tokens is not None:
            kw["max_tokens"] = max_tokens[i]
        else:
            kw["max_new_tokens"] = max_new
        if cfg["greedy"]:        # the reference has no greedy mode (SURVEY.md header): top_k=1 is its deterministic form
            kw.update(temperature=1.0, top_k=1)
        else:
            kw.update(temperature=0.6)
        return SamplingParams(**kw)

    warm = []
    await _consume(llm.generate("Benchmark: ", SamplingParams(max_tokens=10)), warm)       # bench.py:31
    if cfg["workload"] == "shared-prefix":
        # the system prompt has been served once before the measured requests arrive (the reference inserts a sequence
        # into the radix tree when it FINISHES, core/scheduler.py:189): one priming request, in every arm
        await _consume(llm.generate(list(prompts[0][:cfg.get("prefix_len", 2048)]) + [1, 2, 3],
                                    SamplingParams(max_new_tokens=1, ignore_eos=True, temperature=1.0, top_k=1)), warm)
    init_s = time.time() - t_init
    runs = []
    tokens_runs = []
    for r in range(cfg["runs"]):
        sinks = [[] for _ in prompts]
        t = time.time()
        await asyncio.gather(*[_consume(llm.generate(list(p), params(i)), sinks[i]) for i, p in enumerate(prompts)])
        t = time.time() - t
        gen = sum(len(s) for s in sinks)
        ref_total = sum(max_tokens) if max_tokens is not None else gen
        runs.append(dict(wall_s=t, tok_s_reference_formula=ref_total / t, tok_s_generated=gen / t, generated=gen,
                         reference_formula_tokens=ref_total))
        tokens_runs.append(sinks)
    llm.shutdown()
    return dict(init_s=init_s, runs=runs, tokens=tokens_runs[0], tokens_runs=tokens_runs,
                prompt_tokens=sum(len(p) for p in prompts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="0.6b", choices=sorted(MODELS))
    ap.add_argument("--backend", default="b200", choices=["stock", "b200", "oracle"])
    ap.add_argument("--tp", type=int, default=1)
    ap.add_argument("--num-seqs", type=int, default=256)
    ap.add_argument("--max-bs", type=int, default=256)
    ap.add_argument("--context-len", type=int, default=4096)
    ap.add_argument("--workload", default="bench", choices=["bench", "shared-prefix"])
    ap.add_argument("--runs", type=int, default=2, help="timed passes over the workload; the first follows the "
                    "reference bench.py protocol (only a 1-request warm-up before it), later ones are steady state")
    ap.add_argument("--greedy", action="store_true", help="top_k=1, temperature 1: the reference's deterministic form")
    ap.add_argument("--argmax", action="store_true", help="token-identity experiments: plain argmax sampler in every arm")
    ap.add_argument("--enforce-eager", action="store_true")
    ap.add_argument("--fused-qk", action="store_true")
    ap.add_argument("--b200-sampler", action="store_true")
    ap.add_argument("--prefix-cache", action="store_true")
    ap.add_argument("--wire", default="pickle", choices=["pickle", "delta"])
    ap.add_argument("--prefix-len", type=int, default=2048, help="shared-prefix workload: tokens of the common system prompt")
    ap.add_argument("--gen", type=int, default=128, help="shared-prefix workload: new tokens per request")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--init-std", type=float, default=0.02)
    ap.add_argument("--nccl-port", type=int, default=29511)
    ap.add_argument("--device-ids", default=None, help="comma-separated physical GPU ids (default 0..tp-1)")
    ap.add_argument("--stats", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    if not os.path.isdir(os.path.join(REF_DIR, "nanovllmx", "core")):
        raise SystemExit(f"{REF_DIR} is missing or incomplete: run `python baseline/install_ref.py` in the build container")
    cfg = dict(model=a.model, model_dir=write_model_dir(a.model), backend=a.backend, tp=a.tp, num_seqs=a.num_seqs,
               max_bs=a.max_bs, context_len=a.context_len, workload=a.workload, runs=a.runs, greedy=a.greedy or a.argmax, argmax=a.argmax,
               enforce_eager=a.enforce_eager or a.backend == "oracle", fused_qk=a.fused_qk, b200_sampler=a.b200_sampler,
               prefix_cache=a.prefix_cache, wire=a.wire, seed=a.seed, init_std=a.init_std, nccl_port=a.nccl_port,
               verbose=a.verbose, prefix_len=a.prefix_len, gen=a.gen, device_ids=[int(x) for x in a.device_ids.split(",")] if a.device_ids else None)
    if a.stats:
        cfg["stats_path"] = (a.out or "/tmp/ref_engine") + ".steps.json"
    os.environ[CFG_ENV] = json.dumps(cfg)
    if a.backend == "oracle":
        os.environ["TORCHDYNAMO_DISABLE"] = "1"
    apply_patches(cfg)
    res = asyncio.run(run(cfg))
    summary = dict(cfg=cfg, init_s=res["init_s"], runs=res["runs"], prompt_tokens=res["prompt_tokens"])
    if cfg.get("stats_path") and os.path.exists(cfg["stats_path"]):
        with open(cfg["stats_path"]) as f:
            steps = json.load(f)["steps"]
        for mode in ("PREFILL", "DECODE"):
            xs = [s for s in steps if s[0] == mode]
            if xs:
                summary[f"{mode.lower()}_steps"] = dict(n=len(xs), total_s=sum(s[3] for s in xs),
                                                        sched_s=sum(s[4] for s in xs),
                                                        mean_ms=1e3 * sum(s[3] for s in xs) / len(xs))
    print("REF_ENGINE_RESULT " + json.dumps(summary))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(dict(summary, tokens=res["tokens"], tokens_runs=res["tokens_runs"]), f)


if __name__ == "__main__":
    main()
