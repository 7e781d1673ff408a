#!/usr/bin/env python
"""Arbitrate greedy-token divergences between two runs of benchmarks/ref_engine.py (stock FlashInfer arm vs b200 arm).

north_star: "greedy generations must be token-identical on the bench workload".  On random-init weights the top-2
logits of a 151 936-way bf16 head are within a few bf16 ulps of each other at a large fraction of the steps, so two
correct attention implementations that differ in the last bits flip an argmax sooner or later.  This tool makes that
statement checkable instead of eyeballed.  For every sequence whose token lists differ it takes the common prefix
(prompt + identical generated tokens) and runs ONE eager prefill forward of the stock reference model
(baseline/_ref ModelRunner, same seeded weights as the engine runs) per attention flavour:

  fi    the reference's FlashInfer prefill wrapper                       (layers/attention.py:49-53, 106-119)
  b200  this repo's prefill kernel behind AttentionPlan.run               (same stock Attention module, same Triton store)
  fp32  gather + fp32 softmax(QK^T)V on the same bf16 K/V                (the oracle's arithmetic, on the GPU)

and records, from the last position's logits: the two tokens the arms picked, their fp32-flavour logit difference
(the "margin" that decided the flip), the bf16 ulp at that logit magnitude, and max |logits - logits_fp32| for fi
and b200.  tests/test_zz_gpu_ref_engine.py asserts on the report: every divergence is decided by a margin of at most
MARGIN_ULPS bf16 ulps, and b200's logit error against fp32 is not larger than FlashInfer's beyond noise.

  python benchmarks/token_identity.py --stock a.json --b200 b.json --out report.json
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))

MARGIN_ULPS = 4.0      # a flip must be decided by <= this many bf16 ulps (at the magnitude of the top logit)


def bf16_ulp(x: float) -> float:
    return 2.0 ** (math.floor(math.log2(max(abs(x), 1e-30))) - 7)


def first_divergence(a, b):
    n = min(len(a), len(b))
    for i in range(n):
        if a[i] != b[i]:
            return i
    return None if len(a) == len(b) else n


def compare(tok_a, tok_b):
    ident, div = 0, []
    for i, (a, b) in enumerate(zip(tok_a, tok_b)):
        d = first_divergence(a, b)
        if d is None:
            ident += 1
        else:
            div.append((i, d))
    matched = sum((first_divergence(a, b) if first_divergence(a, b) is not None else len(a)) for a, b in zip(tok_a, tok_b))
    return ident, div, matched, sum(len(a) for a in tok_a)


class _Fp32Wrapper:
    """``wrapper`` whose run() is the oracle's arithmetic on the GPU (one sequence, slots 0..L-1, causal)."""
    _sm_scale = None

    def run(self, q, kv):
        import torch
        k_cache, v_cache = kv
        T, Hq, D = q.shape
        Hkv = k_cache.shape[-2]
        K = k_cache.reshape(-1, Hkv, D)[:T].float().repeat_interleave(Hq // Hkv, dim=1).transpose(0, 1)   # [Hq, T, D]
        V = v_cache.reshape(-1, Hkv, D)[:T].float().repeat_interleave(Hq // Hkv, dim=1).transpose(0, 1)
        Q = q.float().transpose(0, 1)
        s = torch.matmul(Q, K.transpose(1, 2)) * self._sm_scale
        s = s.masked_fill(torch.ones(T, T, dtype=torch.bool, device=q.device).triu(1), float("-inf"))
        o = torch.matmul(torch.softmax(s, dim=-1), V)
        return o.transpose(0, 1).contiguous().to(q.dtype)


def arbitrate(cfg, prompts, tok_stock, tok_b200, max_cases=None):
    import torch
    from benchmarks import ref_engine as H
    stock_cfg = dict(cfg, backend="stock", fused_qk=False, b200_sampler=False, prefix_cache=False, wire="pickle")
    stock_cfg.pop("stats_path", None)
    H.apply_patches(stock_cfg)                       # seeded init + stub tokenizer only; attention stays the reference's
    import nanovllmx_b200 as b200
    from nanovllmx.core.common import ForwardBatch, ForwardMode, Sequence
    from nanovllmx.core.model_runner import ModelRunner
    from nanovllmx.distributed.parallel_state import init_distributed_environment, initialize_model_parallel
    from nanovllmx.layers.attention import AttentionMetadata, attention_kv_cache
    torch.set_float32_matmul_precision("highest")
    init_distributed_environment(word_size=1, rank=0, backend="nccl", init_method=f"tcp://127.0.0.1:{cfg['nccl_port'] + 7}")
    initialize_model_parallel(1, 1, 0, 0)
    dev = torch.device("cuda:0")
    runner = ModelRunner(model=cfg["model_dir"], max_bs=1, rank=0, device=dev, enforce_eager=True,
                         context_len=cfg["context_len"])
    runner.load_model()
    runner.initialize_kv_cache(2 * cfg["context_len"])
    mine = b200.AttentionBackend(runner.kv_cache, runner.num_heads, runner.num_kv_heads, runner.head_dim, runner.dtype, dev)
    fp32 = _Fp32Wrapper()
    _, div, matched, total = compare(tok_stock, tok_b200)
    cases = []

    @torch.inference_mode()
    def logits_with(batch, md):
        ids, pos = runner.prepare_input(batch)
        with attention_kv_cache(runner.model, md):
            h = runner.model(ids, pos, None)
        return runner.model.compute_logits(runner.prepare_last_hidden_states(batch, h)).float()[0]

    for n, (i, d) in enumerate(div):
        if max_cases is not None and n >= max_cases:
            break
        prefix = list(prompts[i]) + list(tok_stock[i][:d])
        seq = Sequence(seq_id=f"arb{i}", token_ids=prefix, num_tokens=len(prefix), prompt_len=len(prompts[i]),
                       kv_indices=list(range(len(prefix))), cached_kv_len=0)
        batch = ForwardBatch(forward_mode=ForwardMode.PREFILL, num_seqs=1, seqs=[seq])
        md_fi = runner.attn_backend.build_metadata(batch)
        l_fi = logits_with(batch, md_fi)
        md_b = mine.build_metadata(batch)
        l_b = logits_with(batch, md_b)
        l_32 = logits_with(batch, AttentionMetadata(ForwardMode.PREFILL, runner.kv_cache, md_fi.output_kv_indices, fp32))
        ta = tok_stock[i][d] if d < len(tok_stock[i]) else -1
        tb = tok_b200[i][d] if d < len(tok_b200[i]) else -1
        top = torch.topk(l_32, 2)
        ulp = bf16_ulp(float(top.values[0]))
        margin = abs(float(l_32[ta] - l_32[tb])) if ta >= 0 and tb >= 0 else float("inf")
        cases.append(dict(seq=i, pos=d, prefix_len=len(prefix), tok_stock=ta, tok_b200=tb,
                          fp32_top2=[int(x) for x in top.indices], fp32_top_logit=float(top.values[0]),
                          fp32_top2_margin=float(top.values[0] - top.values[1]), margin_between_picks=margin,
                          bf16_ulp=ulp, margin_ulps=margin / ulp,
                          rank_of_stock_pick=int((l_32 > l_32[ta]).sum()) if ta >= 0 else -1,
                          rank_of_b200_pick=int((l_32 > l_32[tb]).sum()) if tb >= 0 else -1,
                          err_fi=float((l_fi - l_32).abs().max()), err_b200=float((l_b - l_32).abs().max()),
                          rms_fi=float((l_fi - l_32).pow(2).mean().sqrt()), rms_b200=float((l_b - l_32).pow(2).mean().sqrt())))
    return dict(sequences=len(tok_stock), identical_sequences=len(tok_stock) - len(div), generated_tokens=total,
                tokens_in_identical_prefixes=matched, divergences=len(div), arbitrated=len(cases),
                margin_ulps_bound=MARGIN_ULPS, cases=cases)


def summarise(rep):
    cs = rep["cases"]
    if not cs:
        return dict(rep, max_margin_ulps=0.0, mean_rms_fi=0.0, mean_rms_b200=0.0)
    mu = sorted(c["margin_ulps"] for c in cs)
    return dict(rep, max_margin_ulps=mu[-1], median_margin_ulps=mu[len(mu) // 2],
                mean_rms_fi=sum(c["rms_fi"] for c in cs) / len(cs), mean_rms_b200=sum(c["rms_b200"] for c in cs) / len(cs),
                mean_err_fi=sum(c["err_fi"] for c in cs) / len(cs), mean_err_b200=sum(c["err_b200"] for c in cs) / len(cs),
                b200_closer_to_fp32_in=sum(c["rms_b200"] <= c["rms_fi"] for c in cs))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stock", required=True)
    ap.add_argument("--b200", required=True)
    ap.add_argument("--out", default=None)
    ap.add_argument("--max-cases", type=int, default=None)
    a = ap.parse_args()
    A, B = json.load(open(a.stock)), json.load(open(a.b200))
    cfg = A["cfg"]
    for k in ("model", "num_seqs", "workload", "seed", "init_std", "greedy", "context_len"):
        assert A["cfg"][k] == B["cfg"][k], f"the two runs differ in {k}"
    from benchmarks import ref_engine as H
    prompts = (H.bench_workload(cfg["num_seqs"]) if cfg["workload"] == "bench" else H.shared_prefix_workload(cfg["num_seqs"]))[0]
    rep = summarise(arbitrate(cfg, prompts, A["tokens"], B["tokens"], a.max_cases))
    rep["arms"] = dict(stock=A["cfg"]["backend"], b200=B["cfg"]["backend"], model=cfg["model"])
    print("TOKEN_IDENTITY " + json.dumps({k: v for k, v in rep.items() if k != "cases"}))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(rep, f, indent=1)


if __name__ == "__main__":
    main()
