#!/usr/bin/env python
"""Sampler tail timing on one GPU: b200attn::sample next to the reference's torch pipeline restated
(layers/sampler.py:38-55 with _apply_top_k_top_p's full-vocabulary sort, as core/model_runner.py:191-219 always drives
it) and next to the bare softmax + multinomial the engine harness uses.  CUDA events, L2 flushed between iterations."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import nanovllmx_b200 as pkg


def reference_sampler(logits, temps, min_ps, top_ps, top_ks):
    x = logits.to(torch.float)
    x.div_(temps.unsqueeze(1))
    xs, idx = x.sort(dim=-1, descending=False)
    kth = xs.gather(1, (xs.size(1) - top_ks).unsqueeze(1))
    xs.masked_fill_(xs < kth, -float("inf"))
    ps = xs.softmax(dim=-1).cumsum(dim=-1)
    mask = ps <= 1 - top_ps.unsqueeze(1)
    mask[:, -1] = False
    xs.masked_fill_(mask, -float("inf"))
    x = torch.empty_like(xs).scatter_(-1, idx, xs)
    probs = torch.softmax(x, dim=-1)
    x = x.masked_fill_(probs < min_ps.unsqueeze(1) * probs.max(dim=-1, keepdim=True).values, -float("inf"))
    return torch.multinomial(torch.softmax(x, dim=-1, dtype=torch.float), 1).squeeze(-1)


def timeit(fn, flush, iters=20, warm=3):
    for _ in range(warm):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        flush.zero_()
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in ev)
    return ts[len(ts) // 2]


def main():
    dev = torch.device("cuda")
    V = 151936
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    for B in (1, 16, 92, 256):
        g = torch.Generator(device=dev).manual_seed(B)
        logits = (torch.randn(B, V, generator=g, device=dev) * 2).to(torch.bfloat16)
        t = torch.full((B,), 0.6, device=dev)
        kV = torch.full((B,), V, dtype=torch.long, device=dev)
        k50 = torch.full((B,), 50, dtype=torch.long, device=dev)
        p1, p9 = torch.ones(B, device=dev), torch.full((B,), 0.9, device=dev)
        m0, m05 = torch.zeros(B, device=dev), torch.full((B,), 0.05, device=dev)
        row = dict(case="sampler", B=B, V=V, MB=B * V * 2 / 1e6)
        row["b200_greedy_ms"] = timeit(lambda: pkg.ops.sample(logits, None, None, None, None, None, 0, 0), flush)
        row["b200_temperature_ms"] = timeit(lambda: pkg.ops.sample(logits, t, m0, p1, kV, None, 1, 2), flush)
        row["b200_topk50_topp09_minp_ms"] = timeit(lambda: pkg.ops.sample(logits, t, m05, p9, k50, None, 1, 2), flush)
        row["b200_topp09_ms"] = timeit(lambda: pkg.ops.sample(logits, t, m0, p9, kV, None, 1, 2), flush)
        row["torch_argmax_ms"] = timeit(lambda: torch.argmax(logits, -1), flush)
        row["torch_softmax_multinomial_ms"] = timeit(lambda: torch.multinomial(torch.softmax(logits.float() / 0.6, -1), 1), flush)
        row["reference_pipeline_ms"] = timeit(lambda: reference_sampler(logits, t, m0, p1, kV), flush, iters=5, warm=2)
        row["temperature_GBs"] = row["MB"] / row["b200_temperature_ms"]
        row["temperature_frac_measured"] = row["temperature_GBs"] / peak
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
