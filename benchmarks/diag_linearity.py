#!/usr/bin/env python
"""Diagnostic for the linearity property of the full-bench-step prefill test: run-to-run determinism of the kernel and
the size / location of the worst linearity residual for several seeds of v2."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import nanovllmx_b200 as pkg
from oracle.paged_attention_oracle import bench_workload
from tests.util import dev, make_paged_case

plens = [len(p) for p in bench_workload()[0]]
Hq, Hkv, D = 32, 8, 128
c = dev(make_paged_case(plens, Hq, Hkv, q_lens=plens, seed=11))
T, B, sm = sum(plens), len(plens), D ** -0.5
ws = torch.empty(max(pkg.ops.prefill_workspace_bytes(B, T, Hq, Hkv, D), 1 << 20), dtype=torch.uint8, device="cuda")
run = lambda v: pkg.ops.prefill(c["q"], c["k_cache"], v, c["qo_indptr"], c["kv_indptr"], c["kv_indices"], c["last_page_len"], ws, 1, sm, True)
o1 = run(c["v_cache"])
neq = 0
for i in range(6):
    o = run(c["v_cache"])
    if not torch.equal(o, o1):
        neq += 1
        dlt = (o.float() - o1.float()).abs()
        idx = int(dlt.view(T, -1).max(1).values.argmax())
        print(f"run {i}: NOT bit-identical, max diff {float(dlt.max()):.4g} at token row {idx}, rows differing {int((dlt.view(T, -1).max(1).values > 0).sum())}")
print("determinism: ", "all bit-identical" if neq == 0 else f"{neq} of 6 repeats differ")
qo = c["qo_indptr"].cpu().numpy()
o1f = o1.float()
for seed in range(8):
    g = torch.Generator(device="cuda").manual_seed(seed)
    v2 = torch.randn(c["v_cache"].shape, device="cuda", dtype=torch.float32, generator=g).to(torch.bfloat16)
    o2 = run(v2).float()
    o12 = run((c["v_cache"].float() + v2.float()).to(torch.bfloat16)).float()
    bound = 1e-2 + 2.0 ** -8 * (o1f.abs() + o2.abs() + 2 * o12.abs())
    res = (o12 - (o1f + o2)).abs()
    viol = (res - bound).view(T, -1).max(1).values
    w = int(viol.argmax())
    s = int(np.searchsorted(qo, w, side="right") - 1)
    print(f"seed {seed}: worst residual-bound {float(viol.max()):+.4g} (residual {float(res.view(T, -1)[w].max()):.4g}) at row {w} = sequence {s} token {w - qo[s]} of {qo[s + 1] - qo[s]}; violations {int((viol > 0).sum())}")
