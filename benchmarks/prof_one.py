#!/usr/bin/env python
"""Tiny driver for ncu captures: runs ONE kernel family a few times on a fixed case.
  python benchmarks/prof_one.py prefill|prefill_short|decode_cfg5|decode_tp8|decode_short_tp8|decode_small|decode_bench [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nanovllmx_b200 as pkg
from benchmarks.sweep import decode_case, prefill_case, DEV
from oracle.paged_attention_oracle import bench_workload

what = sys.argv[1] if len(sys.argv) > 1 else "prefill"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
if what.startswith("prefill"):
    if what == "prefill":
        c = prefill_case([4096] * 8, [0] * 8, 32, 8)
    else:
        pl = [len(p) for p in bench_workload()[0]][:64]
        c = prefill_case(pl, [0] * 64, 32, 8)
    ws = torch.empty(pkg.ops.prefill_workspace_bytes(c["B"], c["T"], 32, 8, 128), dtype=torch.uint8, device=DEV)
    f = lambda: pkg.ops.prefill(c["q"], c["k"], c["v"], c["qo"], c["indptr"], c["idx"], c["last"], ws, 1, 128 ** -0.5, True)
else:
    c = {"decode_cfg5": lambda: decode_case(64, 32768, 32, 8, 16), "decode_tp8": lambda: decode_case(64, 32768, 4, 1, 16), "decode_tp8_p1": lambda: decode_case(64, 32768, 4, 1, 1), "decode_short_tp8": lambda: decode_case(256, 2048, 4, 1, 1, True),
         "decode_small": lambda: decode_case(8, 1024, 32, 8, 1, True)}.get(what, lambda: decode_case(256, 2048, 16, 8, 1, True))()
    ws = torch.empty(pkg.ops.decode_workspace_bytes(c["B"], c["Hq"], c["Hkv"], 128), dtype=torch.uint8, device=DEV)
    f = lambda: pkg.ops.decode(c["q"], c["k"], c["v"], c["indptr"], c["idx"], c["last"], ws, c["ps"], 128 ** -0.5)
for _ in range(iters):
    f()
torch.cuda.synchronize()
print("ok", what)
