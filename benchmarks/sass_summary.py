#!/usr/bin/env python
"""SASS opcode summary per kernel of libb200attn.so (cuobjdump -sass; runs without a GPU).

  python benchmarks/sass_summary.py [path/to/libb200attn.so] > profiles/rN_sass_opcode_summary.txt

Counts the mnemonics that prove which hardware path a kernel uses (B200_PROFILING.md): UTCHMMA = tcgen05.mma,
LDTM / STTM = tcgen05.ld / st (TMEM), UTCBAR = tcgen05.commit, UTMALDG = cp.async.bulk.tensor (TMA load),
UBLKCP = cp.async.bulk, LDGSTS = cp.async, SYNCS = mbarrier, HMMA = mma.sync, MUFU.EX2, FFMA2 / FADD2 = packed f32x2."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "nano-vllm-x_b200", "libb200attn.so")
OPS = ["UTCHMMA", "LDTM", "STTM", "UTCBAR", "UTMALDG", "UTMAPF", "UBLKCP", "LDGSTS", "SYNCS", "HMMA", "LDSM", "MUFU.EX2",
       "FFMA2", "FADD2", "FFMA", "FMNMX3", "F2FP", "STS", "LDS", "LDG", "STG", "ATOMG", "RED", "USETMAXREG", "ELECT", "NANOSLEEP"]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
kern, counts, total = None, collections.OrderedDict(), collections.Counter()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts.setdefault(kern, collections.Counter())
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m and kern:
        op = m.group(1)
        total[kern] += 1
        for o in OPS:
            if op == o or op.startswith(o + "."):
                counts[kern][o] += 1
                break
print(f"# {os.path.relpath(lib, ROOT)}: SASS instruction counts per kernel (cuobjdump -sass, sm_100a)")
for k, c in counts.items():
    print(f"{k}  [{total[k]} instructions]")
    print("    " + "  ".join(f"{o}={c[o]}" for o in OPS if c[o]))
