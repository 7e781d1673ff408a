#!/usr/bin/env python
"""Warp-role breakdown of an ncu --set full capture of the warp-specialised prefill kernel.

  ncu -i X.ncu-rep --page source --csv --print-source sass > X_src.csv ; python benchmarks/ncu_roles.py X_src.csv

Splits the SASS listing at the USETMAXREG instructions (every role starts with one), then reports per region: warp
state samples, samples on mbarrier-wait sites (SYNCS / NANOSLEEP / the spin branch), executed warp instructions and
the top stall sites."""
import csv
import sys


def main(path, top=8):
    rows = list(csv.reader(open(path)))
    hdr = rows[1]
    data = rows[2:]
    isrc, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    bounds = [0] + [i for i, r in enumerate(data) if "USETMAXREG" in r[isrc]] + [len(data)]
    total = sum(int(r[isamp]) for r in data)
    print(f"{path}: {total} samples, {len(data)} SASS instructions")
    for a, b in zip(bounds[:-1], bounds[1:]):
        reg = data[a:b]
        if not reg:
            continue
        samp = sum(int(r[isamp]) for r in reg)
        waits = sum(int(r[isamp]) for r in reg if any(k in r[isrc] for k in ("SYNCS", "NANOSLEEP", "WARPSYNC")))
        ex = sum(int(r[iex]) for r in reg)
        ops = {}
        for k in ("UTCHMMA", "UTMALDG", "LDGSTS", "MUFU.EX2", "STS", "LDTM", "STTM", "UTCBAR"):
            n = sum(1 for r in reg if k in r[isrc])
            if n:
                ops[k] = n
        print(f"== region @{a:5d}..{b:5d} first={reg[0][isrc].strip()[:40]!r} samples {samp} ({100.0 * samp / max(total, 1):.1f} %) "
              f"in waits {waits} ({100.0 * waits / max(samp, 1):.0f} %) warp-instr {ex} sites {ops}")
        for r in sorted(reg, key=lambda r: -int(r[isamp]))[:top]:
            print(f"      {r[isamp]:>7}  {r[isrc].strip()[:100]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 8)


def reasons(path):
    """Per region: totals of each stall reason (columns stall_* of the source page)."""
    rows = list(csv.reader(open(path)))
    hdr, data = rows[1], rows[2:]
    isrc = hdr.index("Source")
    cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    bounds = [0] + [i for i, r in enumerate(data) if "USETMAXREG" in r[isrc]] + [len(data)]
    for a, b in zip(bounds[:-1], bounds[1:]):
        tot = {h: sum(int(r[i] or 0) for r in data[a:b]) for i, h in cols}
        s = sum(tot.values())
        if s:
            print(f"   region @{a}: " + "  ".join(f"{h[6:]} {100.0 * v / s:.0f}%" for h, v in sorted(tot.items(), key=lambda kv: -kv[1]) if v * 50 > s))


if __name__ == "__main__" and len(sys.argv) > 3 and sys.argv[3] == "reasons":
    reasons(sys.argv[1])
