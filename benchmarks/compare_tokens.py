#!/usr/bin/env python
"""Token-identity check between two `engine_bench.py --greedy --dump-tokens` runs.  Each entry is (token, top-2 logit
margin); a divergence whose margin is within bf16 noise of the logits is a numerical tie, not a kernel bug."""
import json, sys
a, b = json.load(open(sys.argv[1])), json.load(open(sys.argv[2]))
tot = same = seqs_same = 0
div = []
for k in a:
    x, y = a[k], b.get(k, [])
    n = min(len(x), len(y))
    d = next((i for i in range(n) if x[i][0] != y[i][0]), None)
    tot += n
    same += n if d is None else d
    if d is None and len(x) == len(y):
        seqs_same += 1
    elif d is not None:
        div.append(dict(seq=k, at=d, of=len(x), margin_a=x[d][1], margin_b=y[d][1]))
div.sort(key=lambda r: -max(r["margin_a"], r["margin_b"]))
print(json.dumps(dict(sequences=len(a), identical_sequences=seqs_same, tokens_compared=tot, identical_prefix_tokens=same,
                      diverged=len(div), largest_margin_at_divergence=div[0] if div else None,
                      median_margin_at_divergence=sorted(max(r["margin_a"], r["margin_b"]) for r in div)[len(div) // 2] if div else None)))
