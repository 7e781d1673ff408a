"""Generate tests/golden/sampler_golden.npz by RUNNING THE REFERENCE'S OWN Sampler on CPU.

Run in the authoring container only (needs /root/reference):

    TORCHDYNAMO_DISABLE=1 python oracle/gen_golden_sampler.py

Unmodified from /root/reference/nanovllmx/layers/sampler.py (``@torch.compile``; TORCHDYNAMO_DISABLE=1
runs the Python body eagerly), driven as core/model_runner.py:191-219, 262-272 drives it: bf16 logits,
fp32 temperatures / min_ps / top_ps, int64 top_ks with -1 already replaced by vocab_size.  For every
case the global generator is seeded, the reference is run, and the Exp(1) noise ``q`` its
``torch.multinomial`` call consumed is recovered by re-seeding and drawing
``empty(B, V, dtype=float).exponential_(1)`` (the first and only RNG draw of that call).  Logits are
stored as raw bf16 bit patterns; ``q`` as fp32.
"""
import os
import sys

os.environ.setdefault("TORCHDYNAMO_DISABLE", "1")
sys.path.insert(0, "/root/reference")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch


def bits(t):
    return t.contiguous().view(torch.int16).numpy().astype(np.uint16).copy()


CASES = {
    # name: (B, V, seed, logit scale, temperature, top_k, top_p, min_p); scalars are broadcast, lists per row
    "bench_defaults": (6, 4099, 0, 4.0, 0.6, -1, 1.0, 0.0),          # bench.py:29 SamplingParams(temperature=0.6)
    "top_k": (5, 3001, 1, 3.0, 1.0, [1, 2, 17, 300, 3001], 1.0, 0.0),
    "top_p": (5, 2048, 2, 3.0, 0.8, -1, [0.1, 0.5, 0.9, 0.99, 1.0], 0.0),
    "min_p": (4, 2500, 3, 2.0, 1.3, -1, 1.0, [0.0, 0.01, 0.2, 0.9]),
    "all": (6, 5000, 4, 5.0, [0.3, 0.6, 1.0, 1.5, 0.7, 2.0], [50, 5, -1, 1000, 2, 5000],
            [0.9, 0.95, 0.5, 1.0, 0.3, 0.8], [0.0, 0.05, 0.0, 0.1, 0.5, 0.0]),
}


def main():
    from nanovllmx.layers.sampler import Sampler
    from oracle import sampler_oracle as so

    sampler = Sampler()
    out = {}
    for name, (B, V, seed, scale, T, k, p, mp) in CASES.items():
        g = torch.Generator().manual_seed(seed)
        logits = (torch.randn(B, V, generator=g) * scale).to(torch.bfloat16)
        vec = lambda x, dt: torch.tensor(x if isinstance(x, list) else [x] * B, dtype=dt)
        temps, top_ps, min_ps = vec(T, torch.float), vec(p, torch.float), vec(mp, torch.float)
        top_ks = vec(k, torch.long)
        top_ks = torch.where(top_ks == -1, torch.full_like(top_ks, V), torch.clamp(top_ks, max=V))  # model_runner.py:203-206
        torch.manual_seed(1000 + seed)
        ids = sampler(logits.clone(), temps.clone(), min_ps.clone(), top_ps.clone(), top_ks.clone())
        torch.manual_seed(1000 + seed)
        q = torch.empty(B, V, dtype=torch.float).exponential_(1)
        # the restatement must reproduce the reference's own output before it is written anywhere
        mine, _ = so.sample(logits, temps, min_ps.clone(), top_ps, top_ks, q)
        assert torch.equal(mine, ids), (name, mine.tolist(), ids.tolist())
        out[f"{name}.logits_bf16"] = bits(logits)
        out[f"{name}.temperatures"] = temps.numpy()
        out[f"{name}.min_ps"] = min_ps.numpy()
        out[f"{name}.top_ps"] = top_ps.numpy()
        out[f"{name}.top_ks"] = top_ks.numpy()
        out[f"{name}.q"] = q.numpy()
        out[f"{name}.ids"] = ids.numpy()
        print(name, ids.tolist())
    # greedy branch (sampler.py:28-36)
    g = torch.Generator().manual_seed(9)
    logits = (torch.randn(7, 1234, generator=g) * 2).to(torch.bfloat16)
    out["greedy.logits_bf16"] = bits(logits)
    out["greedy.ids"] = sampler(logits).numpy()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "sampler_golden.npz"), **out)


if __name__ == "__main__":
    main()
