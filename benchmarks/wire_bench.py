#!/usr/bin/env python
"""Host-side cost of getting one decode step to the workers: the reference's pickled ForwardBatch
(core/executor.py:86-97 -> mp.Queue pickling, core/worker_client.py:51-52) against core/wire.py's delta message, on
the bench.py workload (256 sequences, prompts 100-1024) at a decode step.  CPU only; prints one JSON line."""
import json
import os
import pickle
import sys
import time
from importlib import import_module

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import random

import nanovllmx_b200 as pkg

wire = import_module("nano-vllm-x_b200.core.wire")


def main():
    random.seed(0)
    seqs, slot = [], 0
    for i in range(256):
        n = random.randint(100, 1024) + 200          # mid-run: prompt + ~200 generated tokens
        s = pkg.Sequence(f"req-{i:032x}", [random.randint(0, 10000) for _ in range(n)], list(range(slot, slot + n)), n - 1)
        s.sampling_params = {"temperature": 0.6, "max_tokens": 2048, "ignore_eos": True}
        slot += n
        seqs.append(s)
    enc, dec = wire.BatchEncoder(), wire.BatchDecoder()
    dec.decode(enc.encode(pkg.ForwardBatch(pkg.ForwardMode.DECODE, 256, seqs)))      # workers know the sequences
    steps = 50
    t_pick = t_delta = 0.0
    b_pick = b_delta = 0
    for _ in range(steps):
        for s in seqs:
            s.cached_kv_len = len(s.kv_indices); s.token_ids.append(1); s.kv_indices.append(slot); slot += 1
        batch = pkg.ForwardBatch(pkg.ForwardMode.DECODE, 256, seqs)
        t0 = time.perf_counter()
        blob = pickle.dumps(batch)
        pickle.loads(blob)
        t1 = time.perf_counter()
        msg = enc.encode(batch)
        got = dec.decode(msg)
        t2 = time.perf_counter()
        t_pick += t1 - t0; t_delta += t2 - t1; b_pick += len(blob); b_delta += len(msg)
    assert got.seqs[7].kv_indices == seqs[7].kv_indices
    print(json.dumps(dict(case="step wire format, 256 seqs decode", tokens_per_seq_avg=sum(len(s.token_ids) for s in seqs) / 256,
                          pickle_bytes_per_step=b_pick // steps, delta_bytes_per_step=b_delta // steps,
                          pickle_ms_per_step=1e3 * t_pick / steps, delta_ms_per_step=1e3 * t_delta / steps,
                          note="encode + decode once; the reference pickles once per worker (tp_size times)")))


if __name__ == "__main__":
    main()
