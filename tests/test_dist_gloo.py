"""CPU, world_size 2, gloo: the multi-GPU partitioning of the path.  TP shards attention by
KV head with NO data-path collective: rank r owns Hkv/tp KV heads + their query heads and a pool holding
only those heads, every rank sees the same page table.  Each rank runs the oracle on its shard; gathering
the shards (test-only collective) must reproduce the unsharded result exactly, and the head partition must
match the reference's model_runner.py:43-44 arithmetic."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import paged_attention_oracle as orc
from tests.util import make_paged_case


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)          # fixed reduction order: two ranks share the host cores, and the exact-equality
                                      # check below must not depend on how the BLAS splits its work that moment
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import nanovllmx_b200 as pkg
    Hq, Hkv = 8, 4
    case = make_paged_case([37, 5, 130], Hq, Hkv, D=64, q_lens=[37, 5, 2], seed=21)   # identical on every rank
    hq_r, hkv_r = pkg.shard_heads(Hq, Hkv, world, rank)
    assert len(hq_r) == Hq // world and len(hkv_r) == max(1, Hkv // world)
    out = orc.paged_attention(case["q"][:, hq_r], case["k_cache"][:, hkv_r], case["v_cache"][:, hkv_r], case["qo_indptr"],
                              case["kv_indptr"], case["kv_indices"], case["last_page_len"], causal=True, return_fp32=True)
    parts = [torch.empty_like(out) for _ in range(world)]
    dist.all_gather(parts, out.contiguous())
    if rank == 0:
        full = orc.paged_attention(case["q"], case["k_cache"], case["v_cache"], case["qo_indptr"], case["kv_indptr"],
                                   case["kv_indices"], case["last_page_len"], causal=True, return_fp32=True)
        got = torch.cat(parts, dim=1)
        q.put((bool(torch.equal(got, full)), float((got - full).abs().max())))
    dist.barrier()
    dist.destroy_process_group()


def test_tp2_head_sharding_needs_no_collective():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    ok, err = q.get(timeout=120)
    [p.join(timeout=120) for p in procs]
    assert ok, f"sharded != unsharded, max abs diff {err}"
    assert [p.exitcode for p in procs] == [0, 0]


def test_shard_heads_matches_reference_arithmetic():
    import nanovllmx_b200 as pkg
    for Hq, Hkv in [(32, 8), (16, 8)]:
        for tp in (1, 2, 4, 8):
            seen_q = []
            for r in range(tp):
                hq, hkv = pkg.shard_heads(Hq, Hkv, tp, r)
                assert len(hq) == Hq // tp and len(hkv) == max(1, Hkv // tp)      # model_runner.py:43-44
                g = Hq // Hkv
                assert all(h // g in hkv for h in hq)                              # q heads stay with their kv head
                seen_q += list(hq)
            assert sorted(seen_q) == list(range(Hq))
    hq, hkv = pkg.shard_heads(16, 2, 4, 3)     # tp > Hkv: kv heads replicated (layers/linear.py:333-338)
    assert list(hkv) == [1] and list(hq) == [12, 13, 14, 15]


def _wire_worker(rank, world, port, q):
    """Rank 0 plays the scheduler: it encodes every step and broadcasts the bytes; every rank (the worker processes of
    core/worker_client.py) decodes them and must end up with the batch rank 0 scheduled."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from importlib import import_module
    import nanovllmx_b200 as pkg
    from tests.test_wire import same_batch, scheduler_run
    wire = import_module("nano-vllm-x_b200.core.wire")
    enc, dec = wire.BatchEncoder(), wire.BatchDecoder()
    ok, steps, prev_done = True, 0, []
    for batch, done in scheduler_run(pkg, False, n=12, seed=3):     # every rank replays the same seeded schedule
        n = torch.zeros(1, dtype=torch.int64)
        if rank == 0:
            for sid in prev_done:
                enc.forget(sid)
            msg = enc.encode(batch)
            n[0] = len(msg)
        prev_done = done
        dist.broadcast(n, 0)
        buf = torch.frombuffer(bytearray(msg), dtype=torch.uint8) if rank == 0 else torch.empty(int(n), dtype=torch.uint8)
        dist.broadcast(buf, 0)
        got = dec.decode(buf.numpy().tobytes())
        try:
            same_batch(got, batch)
        except AssertionError:
            ok = False
        steps += 1
    flags = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(flags, torch.tensor([int(ok)]))
    if rank == 0:
        q.put((all(int(f) for f in flags), steps))
    dist.barrier()
    dist.destroy_process_group()


def test_step_deltas_broadcast_to_two_workers():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_wire_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    ok, steps = q.get(timeout=180)
    [p.join(timeout=120) for p in procs]
    assert ok and steps > 20
    assert [p.exitcode for p in procs] == [0, 0]
