"""CPU: the per-step delta wire format reproduces, on the worker side, exactly the ForwardBatch
the reference would have pickled (core/executor.py:86-97), over a scheduler-shaped run: one prefill batch, decode
steps, sequences finishing, a preemption with re-admission, and a handle being recycled."""
import pickle
import random

import pytest


def same_batch(a, b):
    assert a.forward_mode.name == b.forward_mode.name and a.num_seqs == b.num_seqs == len(b.seqs)
    for x, y in zip(a.seqs, b.seqs):
        assert x.seq_id == y.seq_id and x.cached_kv_len == y.cached_kv_len
        assert list(x.token_ids) == list(y.token_ids) and list(x.kv_indices) == list(y.kv_indices)
        assert getattr(x, "sampling_params", None) == getattr(y, "sampling_params", None)


def scheduler_run(pkg, verify, n=24, seed=0):
    """Yields (batch, finished seq ids after the step), shaped like core/scheduler.py:60-117."""
    rng = random.Random(seed)
    free = list(range(100000 - 1, -1, -1))
    waiting = []
    for i in range(n):
        s = pkg.Sequence(seq_id=f"req-{i}", token_ids=[rng.randint(0, 10000) for _ in range(rng.randint(3, 200))])
        s.sampling_params = {"temperature": 0.6, "max_tokens": len(s.token_ids) + rng.randint(1, 40)}
        waiting.append(s)
    running, step = [], 0
    while waiting or running:
        step += 1
        if waiting:
            take, waiting = waiting[:8], waiting[8:]
            for s in take:
                s.kv_indices.extend(free.pop() for _ in range(len(s.token_ids) - len(s.kv_indices)))
            batch = pkg.ForwardBatch(pkg.ForwardMode.PREFILL, len(take), take)
            running = running + take
        else:
            if step % 7 == 0 and len(running) > 2:           # preemption: core/scheduler.py:130-138
                v = running.pop()
                free.extend(v.kv_indices)
                v.kv_indices = []
                v.cached_kv_len = 0
                waiting.insert(0, v)
            for s in running:
                s.kv_indices.append(free.pop())
            batch = pkg.ForwardBatch(pkg.ForwardMode.DECODE, len(running), list(running))
        done = []
        yield batch, done
        for s in batch.seqs:
            s.cached_kv_len = len(s.kv_indices)
            s.token_ids.append(rng.randint(0, 10000))
            if len(s.token_ids) >= s.sampling_params["max_tokens"]:
                done.append(s.seq_id)
                free.extend(s.kv_indices)
        running = [s for s in running if s.seq_id not in done]


@pytest.mark.parametrize("verify", [False, True])
def test_delta_stream_reproduces_every_batch(pkg, verify):
    from importlib import import_module
    wire = import_module("nano-vllm-x_b200.core.wire")
    enc, dec = wire.BatchEncoder(verify=verify), wire.BatchDecoder()
    delta_bytes = pickle_bytes = steps = 0
    kinds = set()
    prev_done = []
    for batch, done in scheduler_run(pkg, verify):
        for sid in prev_done:                 # the generator fills `done` when it is resumed
            enc.forget(sid)
        prev_done = done
        msg = enc.encode(batch)
        got = dec.decode(msg)
        same_batch(got, batch)
        kinds.add(batch.forward_mode.name)
        delta_bytes += len(msg)
        pickle_bytes += len(pickle.dumps(batch))
        steps += 1
    assert kinds == {"PREFILL", "DECODE"} and steps > 30
    assert delta_bytes * 4 < pickle_bytes, (delta_bytes, pickle_bytes)
    assert len(dec._seqs) <= 24


def test_forget_recycles_handles_and_drops_mirrors(pkg):
    from importlib import import_module
    wire = import_module("nano-vllm-x_b200.core.wire")
    enc, dec = wire.BatchEncoder(), wire.BatchDecoder()
    a = pkg.Sequence("a", [1, 2, 3], [7, 8, 9], 0)
    dec.decode(enc.encode(pkg.ForwardBatch(pkg.ForwardMode.PREFILL, 1, [a])))
    enc.forget("a")
    b = pkg.Sequence("b", [4], [5], 0)
    got = dec.decode(enc.encode(pkg.ForwardBatch(pkg.ForwardMode.PREFILL, 1, [b])))
    assert got.seqs[0].seq_id == "b" and len(dec._seqs) == 1 and enc._next == 1


def test_decoder_rejects_lost_or_corrupt_messages(pkg):
    from importlib import import_module
    wire = import_module("nano-vllm-x_b200.core.wire")
    enc, dec = wire.BatchEncoder(verify=True), wire.BatchDecoder()
    a = pkg.Sequence("a", [1, 2, 3], [7, 8, 9], 0)
    m1 = enc.encode(pkg.ForwardBatch(pkg.ForwardMode.PREFILL, 1, [a]))
    a.cached_kv_len = 3; a.token_ids.append(4); a.kv_indices.append(10)
    m2 = enc.encode(pkg.ForwardBatch(pkg.ForwardMode.DECODE, 1, [a]))
    with pytest.raises(ValueError):
        wire.BatchDecoder().decode(m2)               # delta without the full record
    with pytest.raises(ValueError):
        dec.decode(b"\0" * 32)                       # bad magic
    dec.decode(m1)
    a.cached_kv_len = 4; a.token_ids.append(5); a.kv_indices.append(11)
    m3 = enc.encode(pkg.ForwardBatch(pkg.ForwardMode.DECODE, 1, [a]))
    with pytest.raises(ValueError):
        dec.decode(m3)                               # m2 was skipped: totals disagree
    # a list edited in the middle (neither appended nor caught by the first/last-slot fingerprint): verify checksum
    enc2, dec2 = wire.BatchEncoder(verify=True), wire.BatchDecoder()
    b = pkg.Sequence("b", [1, 2, 3], [7, 8, 9], 0)
    dec2.decode(enc2.encode(pkg.ForwardBatch(pkg.ForwardMode.PREFILL, 1, [b])))
    b.kv_indices[1] = 99; b.token_ids.append(4); b.kv_indices.append(10); b.cached_kv_len = 3
    with pytest.raises(ValueError):
        dec2.decode(enc2.encode(pkg.ForwardBatch(pkg.ForwardMode.DECODE, 1, [b])))


def test_dropped_and_resent_sequences_make_the_worker_forget_its_slot_row(pkg):
    """the worker's SlotTable never heard that a sequence finished.  With the delta wire it does:
    the decoder reports every dropped handle and every full resend, and the binding (core/wire_binding.py) forwards
    that to SlotTable.forget, so a reused seq_id whose middle slots were re-mapped meets no stale row."""
    import torch
    from importlib import import_module
    wire = import_module("nano-vllm-x_b200.core.wire")
    table = pkg.SlotTable(torch.device("cpu"), rows=4, cols=64)
    dropped = []

    def on_drop(seq_id):
        dropped.append(seq_id)
        table.forget(seq_id)

    enc, dec = wire.BatchEncoder(), wire.BatchDecoder(on_drop=on_drop)
    a = pkg.Sequence("req-1", list(range(12)), list(range(100, 112)), 0)
    got = dec.decode(enc.encode(pkg.ForwardBatch(pkg.ForwardMode.PREFILL, 1, [a])))
    assert dropped == ["req-1"]                          # first sight is a full record: nothing to forget yet, harmless
    assert table.pack_step(got.seqs, prefill=True)[1] == 12
    a.cached_kv_len = 12; a.token_ids.append(0); a.kv_indices.append(300)
    got = dec.decode(enc.encode(pkg.ForwardBatch(pkg.ForwardMode.DECODE, 1, [a])))
    assert dropped == ["req-1"] and table.pack_step(got.seqs)[1] == 1        # a delta step reports nothing
    # the scheduler finishes it (wire_binding's finish_sequence hook), then the id comes back with the same length,
    # the same first / last cached slot and tree-owned slots in the middle
    enc.forget("req-1")
    kv = list(a.kv_indices)
    kv[3:9] = range(900, 906)
    b = pkg.Sequence("req-1", list(range(14)), kv + [301], 13)
    before = table.full_resends
    got = dec.decode(enc.encode(pkg.ForwardBatch(pkg.ForwardMode.PREFILL, 1, [b])))
    assert dropped == ["req-1", "req-1", "req-1"]        # the dropped handle, then the full record
    assert "req-1" not in table.row_of
    # the row is gone, so even the O(1) path (prefill=False) has to upload everything
    assert table.pack_step(got.seqs)[1] == 14 and table.full_resends == before + 1
