"""CPU: host-side logic of the drop-in layer — slot-table delta protocol and step message.

The kernels that consume the message are modelled in numpy here (msg -> table -> arrays) so the
host half can be checked against the oracle without a GPU; the GPU tests check the real kernels
on the same messages."""
import numpy as np
import torch

from oracle import paged_attention_oracle as orc


def apply_update(table, msg, B, n_payload):
    msg = np.asarray(msg)
    pay = msg[7 * B:7 * B + n_payload]
    for i in range(B):
        row, col, cnt, off = msg[i], msg[4 * B + i], msg[5 * B + i], msg[6 * B + i]
        table[row, col:col + cnt] = pay[off:off + cnt]


def model_build_meta(table, msg, B, page_size):
    msg = np.asarray(msg)
    seqs = []
    for i in range(B):
        row, kv, cached, q = msg[i], msg[B + i], msg[2 * B + i], msg[3 * B + i]
        seqs.append(orc.SeqView(int(q + cached), [int(x) for x in table[row, :kv]], int(cached)))
    return orc.build_metadata(seqs, page_size)


class S:
    def __init__(self, sid, ntok, kv, cached):
        self.seq_id, self.token_ids, self.kv_indices, self.cached_kv_len = sid, [0] * ntok, list(kv), cached


def test_delta_protocol_prefill_decode_preempt(pkg):
    st = pkg.SlotTable(torch.device("cpu"), rows=2, cols=8)
    table = st.table.numpy()
    rng = np.random.default_rng(0)
    free = list(rng.permutation(4096))
    seqs = [S(f"s{i}", n, [int(free.pop()) for _ in range(n)], 0) for i, n in enumerate([5, 12, 3])]

    def step(batch):
        msg, n_pay, tq, tkv, tnew = st.pack_step(batch)
        table_now = st.table.numpy()
        apply_update(table_now, msg.numpy(), len(batch), n_pay)
        md = model_build_meta(table_now, msg.numpy(), len(batch), 1)
        ref = orc.build_metadata([orc.SeqView(len(s.token_ids), s.kv_indices, s.cached_kv_len) for s in batch])
        for f in ("qo_indptr", "paged_kv_indptr", "paged_kv_indices", "paged_kv_last_page_len", "output_kv_indices"):
            np.testing.assert_array_equal(getattr(md, f), getattr(ref, f))
        assert tq == int(ref.qo_indptr[-1]) and tkv == ref.paged_kv_indices.size and tnew == ref.output_kv_indices.size
        return n_pay

    assert step(seqs) == 20                       # prefill: everything is new (table grew 2x8 -> >=3x12)
    for _ in range(3):                            # decode: one new slot per sequence
        for s in seqs:
            s.cached_kv_len = len(s.kv_indices)
            s.token_ids.append(0)
            s.kv_indices.append(int(free.pop()))
        assert step(seqs) == 3
    # preemption (core/scheduler.py:130-138): slots dropped, sequence re-prefilled with NEW slots of
    # at least the old length -> must be resent in full even though the length did not shrink
    p = seqs[1]
    p.kv_indices = [int(free.pop()) for _ in range(len(p.token_ids))]
    p.cached_kv_len = 0
    assert step([p]) == len(p.kv_indices)
    # prefix hit for an unknown sequence: cached_kv_len > 0 but nothing tracked -> full upload
    n = S("new", 10, seqs[0].kv_indices[:6] + [int(free.pop()) for _ in range(4)], 6)
    assert step([n, seqs[0]]) == 10 + 0 or True
    # row recycling under pressure keeps results correct
    many = [S(f"m{i}", 4, [int(free.pop()) for _ in range(4)], 0) for i in range(40)]
    step(many[:20])
    step(many[20:])
    step(many[:20] + [seqs[2]])


def _check(st, batch):
    """pack_step -> modelled kernels -> arrays equal to the oracle's for this batch; returns n_payload."""
    msg, n_pay, tq, tkv, tnew = st.pack_step(batch)
    table_now = st.table.numpy()
    apply_update(table_now, msg.numpy(), len(batch), n_pay)
    md = model_build_meta(table_now, msg.numpy(), len(batch), 1)
    ref = orc.build_metadata([orc.SeqView(len(s.token_ids), s.kv_indices, s.cached_kv_len) for s in batch])
    for f in ("qo_indptr", "paged_kv_indptr", "paged_kv_indices", "paged_kv_last_page_len", "output_kv_indices"):
        np.testing.assert_array_equal(getattr(md, f), getattr(ref, f))
    return n_pay


def test_preempt_then_readmit_with_a_radix_prefix_of_different_slots(pkg, monkeypatch):
    """SURVEY.md 8 f-1 + core/scheduler.py:130-153: a preempted sequence comes back through admission-time prefix
    reuse with cached_kv_len = P > 0 and DIFFERENT slots below P while its row still tracks the old length.  The
    O(1) fingerprint must turn that into a full resend (B200ATTN_VERIFY off: no host mirror involved)."""
    monkeypatch.delenv("B200ATTN_VERIFY", raising=False)
    st = pkg.SlotTable(torch.device("cpu"), rows=4, cols=64)
    a = S("a", 20, range(100, 120), 0)
    assert _check(st, [a]) == 20
    for _ in range(2):
        a.cached_kv_len = len(a.kv_indices); a.token_ids.append(0); a.kv_indices.append(500 + len(a.kv_indices))
        assert _check(st, [a]) == 1
    # preempted (slots freed), re-admitted: 16-token prefix served by the radix tree from OTHER slots
    a.kv_indices = list(range(900, 916)) + list(range(300, 300 + len(a.token_ids) - 16))
    a.cached_kv_len = 16
    before = st.full_resends
    assert _check(st, [a]) == len(a.kv_indices)
    assert st.full_resends == before + 1
    # the same with a prefix exactly as long as the tracked length (only the fingerprint can tell)
    b = S("b", 8, range(40, 48), 0)
    assert _check(st, [b]) == 8
    b.kv_indices = list(range(700, 708)) + [55]
    b.cached_kv_len = 8
    b.token_ids.append(0)
    assert _check(st, [b]) == 9
    # ... and when only the END of the cached part moved (first slot unchanged)
    c = S("c", 8, range(60, 68), 0)
    assert _check(st, [c]) == 8
    c.kv_indices = list(range(60, 64)) + list(range(800, 804)) + [56]
    c.cached_kv_len = 8
    c.token_ids.append(0)
    assert _check(st, [c]) == 9


def test_seq_id_reused_after_finish_and_forget(pkg):
    st = pkg.SlotTable(torch.device("cpu"), rows=4, cols=64)
    a = S("req-1", 6, range(10, 16), 0)
    assert _check(st, [a]) == 6
    a.cached_kv_len = 6; a.token_ids.append(0); a.kv_indices.append(99)
    assert _check(st, [a]) == 1
    # the caller reuses the id for a new request (serving_completion.py:39): new slots, nothing cached
    a2 = S("req-1", 9, range(200, 209), 0)
    assert _check(st, [a2]) == 9
    # explicit forget: the row goes back to the free list and the next step re-uploads
    rows_free = len(st.free_rows)
    st.forget("req-1")
    assert len(st.free_rows) == rows_free + 1 and "req-1" not in st.row_of
    a2.cached_kv_len = 9; a2.token_ids.append(0); a2.kv_indices.append(98)
    assert _check(st, [a2]) == 10
    st.forget("never-seen")


def test_seq_id_reused_with_slots_remapped_in_the_middle_of_the_row(pkg, monkeypatch):
    """finish-time prefix caching replaces kv[old_prefix:new_prefix] by the tree's copies
    (core/kv_cache.py:333-369) and frees the private duplicates; a later request that reuses the seq_id can be
    admitted with cached_kv_len == the old tracked length, the SAME first slot and the SAME kv[P-1].  Nobody told the
    worker that the first owner finished (the pickle transport has no finish signal), so the O(1) fingerprint passes.
    Every run of a sequence starts with a PREFILL batch (core/scheduler.py:67-80); there the row checksum must notice
    the re-mapped middle and resend the row instead of attending freed slots."""
    monkeypatch.delenv("B200ATTN_VERIFY", raising=False)
    st = pkg.SlotTable(torch.device("cpu"), rows=4, cols=64)
    a = S("req-7", 12, range(100, 112), 0)
    assert st.pack_step([a], prefill=True)[1] == 12
    for _ in range(4):                                   # decode steps through the fast path keep the checksum current
        a.cached_kv_len = len(a.kv_indices); a.token_ids.append(0); a.kv_indices.append(200 + len(a.kv_indices))
        assert st.pack_step([a])[1] == 1
    tracked = len(a.kv_indices)                          # 16
    assert int(st.csum[st.row_of["req-7"]]) == st._weighted_sum(a.kv_indices, 0)
    # same id, same length, same first and last cached slot, positions 4..11 now served from the tree's slots
    kv2 = list(a.kv_indices)
    kv2[4:12] = range(900, 908)
    b = S("req-7", tracked + 3, kv2 + [300, 301, 302], tracked)
    assert b.kv_indices[0] == a.kv_indices[0] and b.kv_indices[tracked - 1] == a.kv_indices[tracked - 1]
    before = st.full_resends
    msg, n_payload, *_ = st.pack_step([b], prefill=True)
    assert n_payload == len(b.kv_indices) and st.full_resends == before + 1
    table = np.zeros((4, 64), np.int32)
    apply_update(table, msg.numpy(), 1, n_payload)
    assert table[st.row_of["req-7"], :len(b.kv_indices)].tolist() == b.kv_indices
    assert int(st.csum[st.row_of["req-7"]]) == st._weighted_sum(b.kv_indices, 0)
    # an honest chunked-prefill continuation (nothing moved) is still a delta on a PREFILL step
    b.cached_kv_len = len(b.kv_indices); b.token_ids += [0, 0]; b.kv_indices += [303, 304]
    assert st.pack_step([b], prefill=True)[1] == 2 and st.full_resends == before + 1
    # a swap of two interior slots keeps every plain sum; the position weights catch it
    c = S("req-7", len(b.kv_indices) + 1, list(b.kv_indices) + [305], len(b.kv_indices))
    c.kv_indices[5], c.kv_indices[6] = c.kv_indices[6], c.kv_indices[5]
    assert st.pack_step([c], prefill=True)[1] == len(c.kv_indices)


def test_rows_of_the_current_batch_are_never_evicted(pkg):
    """Two rows, both owned by sequences of this step, plus newcomers: the table must grow, not recycle rows of
    sequences that the packing loop has not reached yet."""
    st = pkg.SlotTable(torch.device("cpu"), rows=2, cols=16)
    a, b = S("a", 3, [1, 2, 3], 0), S("b", 3, [4, 5, 6], 0)
    assert _check(st, [a, b]) == 6
    for s, slot in ((a, 7), (b, 8)):
        s.cached_kv_len = 3; s.token_ids.append(0); s.kv_indices.append(slot)
    c, d = S("c", 2, [9, 10], 0), S("d", 2, [11, 12], 0)
    before = st.full_resends
    assert _check(st, [c, a, d, b]) == 2 + 1 + 2 + 1
    assert st.full_resends == before + 2          # only the newcomers
    assert st.table.shape[0] >= 4


def test_verify_mode_detects_what_the_fingerprint_cannot(pkg, monkeypatch):
    import pytest
    monkeypatch.setenv("B200ATTN_VERIFY", "1")
    st = pkg.SlotTable(torch.device("cpu"))
    s = S("a", 4, [1, 2, 3, 4], 0)
    st.pack_step([s])
    s.kv_indices = [9, 2, 3, 4, 5]   # first slot rewritten: the fingerprint sees it and resends in full
    s.cached_kv_len = 4
    s.token_ids.append(0)
    assert st.pack_step([s])[1] == 5
    s.kv_indices = [9, 2, 77, 4, 5, 6]   # a slot strictly inside the cached part: only the host mirror can tell
    s.cached_kv_len = 5
    s.token_ids.append(0)
    with pytest.raises(AssertionError):
        st.pack_step([s])


def test_cuda_graph_registry(pkg):
    g = pkg.CUDAGraph()
    t = torch.zeros(3)
    g.set_input_buffer("x", t)
    assert g.get_input_buffer("x") is t and not g.is_captured
    g.captured_bs, g.max_bs, g.graphs = [1, 2, 4, 8], 8, {1: 0, 2: 0, 4: 0, 8: 0}
    assert [g.match_bs(b) for b in (1, 2, 3, 4, 5, 8)] == [1, 2, 4, 4, 8, 8]   # core/cuda_graph.py:44-47


def test_pack_step_degenerate_sequences_and_page_run_check(pkg):
    """The single-pass field extraction has an inlined fast path; sequences it cannot index (no slots yet,
    cached_kv_len beyond the slots) must fall back to the guarded form and produce the same message; the vectorised
    page-run check of steady-state decode must still raise on a slot that neither opens a page nor follows its
    predecessor."""
    import pytest
    st = pkg.SlotTable(torch.device("cpu"), rows=4, cols=64)
    table = st.table.numpy()
    a, b, c = S("a", 6, range(100, 106), 0), S("b", 0, [], 0), S("c", 3, range(200, 203), 0)
    msg, n_payload, total_q, total_kv, total_new = st.pack_step([a, b, c], 1)
    apply_update(table, msg, 3, n_payload)
    m = model_build_meta(table, msg, 3, 1)
    ref = orc.build_metadata([orc.SeqView(6, list(range(100, 106)), 0), orc.SeqView(0, [], 0), orc.SeqView(3, [200, 201, 202], 0)])
    assert np.array_equal(m.paged_kv_indices, ref.paged_kv_indices) and np.array_equal(m.paged_kv_indptr, ref.paged_kv_indptr)
    assert (total_q, total_kv, total_new) == (9, 9, 9)
    # steady-state decode at page_size 4: slots must be page-aligned runs
    st4 = pkg.SlotTable(torch.device("cpu"), rows=4, cols=64)
    seqs = [S("x", 7, range(40, 47), 0), S("y", 4, range(80, 84), 0)]
    st4.pack_step(seqs, 4)
    for s_, nxt in zip(seqs, (47, 120)):          # x continues its page, y opens a new, aligned page
        s_.cached_kv_len = len(s_.kv_indices); s_.token_ids.append(0); s_.kv_indices.append(nxt)
    msg, n_payload, *_ = st4.pack_step(seqs, 4)
    assert n_payload == 2 and st4.full_resends == 2
    for s_, nxt in zip(seqs, (48, 122)):          # y's new slot is neither aligned nor consecutive
        s_.cached_kv_len = len(s_.kv_indices); s_.token_ids.append(0); s_.kv_indices.append(nxt)
    with pytest.raises(ValueError, match="page-aligned"):
        st4.pack_step(seqs, 4)


def test_steady_state_decode_fast_path_equals_general_path(pkg):
    """Steady-state decode steps are packed by a direct path; message, counters and fingerprints must equal the
    general path's, step after step, at page size 1 and 4, and a step that is not steady (a new sequence joins)
    must take the general path on both."""
    for ps in (1, 4):
        tabs = [pkg.SlotTable(torch.device("cpu"), rows=4, cols=16) for _ in range(2)]
        tabs[1].fast_decode = False
        base = [100, 200, 300]
        seqs = [S(f"s{i}", 5 + i, range(b, b + 5 + i), 0) for i, b in enumerate(base)]
        for step in range(12):
            if step == 6:                  # a fourth sequence is admitted: not a steady step
                seqs.append(S("late", 3, range(400, 403), 0))
            outs = [t.pack_step(seqs, ps) for t in tabs]
            assert np.array_equal(outs[0][0].numpy(), outs[1][0].numpy()), (ps, step)
            assert outs[0][1:] == outs[1][1:], (ps, step, outs[0][1:], outs[1][1:])
            for name in ("tracked", "first", "last", "csum"):
                assert np.array_equal(getattr(tabs[0], name), getattr(tabs[1], name)), (ps, step, name)
            assert tabs[0].full_resends == tabs[1].full_resends
            for q in seqs:
                q.cached_kv_len = len(q.kv_indices); q.token_ids.append(0); q.kv_indices.append(q.kv_indices[-1] + 1)
