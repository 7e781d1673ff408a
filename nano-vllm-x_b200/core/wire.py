"""Scheduler -> worker wire format for one engine step.

The reference pickles the whole ``ForwardBatch`` — every sequence with ALL its ``token_ids`` and
``kv_indices`` — into every worker's queue on every step (core/executor.py:86-97 ``submit`` ->
core/worker_client.py:51-52 ``send_request``): O(sum L) bytes and pickling work per step per worker,
about 1.5 MB per decode step at 256 sequences x 1.3 k tokens.  Once the attention metadata is built
on the device from a slot table that itself only receives the NEW slots (layers/attention.py
``SlotTable.pack_step``), this transfer is the remaining O(sum L) term of a step.

Here one step travels as a flat little-endian int32 message holding only what the worker does not
know yet: for a sequence first seen, its id string, sampling parameters, tokens and slots; for a
known one, the tokens and slots appended since the last step (decode: one token, one slot).  The
worker keeps mirror ``Sequence`` objects and hands the attention backend a ``ForwardBatch`` whose
fields equal the scheduler's, so ``AttentionBackend.build_metadata`` and everything behind it are
unchanged.  The same bytes go to every tensor-parallel rank (the reference sends the same batch to
all workers too, core/executor.py:94-95).

Contract (the reference scheduler satisfies it): between two steps a live sequence's ``token_ids``
and ``kv_indices`` only grow by appending, except on preemption (core/scheduler.py:130-138 clears
``kv_indices`` and resets ``cached_kv_len``) and prefix re-mapping at admission.  The scheduler calls
``BatchEncoder.forget(seq_id)`` at those points and when a sequence finishes; as a safety net the
encoder also resends in full whenever a list shrank, ``cached_kv_len`` went backwards, or the O(1)
fingerprint (first slot, last slot sent) changed.  ``verify=True`` adds an O(L) checksum of both
lists per sequence (tests, debugging).

Message layout (int32 words):
  [0] magic 0x42325731 ("B2W1")  [1] mode (0 PREFILL, 1 DECODE)  [2] B  [3] n_drop  [4] flags (1 = checksums)
  n_drop words   : handles of sequences the worker may forget
  B records of 8 : handle, kind (0 delta, 1 full), n_tok, n_kv, cached_kv_len, total tokens, total kv, blob words
  payload        : per record, in order: blob (full only: pickled (seq_id, sampling_params), word padded),
                   n_tok token ids, n_kv slot ids, [2 checksum words]
"""
from __future__ import annotations

import pickle

import numpy as np

from .common import ForwardBatch, ForwardMode, Sequence, mode_name

MAGIC = 0x42325731
_REC = 8


def _checksum(xs) -> int:
    a = np.asarray(xs, dtype=np.int64)
    return int((a * (np.arange(a.size, dtype=np.int64) % 65521 + 1)).sum() & 0x7FFFFFFF) if a.size else 0


class BatchEncoder:
    """Scheduler side.  One encoder per worker group (all ranks receive the same bytes)."""

    def __init__(self, verify: bool = False):
        self.verify = verify
        self._known: dict = {}          # seq_id -> [handle, tokens sent, kv sent, cached_kv_len sent, first slot, last slot sent]
        self._next = 0
        self._free: list[int] = []
        self._drop: list[int] = []

    def forget(self, seq_id) -> None:
        """Call when a sequence finishes (core/scheduler.py:183-190) or is preempted (:130-138): its handle is
        recycled and the worker drops its mirror with the next message; a preempted sequence is resent in full
        when it is scheduled again."""
        st = self._known.pop(seq_id, None)
        if st is not None:
            self._drop.append(st[0])

    def encode(self, batch) -> bytes:
        seqs = batch.seqs
        B = len(seqs)
        head = [MAGIC, 0 if mode_name(batch.forward_mode) == "PREFILL" else 1, B, len(self._drop), int(self.verify)]
        head.extend(self._drop)
        self._free.extend(self._drop)
        self._drop = []
        recs: list[int] = []
        payload: list[int] = []          # plain ints; ONE array conversion per message
        known, verify = self._known, self.verify
        for s in seqs:
            toks, kv, cached = s.token_ids, s.kv_indices, s.cached_kv_len
            nt, nk = len(toks), len(kv)
            st = known.get(s.seq_id)
            if (st is None or nt < st[1] or nk < st[2] or cached < st[3]
                    or (st[2] > 0 and (kv[0] != st[4] or kv[st[2] - 1] != st[5]))):
                if st is None:
                    handle = self._free.pop() if self._free else self._next
                    if handle == self._next:
                        self._next += 1
                    st = known[s.seq_id] = [handle, 0, 0, 0, -1, -1]
                blob = pickle.dumps((s.seq_id, getattr(s, "sampling_params", None)), protocol=pickle.HIGHEST_PROTOCOL)
                blob += b"\0" * (-len(blob) % 4)
                words = np.frombuffer(blob, dtype="<i4").tolist()
                recs += (st[0], 1, nt, nk, cached, nt, nk, len(words))
                payload += words
                payload += toks
                payload += kv
            else:
                dt, dk = nt - st[1], nk - st[2]
                recs += (st[0], 0, dt, dk, cached, nt, nk, 0)
                if dt == 1:
                    payload.append(toks[-1])
                elif dt:
                    payload += toks[st[1]:]
                if dk == 1:
                    payload.append(kv[-1])
                elif dk:
                    payload += kv[st[2]:]
            if verify:
                payload.append(_checksum(toks))
                payload.append(_checksum(kv))
            st[1], st[2], st[3] = nt, nk, cached
            if nk:
                st[4], st[5] = kv[0], kv[-1]
        return np.array(head + recs + payload, dtype="<i4").tobytes()


class BatchDecoder:
    """Worker side: applies a message to the mirror sequences and returns the step's ForwardBatch."""

    def __init__(self, sequence_cls=Sequence, batch_cls=ForwardBatch, mode_cls=ForwardMode, on_drop=None):
        """``on_drop(seq_id)`` is called whenever the worker must stop trusting what it holds for a sequence: its
        handle was dropped by the scheduler (finish / preempt) or the scheduler resent it in full.  The binding uses
        it to call ``SlotTable.forget`` so that a reused seq_id can never meet a stale device row."""
        self._seqs: dict[int, object] = {}
        self._sequence_cls, self._batch_cls, self._mode_cls = sequence_cls, batch_cls, mode_cls
        self._on_drop = on_drop

    def decode(self, data) -> ForwardBatch:
        m = np.frombuffer(data, dtype="<i4")
        if m.size < 5 or int(m[0]) != MAGIC:
            raise ValueError("not a b200 step message (bad magic)")
        mode = self._mode_cls.PREFILL if int(m[1]) == 0 else self._mode_cls.DECODE
        B, n_drop, verify = int(m[2]), int(m[3]), bool(m[4] & 1)
        for h in m[5:5 + n_drop].tolist():
            gone = self._seqs.pop(h, None)
            if gone is not None and self._on_drop is not None:
                self._on_drop(gone.seq_id)
        pos = 5 + n_drop
        w = m.tolist()                     # one conversion; the loop below touches plain ints only
        rec_end = pos + _REC * B
        if rec_end > len(w):
            raise ValueError("truncated step message")
        recs = [w[i:i + _REC] for i in range(pos, rec_end, _REC)]
        pos = rec_end
        seqs = []
        for handle, kind, n_tok, n_kv, cached, tot_tok, tot_kv, blob_words in recs:
            if kind == 1:
                seq_id, sp = pickle.loads(m[pos:pos + blob_words].tobytes())
                pos += blob_words
                if self._on_drop is not None:
                    self._on_drop(seq_id)        # a full record restarts the sequence: drop whatever row it had
                s = self._sequence_cls(seq_id=seq_id)
                if sp is not None:
                    s.sampling_params = sp
                s.token_ids = w[pos:pos + n_tok]
                pos += n_tok
                s.kv_indices = w[pos:pos + n_kv]
                pos += n_kv
                self._seqs[handle] = s
            else:
                s = self._seqs.get(handle)
                if s is None:
                    raise ValueError(f"delta for unknown sequence handle {handle} (a message was lost or reordered)")
                if n_tok == 1:
                    s.token_ids.append(w[pos])
                elif n_tok:
                    s.token_ids.extend(w[pos:pos + n_tok])
                pos += n_tok
                if n_kv == 1:
                    s.kv_indices.append(w[pos])
                elif n_kv:
                    s.kv_indices.extend(w[pos:pos + n_kv])
                pos += n_kv
            s.cached_kv_len = cached
            if len(s.token_ids) != tot_tok or len(s.kv_indices) != tot_kv:
                raise ValueError(f"sequence {s.seq_id!r} out of sync: have {len(s.token_ids)} tokens / "
                                 f"{len(s.kv_indices)} slots, scheduler has {tot_tok} / {tot_kv}")
            if verify:
                if _checksum(s.token_ids) != w[pos] or _checksum(s.kv_indices) != w[pos + 1]:
                    raise ValueError(f"sequence {s.seq_id!r} checksum mismatch (a list changed other than by appending)")
                pos += 2
            seqs.append(s)
        if pos != len(w):
            raise ValueError("trailing words in step message")
        return self._batch_cls(forward_mode=mode, num_seqs=B, seqs=seqs)
