"""Drop-in replacement for nanovllmx/layers/attention.py (reference :17-373).

Same public surface — ``AttentionMetadata``, ``AttentionBackend`` (``build_metadata``,
``prepare_for_cuda_graph_capture``, ``build_metadata_for_cuda_graph_capture``,
``build_metadata_for_cuda_graph_replay``), ``attention_kv_cache``, ``Attention`` — and the same
named CUDA-graph buffers (``paged_kv_indptr_buffer``, ``paged_kv_indices_buffer``,
``paged_kv_last_page_len_buffer``, ``output_kv_indices_buffer``; reference :158-168) with
bit-identical contents.  What changes underneath:

* no FlashInfer, no Triton, no host ``plan()``: the ``wrapper`` field of the metadata holds an
  ``AttentionPlan`` (device index arrays + workspace) consumed by the b200attn CUDA kernels;
* the index arrays are built ON THE DEVICE from a device-resident slot table.  Per step the host
  sends one packed int32 message (rows, lengths and only the NEW slot ids) instead of the
  reference's 2*B list->tensor H2D copies, B ``.item()`` syncs (reference :92-101) and O(sum L)
  Python list traversal;
* nothing in this file synchronises with the device.
"""
from __future__ import annotations

import os
from contextlib import contextmanager
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch
import torch.nn as nn

from .. import ops
from ..core.common import mode_name
from ..core.kv_cache import KVCachePool


# =======================================================================================================
# plan object stored in AttentionMetadata.wrapper
# =======================================================================================================
class AttentionPlan:
    """What FlashInfer's wrapper object is to the reference (:49-59): the per-step state the
    attention kernels read.  All arrays live on the device; ``total_q`` / ``num_seqs`` are host
    ints known without a sync."""

    def __init__(self, mode: str, num_seqs: int, total_q: int, page_size: int,
                 qo_indptr: Optional[torch.Tensor], paged_kv_indptr: torch.Tensor,
                 paged_kv_indices: torch.Tensor, paged_kv_last_page_len: torch.Tensor,
                 workspace: torch.Tensor, num_qo_heads: int, num_kv_heads: int, head_dim: int):
        self.mode = mode
        self.num_seqs = num_seqs
        self.total_q = total_q
        self.page_size = page_size
        self.qo_indptr = qo_indptr
        self.paged_kv_indptr = paged_kv_indptr
        self.paged_kv_indices = paged_kv_indices
        self.paged_kv_last_page_len = paged_kv_last_page_len
        self.workspace = workspace
        self.num_qo_heads = num_qo_heads
        self.num_kv_heads = num_kv_heads
        self.head_dim = head_dim
        self._sm_scale = head_dim ** -0.5   # overwritten by Attention.forward, as the reference does (:361,:366)

    def run(self, q: torch.Tensor, paged_kv_cache) -> torch.Tensor:
        """q: [T, Hq, D]; paged_kv_cache = (k_cache, v_cache), each [num_tokens, 1, Hkv, D] or
        [num_tokens, Hkv, D].  Returns [T, Hq, D] (same call shape as FlashInfer's ``run``)."""
        k_cache, v_cache = paged_kv_cache
        k_cache = k_cache.view(-1, self.num_kv_heads, self.head_dim)
        v_cache = v_cache.view(-1, self.num_kv_heads, self.head_dim)
        if self.mode == "DECODE":
            return ops.call("decode", q, k_cache, v_cache, self.paged_kv_indptr, self.paged_kv_indices,
                              self.paged_kv_last_page_len, self.workspace, self.page_size, self._sm_scale)
        return ops.call("prefill", q, k_cache, v_cache, self.qo_indptr, self.paged_kv_indptr, self.paged_kv_indices,
                           self.paged_kv_last_page_len, self.workspace, self.page_size, self._sm_scale, True)


@dataclass
class AttentionMetadata:
    forward_mode: object            # ForwardMode (reference's or ours)
    kv_cache: KVCachePool
    output_kv_indices: torch.Tensor
    wrapper: AttentionPlan


# =======================================================================================================
# device-resident slot table
# =======================================================================================================
class SlotTable:
    """``table[row, pos]`` = KV slot id of token ``pos`` of the live sequence in ``row``.

    Workers receive freshly pickled ``Sequence`` objects every step (reference
    core/executor.py:94-95), so rows are keyed by ``seq_id``.  Delta contract: a row is trusted only
    when the sequence continues exactly where this table left it, which is checked in O(1) per
    sequence from a fingerprint kept on the host:

        tracked == seq.cached_kv_len > 0  and  kv[0] == first slot sent  and  kv[tracked-1] == last slot sent

    Then only ``kv[tracked:]`` (one slot per decode step) is uploaded.  Anything else is resent in
    full: an unknown ``seq_id``; a preempted sequence (back with ``cached_kv_len == 0``,
    core/scheduler.py:130-138); a sequence re-admitted with a radix-cache prefix whose slots differ
    from the ones this row holds (``cached_kv_len = P`` with new prefix slots); a
    ``seq_id`` reused after its first owner finished (request ids are caller-supplied,
    entrypoints/openai/serving_completion.py:39).  ``forget(seq_id)`` drops a row explicitly.

    The three-slot fingerprint cannot see slots that changed in the MIDDLE of a row (finish-time prefix
    caching replaces ``kv[old_prefix:new_prefix]`` by the tree's copies, core/kv_cache.py:333-369; a later
    request with the same id can then come back with the same length, first slot and last slot).  The
    reference only re-maps slots between a sequence's runs, and every run starts with a PREFILL batch
    (core/scheduler.py:67-80 admits from ``waiting`` only), so PREFILL steps additionally compare a
    position-weighted checksum of ``kv[:tracked]`` with the one accumulated while the row was uploaded
    (O(L) per trusted sequence, on the step that is O(L^2) on the device anyway); DECODE steps keep the O(1) check.
    ``B200ATTN_VERIFY=1`` additionally keeps a full host mirror and checks every step (debug aid;
    O(sum L) like the reference).
    """

    RING = 4            # pinned staging buffers in flight

    def __init__(self, device: torch.device, rows: int = 64, cols: int = 1024):
        self.device = device
        self.table = torch.zeros((rows, cols), dtype=torch.int32, device=device)
        self.row_of: dict = {}
        self.tracked = np.zeros(rows, dtype=np.int64)
        self.first = np.full(rows, -1, dtype=np.int64)
        self.last = np.full(rows, -1, dtype=np.int64)
        self.csum = np.zeros(rows, dtype=np.int64)      # sum(slot[p] * (p % 65521 + 1)) over the tracked prefix, mod 2^64
        self.last_used = np.zeros(rows, dtype=np.int64)
        self.owner: list = [None] * rows
        self.free_rows = list(range(rows - 1, -1, -1))
        self.step = 0
        self.verify = os.environ.get("B200ATTN_VERIFY", "0") == "1"
        self._mirror: dict = {}
        self._pin: list = [None] * self.RING
        self._dev: list = [None] * self.RING
        self._evt: list = [None] * self.RING
        self.last_msg_bytes = 0
        self.full_resends = 0           # sequences uploaded from position 0 (statistics / tests)
        self.fast_decode = True         # direct packing of steady-state decode steps (tests switch it off to compare)
        self._iota = None

    # -- capacity ----------------------------------------------------------------------------------
    def _grow(self, rows: int, cols: int) -> None:
        old = self.table
        rows = max(rows, old.shape[0])
        cols = max(cols, old.shape[1])
        if rows == old.shape[0] and cols == old.shape[1]:
            return
        new = torch.zeros((rows, cols), dtype=torch.int32, device=self.device)
        new[:old.shape[0], :old.shape[1]].copy_(old)
        self.table = new
        extra = rows - old.shape[0]
        if extra > 0:
            self.free_rows = list(range(rows - 1, old.shape[0] - 1, -1)) + self.free_rows
            self.tracked = np.concatenate([self.tracked, np.zeros(extra, np.int64)])
            self.first = np.concatenate([self.first, np.full(extra, -1, np.int64)])
            self.last = np.concatenate([self.last, np.full(extra, -1, np.int64)])
            self.csum = np.concatenate([self.csum, np.zeros(extra, np.int64)])
            self.last_used = np.concatenate([self.last_used, np.zeros(extra, np.int64)])
            self.owner += [None] * extra

    def reserve(self, rows: int, cols: int) -> None:
        self._grow(rows, cols)

    def forget(self, seq_id) -> None:
        """Drop the row of ``seq_id`` (finish / abort / preempt hooks may call this; it is never required
        for correctness: a stale row fails the fingerprint check and is resent)."""
        row = self.row_of.pop(seq_id, None)
        if row is not None:
            self._mirror.pop(seq_id, None)
            self.owner[row] = None
            self.tracked[row] = 0
            self.first[row] = self.last[row] = -1
            self.csum[row] = 0
            self.free_rows.append(row)

    def _take_rows(self, need: int) -> None:
        """Make ``need`` rows available.  Rows of sequences in the current step carry ``last_used == step``
        (set before any row is taken) and are never evicted."""
        if len(self.free_rows) >= need:
            return
        stale = [r for r in range(len(self.owner)) if self.owner[r] is not None and self.last_used[r] != self.step]
        if len(stale) + len(self.free_rows) < need:
            self._grow(max(2 * self.table.shape[0], self.table.shape[0] + need), self.table.shape[1])
            return
        stale.sort(key=lambda r: self.last_used[r])
        for r in stale[:max(need - len(self.free_rows), len(stale) // 2)]:
            self.forget(self.owner[r])

    # -- staging ------------------------------------------------------------------------------------
    def _upload(self, msg: np.ndarray) -> torch.Tensor:
        """One H2D copy of the packed message through a ring of pinned buffers (no allocation per step)."""
        n = msg.size
        if self.device.type != "cuda":
            return torch.from_numpy(msg)
        i = self.step % self.RING
        if self._pin[i] is None or self._pin[i].numel() < n:
            cap = max(4096, 1 << (n - 1).bit_length())
            self._pin[i] = torch.empty(cap, dtype=torch.int32).pin_memory()
            self._dev[i] = torch.empty(cap, dtype=torch.int32, device=self.device)
            self._evt[i] = torch.cuda.Event()
        else:
            self._evt[i].synchronize()          # the copy that last used this buffer (RING steps ago) is long done
        self._pin[i].numpy()[:n] = msg
        dev = self._dev[i][:n]
        dev.copy_(self._pin[i][:n], non_blocking=True)
        self._evt[i].record()
        return dev

    @staticmethod
    def _check_pages(seqs, payload, start, cnt, woff, ps: int) -> None:
        for i in np.flatnonzero(cnt):
            a, c, st = int(woff[i]), int(cnt[i]), int(start[i])
            arr = payload[a:a + c].astype(np.int64)
            pos = st + np.arange(c)
            head = pos % ps == 0
            prev = np.empty(c, dtype=np.int64)
            prev[1:] = arr[:-1]
            prev[0] = seqs[i].kv_indices[st - 1] if st > 0 else -1
            if not (np.all(arr[head] % ps == 0) and np.all(arr[~head] == prev[~head] + 1)):
                raise ValueError(
                    f"AttentionBackend(page_size={ps}): the KV slots of sequence {seqs[i].seq_id!r} are not "
                    f"page-aligned runs of {ps} consecutive slots; the reference's token-granular KVCacheAllocator "
                    "(core/kv_cache.py:101-117) needs page_size=1")

    @staticmethod
    def _weighted_sum(slots, start: int) -> int:
        """sum(slots[j] * ((start + j) % 65521 + 1)) as a wrapped int64 (the row checksum's increment)."""
        a = np.asarray(slots, dtype=np.int64)
        if a.size == 0:
            return 0
        with np.errstate(over="ignore"):
            return int((a * ((np.arange(start, start + a.size, dtype=np.int64) % 65521) + 1)).sum())

    # -- per step ----------------------------------------------------------------------------------
    def pack_step(self, seqs, page_size: int = 1, prefill: bool = False) -> tuple[torch.Tensor, int, int, int, int]:
        """Returns (device message, n_payload, total_q, total_kv, total_new).

        ``prefill=True`` (a PREFILL batch: the only place a sequence can come back with re-mapped slots) makes every
        trusted row prove itself with the O(L) checksum described in the class docstring; a mismatch is a full resend.

        ``page_size > 1`` (an extension: the reference is token granular, layers/attention.py:66) needs every
        page of a sequence to be ``page_size`` consecutive, page-aligned slots; the newly uploaded slots are
        checked here (O(new) per step) and a violation raises instead of attending the wrong rows.

        Message layout (int32): rows[B] kv_len[B] cached[B] q_len[B] write_col[B] write_cnt[B]
        write_off[B] payload[n_payload] (include/b200attn.h, b200_slot_table_update)."""
        self.step += 1
        step = self.step
        B = len(seqs)
        row_of = self.row_of
        # ONE pass over the Python objects: row, lengths and the three slots the O(1) checks need
        # (first, last, and kv[cached_kv_len - 1]: the slot the fingerprint compares, since a trusted row has
        # tracked == cached_kv_len)
        def fields(q):
            kv, c = q.kv_indices, q.cached_kv_len
            n = len(kv)
            return (row_of.get(q.seq_id, -1), n, c, len(q.token_ids), kv[0] if n else -1, kv[-1] if n else -1,
                    kv[c - 1] if 0 < c <= n else -1)
        get = row_of.get
        try:      # common case inlined (no call per sequence): every sequence owns slots and 0 < cached_kv_len <= len(kv)
            arr = np.array([(get(q.seq_id, -1), len(q.kv_indices), q.cached_kv_len, len(q.token_ids), q.kv_indices[0],
                             q.kv_indices[-1], q.kv_indices[q.cached_kv_len - 1] if q.cached_kv_len > 0 else -1)
                            for q in seqs], dtype=np.int64).reshape(B, 7)
        except IndexError:
            arr = np.array([fields(q) for q in seqs], dtype=np.int64).reshape(B, 7)
        rows, kvl, cach, ntok = arr[:, 0].copy(), arr[:, 1], arr[:, 2], arr[:, 3]
        kv_first, kv_last, kv_prev = arr[:, 4], arr[:, 5], arr[:, 6]
        known = rows >= 0
        self.last_used[rows[known]] = step
        # O(1) fingerprint per known sequence
        trusted = known.copy()
        if known.any():
            r = np.where(known, rows, 0)
            t = self.tracked[r]
            trusted &= (t == cach) & (t > 0) & (kvl >= t) & (kv_first == self.first[r]) & (kv_prev == self.last[r])
            if prefill:
                for i in np.flatnonzero(trusted):
                    n = int(self.tracked[rows[i]])
                    kv = seqs[i].kv_indices
                    if np.int64(self._weighted_sum(kv if n == len(kv) else kv[:n], 0)) != self.csum[rows[i]]:
                        trusted[i] = False          # same ends, different middle: resend the row
        if self.fast_decode and B and not self.verify and bool(trusted.all()) and bool((kvl == cach + 1).all()):
            # Steady-state decode (every sequence known and trusted, exactly one new slot each): the message and the
            # bookkeeping of the general path below, written out directly (a dozen numpy calls instead of ~40; at the
            # TP8 shapes the device step is 0.18 ms and this packing is most of the host's share of a step).
            if page_size > 1:
                head = cach % page_size == 0
                if not bool(np.all(np.where(head, kv_last % page_size == 0, kv_last == self.last[rows] + 1))):
                    self._check_pages(seqs, kv_last.astype(np.int32), cach, np.ones(B, np.int64), np.arange(B), page_size)
            max_len = int(kvl.max())
            if max_len > self.table.shape[1]:
                self._grow(self.table.shape[0], max(max_len, 2 * self.table.shape[1]))
            if self._iota is None or self._iota.size < B:
                self._iota = np.arange(max(B, 256), dtype=np.int64)
            qls = ntok - cach
            msg = np.empty(8 * B, dtype=np.int32)
            m = msg.reshape(8, B)
            m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7] = rows, kvl, cach, qls, cach, 1, self._iota[:B], kv_last
            self.last[rows] = kv_last
            self.tracked[rows] = kvl
            self.csum[rows] += kv_last * (cach % 65521 + 1)         # the new slot sits at position cach (wraps like _weighted_sum)
            dev = self._upload(msg)
            self.last_msg_bytes = msg.nbytes
            return dev, B, int(qls.sum()), int(kvl.sum()), B
        n_unknown = int(B - known.sum())
        if n_unknown:
            self._take_rows(n_unknown)
            for i in np.flatnonzero(~known):
                sid = seqs[i].seq_id
                row = row_of.get(sid)           # the same seq_id twice in one batch shares a row
                if row is None:
                    row = self.free_rows.pop()
                    row_of[sid] = row
                    self.owner[row] = sid
                rows[i] = row
            self.last_used[rows] = step
        start = np.where(trusted, cach, 0)
        cnt = np.maximum(kvl - start, 0)
        woff = np.zeros(B, dtype=np.int64)
        if B > 1:
            np.cumsum(cnt[:-1], out=woff[1:])
        n_payload = int(cnt.sum())
        payload = np.empty(n_payload, dtype=np.int32)
        # row checksums: a row (re)sent from position 0 starts over, a delta adds to what the row has
        if n_payload == B and bool((cnt == 1).all()):            # steady-state decode: one new slot per sequence
            payload[:] = kv_last
            self.csum[rows] = np.where(start == 0, 0, self.csum[rows]) + kv_last * (start % 65521 + 1)
        else:
            self.csum[rows[start == 0]] = 0
            for i in np.flatnonzero(cnt):
                a, c = int(woff[i]), int(cnt[i])
                kv = seqs[i].kv_indices
                payload[a:a + c] = kv if c == len(kv) else kv[len(kv) - c:]
                base = self.csum[rows[i]] if start[i] else np.int64(0)
                with np.errstate(over="ignore"):
                    self.csum[rows[i]] = base + np.int64(self._weighted_sum(payload[a:a + c], int(start[i])))
        self.full_resends += int(((start == 0) & (kvl > 0)).sum())
        if page_size > 1 and n_payload:
            if n_payload == B and bool(trusted.all()) and bool((cnt == 1).all()):
                # steady-state decode, vectorised: the one new slot of every sequence either opens a page or follows
                # the slot this table saw last (the fingerprint just verified that it is kv[start - 1])
                head = start % page_size == 0
                new = payload.astype(np.int64)
                if not bool(np.all(np.where(head, new % page_size == 0, new == self.last[rows] + 1))):
                    self._check_pages(seqs, payload, start, cnt, woff, page_size)     # raises with the sequence id
            else:
                self._check_pages(seqs, payload, start, cnt, woff, page_size)
        if self.verify:
            for i, s in enumerate(seqs):
                mir = self._mirror.get(s.seq_id, [])
                st = int(start[i])
                assert mir[:st] == list(s.kv_indices[:st]), f"slot table contract violated for seq {s.seq_id}"
                self._mirror[s.seq_id] = list(s.kv_indices)
        # fingerprints for the next step
        nz = kvl > 0
        if nz.any():
            rr = rows[nz]
            self.first[rr] = kv_first[nz]
            self.last[rr] = kv_last[nz]
        self.tracked[rows] = kvl
        max_len = int(kvl.max()) if B else 0
        if max_len > self.table.shape[1]:      # (the copy of the old contents is ordered before this step's update)
            self._grow(self.table.shape[0], max(max_len, 2 * self.table.shape[1]))
        qls = ntok - cach
        msg = np.concatenate([rows, kvl, cach, qls, start, cnt, woff]).astype(np.int32)
        msg = np.concatenate([msg, payload]) if n_payload else msg
        dev = self._upload(msg)
        self.last_msg_bytes = msg.nbytes
        total_new = int(np.maximum(kvl - cach, 0).sum())
        return dev, n_payload, int(qls.sum()), int(kvl.sum()), total_new


# =======================================================================================================
# backend
# =======================================================================================================
class AttentionBackend:
    """Signature of the reference constructor (:26-34); ``page_size`` is an extension (the
    reference hard-codes 1, :66) used by the long-context page-16 configuration."""

    def __init__(self, kv_cache: KVCachePool, num_qo_heads: int, num_kv_heads: int, head_dim: int,
                 dtype: torch.dtype, device: torch.device, page_size: int = 1):
        self.kv_cache = kv_cache
        self.num_qo_heads = num_qo_heads
        self.num_kv_heads = num_kv_heads
        self.head_dim = head_dim
        self.dtype = dtype
        self.device = torch.device(device)
        self.page_size = page_size
        assert page_size >= 1 and (page_size & (page_size - 1)) == 0, "page_size must be a power of two"
        if self.device.type != "cuda":
            raise RuntimeError("AttentionBackend needs a CUDA device: the b200attn path has no CPU fallback")
        self.slot_table = SlotTable(self.device)
        self._workspace = torch.empty(0, dtype=torch.uint8, device=self.device)
        self._graph_plans: dict[int, AttentionPlan] = {}
        self._graph_scratch: Optional[torch.Tensor] = None

    # -- helpers -----------------------------------------------------------------------------------
    def _ensure_workspace(self, bs: int, total_q: int) -> torch.Tensor:
        need = max(ops.decode_workspace_bytes(max(bs, 1), self.num_qo_heads, self.num_kv_heads, self.head_dim),
                   ops.prefill_workspace_bytes(max(bs, 1), max(total_q, 1), self.num_qo_heads, self.num_kv_heads,
                                               self.head_dim))
        if self._workspace.numel() < need:
            # plans captured in CUDA graphs keep a reference to the buffer they were captured with
            self._workspace = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._workspace

    def _pages(self, n_tokens_total_upper: int, B: int) -> int:
        ps = self.page_size
        return n_tokens_total_upper if ps == 1 else (n_tokens_total_upper + ps - 1) // ps + B

    # -- eager ---------------------------------------------------------------------------------------
    def build_metadata(self, batch) -> AttentionMetadata:
        """Reference :62-149.  Produces the same arrays (bit-exact at page_size 1)."""
        seqs = batch.seqs
        B = len(seqs)
        mode = mode_name(batch.forward_mode)
        st = self.slot_table
        ps = self.page_size
        msg, n_payload, total_q, total_kv, total_new = st.pack_step(seqs, ps, prefill=(mode == "PREFILL"))
        if ps == 1:
            n_pages = total_kv
        else:
            n_pages = sum((len(s.kv_indices) + ps - 1) // ps for s in seqs)
        dev = self.device
        # one allocation for the six arrays of a step; every array starts on a 16-byte boundary
        nb, n64 = (B + 4) & ~3, (2 * total_new + 3) & ~3
        buf = torch.empty(n64 + 4 * nb + n_pages, dtype=torch.int32, device=dev)
        out_slots = buf[:2 * total_new].view(torch.int64)
        qo_indptr, kv_indptr = buf[n64:n64 + B + 1], buf[n64 + nb:n64 + nb + B + 1]
        scratch, last_page_len = buf[n64 + 2 * nb:n64 + 2 * nb + B + 1], buf[n64 + 3 * nb:n64 + 3 * nb + B]
        kv_indices = buf[n64 + 4 * nb:n64 + 4 * nb + n_pages]
        if B > 0:
            ops.call("step_meta", st.table, msg, B, n_payload, B, ps, False, qo_indptr, kv_indptr, kv_indices, last_page_len,
                          out_slots, scratch)
        else:
            qo_indptr.zero_()
            kv_indptr.zero_()
        plan = AttentionPlan(mode, B, total_q, ps, qo_indptr, kv_indptr, kv_indices, last_page_len,
                             self._ensure_workspace(B, total_q), self.num_qo_heads, self.num_kv_heads,
                             self.head_dim)
        return AttentionMetadata(forward_mode=batch.forward_mode, kv_cache=self.kv_cache,
                                 output_kv_indices=out_slots, wrapper=plan)

    # -- CUDA graphs -----------------------------------------------------------------------------------
    def prepare_for_cuda_graph_capture(self, graph, max_bs: int, context_len: int) -> None:
        """Reference :151-168: same four named buffers, same dtypes, shapes and initial contents."""
        dev = self.device
        graph.set_input_buffer("paged_kv_indptr_buffer",
                               torch.arange(0, max_bs + 1, dtype=torch.int32, device=dev) * context_len)
        graph.set_input_buffer("paged_kv_indices_buffer",
                               torch.zeros((max_bs * context_len,), dtype=torch.int32, device=dev))
        graph.set_input_buffer("paged_kv_last_page_len_buffer",
                               torch.ones((max_bs,), dtype=torch.int32, device=dev))
        graph.set_input_buffer("output_kv_indices_buffer",
                               torch.zeros((max_bs,), dtype=torch.int64, device=dev))
        self._graph_scratch = torch.zeros(max_bs + 1, dtype=torch.int32, device=dev)
        self.slot_table.reserve(2 * max_bs, context_len)
        # one workspace, sized for the largest captured batch, shared by all graphs
        self._graph_workspace = torch.empty(
            ops.decode_workspace_bytes(max_bs, self.num_qo_heads, self.num_kv_heads, self.head_dim),
            dtype=torch.uint8, device=dev)
        self._graph_context_len = context_len

    def build_metadata_for_cuda_graph_capture(self, graph, bs: int, context_len: int) -> AttentionMetadata:
        """Reference :171-216: metadata over slices of the static buffers (worst case: every
        sequence ``context_len`` long, which is what the initial indptr encodes)."""
        indptr = graph.get_input_buffer("paged_kv_indptr_buffer")[:bs + 1]
        indices = graph.get_input_buffer("paged_kv_indices_buffer")[:bs * context_len]
        last = graph.get_input_buffer("paged_kv_last_page_len_buffer")[:bs]
        out_idx = graph.get_input_buffer("output_kv_indices_buffer")[:bs]
        from ..core.common import ForwardMode
        plan = AttentionPlan("DECODE", bs, bs, self.page_size, None, indptr, indices, last,
                             self._graph_workspace, self.num_qo_heads, self.num_kv_heads, self.head_dim)
        self._graph_plans[bs] = plan
        return AttentionMetadata(forward_mode=ForwardMode.DECODE, kv_cache=self.kv_cache,
                                 output_kv_indices=out_idx, wrapper=plan)

    def build_metadata_for_cuda_graph_replay(self, graph, batch, padded_bs: int) -> AttentionMetadata:
        """Reference :219-295: refill the static buffers for this step (padding sequences have
        kv_len 1 -> slot 0, slot map -1), with two kernels instead of 2*B copies + a host plan."""
        seqs = batch.seqs
        bs = len(seqs)
        st = self.slot_table
        msg, n_payload, _tq, total_kv, _tn = st.pack_step(seqs, self.page_size)
        indptr_buf = graph.get_input_buffer("paged_kv_indptr_buffer")
        indices_buf = graph.get_input_buffer("paged_kv_indices_buffer")
        last_buf = graph.get_input_buffer("paged_kv_last_page_len_buffer")
        out_buf = graph.get_input_buffer("output_kv_indices_buffer")
        # the reference never rewrites last_page_len on replay (stays 1 at page_size 1); for
        # page_size > 1 it must be rebuilt
        ops.call("step_meta", st.table, msg, bs, n_payload, padded_bs, self.page_size, True, None, indptr_buf, indices_buf,
                      None if self.page_size == 1 else last_buf, out_buf, self._graph_scratch)
        plan = self._graph_plans[padded_bs]
        return AttentionMetadata(forward_mode=batch.forward_mode, kv_cache=self.kv_cache,
                                 output_kv_indices=out_buf[:padded_bs], wrapper=plan)


# =======================================================================================================
# module side
# =======================================================================================================
@contextmanager
def attention_kv_cache(model: nn.Module, metadata: AttentionMetadata):
    """Reference :297-306."""
    mods = [m for m in model.modules() if isinstance(m, Attention)]
    for m in mods:
        m.set_attention_metadata(metadata)
    yield
    for m in mods:
        m.set_attention_metadata(None)


class Attention(nn.Module):
    """Reference :310-373: store this step's K/V into the pool, then paged attention."""

    def __init__(self, num_heads: int, head_dim: int, scaling: Optional[float], num_kv_heads: int, layer_id: int):
        super().__init__()
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.num_kv_heads = num_kv_heads
        self.num_kv_groups = num_heads // num_kv_heads
        self.scaling = self.head_dim ** -0.5 if scaling is None else scaling
        self.layer_id = layer_id
        self.attention_metadata: Optional[AttentionMetadata] = None

    def set_attention_metadata(self, metadata: Optional[AttentionMetadata]) -> None:
        self.attention_metadata = metadata

    def forward_fused_qkv(self, qkv: torch.Tensor, positions: torch.Tensor, q_norm_weight: torch.Tensor,
                          k_norm_weight: torch.Tensor, eps: float, cos_sin_cache: torch.Tensor) -> torch.Tensor:
        """what Qwen3Attention.forward does between qkv_proj and o_proj (models/qwen3.py:123-134:
        split, q_norm, k_norm, rotary_emb, then Attention.forward with its slot write, reference :347-352) with ONE
        kernel ahead of the attention kernel.  q is normalised/rotated in place inside ``qkv``; K (normalised,
        rotated, the reference's interleaved layout) and V go straight to the pool rows ``output_kv_indices``."""
        md = self.attention_metadata
        assert md is not None
        Hq, Hkv, D = self.num_heads, self.num_kv_heads, self.head_dim
        qkv = qkv.view(-1, (Hq + 2 * Hkv) * D)
        k_cache, v_cache = md.kv_cache.get_kv_cache(self.layer_id)
        q = qkv[:, : Hq * D]
        ops.call("qk_norm_rope_store", qkv, positions.reshape(-1), q_norm_weight, k_norm_weight, eps, cos_sin_cache, q, None,
                               k_cache, v_cache, md.output_kv_indices, Hq, Hkv, D)
        return self.forward(q, None, None)

    def forward(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, save_kv_cache: bool = True):
        md = self.attention_metadata
        assert md is not None
        if q.dim() != 2 or q.stride(-1) != 1 or q.stride(0) % 8 or q.data_ptr() % 16:
            q = q.contiguous()      # the kernels take a row stride: a column slice of the qkv buffer is read in place
        if k is not None:
            assert v is not None
            if save_kv_cache:
                md.kv_cache.set_kv_cache(self.layer_id, md.output_kv_indices,
                                         k.view(-1, self.num_kv_heads, self.head_dim),
                                         v.view(-1, self.num_kv_heads, self.head_dim))
        q = q.unflatten(-1, (self.num_heads, self.head_dim))
        k_cache, v_cache = md.kv_cache.get_kv_cache(self.layer_id)
        name = mode_name(md.forward_mode)
        if name not in ("PREFILL", "DECODE"):
            raise NotImplementedError(name)
        md.wrapper._sm_scale = self.scaling
        o = md.wrapper.run(q, (k_cache, v_cache))
        return o.view(-1, self.num_heads * self.head_dim)
