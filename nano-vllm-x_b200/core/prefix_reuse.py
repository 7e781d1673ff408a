"""Admission-time prefix reuse for the reference scheduler.

In the reference snapshot the radix tree is only written (``KVCacheManager.cache_sequence`` at finish,
core/kv_cache.py:333-369) and never consulted when a sequence is admitted: new sequences always start with
``kv_indices = []`` / ``cached_kv_len = 0`` (core/engine.py:62-68, core/scheduler.py:67-74), so the
prefix-page prefill path of the attention kernels never runs under the shipped engine.  ``install()`` applies
the smallest change that makes BASELINE.json configs[3] ("radix-cache hits drive prefix-page prefill") real,
as wrappers around the reference's own classes — nothing in the reference tree is edited:

1. ``Scheduler.alloc_kv_slots`` (core/scheduler.py:119-128), called for every sequence of the prefill loop
   (:67-74): a sequence that owns no slots yet first asks ``RadixTree.match_prefix(token_ids[:-1])``
   (core/kv_cache.py:159; the last token is always computed, it produces the logits), adopts the matched slots
   as ``kv_indices[:P]`` with ``cached_kv_len = P``, pins the matched path (``inc_ref``) and records
   ``(P, node)`` in ``KVCacheManager.unfinished_sequences`` — the bookkeeping ``cache_sequence`` already reads.
2. ``Scheduler.free_kv_slots`` (core/scheduler.py:130-138, preemption): the first ``P`` slots belong to the tree,
   only the private tail is freed and the pin is dropped.  (Reached through ``Scheduler.preempt`` for a sequence
   that is still in ``running``.  The stock memory-pressure path, core/scheduler.py:98-101, pops the victim from
   ``running`` BEFORE calling ``preempt``, which then returns at its ``if not seq in self.running`` guard (:146-147):
   in the shipped engine that preemption is a no-op and the victim simply leaves the queue.  That is reference
   behaviour outside this row and is left as it is.)  If the pool cannot supply the private tail at admission,
   the adoption of step 1 is undone before the error propagates.
3. ``KVCacheManager.cache_sequence`` (core/kv_cache.py:333-369), three fixes the hook exposes:
   * tokens without KV are not inserted: at finish ``token_ids`` is one longer than ``kv_indices``
     (core/scheduler.py:161-171 appends the sampled token), and ``RadixTree.insert`` slices key and value by the
     same prefix length (:193-194), so the stock call stores nodes whose value is shorter than their key and a
     later match would return slots shifted against token positions; slots allocated but not yet computed
     (``kv_indices`` beyond ``cached_kv_len``: a sequence finished between scheduling and its update) go back
     to the allocator instead of into the tree;
   * the duplicate-slot leak: slots in ``[old_prefix_len, new_prefix_len)`` duplicate slots the tree already
     holds and are freed only ``if old_prefix_len != 0`` (:353-357); a sequence first cached at finish
     (``old_prefix_len == 0``, every sequence of the stock engine) leaks them;
   * ``RadixTree.match_prefix`` / ``insert`` read ``key[0]`` before checking for an empty key (:161, :189).

The attention side needs nothing new: ``AttentionBackend.build_metadata`` already takes
``seqlens_q = len(token_ids) - cached_kv_len`` (layers/attention.py:67-71) and the slot table re-uploads a row whose
prefix slots changed (``SlotTable`` fingerprint).
"""
from __future__ import annotations

_INSTALLED = False


def install() -> None:
    """Patch ``nanovllmx.core.scheduler.Scheduler`` and ``nanovllmx.core.kv_cache`` in this process.
    Idempotent.  Call it wherever the engine process imports the reference (the scheduler lives in the engine
    process: core/engine.py:43-46)."""
    global _INSTALLED
    if _INSTALLED:
        return
    import nanovllmx.core.kv_cache as K
    import nanovllmx.core.scheduler as S
    from nanovllmx.core.common import SequenceStatus

    orig_match, orig_insert = K.RadixTree.match_prefix, K.RadixTree.insert

    def match_prefix(self, key):
        if len(key) == 0:                               # core/kv_cache.py:161 reads key[0] unguarded
            return [], self.root
        return orig_match(self, key)

    def insert(self, key, value):
        if len(key) == 0:                               # core/kv_cache.py:189
            return 0, self.root
        return orig_insert(self, key, value)

    def cache_sequence(self, seq):
        """core/kv_cache.py:333-369 with the three fixes listed in the module docstring."""
        tree = self.radix_tree
        # only tokens whose KV row has been COMPUTED: a sequence finished or aborted between schedule() and
        # update_sequence() (abort while a pipeline-parallel step is in flight, core/engine.py step_with_pp) owns
        # slots beyond cached_kv_len that were never written; they must not become valid KV in the tree
        n = min(len(seq.kv_indices), seq.cached_kv_len)
        if seq.status == SequenceStatus.FINISHED and n < len(seq.kv_indices):
            self.kv_cache_allocator.free(seq.kv_indices[n:])
            del seq.kv_indices[n:]
        key = seq.token_ids[:n]
        old_prefix_len, old_last_node = self.unfinished_sequences.get(seq.seq_id, (0, tree.root))
        new_prefix_len, new_last_node = tree.insert(key, seq.kv_indices[:n])
        new_indices, _ = tree.match_prefix(key)
        if new_prefix_len > old_prefix_len:             # our copies of rows the tree already holds: free ours, adopt the tree's
            self.kv_cache_allocator.free(seq.kv_indices[old_prefix_len:new_prefix_len])
            seq.kv_indices[old_prefix_len:new_prefix_len] = new_indices[old_prefix_len:new_prefix_len]
        self.unfinished_sequences[seq.seq_id] = (len(new_indices), new_last_node)
        tree.dec_ref(old_last_node)
        tree.inc_ref(new_last_node)
        if seq.status == SequenceStatus.FINISHED:
            tree.dec_ref(new_last_node)
            self.unfinished_sequences.pop(seq.seq_id, None)
            seq.kv_indices.clear()                      # the tree owns the rows now
            seq.cached_kv_len = 0

    orig_alloc = S.Scheduler.alloc_kv_slots

    def alloc_kv_slots(self, seq):
        km = self.kv_manager
        if not seq.kv_indices and seq.cached_kv_len == 0 and len(seq.token_ids) > 1 \
                and seq.seq_id not in km.unfinished_sequences:
            indices, node = km.radix_tree.match_prefix(seq.token_ids[:-1])
            if indices:
                seq.kv_indices.extend(indices)
                seq.cached_kv_len = len(indices)
                km.radix_tree.inc_ref(node)
                km.unfinished_sequences[seq.seq_id] = (len(indices), node)
                try:
                    orig_alloc(self, seq)
                except Exception:
                    # pool full at admission (core/scheduler.py:71 has no try): give the prefix back, or the pinned
                    # path could never be evicted again and a retry would find the sequence half admitted
                    km.radix_tree.dec_ref(node)
                    km.unfinished_sequences.pop(seq.seq_id, None)
                    seq.kv_indices.clear()
                    seq.cached_kv_len = 0
                    raise
                return
        orig_alloc(self, seq)

    def free_kv_slots(self, seq):
        if not seq.kv_indices:
            return
        km = self.kv_manager
        owned_by_tree = 0
        entry = km.unfinished_sequences.pop(seq.seq_id, None)
        if entry is not None:
            owned_by_tree, node = entry
            km.radix_tree.dec_ref(node)
        km.free_slots(seq.kv_indices[owned_by_tree:])
        seq.kv_indices.clear()
        seq.cached_kv_len = 0

    K.RadixTree.match_prefix = match_prefix
    K.RadixTree.insert = insert
    K.KVCacheManager.cache_sequence = cache_sequence
    S.Scheduler.alloc_kv_slots = alloc_kv_slots
    S.Scheduler.free_kv_slots = free_kv_slots
    _INSTALLED = True
