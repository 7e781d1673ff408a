"""CPU: SURVEY.md 8 f-1 against the reference's OWN Scheduler / KVCacheManager / RadixTree (baseline/_ref), patched
by nanovllmx_b200.core.prefix_reuse.install(): admission-time radix hits, the cache_sequence fixes, preemption with a
pinned prefix, and a slot-conservation invariant under a random workload (no leak, no double free)."""
import os
import random
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")


@pytest.fixture(scope="module")
def ref():
    if not os.path.isdir(os.path.join(REF, "nanovllmx", "core")):
        pytest.skip("baseline/_ref not installed (python baseline/install_ref.py; needs /root/reference)")
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import nanovllmx_b200  # noqa: F401  (alias module)
    from nanovllmx_b200.core import prefix_reuse
    prefix_reuse.install()
    import nanovllmx.core.scheduler as S
    import nanovllmx.core.common as C
    return S, C


def _tree_slots(tree):
    out = []

    def walk(n):
        out.extend(n.value)
        for c in n.children.values():
            walk(c)
    walk(tree.root)
    return out


def _check_conservation(sched, size):
    """Every slot is in exactly one place: the free list, the radix tree, or the private tail of a live sequence."""
    km = sched.kv_manager
    free = list(km.kv_cache_allocator.free_slots)
    tree = _tree_slots(km.radix_tree)
    private = []
    for s in list(sched.running) + list(sched.waiting):
        p = km.unfinished_sequences.get(s.seq_id, (0, None))[0]
        private.extend(s.kv_indices[p:])
    allslots = free + tree + private
    assert len(allslots) == len(set(allslots)), "a slot is owned twice"
    assert sorted(allslots) == list(range(size)), f"{size - len(allslots)} slots leaked"


def _run_step(sched, C, next_token):
    """One engine step without a model (core/engine.py:81-100, 119-141): schedule, 'sample', update, finish."""
    batch = sched.schedule()
    if batch is None:
        return None
    for s in batch.seqs:
        sched.update_sequence(s, next_token(s))
        sp = s.sampling_params
        if sp.max_new_tokens and s.num_tokens >= s.prompt_len + sp.max_new_tokens:
            sched.finish_sequence(s)
    return batch


def _seq(C, sid, toks, new):
    return C.Sequence(seq_id=sid, token_ids=list(toks), num_tokens=len(toks), prompt_len=len(toks),
                      sampling_params=C.SamplingParams(max_new_tokens=new, ignore_eos=True))


def test_admission_hits_the_radix_tree_and_positions_stay_aligned(ref):
    S, C = ref
    size = 256
    sched = S.Scheduler(kv_cache_size=size, max_bs=8)
    a = _seq(C, "a", range(100, 140), 3)
    sched.add_sequence(a)
    b = _run_step(sched, C, lambda s: 7)
    assert b.forward_mode == C.ForwardMode.PREFILL and a.cached_kv_len == 40
    a_slots = list(a.kv_indices)
    while sched.has_unfinished_sequences():
        _run_step(sched, C, lambda s: 7)
    _check_conservation(sched, size)
    # the tree holds exactly the tokens that HAD a KV row: 40 prompt + 2 generated (the 3rd was sampled, never computed)
    tree = sched.kv_manager.radix_tree
    idx, _ = tree.match_prefix(list(range(100, 140)) + [7, 7, 7, 7])
    assert len(idx) == 42 and idx[:40] == a_slots[:40]
    # a new request sharing 30 prompt tokens: admitted with those 30 slots, computes only the rest
    c = _seq(C, "c", list(range(100, 130)) + [1, 2, 3, 4, 5], 2)
    sched.add_sequence(c)
    batch = sched.schedule()
    assert batch.forward_mode == C.ForwardMode.PREFILL
    assert c.cached_kv_len == 30 and c.kv_indices[:30] == a_slots[:30] and len(c.kv_indices) == 35
    assert len(set(c.kv_indices[30:]) & set(a_slots)) == 0
    _check_conservation(sched, size)
    # a request that IS a cached prefix entirely: the last token is still computed (it produces the logits)
    d = _seq(C, "d", range(100, 120), 1)
    sched.add_sequence(d)
    for s in batch.seqs:
        sched.update_sequence(s, 9)
    sched.schedule()
    assert d.cached_kv_len == 19 and len(d.kv_indices) == 20
    # single-token and empty-prefix prompts do not trip the key[0] reads (core/kv_cache.py:161, :189)
    e = _seq(C, "e", [5], 1)
    sched.add_sequence(e)
    sched.update_sequence(d, 9)
    sched.finish_sequence(d)
    sched.schedule()
    assert e.cached_kv_len == 0 and len(e.kv_indices) == 1
    _check_conservation(sched, size)


def test_duplicate_slots_of_sequences_first_cached_at_finish_are_freed(ref):
    """core/kv_cache.py:353-357 frees the duplicate range only `if old_prefix_len != 0`: two sequences with a common
    prefix that were admitted before either finished (the worked example in the reference's own comment, :346-352)."""
    S, C = ref
    size = 64
    sched = S.Scheduler(kv_cache_size=size, max_bs=8)
    s1, s2 = _seq(C, "s1", [1, 2, 3, 4], 2), _seq(C, "s2", [1, 2, 3, 5], 2)
    sched.add_sequence(s1)
    sched.add_sequence(s2)
    while sched.has_unfinished_sequences():
        _run_step(sched, C, lambda s: 42)
    _check_conservation(sched, size)                      # the stock code leaks 3 slots here
    tree = sched.kv_manager.radix_tree
    i1, _ = tree.match_prefix([1, 2, 3, 4])
    i2, _ = tree.match_prefix([1, 2, 3, 5])
    assert i1[:3] == i2[:3] and i1[3] != i2[3]            # "[1,2,3,5] -> [10,11,12,23]"


def test_preemption_with_a_pinned_prefix_and_random_workload_conserve_slots(ref):
    S, C = ref
    size = 300
    rng = random.Random(0)
    sched = S.Scheduler(kv_cache_size=size, max_bs=6)
    base = [rng.randrange(50) for _ in range(40)]
    n = 0
    for step in range(400):
        if rng.random() < 0.25 and len(sched.waiting) < 4:
            cut = rng.randrange(0, 40)
            toks = base[:cut] + [rng.randrange(50) for _ in range(rng.randrange(1, 30))]
            sched.add_sequence(_seq(C, f"r{n}", toks, rng.randrange(1, 25)))
            n += 1
        if rng.random() < 0.05 and sched.running:
            victim = rng.choice(list(sched.running))
            if victim.seq_id not in sched.scheduled:
                sched.preempt(victim)                      # core/scheduler.py:140-153
                assert victim.kv_indices == [] and victim.cached_kv_len == 0
        try:
            _run_step(sched, C, lambda s: rng.randrange(50))
        except RuntimeError:
            pass                                           # pool exhausted at admission: stock behaviour (scheduler.py:71)
        if step % 10 == 0:
            _check_conservation(sched, size)
    while sched.has_unfinished_sequences():
        _run_step(sched, C, lambda s: 3)
    _check_conservation(sched, size)
    assert n > 50


def test_failed_admission_gives_the_adopted_prefix_back(ref):
    """alloc_kv_slots adopts and pins the matched prefix before the stock allocation; when the pool
    cannot supply the private tail the RuntimeError must leave nothing behind (no pin, no half-admitted sequence)."""
    S, C = ref
    size = 48
    sched = S.Scheduler(kv_cache_size=size, max_bs=8)
    a = _seq(C, "a", range(100, 120), 2)
    sched.add_sequence(a)
    while sched.has_unfinished_sequences():
        _run_step(sched, C, lambda s: 7)
    km = sched.kv_manager
    tree = km.radix_tree
    _, node = tree.match_prefix(list(range(100, 120)))
    ref_before = node.ref_count if hasattr(node, "ref_count") else None
    big = _seq(C, "big", list(range(100, 120)) + list(range(1000, 1000 + size)), 1)   # shares 20 tokens, needs > pool
    with pytest.raises(RuntimeError):
        sched.alloc_kv_slots(big)
    assert big.kv_indices == [] and big.cached_kv_len == 0
    assert "big" not in km.unfinished_sequences
    if ref_before is not None:
        assert node.ref_count == ref_before
    # (the failed attempt may have evicted unpinned leaves, core/kv_cache.py:322-326: the free list can only have grown)
    _check_conservation(sched, size)
    # the same request can be admitted once there is room: nothing was left half done
    ok = _seq(C, "ok", list(range(100, 120)) + [1, 2, 3], 1)
    sched.alloc_kv_slots(ok)
    assert ok.cached_kv_len == 20 and len(ok.kv_indices) == 23


def test_slots_scheduled_but_never_computed_do_not_enter_the_tree(ref):
    """a sequence finished between schedule() and update_sequence() (abort with a step in flight) has
    kv_indices longer than cached_kv_len.  Those rows were never written; cache_sequence must free them, not cache them."""
    S, C = ref
    size = 64
    sched = S.Scheduler(kv_cache_size=size, max_bs=8)
    a = _seq(C, "a", range(10, 30), 8)
    sched.add_sequence(a)
    _run_step(sched, C, lambda s: 5)                      # prefill computed: cached_kv_len 20, 21 tokens
    batch = sched.schedule()                              # decode step scheduled: slot 21 allocated, not computed
    assert batch.seqs == [a] and len(a.kv_indices) == 21 and a.cached_kv_len == 20
    sched.finish_sequence(a)                              # aborted before the step's result came back
    _check_conservation(sched, size)
    idx, _ = sched.kv_manager.radix_tree.match_prefix(list(range(10, 30)) + [5, 5])
    assert len(idx) == 20, "an uncomputed slot was cached as valid KV"
    assert len(_tree_slots(sched.kv_manager.radix_tree)) == 20
