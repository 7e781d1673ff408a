"""GPU parity of b200attn::sample through the C ABI against the sampler oracle and the fixture
made by the reference's own Sampler.  The draw is argmax(score); an id may differ from the oracle's only on a
floating-point near-tie of the two scores (tolerance written below), never on the kept set."""
import os

import numpy as np
import pytest
import torch

from oracle import sampler_oracle as so
from tests.conftest import ROOT
from tests.test_sampler_oracle import CASES, load_case

pytestmark = pytest.mark.gpu
NEAR_TIE = 1e-4      # relative gap of the oracle's two scores below which either id is accepted


@pytest.fixture(scope="module")
def sgold():
    return np.load(os.path.join(ROOT, "tests", "golden", "sampler_golden.npz"))


def cu(t):
    return None if t is None else t.cuda()


def run(pkg, logits, temps=None, min_ps=None, top_ps=None, top_ks=None, q=None, seed=0, offset=0):
    return pkg.ops.sample(cu(logits), cu(temps), cu(min_ps), cu(top_ps), cu(top_ks), cu(q), seed, offset).cpu()


def assert_ids(got, want, scores, what, logits=None):
    """ids equal, or a floating-point near-tie of the two scores, or the reference's winner is one of the ties AT the
    cut value (its unstable sort decides which of those survive, see boundary_probe; needs ``logits``)."""
    for b, (g, w) in enumerate(zip(got.tolist(), want.tolist())):
        if g == w:
            continue
        sg, sw = scores[b, g].item(), scores[b, w].item()
        if sg > 0 and abs(sg - sw) <= NEAR_TIE * abs(sw):
            continue
        kept = scores[b] > 0
        if logits is not None and logits[b, w].float() == logits[b][kept].float().min():
            continue
        raise AssertionError(f"{what}: row {b} got {g} (score {sg}) want {w} (score {sw}); kept {int(kept.sum())}")


def boundary_probe(pkg, logits, temps, min_ps, top_ps, top_ks):
    """Pins the kept set's lower boundary per row.  Every cut the sampler makes is monotone in ascending (value,
    index) order, so two membership probes decide it: the oracle's first kept token must be kept and its last removed
    token must be removed.  A probe gives token j the noise 1e-37 and everything else 1e37 (a 170-nat head start, more
    than any scaled-logit range used here), so the kernel returns j iff j is in its kept set."""
    B, V = logits.shape
    x = logits.float() / (temps.unsqueeze(1) if temps is not None else 1.0)
    assert (x.max(-1).values - x.min(-1).values).max() < 150
    kept = so.masked_logits(logits, temps, None if min_ps is None else min_ps.clone(), top_ps, top_ks) > -float("inf")
    idx = torch.arange(V)
    first_kept, last_removed = [], []
    for b in range(B):
        kv = torch.where(kept[b], x[b], torch.full_like(x[b], float("inf")))
        # Ties at the cut value: the reference's sort is not stable (sampler.py:65), so WHICH of them survive is
        # implementation-defined there; HOW MANY survive is not.  The build removes lowest indices first.
        ties = idx[x[b] == kv.min()]
        n_kept = int(kept[b][ties].sum())
        kept[b][ties] = False
        kept[b][ties[len(ties) - n_kept:]] = True
        kv = torch.where(kept[b], x[b], torch.full_like(x[b], float("inf")))
        first_kept.append(int(idx[kv == kv.min()][0]))
        if kept[b].all():
            last_removed.append(-1)
        else:
            rv = torch.where(~kept[b], x[b], torch.full_like(x[b], -float("inf")))
            last_removed.append(int(idx[rv == rv.max()][-1]))
    rows = torch.arange(B)
    q = torch.full((B, V), 1e37)
    q[rows, torch.tensor(first_kept)] = 1e-37
    got = run(pkg, logits, temps, min_ps, top_ps, top_ks, q)
    assert got.tolist() == first_kept, f"first kept token dropped: got {got.tolist()} want {first_kept}; kept {kept.sum(-1).tolist()}"
    probe = [j if j >= 0 else f for j, f in zip(last_removed, first_kept)]
    q = torch.full((B, V), 1e37)
    q[rows, torch.tensor(probe)] = 1e-37
    got = run(pkg, logits, temps, min_ps, top_ps, top_ks, q)
    for b in range(B):
        if last_removed[b] >= 0:
            assert int(got[b]) != last_removed[b], f"row {b}: token {last_removed[b]} should be removed (kept {int(kept[b].sum())} of {V})"


@pytest.mark.parametrize("name", CASES)
def test_golden_reference_sampler(pkg, sgold, name):
    logits, temps, min_ps, top_ps, top_ks, q, ids = load_case(sgold, name)
    _, scores = so.sample(logits, temps, min_ps.clone(), top_ps, top_ks, q)
    assert_ids(run(pkg, logits, temps, min_ps, top_ps, top_ks, q), ids, scores, name, logits)
    boundary_probe(pkg, logits, temps, min_ps, top_ps, top_ks)


def test_golden_greedy(pkg, sgold):
    logits = torch.from_numpy(sgold["greedy.logits_bf16"].astype(np.int16)).view(torch.bfloat16)
    assert torch.equal(run(pkg, logits), torch.from_numpy(sgold["greedy.ids"]))
    # temperature <= 0 rows are greedy as well (the build's definition; the reference divides by zero)
    t = torch.zeros(logits.shape[0])
    assert torch.equal(run(pkg, logits, t, None, None, None, torch.ones(logits.shape)), torch.from_numpy(sgold["greedy.ids"]))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("V", [1, 7, 1000, 4097, 151936])
def test_random_rows_vs_oracle(pkg, dtype, V):
    B = 5 if V > 100000 else 9
    g = torch.Generator().manual_seed(V)
    logits = (torch.randn(B, V, generator=g) * 3).to(dtype)
    temps = torch.rand(B, generator=g) * 1.5 + 0.3
    top_ks = torch.randint(1, max(2, V // 3 + 2), (B,), generator=g).clamp(max=V)
    top_ks[0] = V
    top_ps = torch.rand(B, generator=g).clamp(0.05, 1.0)
    top_ps[1 % B] = 1.0
    min_ps = torch.rand(B, generator=g) * 0.2
    min_ps[2 % B] = 0.0
    q = torch.empty(B, V).exponential_(1, generator=g)
    for kw in (dict(temps=temps), dict(temps=temps, top_ks=top_ks), dict(temps=temps, top_ps=top_ps),
               dict(temps=temps, min_ps=min_ps), dict(temps=temps, top_ks=top_ks, top_ps=top_ps, min_ps=min_ps),
               dict(top_ks=top_ks), dict(top_ps=top_ps)):
        a = (logits, kw.get("temps"), kw.get("min_ps"), kw.get("top_ps"), kw.get("top_ks"))
        want, scores = so.sample(a[0], a[1], None if a[2] is None else a[2].clone(), a[3], a[4], q)
        assert_ids(run(pkg, *a, q), want, scores, f"V={V} {sorted(kw)}", logits)
        boundary_probe(pkg, *a)


def test_ties_cut_by_index(pkg):
    """All-equal and few-valued rows: the top-p cut falls inside a run of ties and must remove the lowest indices
    first (stable ascending sort, sampler.py:65); top-k keeps every tie of the k-th value (:72)."""
    V = 5003      # prime: (1 - p) * V is never an integer, so the fp32 cumsum of the oracle cannot sit on the threshold
    rows = [torch.zeros(V), torch.arange(V).remainder(2).float(), torch.arange(V).remainder(5).float()]
    logits = torch.stack(rows).to(torch.bfloat16)
    B = logits.shape[0]
    temps = torch.ones(B)
    for p in (0.9, 0.5, 0.123, 0.0):
        boundary_probe(pkg, logits, temps, None, torch.full((B,), p), None)
    for k in (1, 2, 1000, 2501):
        boundary_probe(pkg, logits, temps, None, None, torch.full((B,), k, dtype=torch.long))
    boundary_probe(pkg, logits, temps, torch.full((B,), 0.5), torch.full((B,), 0.7), torch.full((B,), 3000, dtype=torch.long))


def test_greedy_lowest_index_and_top_k_1(pkg):
    g = torch.Generator().manual_seed(3)
    logits = torch.randn(6, 30000, generator=g).to(torch.bfloat16)
    logits[0, 17] = logits[0, 29000] = 9.0
    assert torch.equal(run(pkg, logits), torch.argmax(logits.float(), -1))
    assert run(pkg, logits)[0] == 17
    q = torch.empty(6, 30000).exponential_(1, generator=g)
    got = run(pkg, logits, torch.full((6,), 0.7), None, None, torch.ones(6, dtype=torch.long), q)
    assert (logits.float().gather(1, got.unsqueeze(1)).squeeze(1) == logits.float().max(-1).values).all()


def test_seed_mode_matches_oracle_stream_and_is_deterministic(pkg):
    B, V = 6, 10007
    g = torch.Generator().manual_seed(8)
    logits = (torch.randn(B, V, generator=g) * 2).to(torch.bfloat16)
    temps = torch.full((B,), 0.6)
    top_ks = torch.tensor([V, 50, 50, 1000, V, 3])
    a = run(pkg, logits, temps, None, None, top_ks, None, seed=42, offset=5)
    want, scores = so.sample(logits, temps, None, None, top_ks, so.seed_noise(42, 5, B, V))
    assert_ids(a, want, scores, "seed mode")
    assert torch.equal(a, run(pkg, logits, temps, None, None, top_ks, None, seed=42, offset=5))
    others = [run(pkg, logits, temps, None, None, top_ks, None, seed=42, offset=o) for o in range(6, 12)]
    assert any(not torch.equal(a, o) for o in others)


def test_seed_mode_distribution(pkg):
    """20 000 draws from one 64-token distribution: chi-square against softmax(logits / T)."""
    V, N, T = 64, 20000, 0.8
    g = torch.Generator().manual_seed(1)
    row = torch.randn(V, generator=g) * 1.5
    logits = row.repeat(N, 1)
    ids = run(pkg, logits, torch.full((N,), T), None, None, None, None, seed=7, offset=1)
    p = torch.softmax(row / T, -1).double().numpy()
    cnt = np.bincount(ids.numpy(), minlength=V).astype(np.float64)
    chi2 = ((cnt - N * p) ** 2 / (N * p)).sum()
    assert chi2 < 125.0, chi2        # 63 dof: P(chi2 > 125) < 1e-5


def test_sampler_module_equals_torch_pipeline(pkg):
    """Sampler(torch_noise=True) consumes the generator exactly as torch.multinomial does, so for the same CUDA
    generator state it returns the reference pipeline's ids (softmax(logits / T) -> multinomial, sampler.py:38-55)."""
    B, V = 32, 151936
    g = torch.Generator().manual_seed(2)
    logits = (torch.randn(B, V, generator=g) * 2).to(torch.bfloat16).cuda()
    temps = torch.full((B,), 0.6, device="cuda")
    torch.manual_seed(123)
    probs = torch.softmax(logits.float() / temps.unsqueeze(1), -1)
    want = torch.multinomial(probs, 1).squeeze(-1)
    torch.manual_seed(123)
    q = torch.empty_like(probs).exponential_(1)
    torch.manual_seed(123)
    got = pkg.Sampler(torch_noise=True)(logits, temps)
    assert_ids(got.cpu(), want.cpu(), (probs / q).cpu(), "Sampler module")
    assert pkg.Sampler()(logits).equal(torch.argmax(logits, -1))


def test_errors(pkg):
    with pytest.raises(RuntimeError):
        pkg.ops.sample(torch.zeros(2, 8), None, None, None, None, None, 0, 0)
    assert pkg.ops.sample(torch.zeros(0, 8, device="cuda"), None, None, None, None, None, 0, 0).numel() == 0
