"""CPU: the sampler oracle against the fixture produced by the reference's own Sampler,
against torch.multinomial itself, and its Philox stream against the Random123 known-answer vectors."""
import os

import numpy as np
import pytest
import torch

from oracle import sampler_oracle as so
from tests.conftest import ROOT

CASES = ["bench_defaults", "top_k", "top_p", "min_p", "all"]


@pytest.fixture(scope="module")
def sgold():
    return np.load(os.path.join(ROOT, "tests", "golden", "sampler_golden.npz"))


def load_case(g, name):
    logits = torch.from_numpy(g[f"{name}.logits_bf16"].astype(np.int16)).view(torch.bfloat16)
    t = lambda k: torch.from_numpy(g[f"{name}.{k}"])
    return logits, t("temperatures"), t("min_ps"), t("top_ps"), t("top_ks"), t("q"), t("ids")


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_reference_sampler(sgold, name):
    logits, temps, min_ps, top_ps, top_ks, q, ids = load_case(sgold, name)
    mine, scores = so.sample(logits, temps, min_ps.clone(), top_ps, top_ks, q)
    assert torch.equal(mine, ids)
    assert torch.isfinite(scores.max(dim=-1).values).all()


def test_oracle_greedy_branch(sgold):
    logits = torch.from_numpy(sgold["greedy.logits_bf16"].astype(np.int16)).view(torch.bfloat16)
    ids, _ = so.sample(logits)
    assert torch.equal(ids, torch.from_numpy(sgold["greedy.ids"]))


def test_multinomial_is_the_exponential_race():
    """The identity the explicit-noise formulation rests on: torch.multinomial(p, 1) ==
    argmax(p / q), q the Exp(1) tensor drawn first from the same generator state."""
    g = torch.Generator().manual_seed(5)
    p = torch.softmax(torch.randn(16, 3000, generator=g) * 3, -1)
    torch.manual_seed(77)
    a = torch.multinomial(p, 1).squeeze(-1)
    torch.manual_seed(77)
    q = torch.empty_like(p).exponential_(1)
    assert torch.equal(a, torch.argmax(p / q, -1))


def test_philox_known_answers():
    kat = [([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
           ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
           ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
            [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1])]
    for c, k, want in kat:
        got = so.philox4x32_10(np.array(c, np.uint32), np.array(k, np.uint32))
        assert [int(x) for x in got] == want


def test_seed_noise_is_exponential():
    q = so.seed_noise(seed=123, offset=7, B=4, V=50001).numpy().astype(np.float64)
    assert q.shape == (4, 50001) and (q > 0).all() and np.isfinite(q).all()
    assert abs(q.mean() - 1.0) < 0.02 and abs(q.var() - 1.0) < 0.05
    assert not np.array_equal(q[0], q[1])
    q2 = so.seed_noise(seed=123, offset=8, B=4, V=50001).numpy()
    assert not np.array_equal(q[0].astype(np.float32), q2[0])
