"""GPU: the fused q/k RMSNorm + RoPE + slot write kernel through the C ABI, against the oracle and
against outputs of the reference's own RMSNorm / RotaryEmbedding modules (tests/golden/qk_norm_rope_golden.npz)."""
import numpy as np
import pytest
import torch

from oracle import paged_attention_oracle as orc
from tests.test_qknorm_oracle import CASES, load_case, ulp_diff

pytestmark = pytest.mark.gpu

EPS = 1e-6


def _bits(t):
    return t.detach().cpu().contiguous().view(torch.int16)


def _cos_sin_cache(max_pos, D=128, base=1e6):
    """layers/rotary_embedding.py:53-71 (fp32 table, [cos | sin])."""
    inv = 1.0 / (base ** (torch.arange(0, D, 2, dtype=torch.float) / D))
    fr = torch.einsum("i,j -> ij", torch.arange(max_pos, dtype=torch.float), inv)
    return torch.cat((fr.cos(), fr.sin()), dim=-1)


def _check_rotated(mine, ref, T, D=128):
    """The kernel and the oracle differ only through the reduction order of mean(x^2) and rsqrt's last ulp: a flip of
    one bf16 rounding of the normalised value, bounded like the golden test (2^-7 of the rotation pair norm)."""
    d = ulp_diff(mine, ref)
    a = (mine.float() - ref.float()).abs().reshape(T, -1, D // 2, 2)
    m = ref.float().reshape(T, -1, D // 2, 2).pow(2).sum(-1, keepdim=True).sqrt()
    assert bool((a <= m * 2.0 ** -7 + 1e-30).all())
    assert float((d == 0).float().mean()) > 0.98, float((d == 0).float().mean())


@pytest.mark.parametrize("T,Hq,Hkv", [(1, 16, 8), (37, 16, 8), (300, 32, 8), (19, 4, 1), (64, 8, 2), (257, 16, 4),
                                      (5000, 32, 8)])
@pytest.mark.parametrize("in_place", [True, False])
def test_fused_vs_oracle(pkg, T, Hq, Hkv, in_place):
    D, max_pos, pool = 128, 4096, T + 50
    g = torch.Generator().manual_seed(T * 131 + Hq)
    qkv = (torch.randn((T, (Hq + 2 * Hkv) * D), generator=g) * 2.0).to(torch.bfloat16)
    qw = (1.0 + 0.25 * torch.randn(D, generator=g)).to(torch.bfloat16)
    kw = (1.0 + 0.25 * torch.randn(D, generator=g)).to(torch.bfloat16)
    pos = torch.randint(0, max_pos, (T,), generator=g)
    cache = _cos_sin_cache(max_pos, D)
    slots = torch.from_numpy(np.random.default_rng(T).permutation(pool)[:T].astype(np.int64))
    slots[::7] = -1
    init = torch.randn((pool, Hkv, D), generator=g).to(torch.bfloat16)
    kc_ref, vc_ref = init.clone(), init.clone()
    q_ref, k_ref = orc.qk_norm_rope_store(qkv, pos, qw, kw, EPS, cache, kc_ref, vc_ref, slots, Hq, Hkv, D)

    qkv_d = qkv.cuda()
    kc, vc = init.clone().cuda(), init.clone().cuda()
    q_out = qkv_d[:, : Hq * D] if in_place else torch.empty((T, Hq * D), dtype=torch.bfloat16, device="cuda")
    k_out = None if in_place else torch.empty((T, Hkv * D), dtype=torch.bfloat16, device="cuda")
    pkg.ops.qk_norm_rope_store(qkv_d, pos.cuda(), qw.cuda(), kw.cuda(), EPS, cache.cuda(), q_out, k_out, kc, vc,
                               slots.cuda(), Hq, Hkv, D)
    torch.cuda.synchronize()
    _check_rotated(q_out.cpu(), q_ref, T)
    if k_out is not None:
        _check_rotated(k_out.cpu(), k_ref, T)
    # V rows: bit-exact copies; skipped slots and untouched pool rows keep their old bits
    assert torch.equal(_bits(vc), _bits(vc_ref))
    live = slots >= 0
    if int(live.sum()):
        _check_rotated(kc.cpu()[slots[live]].reshape(int(live.sum()), -1), kc_ref[slots[live]].reshape(int(live.sum()), -1), int(live.sum()))
    untouched = torch.ones(pool, dtype=torch.bool)
    untouched[slots[live]] = False
    assert torch.equal(_bits(kc.cpu()[untouched]), _bits(init[untouched]))
    if not in_place:      # the input buffer is not modified when q_out is a separate buffer
        assert torch.equal(_bits(qkv_d), _bits(qkv))
    else:                 # in place: only the q columns change
        assert torch.equal(_bits(qkv_d[:, Hq * D:]), _bits(qkv[:, Hq * D:]))


@pytest.mark.parametrize("name", CASES)
def test_fused_vs_reference_modules_golden(pkg, name):
    c = load_case(name)
    T, Hq, Hkv, D = c["T"], c["Hq"], c["Hkv"], c["D"]
    kc = torch.zeros((64, Hkv, D), dtype=torch.bfloat16, device="cuda")
    vc = torch.zeros_like(kc)
    slots = torch.arange(T, dtype=torch.int64)
    q_out = torch.empty((T, Hq * D), dtype=torch.bfloat16, device="cuda")
    k_out = torch.empty((T, Hkv * D), dtype=torch.bfloat16, device="cuda")
    pkg.ops.qk_norm_rope_store(c["qkv"].cuda(), c["pos"].cuda(), c["q_weight"].cuda(), c["k_weight"].cuda(), c["eps"],
                               c["cache"].cuda(), q_out, k_out, kc, vc, slots.cuda(), Hq, Hkv, D)
    torch.cuda.synchronize()
    # same bound as the oracle's own golden test: the reference modules ran eagerly on CPU (three bf16 roundings)
    for mine, ref in ((q_out.cpu(), c["q_out"]), (k_out.cpu(), c["k_out"])):
        a = (mine.float() - ref.float()).abs().reshape(T, -1, D // 2, 2)
        m = ref.float().reshape(T, -1, D // 2, 2).pow(2).sum(-1, keepdim=True).sqrt()
        assert bool((a <= m * 2.0 ** -7 + 1e-30).all())
        assert float((ulp_diff(mine, ref) == 0).float().mean()) > 0.70
    assert torch.equal(_bits(kc[:T].reshape(T, -1)), _bits(k_out))
    assert torch.equal(_bits(vc[:T].reshape(T, -1)), _bits(c["qkv"][:, (Hq + Hkv) * D:]))


def test_fused_path_feeds_attention(pkg):
    """Attention.forward_fused_qkv == q_norm/k_norm/rotary_emb/Attention.forward of the reference (models/qwen3.py:
    123-134) on one decode step: the pool rows and the attention output agree with the oracle pipeline."""
    from tests.util import check_close
    Hq, Hkv, D, B, L = 16, 8, 128, 5, 70
    g = torch.Generator().manual_seed(7)
    pool = B * L + 33
    cache = _cos_sin_cache(4096, D)
    kc0 = torch.randn((1, pool, Hkv, D), generator=g).to(torch.bfloat16)
    vc0 = torch.randn((1, pool, Hkv, D), generator=g).to(torch.bfloat16)
    perm = np.random.default_rng(3).permutation(pool)[: B * L].astype(np.int64).reshape(B, L)
    qkv = (torch.randn((B, (Hq + 2 * Hkv) * D), generator=g)).to(torch.bfloat16)
    qw = (1.0 + 0.1 * torch.randn(D, generator=g)).to(torch.bfloat16)
    kw = (1.0 + 0.1 * torch.randn(D, generator=g)).to(torch.bfloat16)
    pos = torch.full((B,), L - 1, dtype=torch.int64)
    new_slots = torch.from_numpy(perm[:, -1].copy())
    # oracle pipeline
    kc_ref, vc_ref = kc0[0].clone(), vc0[0].clone()
    q_ref, _ = orc.qk_norm_rope_store(qkv, pos, qw, kw, EPS, cache, kc_ref, vc_ref, new_slots, Hq, Hkv, D)
    indptr = (np.arange(B + 1) * L).astype(np.int32)
    o_ref = orc.decode_attention(q_ref.view(B, Hq, D), kc_ref, vc_ref, indptr, perm.reshape(-1).astype(np.int32),
                                 np.ones(B, np.int32), 1)

    class Seq:
        def __init__(self, i):
            self.seq_id, self.token_ids, self.kv_indices, self.cached_kv_len = f"s{i}", [0] * L, perm[i].tolist(), L - 1

    class Batch:
        forward_mode = pkg.ForwardMode.DECODE
        seqs = [Seq(i) for i in range(B)]
        num_seqs = B

    kv = pkg.KVCachePool(torch.bfloat16, torch.device("cuda"), pool, 1, Hkv, D)
    kv.k_cache.copy_(kc0); kv.v_cache.copy_(vc0)
    backend = pkg.AttentionBackend(kv, Hq, Hkv, D, torch.bfloat16, torch.device("cuda"))
    attn = pkg.Attention(Hq, D, None, Hkv, 0)
    md = backend.build_metadata(Batch)
    attn.set_attention_metadata(md)
    o = attn.forward_fused_qkv(qkv.cuda(), pos.cuda(), qw.cuda(), kw.cuda(), EPS, cache.cuda())
    torch.cuda.synchronize()
    check_close(o.view(B, Hq, D), o_ref.view(B, Hq, D) if hasattr(o_ref, "view") else torch.as_tensor(o_ref), "fused->decode")
    assert torch.equal(_bits(kv.v_cache[0]), _bits(vc_ref))
