"""Shared builders for the parity tests (seeded, identical inputs for oracle and CUDA path)."""
import numpy as np
import torch

from oracle import paged_attention_oracle as orc

ATOL = 2e-2   # north_star: max abs error <= 2e-2 (bf16 outputs)
RTOL = 1e-2   # north_star: fp32-accumulate relative error <= 1e-2 (Frobenius, vs fp32 oracle)


def make_paged_case(kv_lens, Hq, Hkv, D=128, page_size=1, seed=0, q_lens=None, extra_slots=64,
                    dtype=torch.bfloat16, scale=1.0):
    """Random pool + fragmented page table.  Returns dict of CPU tensors / numpy arrays."""
    g = torch.Generator().manual_seed(seed)
    rng = np.random.default_rng(seed)
    B = len(kv_lens)
    pages = [(-(-L // page_size)) for L in kv_lens]
    n_pages = sum(pages)
    pool_pages = n_pages + extra_slots
    perm = rng.permutation(pool_pages)[:n_pages].astype(np.int32)
    kv_indptr = np.zeros(B + 1, np.int32)
    kv_indptr[1:] = np.cumsum(pages)
    last = np.array([(L % page_size) or page_size for L in kv_lens], np.int32)
    num_slots = pool_pages * page_size
    k_cache = (torch.randn((num_slots, Hkv, D), generator=g) * scale).to(dtype)
    v_cache = (torch.randn((num_slots, Hkv, D), generator=g) * scale).to(dtype)
    if q_lens is None:
        q_lens = [1] * B
    qo_indptr = np.zeros(B + 1, np.int32)
    qo_indptr[1:] = np.cumsum(q_lens)
    q = (torch.randn((int(qo_indptr[-1]), Hq, D), generator=g) * scale).to(dtype)
    return dict(q=q, k_cache=k_cache, v_cache=v_cache, qo_indptr=qo_indptr, kv_indptr=kv_indptr,
                kv_indices=perm, last_page_len=last, page_size=page_size, B=B, Hq=Hq, Hkv=Hkv, D=D)


def token_slots(case, i):
    return orc._token_slots(case["kv_indptr"], case["kv_indices"], case["last_page_len"], i, case["page_size"])


def check_close(out_bf16: torch.Tensor, ref_fp32: torch.Tensor, what=""):
    out = out_bf16.detach().float().cpu()
    ref = ref_fp32.detach().float().cpu()
    assert out.shape == ref.shape, (out.shape, ref.shape)
    assert torch.isfinite(out).all(), f"{what}: non-finite output"
    max_abs = (out - ref).abs().max().item() if out.numel() else 0.0
    denom = ref.norm().item()
    rel = (out - ref).norm().item() / denom if denom > 0 else 0.0
    assert max_abs <= ATOL, f"{what}: max abs err {max_abs:.4g} > {ATOL}"
    assert rel <= RTOL, f"{what}: rel err {rel:.4g} > {RTOL}"
    return max_abs, rel


def dev(case, device="cuda"):
    out = {}
    for k, v in case.items():
        if isinstance(v, torch.Tensor):
            out[k] = v.to(device)
        elif isinstance(v, np.ndarray):
            out[k] = torch.from_numpy(v).to(device)
        else:
            out[k] = v
    return out
