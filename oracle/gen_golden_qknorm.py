"""Generate tests/golden/qk_norm_rope_golden.npz by RUNNING THE REFERENCE'S OWN MODULES on CPU.

Run in the authoring container only (needs /root/reference):

    TORCHDYNAMO_DISABLE=1 python oracle/gen_golden_qknorm.py

Unmodified from /root/reference/nanovllmx: layers/layernorm.py RMSNorm, layers/rotary_embedding.py
RotaryEmbedding (both ``@torch.compile``; TORCHDYNAMO_DISABLE=1 runs their Python bodies eagerly), driven
exactly as models/qwen3.py:122-133 drives them (split, per-head view, q_norm / k_norm, rotary_emb).
The fixture stores inputs and the reference's (q, k) outputs as raw bf16 bit patterns (uint16).
"""
import os
import sys

os.environ.setdefault("TORCHDYNAMO_DISABLE", "1")
sys.path.insert(0, "/root/reference")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch


def bits(t):
    return t.contiguous().view(torch.int16).numpy().astype(np.uint16).copy()


def main():
    from nanovllmx.layers.layernorm import RMSNorm
    from nanovllmx.layers.rotary_embedding import RotaryEmbedding

    out = {}
    D, eps, max_pos, theta = 128, 1e-6, 4096, 1e6
    torch.set_default_dtype(torch.bfloat16)                  # core/model_runner.py:123
    rot = RotaryEmbedding(D, rotary_dim=D, max_position_embeddings=max_pos, base=theta)
    out["max_pos"] = np.array([max_pos], np.int64)
    assert rot.cos_sin_cache.dtype == torch.float32          # built from explicit dtype=torch.float aranges
    for name, (T, Hq, Hkv, seed) in {"tp1_06b": (37, 16, 8, 0), "tp8_8b": (19, 4, 1, 1), "tp1_8b": (5, 32, 8, 2)}.items():
        g = torch.Generator().manual_seed(seed)
        qn, kn = RMSNorm(D, eps=eps), RMSNorm(D, eps=eps)
        with torch.no_grad():
            qn.weight.copy_((1.0 + 0.25 * torch.randn(D, generator=g, dtype=torch.float32)).to(torch.bfloat16))
            kn.weight.copy_((1.0 + 0.25 * torch.randn(D, generator=g, dtype=torch.float32)).to(torch.bfloat16))
        qkv = (torch.randn((T, (Hq + 2 * Hkv) * D), generator=g, dtype=torch.float32) * 2.0).to(torch.bfloat16)
        pos = torch.randint(0, max_pos, (T,), generator=g)
        q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
        with torch.no_grad():
            qh = qn(q.view(T, Hq, D)).view(q.shape)          # models/qwen3.py:125-128
            kh = kn(k.view(T, Hkv, D)).view(k.shape)         # :129-132
            qo, ko = rot(pos, qh, kh)                        # :133
        out[f"{name}.shape"] = np.array([T, Hq, Hkv, D], np.int64)
        out[f"{name}.qkv"] = bits(qkv)
        out[f"{name}.pos"] = pos.numpy().astype(np.int64)
        out[f"{name}.cos_sin_rows"] = rot.cos_sin_cache[pos].float().numpy().copy()   # rows of the fp32 table actually used
        out[f"{name}.q_weight"] = bits(qn.weight.detach())
        out[f"{name}.k_weight"] = bits(kn.weight.detach())
        out[f"{name}.q_out"] = bits(qo)
        out[f"{name}.k_out"] = bits(ko)
    out["eps"] = np.array([eps], np.float64)
    dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "qk_norm_rope_golden.npz")
    np.savez_compressed(dst, **out)
    print("wrote", dst, {k: v.shape for k, v in out.items() if k.endswith("q_out")})


if __name__ == "__main__":
    main()
