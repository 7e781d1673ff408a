"""Drop-in for nanovllmx/layers/sampler.py (reference :4-55): same class name and ``forward`` signature.

The reference divides by the temperature, SORTS the whole [B, V] logit matrix for top-k / top-p
(:58-90; core/model_runner.py:191-219 always sends top_k / top_p, so the sort runs on every step),
takes three softmaxes and calls ``torch.multinomial``.  Here all of it is one launch of
``b200attn::sample`` (csrc/sample.cu): thresholds by radix select, the draw as the same
``argmax(probs / q)`` race ``torch.multinomial`` runs.

Randomness: by default q comes from the kernel's own Philox stream, keyed by ``seed`` and a call
counter (no host sync, CUDA-graph safe when the counter is advanced by the caller).  With
``torch_noise=True`` q is drawn by ``torch.Tensor.exponential_`` from the given / default generator,
which is exactly the draw ``torch.multinomial`` would make — the samples then equal the reference's
for the same generator state.
"""
from __future__ import annotations

import torch
from torch import nn

from .. import ops


class Sampler(nn.Module):
    def __init__(self, seed: int = 0, torch_noise: bool = False, generator: torch.Generator | None = None):
        super().__init__()
        self.seed = seed
        self.calls = 0
        self.torch_noise = torch_noise
        self.generator = generator

    def forward(self, logits: torch.Tensor, temperatures: torch.Tensor | None = None,
                min_ps: torch.Tensor | None = None, top_ps: torch.Tensor | None = None,
                top_ks: torch.Tensor | None = None) -> torch.Tensor:
        if not logits.is_cuda:
            raise RuntimeError("b200attn Sampler runs on CUDA tensors only (there is no CPU fallback)")
        greedy = temperatures is None and min_ps is None and top_ps is None and top_ks is None   # reference :28-33
        noise = None
        if self.torch_noise and not greedy:
            noise = torch.empty(logits.shape, dtype=torch.float32, device=logits.device).exponential_(1, generator=self.generator)
        if logits.dtype not in (torch.bfloat16, torch.float32):
            logits = logits.float()
        if logits.stride(-1) != 1:
            logits = logits.contiguous()
        self.calls += 1
        return ops.sample(logits, temperatures, min_ps, top_ps, top_ks, noise, self.seed, self.calls)
