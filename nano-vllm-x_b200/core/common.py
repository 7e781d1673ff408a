"""The data types the attention path reads, mirroring core/common.py:44-65 of the reference.

Inside the reference engine the reference's own ``Sequence`` / ``ForwardBatch`` /
``ForwardMode`` objects are passed in; the host code here is duck-typed on the field names
(``token_ids``, ``kv_indices``, ``cached_kv_len``, ``seq_id``; ``forward_mode``, ``num_seqs``,
``seqs``) and compares modes by ``.name`` so either set of classes works.
"""
from __future__ import annotations

import enum
from dataclasses import dataclass, field


class ForwardMode(enum.Enum):
    PREFILL = enum.auto()
    DECODE = enum.auto()


@dataclass
class Sequence:
    seq_id: str
    token_ids: list = field(default_factory=list)
    kv_indices: list = field(default_factory=list)
    cached_kv_len: int = 0


@dataclass
class ForwardBatch:
    forward_mode: ForwardMode
    num_seqs: int
    seqs: list


def mode_name(mode) -> str:
    return getattr(mode, "name", str(mode))
