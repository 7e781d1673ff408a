"""Per-batch-size CUDA graph registry with named static buffers.

Same surface as the reference's core/cuda_graph.py:5-56 (set/get_input_buffer,
set/get_output_buffer, capture(bs) context manager, replay(bs), match_bs(bs), clear()) so
``AttentionBackend.prepare_for_cuda_graph_capture`` & co. can be driven identically; the
reference's own CUDAGraph object works just as well (duck-typed on those methods).
"""
from __future__ import annotations

import bisect
from contextlib import contextmanager

import torch


class CUDAGraph:
    def __init__(self):
        self.graphs: dict[int, "torch.cuda.CUDAGraph"] = {}
        self.input_buffers: dict[str, torch.Tensor] = {}
        self.output_buffers: dict[str, torch.Tensor] = {}
        self.captured_bs: list[int] = []
        self.graph_pool = None
        self.max_bs = 0

    @property
    def is_captured(self) -> bool:
        return bool(self.graphs)

    def set_input_buffer(self, name: str, tensor: torch.Tensor) -> None:
        self.input_buffers[name] = tensor

    def get_input_buffer(self, name: str) -> torch.Tensor:
        return self.input_buffers[name]

    def set_output_buffer(self, name: str, tensor: torch.Tensor) -> None:
        self.output_buffers[name] = tensor

    def get_output_buffer(self, name: str) -> torch.Tensor:
        return self.output_buffers[name]

    @contextmanager
    def capture(self, bs: int):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, self.graph_pool):
            yield
        if self.graph_pool is None:
            self.graph_pool = g.pool()
        if bs not in self.graphs:
            bisect.insort(self.captured_bs, bs)
        self.graphs[bs] = g
        self.max_bs = max(self.max_bs, bs)

    def replay(self, bs: int) -> None:
        self.graphs[bs].replay()

    def match_bs(self, bs: int) -> int:
        assert self.graphs and bs <= self.max_bs
        return self.captured_bs[bisect.bisect_left(self.captured_bs, bs)]

    def clear(self) -> None:
        self.graphs.clear()
        self.input_buffers.clear()
        self.output_buffers.clear()
        self.captured_bs.clear()
        self.graph_pool = None
        self.max_bs = 0
