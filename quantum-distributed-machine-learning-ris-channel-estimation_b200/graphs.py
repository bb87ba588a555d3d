"""Replay of a recorded step: ``GraphedCall`` records a function of CUDA tensors once into a CUDA graph and
replays it for every later call with inputs of the same shapes.

At the reference's sizes (256 to 4,096 samples of a 6-qubit circuit per call, Runner_P128_...py:341-352) a
forward + backward through the module surface is about twenty kernels of a few microseconds each: the host's
per-launch cost, not the GPU, bounds the step.  Recording removes that cost without changing a single kernel or
its arguments' meaning.  ``runner.GraphedQSCStep`` is the same idea specialised to the whole optimiser step.
"""
from __future__ import annotations

from typing import Callable, Sequence

import torch


class GraphedCall:
    """``GraphedCall(fn, example_inputs)``: ``fn(*tensors) -> tensor or tuple of tensors``.

    The first ``warmup`` calls run eagerly on a side stream (real calls: allocator and library warm-up), the
    next one is recorded, every later one copies its inputs into the recording's fixed input buffers and replays.
    Returned tensors are the recording's fixed output buffers: read or copy them before the next call.  ``fn``
    must launch the same kernels every time (no data-dependent Python control flow, no host synchronisation) and
    may use torch's random generator (its state is handled by the recording).  Inputs of another shape raise.
    """

    def __init__(self, fn: Callable, example_inputs: Sequence[torch.Tensor], warmup: int = 3):
        self.fn, self.warmup, self.calls = fn, warmup, 0
        self.static = [t.detach().clone().requires_grad_(t.requires_grad) for t in example_inputs]
        self.graph = None
        self.out = None
        self.stream = torch.cuda.Stream(device=self.static[0].device)

    def __call__(self, *inputs: torch.Tensor):
        if len(inputs) != len(self.static) or any(i.shape != s.shape or i.dtype != s.dtype for i, s in zip(inputs, self.static)):
            raise ValueError("GraphedCall: inputs differ in number, shape or dtype from the recorded ones")
        with torch.no_grad():
            for s, i in zip(self.static, inputs):
                s.copy_(i, non_blocking=True)
        self.calls += 1
        if self.graph is not None:
            self.graph.replay()
            return self.out
        cur = torch.cuda.current_stream()
        self.stream.wait_stream(cur)
        if self.calls <= self.warmup:
            with torch.cuda.stream(self.stream):
                out = self.fn(*self.static)
            cur.wait_stream(self.stream)
            return out
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=self.stream):
            self.out = self.fn(*self.static)
        self.graph.replay()              # recording does not execute: run what was recorded
        return self.out
