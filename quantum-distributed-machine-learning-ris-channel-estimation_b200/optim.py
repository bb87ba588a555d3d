"""Fused optimiser tail: gradient pruning + AdamW as one kernel over flat buffers.

Replaces ``model.apply_gradient_pruning()`` + ``optimizer.step()`` (Runner_P128_...py:367-369;
Estimators_...py:205-228) with ``qnn_pruned_adamw_step``: same arithmetic as ``torch.optim.AdamW``,
no per-parameter Python loop, no host synchronisation (the pruned count stays on the device until read).
"""
from __future__ import annotations

from typing import Iterable, Optional

import torch

from . import _lib
from .parallel import FlatGradBucket


class FlatPrunedAdamW:
    """AdamW(lr, betas, eps, weight_decay) with optional |g| <= threshold pruning, on flat storage.

    Parameters are re-pointed at slices of one flat fp32 buffer (the ``nn.Parameter`` objects, and so
    ``state_dict`` keys, are unchanged); gradients live in a ``FlatGradBucket`` so a data-parallel step
    is: ``zero_grad() -> backward -> step()`` with exactly one all-reduce and one optimiser kernel.
    """

    def __init__(self, params: Iterable[torch.nn.Parameter], lr=1e-3, betas=(0.9, 0.999), eps=1e-8,
                 weight_decay=0.01, prune_threshold: Optional[float] = None, process_group=None,
                 device_step_count: bool = False):
        self.params = [p for p in params if p.requires_grad]
        if not self.params or not self.params[0].is_cuda:
            raise _lib.QnnError("FlatPrunedAdamW needs CUDA parameters (its step is a CUDA kernel; no CPU fallback)")
        dev = self.params[0].device
        self.flat = torch.cat([p.detach().reshape(-1).to(torch.float32) for p in self.params]).contiguous()
        off = 0
        for p in self.params:
            p.data = self.flat[off:off + p.numel()].view_as(p)
            off += p.numel()
        self.bucket = FlatGradBucket(self.params)
        self.exp_avg = torch.zeros_like(self.flat)
        self.exp_avg_sq = torch.zeros_like(self.flat)
        self.pruned = torch.zeros(1, dtype=torch.int64, device=dev)
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.prune_threshold = prune_threshold
        self.process_group = process_group
        self.t = 0
        # step count on the device: no argument of step() changes between steps, so a whole training step can
        # be recorded into a CUDA graph and replayed (runner.GraphedQSCStep)
        self.steps_done = torch.zeros(1, dtype=torch.int64, device=dev) if device_step_count else None

    def zero_grad(self) -> None:
        self.bucket.zero()

    def step(self, local_rows: Optional[int] = None) -> None:
        """``local_rows``: this rank's sample count for the step when data-parallel shards may be unequal (the
        last batch of an epoch): the bucket then averages weighted by rows (``FlatGradBucket.all_reduce``)."""
        self.bucket.all_reduce(group=self.process_group, local_rows=local_rows)
        self.t += 1
        dev = self.flat.device.index
        thr = -1.0 if self.prune_threshold is None else float(self.prune_threshold)
        if self.steps_done is not None:
            with torch.cuda.device(dev):
                rc = _lib.lib().qnn_pruned_adamw_step_counted(
                    self.flat.data_ptr(), self.bucket.flat.data_ptr(), self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(),
                    self.flat.numel(), self.lr, self.betas[0], self.betas[1], self.eps, self.weight_decay,
                    self.steps_done.data_ptr(), thr, self.pruned.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
            _lib.check(rc, "qnn_pruned_adamw_step_counted")
            return
        with torch.cuda.device(dev):
            rc = _lib.lib().qnn_pruned_adamw_step(
                self.flat.data_ptr(), self.bucket.flat.data_ptr(), self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(),
                self.flat.numel(), self.lr, self.betas[0], self.betas[1], self.eps, self.weight_decay, self.t, thr,
                self.pruned.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(rc, "qnn_pruned_adamw_step")

    def pruned_fraction(self) -> float:
        """Pruned entries / all entries since construction (one device -> host read)."""
        steps = self.t if self.steps_done is None else int(self.steps_done.item())   # replays do not pass through step()
        return float(self.pruned.item()) / max(1, steps * self.flat.numel())
