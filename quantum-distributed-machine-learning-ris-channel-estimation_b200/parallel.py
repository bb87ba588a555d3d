"""Data-parallel plumbing: one process per GPU, batch sharded across ranks, ONE all-reduce per step.

The reference never runs the QNN on more than one device (Runner_P128_...py:313-316) and uses
``torch.nn.DataParallel`` only for the classical estimators (Run:144-148).  Here every rank takes a
contiguous slice of the batch; the only cross-sample quantity of the hot path, the parameter gradient
(a few KB), lives in one flat fp32 bucket that is all-reduced once per step over NCCL/NVLink
(``gloo`` on CPU for the host-logic tests).  No statevector is ever split across GPUs.
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Rank r of N takes rows [r*ceil(n/N), min((r+1)*ceil(n/N), n)) -- contiguous, ragged tail last."""
    per = -(-n_items // world)
    lo = min(rank * per, n_items)
    return lo, min(lo + per, n_items)


class FlatGradBucket:
    """All parameter gradients as views into one contiguous fp32 buffer.

    ``p.grad`` of every parameter is pointed at its slice, so autograd accumulates straight into the
    bucket and a step needs exactly one collective.  ``zero()`` replaces ``optimizer.zero_grad()``.
    """

    def __init__(self, params: Iterable[torch.nn.Parameter]):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("no trainable parameters")
        dev, total = self.params[0].device, sum(p.numel() for p in self.params)
        # one spare element behind the gradients carries this rank's row count through the same all-reduce
        self._buf = torch.zeros(total + 1, dtype=torch.float32, device=dev)
        self.flat = self._buf[:total]
        off = 0
        self.slices = []
        for p in self.params:
            if p.dtype != torch.float32 or p.device != dev:
                raise ValueError("FlatGradBucket needs float32 parameters on one device")
            view = self.flat[off:off + p.numel()].view_as(p)
            p.grad = view
            self.slices.append((off, off + p.numel()))
            off += p.numel()

    def zero(self) -> None:
        self.flat.zero_()
        for p, (a, b) in zip(self.params, self.slices):   # re-attach if something replaced .grad
            if p.grad is None or p.grad.data_ptr() != self.flat[a:b].data_ptr():
                p.grad = self.flat[a:b].view_as(p)

    def all_reduce(self, group: Optional[dist.ProcessGroup] = None, average: bool = True,
                   local_rows: Optional[int] = None) -> None:
        """Sum over ranks (then / world if ``average``): with equal shard sizes and per-shard mean
        losses this reproduces the single-process gradient of the global-batch mean loss.

        ``local_rows``: this rank's number of samples when shards are NOT equal (ragged tail of
        ``shard_range``, last batch of an epoch).  Each rank's mean-loss gradient is then weighted by its
        row count and the sum divided by the global count, which travels in the same collective (no host
        round trip): the result is the gradient of the mean over the global batch whatever the split.
        That holds for models whose samples are independent.  Where the forward couples samples of different
        ranks (``QuantumNATHead(process_group=...)``: shared batch statistics), scale the loss by the global
        row count on every rank instead (``sum / B_global``) and call ``all_reduce(average=False)``."""
        if not (dist.is_available() and dist.is_initialized()):
            return
        world = dist.get_world_size(group)
        if world == 1:
            return
        if self.flat.is_cuda:
            with torch.cuda.nvtx.range("allreduce/flat-gradient-bucket"):
                self._all_reduce(group, average, local_rows, world)
        else:
            self._all_reduce(group, average, local_rows, world)

    def _all_reduce(self, group, average, local_rows, world) -> None:
        if local_rows is not None and average:
            self.flat.mul_(float(local_rows))
            self._buf[-1] = float(local_rows)
            dist.all_reduce(self._buf, op=dist.ReduceOp.SUM, group=group)
            self.flat.div_(self._buf[-1])
            return
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
        if average:
            self.flat.mul_(1.0 / world)
