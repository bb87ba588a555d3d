"""QuantumNAT post-measurement pipeline as a PyTorch module (libqnn_b200.so: qnat_*).

Not in the reference (its "QuantumNAT" is weight noise, Estimators_QuantumNAT_onchipQNN.py:176-196,
which ``estimators.QSC_P128`` keeps); BASELINE.json's north_star asks for normalisation, noise
injection and quantisation of the measured expectation values, defined in SURVEY.md section 5.9 and
fixed by ``oracle/qnat_oracle.py``.  Noise is drawn by torch (caller's generator); kernels own no RNG.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch
from torch import nn
from torch.autograd.function import once_differentiable

from . import _lib
from .qlayer import _ptr, _require_cuda, _workspace


def _params(sigma, lo, hi, eps, levels, normalise, quantise):
    return _lib.QnatParams(float(sigma), float(lo), float(hi), float(eps), int(levels),
                           int(bool(normalise)), int(bool(quantise)), 0)


class _QnatFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y, noise, cfg, group):
        _require_cuda(y, "QuantumNAT input")
        lib = _lib.lib()
        yc = y.detach().to(torch.float32).contiguous()
        B, n = yc.shape
        p = _params(cfg["sigma"] if noise is not None else 0.0, cfg["lo"], cfg["hi"], cfg["eps"],
                    cfg["levels"], cfg["normalise"], cfg["quantise"])
        dev = yc.device.index
        out = torch.empty_like(yc)
        yn = torch.empty_like(yc)
        stats = torch.empty(2 * n, dtype=torch.float32, device=yc.device)
        # [sum y (n), sum y^2 (n), row count]: the count rides along with the sums through the all-reduce, so
        # ranks with unequal shard sizes (ragged tail of parallel.shard_range, last batch of an epoch) all
        # normalise with the true global count -- read on the device (B_total = 0), no host round trip
        sums = torch.zeros(2 * n + 1, dtype=torch.float64, device=yc.device)
        B_total = B
        nz = None if noise is None else noise.detach().to(torch.float32).contiguous()
        with torch.cuda.device(dev), torch.cuda.nvtx.range("quantumnat/forward"):
            st = torch.cuda.current_stream(dev).cuda_stream
            if cfg["normalise"]:
                ws = _workspace(lib.qnat_workspace_bytes(B, n), yc.device)
                _lib.check(lib.qnat_batch_sums(_ptr(yc), _ptr(sums), B, n, _ptr(ws), ws.numel(), st), "qnat_batch_sums")
                if group is not None:
                    import torch.distributed as dist
                    sums[2 * n] = float(B)
                    dist.all_reduce(sums, group=group)
                    B_total = 0
            _lib.check(lib.qnat_apply(_ptr(yc), _ptr(nz), _ptr(sums), B_total, _ptr(out), _ptr(yn), _ptr(stats),
                                      B, n, C.byref(p), st), "qnat_apply")
        ctx.save_for_backward(yc, yn, stats, sums)
        ctx.p, ctx.group, ctx.B_total = p, group, B_total
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, gout):
        lib = _lib.lib()
        yc, yn, stats, sums = ctx.saved_tensors
        B, n = yc.shape
        g = gout.detach().to(torch.float32).contiguous()
        gy = torch.empty_like(yc)
        bsums = torch.zeros(2 * n + 1, dtype=torch.float64, device=yc.device)
        dev = yc.device.index
        p = ctx.p
        with torch.cuda.device(dev), torch.cuda.nvtx.range("quantumnat/backward"):
            st = torch.cuda.current_stream(dev).cuda_stream
            if p.normalise:
                ws = _workspace(lib.qnat_workspace_bytes(B, n), yc.device)
                _lib.check(lib.qnat_backward_sums(_ptr(g), _ptr(yc), _ptr(yn), _ptr(stats), _ptr(bsums), B, n,
                                                  C.byref(p), _ptr(ws), ws.numel(), st), "qnat_backward_sums")
                if ctx.group is not None:
                    import torch.distributed as dist
                    dist.all_reduce(bsums, group=ctx.group)
                    bsums[2 * n] = sums[2 * n]          # the global row count of the forward pass
            _lib.check(lib.qnat_backward_apply(_ptr(g), _ptr(yc), _ptr(yn), _ptr(stats), _ptr(bsums), ctx.B_total,
                                               _ptr(gy), B, n, C.byref(p), st), "qnat_backward_apply")
        return gy, None, None, None


class QuantumNATHead(nn.Module):
    """normalise -> (+ sigma * noise while training) -> quantise, on (B, n_qubits) expectation values.

    Batch statistics are taken over the batch of the call, in training and in evaluation alike.  With
    ``process_group`` set (data-parallel training) the 2n sums are all-reduced so every rank normalises
    with the statistics of the global batch, and the backward sums likewise.
    """

    def __init__(self, sigma: float = 0.01, lo: float = -2.0, hi: float = 2.0, levels: int = 5,
                 normalise: bool = True, quantise: bool = True, eps: float = 1e-10, process_group=None):
        super().__init__()
        self.cfg = dict(sigma=sigma, lo=lo, hi=hi, levels=levels, normalise=normalise, quantise=quantise, eps=eps)
        self.process_group = process_group

    def forward(self, y: torch.Tensor, noise: Optional[torch.Tensor] = None) -> torch.Tensor:
        if noise is None and self.training and self.cfg["sigma"] != 0.0:
            noise = torch.randn_like(y)
        if not self.training:
            noise = None
        return _QnatFunction.apply(y, noise, self.cfg, self.process_group)

    def extra_repr(self) -> str:
        return ", ".join(f"{k}={v}" for k, v in self.cfg.items())
