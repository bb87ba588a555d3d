"""Angle head of the scenario classifier as a PyTorch function (libqnn_b200.so: qnn_head_*).

The reference ends its ``preprocess`` stack with ``nn.Linear(32*(h/4)*(w/4), n_qubits)`` and ``nn.Tanh()``
(Estimators_QuantumNAT_onchipQNN.py:161-162); their output is the encoder angles of the circuit.
``linear_tanh(f, weight, bias)`` computes the same ``tanh(f @ weight.T + bias)`` in one kernel forward and
one (plus a small deterministic reduction) backward, instead of addmm + tanh and tanh_backward + two GEMMs +
a column sum.  ``estimators.QSC_P128`` calls it with the weights of its unchanged ``preprocess[7]`` module,
so the ``state_dict`` keys stay the reference's.
"""
from __future__ import annotations

import torch
from torch.autograd.function import once_differentiable

from . import _lib
from .qlayer import _ptr, _require_cuda, _workspace

MAX_OUTPUTS = 16          # limits of qnn_head_forward (include/qnn_b200.h)
MAX_FEATURES = 4096
MAX_WEIGHTS = 49152
# Below this batch size the step is host-bound and the stock modules' two launches cost the same as this
# wrapper's (measured on B200, profiles/experiments/head_rate.py: 196 vs 162 us per forward+backward at
# B = 256, 225 vs 240 us at B = 65,536), so the classifier only routes large batches here.
MIN_BATCH = 8192


def supported(n_features: int, n_outputs: int) -> bool:
    return 0 < n_outputs <= MAX_OUTPUTS and 0 < n_features <= MAX_FEATURES and n_features * n_outputs <= MAX_WEIGHTS


class _LinearTanh(torch.autograd.Function):
    @staticmethod
    def forward(ctx, f, weight, bias):
        _require_cuda(f, "head features")
        lib = _lib.lib()
        fc = f.detach().to(torch.float32).contiguous()
        wc = weight.detach().to(torch.float32).contiguous()
        bc = bias.detach().to(torch.float32).contiguous()
        B, K = fc.shape
        n = wc.shape[0]
        if wc.shape[1] != K or bc.numel() != n:
            raise _lib.QnnError(f"linear_tanh: features {tuple(fc.shape)}, weight {tuple(wc.shape)}, bias {tuple(bc.shape)} do not fit")
        x = torch.empty(B, n, dtype=torch.float32, device=fc.device)
        if B > 0:
            dev = fc.device.index
            with torch.cuda.device(dev):
                st = torch.cuda.current_stream(dev).cuda_stream
                _lib.check(lib.qnn_head_forward(_ptr(fc), _ptr(wc), _ptr(bc), _ptr(x), B, K, n, st), "qnn_head_forward")
        ctx.save_for_backward(fc, wc, x)
        ctx.need_gf = f.requires_grad
        return x

    @staticmethod
    @once_differentiable
    def backward(ctx, gx):
        lib = _lib.lib()
        fc, wc, x = ctx.saved_tensors
        B, K = fc.shape
        n = wc.shape[0]
        gxc = gx.detach().to(torch.float32).contiguous()
        gf = torch.empty_like(fc) if ctx.need_gf else None
        gw = torch.zeros_like(wc)
        gb = torch.zeros(n, dtype=torch.float32, device=fc.device)
        if B > 0:
            dev = fc.device.index
            with torch.cuda.device(dev):
                st = torch.cuda.current_stream(dev).cuda_stream
                ws = _workspace(lib.qnn_head_workspace_bytes(B, K, n), fc.device)
                _lib.check(lib.qnn_head_backward(_ptr(fc), _ptr(wc), _ptr(x), _ptr(gxc), _ptr(gf), _ptr(gw), _ptr(gb),
                                                 B, K, n, _ptr(ws), ws.numel(), st), "qnn_head_backward")
        return gf, gw, gb


def linear_tanh(f: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor) -> torch.Tensor:
    """``tanh(f @ weight.T + bias)`` on the GPU kernels; ``f [B, K]``, ``weight [n, K]``, ``bias [n]``."""
    return _LinearTanh.apply(f, weight, bias)
