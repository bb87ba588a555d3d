#!/usr/bin/env python
"""Fused angle head (qnn_head_*) against the stock Linear + Tanh modules, forward + backward, whole
preprocess tail only (features precomputed)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import qris_b200  # noqa: E402

lin = torch.nn.Linear(256, 6).cuda()
for B in (256, 2304, 65536):
    f = torch.randn(B, 256, device="cuda", requires_grad=True)
    go = torch.randn(B, 6, device="cuda")

    def fused():
        x = qris_b200.linear_tanh(f, lin.weight, lin.bias)
        return torch.autograd.grad((x * go).sum(), [f, lin.weight, lin.bias])

    def stock():
        x = torch.tanh(lin(f))
        return torch.autograd.grad((x * go).sum(), [f, lin.weight, lin.bias])

    for name, fn in (("fused", fused), ("stock", stock), ("fused", fused), ("stock", stock)):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            fn()
        e1.record(); torch.cuda.synchronize()
        print(f"B={B} {name}: {e0.elapsed_time(e1) / 50 * 1e3:.1f} us per forward+backward")
