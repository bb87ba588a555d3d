#!/usr/bin/env python
"""Minimal driver for ncu: a few forward + backward calls of one workload, nothing else.

    python profiles/run_hot_path.py [cfg3|cfg2|ref10] [iters] [batch]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import qris_b200  # noqa: E402
from bench import WORKLOADS  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
kind, n, L, B, qnat, desc = WORKLOADS[wl]
if len(sys.argv) > 3:
    B = int(sys.argv[3])
circ = qris_b200.reference_circuit(n, L) if kind == "ref" else qris_b200.u3cu3_circuit(n, L)
torch.manual_seed(0)
x = (2 * torch.rand(B, n, device="cuda") - 1)
w = 2 * torch.pi * torch.rand(circ.n_weights, device="cuda")
gy = torch.randn(B, n, device="cuda")
def timed(fn, reps=5):
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); out = fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts), out

for _ in range(iters):
    y = qris_b200.qnn_forward_raw(circ, x, w)
    gx, gw = qris_b200.qnn_backward_raw(circ, x, w, gy)
    _, state = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
    qris_b200.qnn_backward_raw(circ, x, w, gy, state=state)
torch.cuda.synchronize()
f0, _ = timed(lambda: qris_b200.qnn_forward_raw(circ, x, w))
b0, (gx, gw) = timed(lambda: qris_b200.qnn_backward_raw(circ, x, w, gy))
f1, (_, state) = timed(lambda: qris_b200.qnn_forward_raw(circ, x, w, save_state=True))
b1, (gx2, gw2) = timed(lambda: qris_b200.qnn_backward_raw(circ, x, w, gy, state=state))
plan = qris_b200._lib.lib().qnn_circuit_plan(circ.handle(0), B).decode()
print(f"{wl} B={B} plan={plan} (best of 5): recompute fwd {f0:.3f} + bwd {b0:.3f} ms | saved-state fwd {f1:.3f} + bwd {b1:.3f} = {f1 + b1:.3f} ms | "
      f"max|dgw| {float((gw2 - gw).abs().max()):.1e} max|dgx| {float((gx2 - gx).abs().max()):.1e}")
