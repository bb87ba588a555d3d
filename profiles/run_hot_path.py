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
for _ in range(iters):
    y = qris_b200.qnn_forward_raw(circ, x, w)
    gx, gw = qris_b200.qnn_backward_raw(circ, x, w, gy)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
ev[0].record(); y = qris_b200.qnn_forward_raw(circ, x, w); ev[1].record()
gx, gw = qris_b200.qnn_backward_raw(circ, x, w, gy); ev[2].record()
torch.cuda.synchronize()
for _ in range(2):
    _, state = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
    qris_b200.qnn_backward_raw(circ, x, w, gy, state=state)
del state
torch.cuda.synchronize()
ev2 = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
ev2[0].record(); y2, state = qris_b200.qnn_forward_raw(circ, x, w, save_state=True); ev2[1].record()
gx2, gw2 = qris_b200.qnn_backward_raw(circ, x, w, gy, state=state); ev2[2].record()
torch.cuda.synchronize()
print(f"   with saved state: fwd {ev2[0].elapsed_time(ev2[1]):.3f} ms, bwd {ev2[1].elapsed_time(ev2[2]):.3f} ms, "
      f"max|dgw| vs recompute {float((gw2 - gw).abs().max()):.2e}, max|dgx| {float((gx2 - gx).abs().max()):.2e}")
print(f"{wl} B={B}: fwd {ev[0].elapsed_time(ev[1]):.3f} ms, bwd {ev[1].elapsed_time(ev[2]):.3f} ms, "
      f"plan {qris_b200._lib.lib().qnn_circuit_plan(circ.handle(0), B).decode()}")
