#!/usr/bin/env python
"""ncu driver: one saved-state forward, then saved-state backward calls of one workload (cfg3 by default)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import qris_b200  # noqa: E402
from bench import WORKLOADS  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
kind, n, L, B, qnat, desc = WORKLOADS[wl]
circ = qris_b200.reference_circuit(n, L) if kind == "ref" else qris_b200.u3cu3_circuit(n, L)
torch.manual_seed(0)
x = (2 * torch.rand(B, n, device="cuda") - 1)
w = 2 * torch.pi * torch.rand(circ.n_weights, device="cuda")
gy = torch.randn(B, n, device="cuda")
_, state = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
for _ in range(2):
    qris_b200.qnn_backward_raw(circ, x, w, gy, state=state)
torch.cuda.synchronize()
print("ok")
