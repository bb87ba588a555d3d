#!/usr/bin/env python
"""ncu driver: saved-state forward + two saved-state backward calls of the U3+CU3 circuit at n qubits.

    python profiles/experiments/prof_n.py n [L] [B] [variant: wide | narrow | 16-per-thread]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import qris_b200  # noqa: E402

n = int(sys.argv[1]); L = int(sys.argv[2]) if len(sys.argv) > 2 else 3
B = int(sys.argv[3]) if len(sys.argv) > 3 else max(148, (1 << 23) >> n)
circ = qris_b200.u3cu3_circuit(n, L)
if len(sys.argv) > 4:
    circ = circ.pinned(sys.argv[4])
torch.manual_seed(0)
x = (2 * torch.rand(B, n, device="cuda") - 1)
w = 2 * torch.pi * torch.rand(circ.n_weights, device="cuda")
gy = torch.randn(B, n, device="cuda")
ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
_, state = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
qris_b200.qnn_backward_raw(circ, x, w, gy, state=state)
ev[0].record(); _, state = qris_b200.qnn_forward_raw(circ, x, w, save_state=True); ev[1].record()
qris_b200.qnn_backward_raw(circ, x, w, gy, state=state); ev[2].record()
torch.cuda.synchronize()
print(f"n={n} L={L} B={B}: saved-state fwd {ev[0].elapsed_time(ev[1]):.3f} ms, bwd {ev[1].elapsed_time(ev[2]):.3f} ms")
