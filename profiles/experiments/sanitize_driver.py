#!/usr/bin/env python
"""Small forward + backward runs of every kernel geometry, for compute-sanitizer (one tool per gpurun call):

    compute-sanitizer --tool memcheck|racecheck|synccheck python profiles/experiments/sanitize_driver.py

n = 8 (a warp carries whole samples), n = 12 (a sample spans warps of one CTA: named barriers), n = 14 and 16
(a sample spans a thread-block cluster: cluster barriers, alternating exchange buffers), the amplitude encoder,
and the generic family; plain pair and saved-state pair; two calls each so that a second sample round reuses the
exchange buffers.  Results are checked for finiteness only (parity is the test-suite's job)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import qris_b200  # noqa: E402

if "--checked" in sys.argv:
    # the self-checking build (make -C .../csrc checked): every exchange address is checked on the device and a
    # violation prints "QNN_VM_CHECKS: violation ..." (grep the output for it)
    sys.argv.remove("--checked")
    qris_b200._lib.LIB_PATH = os.path.join(os.path.dirname(qris_b200._lib.LIB_PATH), "libqnn_b200_checked.so")
    print("using", qris_b200._lib.LIB_PATH, flush=True)

cases = [("u3cu3 n=8", qris_b200.u3cu3_circuit(8, 2), 9), ("ref n=6", qris_b200.reference_circuit(6, 3), 37),
         ("u3cu3 n=12", qris_b200.u3cu3_circuit(12, 2), 5), ("u3cu3 n=14", qris_b200.u3cu3_circuit(14, 2), 3),
         ("u3cu3 n=16", qris_b200.u3cu3_circuit(16, 2), 2), ("amplitude n=8", qris_b200.u3cu3_circuit(8, 1, encoder="amplitude"), 7),
         ("generic n=7", qris_b200.u3cu3_circuit(7, 1).generic(), 5), ("generic n=12", qris_b200.reference_circuit(12, 1).generic(), 2),
         ("narrow n=6", qris_b200.reference_circuit(6, 3).pinned("narrow"), 301), ("narrow n=3", qris_b200.u3cu3_circuit(3, 2), 70),
         ("narrow n=8", qris_b200.u3cu3_circuit(8, 3).pinned("narrow"), 37), ("wide n=6", qris_b200.reference_circuit(6, 3).pinned("wide"), 301),
         ("16-per-thread n=10", qris_b200.u3cu3_circuit(10, 2).pinned("16-per-thread"), 9),
         ("16-per-thread n=12", qris_b200.reference_circuit(12, 2).pinned("16-per-thread"), 3),
         ("ref n=10", qris_b200.reference_circuit(10, 3), 70), ("u3cu3 n=13", qris_b200.u3cu3_circuit(13, 3), 20),
         ("ref n=15", qris_b200.reference_circuit(15, 2), 40), ("u3cu3 n=16 L=8", qris_b200.u3cu3_circuit(16, 8), 40),
         ("ref n=16", qris_b200.reference_circuit(16, 3), 37), ("deep n=14 (L2 exchanges only)", qris_b200.u3cu3_circuit(14, 20), 3)]
only = sys.argv[1:] if len(sys.argv) > 1 else None
for name, circ, B in cases:
    if only and not any(o in name for o in only):
        continue
    torch.manual_seed(0)
    x = (2 * torch.rand(B, circ.n_inputs, device="cuda") - 1)
    w = 2 * torch.pi * torch.rand(circ.n_weights, device="cuda")
    gy = torch.randn(B, circ.n_qubits, device="cuda")
    for rep in range(2):
        y = qris_b200.qnn_forward_raw(circ, x, w)
        gx, gw = qris_b200.qnn_backward_raw(circ, x, w, gy)
        y2, st = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
        gx2, gw2 = qris_b200.qnn_backward_raw(circ, x, w, gy, state=st)
    torch.cuda.synchronize()
    ok = all(torch.isfinite(t).all().item() for t in (y, gx, gw, y2, gx2, gw2))
    print(f"{name}: B={B} plan={qris_b200._lib.lib().qnn_circuit_plan(circ.handle(0), B).decode()} finite={ok}", flush=True)
    assert ok
print("sanitize driver done")
