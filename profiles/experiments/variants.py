#!/usr/bin/env python
"""Same-box A/B of the kernel variants (amplitudes per thread): saved-state forward + backward, CUDA events,
median of 20 calls after 5 warm-up calls.

    python profiles/experiments/variants.py
"""
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import qris_b200  # noqa: E402


def timed(circ, B, reps=20):
    torch.manual_seed(0)
    x = (2 * torch.rand(B, circ.n_inputs, device="cuda") - 1)
    w = 2 * torch.pi * torch.rand(circ.n_weights, device="cuda")
    gy = torch.randn(B, circ.n_qubits, device="cuda")
    f, b = [], []
    for it in range(reps + 5):
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record(); _, st = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
        e[1].record(); qris_b200.qnn_backward_raw(circ, x, w, gy, state=st)
        e[2].record(); torch.cuda.synchronize()
        if it >= 5:
            f.append(e[0].elapsed_time(e[1])); b.append(e[1].elapsed_time(e[2]))
    return statistics.median(f), statistics.median(b)


def graphed(circ, B, reps=50):
    """The same pair recorded once into a CUDA graph and replayed: device time without the host's per-call cost."""
    torch.manual_seed(0)
    x = (2 * torch.rand(B, circ.n_inputs, device="cuda") - 1)
    w = 2 * torch.pi * torch.rand(circ.n_weights, device="cuda")
    gy = torch.randn(B, circ.n_qubits, device="cuda")
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            _, st = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
            qris_b200.qnn_backward_raw(circ, x, w, gy, state=st)
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        _, st = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
        qris_b200.qnn_backward_raw(circ, x, w, gy, state=st)
    for _ in range(5):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def line(name, circ, B):
    plan = qris_b200._lib.lib().qnn_circuit_plan(circ.handle(0), B).decode()
    f, b = timed(circ, B)
    gms = graphed(circ, B)
    print(f"{name:34s} B={B:6d} {plan:20s} eager fwd {f * 1e3:8.1f} us  bwd {b * 1e3:8.1f} us  total {(f + b) * 1e3:8.1f} us "
          f"{B / (f + b) / 1e3:8.2f} M/s | graph replay {gms * 1e3:8.1f} us {B / gms / 1e3:9.2f} M samples/s", flush=True)


ref6 = qris_b200.reference_circuit(6, 3)
for B in (256, 1024, 2304, 4096, 16384, 65536):
    for v in ("wide", "narrow"):
        line(f"reference n=6 L=3 [{v}]", ref6.pinned(v), B)
for n in (4, 8):
    c = qris_b200.reference_circuit(n, 3)
    for B in (256, 4096, 65536):
        for v in (("narrow",) if n < 5 else ("wide", "narrow")):
            line(f"reference n={n} L=3 [{v}]", c.pinned(v), B)
cfg3 = qris_b200.u3cu3_circuit(10, 3)
for B in (8192, 65536):
    for v in ("wide", "16-per-thread"):
        line(f"U3+CU3 n=10 L=3 [{v}]", cfg3.pinned(v), B)
for n in (8, 9, 12):
    c = qris_b200.u3cu3_circuit(n, 3)
    B = (1 << 26) >> n
    for v in ("wide", "16-per-thread"):
        line(f"U3+CU3 n={n} L=3 [{v}]", c.pinned(v), B)
