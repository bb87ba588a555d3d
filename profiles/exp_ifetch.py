"""Experiment: is the VM limited by instruction fetch?  Same work, one body vs five bodies."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, qris_b200
Op = qris_b200.Op
n, B, reps = 10, 65536, 200
def circ(wires):
    ops = [Op("RY", (wires[i % len(wires)],), (("w", i),)) for i in range(reps)]
    return qris_b200.Circuit(n, ops, n_inputs=n, n_weights=reps)
x = torch.zeros(B, n, device="cuda"); gy = torch.randn(B, n, device="cuda")
for name, wires in (("one body", [0]), ("five bodies", [0, 1, 2, 3, 4])):
    c = circ(wires); w = torch.rand(reps, device="cuda")
    for _ in range(2):
        qris_b200.qnn_forward_raw(c, x, w); qris_b200.qnn_backward_raw(c, x, w, gy)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    ev[0].record(); qris_b200.qnn_forward_raw(c, x, w); ev[1].record(); qris_b200.qnn_backward_raw(c, x, w, gy); ev[2].record()
    torch.cuda.synchronize()
    f, b = ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2])
    # REAL fwd: 4 instr/amp -> 128 per lane per op; 65536 warps-samples
    fi = reps * 128 * B; bi = reps * (128 + 320) * B
    print(f"{name}: fwd {f:.3f} ms ({fi / (f * 1e-3) / (148 * 4 * 1.965e9) * 100:.0f}% issue), bwd {b:.3f} ms ({bi / (b * 1e-3) / (148 * 4 * 1.965e9) * 100:.0f}% issue)")
