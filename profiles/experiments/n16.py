"""Experiment: n = 11..16 on the register-resident VM (one CTA group up to n = 13, thread-block clusters
above) against the generic family (QNN_DISABLE_VM=1)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch, qris_b200
from bench import algorithmic_flops
for n, L, B in ((10, 3, 16384), (11, 3, 8192), (12, 3, 4096), (13, 3, 2048), (14, 3, 1184), (15, 3, 592), (16, 3, 296), (16, 8, 296)):
    for disable in ("0", "1"):
        os.environ["QNN_DISABLE_VM"] = disable
        c = qris_b200.u3cu3_circuit(n, L)
        x = 2 * torch.rand(B, n, device="cuda") - 1; w = 6.28 * torch.rand(c.n_weights, device="cuda"); gy = torch.randn(B, n, device="cuda")
        for _ in range(2):
            qris_b200.qnn_forward_raw(c, x, w); qris_b200.qnn_backward_raw(c, x, w, gy)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        ev[0].record(); qris_b200.qnn_forward_raw(c, x, w); ev[1].record(); qris_b200.qnn_backward_raw(c, x, w, gy); ev[2].record()
        _, state = qris_b200.qnn_forward_raw(c, x, w, save_state=True); ev[3].record()
        qris_b200.qnn_backward_raw(c, x, w, gy, state=state); ev[4].record()
        torch.cuda.synchronize()
        f, b = ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2])
        fs, bs = ev[2].elapsed_time(ev[3]), ev[3].elapsed_time(ev[4])
        ff, fb = algorithmic_flops([(o.kind, o.wires, o.params) for o in c.ops], n)
        plan = qris_b200._lib.lib().qnn_circuit_plan(c.handle(0), B).decode()
        pct = lambda t: (ff + fb) * B / (t * 1e-3) / 74.45e12 * 100
        print(f"n={n} L={L} B={B} plan={plan}: recompute fwd {f:.2f} + bwd {b:.2f} ms = {B / ((f + b) * 1e-3):.0f} samples/s, {pct(f + b):.1f}% | "
              f"saved state fwd {fs:.2f} + bwd {bs:.2f} ms = {B / ((fs + bs) * 1e-3):.0f} samples/s, {pct(fs + bs):.1f}% of FP32 roofline")
