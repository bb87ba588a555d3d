"""Experiment: time per micro-op kind of the VM (cycles per warp-batch), from synthetic circuits."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, qris_b200, collections
Op = qris_b200.Op
n, B = 10, 65536
names={0:'END',1:'ROT',2:'ROTX',4:'CUL',5:'CUP',6:'DIAG',7:'DIAGR',8:'XCHG',9:'ROT5',10:'ROTX5',11:'CHAIN'}
def build(kind, reps):
    ops, nw = [], 0
    def w():
        nonlocal nw; nw += 1; return ("w", nw - 1)
    for r in range(reps):
        if kind == "rot5":
            ops += [Op("RY", (q,), (w(),)) for q in range(5)]
        elif kind == "chain":
            ops += [Op("CU3", (q, q + 1), (w(), w(), w())) for q in range(4)]
        elif kind == "cup":
            ops += [Op("CU3", (9, 0), (w(), w(), w()))]
        elif kind == "diag+rot":
            ops += [Op("RZ", (q,), (w(),)) for q in range(10)] + [Op("RY", (0,), (w(),))]
        elif kind == "diagreg+rot":
            ops += [Op("RZ", (q,), (w(),)) for q in range(5)] + [Op("RY", (0,), (w(),))]
        elif kind == "xchg+2rot":
            ops += [Op("RY", (0,), (w(),)), Op("RY", (5,), (w(),))]
        elif kind == "xchg+rot5":
            ops += [Op("RY", (q + 5 * (r % 2),), (w(),)) for q in range(5)]
        elif kind == "cup9":
            ops += [Op("RY", (q,), (w(),)) for q in range(5)] + [Op("CU3", (9, 0), (w(), w(), w()))]
        elif kind == "rot":
            ops += [Op("RY", (0,), (w(),))]
    return qris_b200.Circuit(n, ops, n_inputs=n, n_weights=nw)
x = torch.zeros(B, n, device="cuda"); gy = torch.randn(B, n, device="cuda")
for kind, reps in (("rot", 100), ("rot5", 40), ("chain", 40), ("cup", 60), ("diag+rot", 40), ("diagreg+rot", 40), ("xchg+rot5", 40), ("cup9", 40)):
    c = build(kind, reps); w = 6.28 * torch.rand(c.n_weights, device="cuda")
    d = c.describe(); cnt = collections.Counter(names[o[0]] for o in d["ops"])
    for _ in range(2):
        qris_b200.qnn_forward_raw(c, x, w); qris_b200.qnn_backward_raw(c, x, w, gy)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    ev[0].record(); qris_b200.qnn_forward_raw(c, x, w); ev[1].record(); qris_b200.qnn_backward_raw(c, x, w, gy); ev[2].record()
    torch.cuda.synchronize()
    f, b = ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2])
    # cycles per SMSP per (warp-batch, repetition): time * clock / (warp-batches per SMSP) / reps
    per = lambda ms: ms * 1e-3 * 1.965e9 / (B / (148 * 4)) / reps
    print(f"{kind:12s} reps={reps:3d} ops={dict(cnt)}  fwd {f:.3f} ms = {per(f):6.0f} cyc/rep   bwd-minus-fwd {b - f:.3f} ms = {per(b - f):6.0f} cyc/rep")
