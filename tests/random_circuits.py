"""Seeded random circuits over the whole gate set (test infrastructure): irregular gate lists with re-used input
columns, shared weights, constants, CNOT / CZ / CU3 between arbitrary wire pairs."""
import numpy as np

import qris_b200


def random_circuit(n, n_gates, seed, n_weights=None, vm_only=True):
    """vm_only: keep to what the register-resident VM takes (input-sourced angles only on RY; else RX/RZ/U3/CU3
    angles may also come from the input row, which the generic family handles)."""
    rng = np.random.default_rng(seed)
    Op = qris_b200.Op
    n_weights = n_weights or max(4, n_gates // 2)
    ops = [Op("RY", (i,), (("x", i),)) for i in range(n)]                 # angle embedding first, as the reference does

    def src(allow_x):
        r = rng.random()
        if allow_x and r < 0.2:
            return ("x", int(rng.integers(n)))
        if r < 0.85:
            return ("w", int(rng.integers(n_weights)))
        return ("c", float(rng.uniform(-3, 3)))
    for _ in range(n_gates):
        kind = rng.choice(["RY", "RZ", "RX", "U3", "CNOT", "CZ", "CU3"], p=[0.2, 0.15, 0.1, 0.2, 0.15, 0.05, 0.15]) if n > 1 else \
            rng.choice(["RY", "RZ", "RX", "U3"])
        if kind in ("RY", "RZ", "RX"):
            ops.append(Op(str(kind), (int(rng.integers(n)),), (src(kind == "RY" or not vm_only),)))
        elif kind == "U3":
            ops.append(Op("U3", (int(rng.integers(n)),), tuple(src(not vm_only) for _ in range(3))))
        else:
            a, b = (int(v) for v in rng.choice(n, size=2, replace=False))
            ops.append(Op(str(kind), (a, b), tuple(src(not vm_only) for _ in range(3)) if kind == "CU3" else ()))
    used = sorted({int(v) for op in ops for k, v in op.params if k == "w"})
    remap = {w: i for i, w in enumerate(used)}                             # every weight index is used
    ops = [Op(op.kind, op.wires, tuple((k, remap[int(v)]) if k == "w" else (k, v) for k, v in op.params)) for op in ops]
    return qris_b200.Circuit(n, ops, n_inputs=n, n_weights=max(len(used), 1))
