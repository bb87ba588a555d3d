"""Generates tests/golden/qnn_golden.npz from oracle/qnn_oracle.py (complex128).

The reference itself cannot produce fixtures: PennyLane is not installed and not obtainable here
(SURVEY.md F3), and the reference ships none.  These vectors therefore freeze the ORACLE (pinned by
the analytic tests in tests/test_oracle.py), so that the oracle and the kernels cannot drift together
unnoticed.  Re-run from the repository root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import qnn_oracle as O  # noqa: E402

CASES = {
    # name: (builder, n, L, encoder, B)
    "ref_n6_L3": ("ref", 6, 3, "none", 16),       # the reference's shipped shape (Est:108)
    "ref_n4_L3": ("ref", 4, 3, "none", 8),        # Loss-curve variants (4 / 8 "bits")
    "ref_n8_L3": ("ref", 8, 3, "none", 8),
    "ref_n10_L3": ("ref", 10, 3, "none", 4),
    "u3cu3_n5_L2": ("u3", 5, 2, "none", 8),
    "u3cu3_n10_L3": ("u3", 10, 3, "none", 4),     # BASELINE config 3 circuit
    "u3cu3_amp_n4_L2": ("u3amp", 4, 2, "amplitude", 8),
    "u3cu3_n12_L2": ("u3", 12, 2, "none", 2),
    "u3cu3_n16_L8": ("u3", 16, 8, "none", 2),     # BASELINE config 4's exact circuit (768 weights, 512 KiB per state)
}


def build(kind, n, L):
    if kind == "ref":
        return O.ref_circuit_ops(n, L)
    return O.u3cu3_circuit_ops(n, L, angle_encoder=(kind == "u3"))


def main():
    out = {}
    for name, (kind, n, L, enc, B) in CASES.items():
        ops = build(kind, n, L)
        nw = O.n_weights_of(ops)
        g = torch.Generator().manual_seed(1234 + n * 10 + L)
        n_in = n if enc == "none" else (1 << n)
        x = (2 * torch.rand(B, n_in, generator=g) - 1).float()
        w = (2 * np.pi * torch.rand(nw, generator=g)).float()
        gy = torch.randn(B, n, generator=g).float()
        y, gx, gw = O.forward_and_grads(ops, n, x, w, gy, encoder=enc)
        y2, gx2, gw2 = O.adjoint_vjp_numpy(ops, n, x.numpy(), w.numpy(), gy.numpy(), encoder=enc)
        assert np.abs(y.numpy() - y2).max() < 1e-12 and np.abs(gx.numpy() - gx2).max() < 1e-11
        assert np.abs(gw.numpy() - gw2).max() < 1e-10
        for k, v in dict(x=x, w=w, gy=gy, y=y, gx=gx, gw=gw).items():
            out[f"{name}/{k}"] = v.numpy()
        out[f"{name}/meta"] = np.array([n, L, B, 0 if enc == "none" else 1], dtype=np.int64)
        print(name, "ok")
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "qnn_golden.npz"), **out)


if __name__ == "__main__":
    main()
