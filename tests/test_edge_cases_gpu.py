"""GPU: edge cases of the hot path through every kernel variant: circuits without weights, gates that use no input
column, 32 re-uploaded input columns, million-sample batches, a 16-qubit batch larger than the cluster count."""
import numpy as np
import pytest
import torch

import qris_b200
from helpers import assert_close, oracle_ops, seeded_inputs
from oracle import qnn_oracle as O

pytestmark = pytest.mark.gpu
Op = qris_b200.Op


def check(circ, B, seed=0, tag=""):
    x, w, gy = seeded_inputs(circ, B, seed)
    xd, wd, gd = x.cuda(), w.cuda(), gy.cuda()
    y, st = qris_b200.qnn_forward_raw(circ, xd, wd, save_state=True)
    gx, gw = qris_b200.qnn_backward_raw(circ, xd, wd, gd, state=st)
    idx = torch.tensor(sorted({0, B // 3, B - 1}))
    yo, gxo, _ = O.forward_and_grads(oracle_ops(circ), circ.n_qubits, x[idx], w, gy[idx])
    assert_close(y[idx.cuda()].cpu().numpy(), yo.numpy(), what=tag + " y")
    assert_close(gx[idx.cuda()].cpu().numpy(), gxo.numpy(), scale=float(gxo.abs().max()) + 1e-30, what=tag + " gx")
    assert torch.isfinite(gw).all() and torch.isfinite(y).all() and torch.isfinite(gx).all()


@pytest.mark.parametrize("n", [3, 6, 9])
def test_circuit_without_weights(n):
    ops = [Op("RY", (i,), (("x", i),)) for i in range(n)] + [Op("CNOT", (i, (i + 1) % n)) for i in range(n)]
    c = qris_b200.Circuit(n, ops, n_inputs=n, n_weights=0)
    x, w, gy = seeded_inputs(c, 50, 1)
    y = qris_b200.qnn_forward_raw(c, x.cuda(), torch.zeros(0, device="cuda"))
    gx, gw = qris_b200.qnn_backward_raw(c, x.cuda(), torch.zeros(0, device="cuda"), gy.cuda())
    yo, gxo, _ = O.forward_and_grads(oracle_ops(c), n, x, torch.zeros(0), gy)
    assert_close(y.cpu().numpy(), yo.numpy(), what="y")
    assert_close(gx.cpu().numpy(), gxo.numpy(), scale=float(gxo.abs().max()), what="gx")
    assert gw.numel() == 0


@pytest.mark.parametrize("variant", [None, "wide", "narrow", "generic", "16-per-thread"])
def test_gates_that_use_no_input_column(variant):
    ops = [Op("RX", (0,), (("c", 0.3),)), Op("CZ", (0, 1)), Op("U3", (2,), (("c", 0.1), ("c", 0.2), ("w", 0))),
           Op("CU3", (2, 3), (("w", 1), ("c", 0.5), ("w", 0)))]
    c = qris_b200.Circuit(5, ops, n_inputs=5, n_weights=2)
    cc = c if variant is None else c.pinned(variant)
    x, w, gy = seeded_inputs(cc, 33, 2)
    y = qris_b200.qnn_forward_raw(cc, x.cuda(), w.cuda())
    gx, gw = qris_b200.qnn_backward_raw(cc, x.cuda(), w.cuda(), gy.cuda())
    yo, gxo, gwo = O.forward_and_grads(oracle_ops(cc), 5, x, w, gy)
    assert_close(y.cpu().numpy(), yo.numpy(), what="y")
    assert_close(gw.cpu().numpy(), gwo.numpy(), scale=float(gwo.abs().max()), what="gw")
    assert gx.abs().max() == 0


def test_thirty_two_input_columns_reuploaded():
    n = 6
    ops = [Op("RY", (i % n,), (("x", i),)) for i in range(32)] + [Op("CNOT", (i, (i + 1) % n)) for i in range(n)] + \
          [Op("RY", (i,), (("w", i),)) for i in range(n)]
    c = qris_b200.Circuit(n, ops, n_inputs=32, n_weights=n)
    check(c, 100, 3, "narrow")
    check(c.pinned("wide"), 100, 3, "wide")


@pytest.mark.parametrize("make,B", [(lambda: qris_b200.reference_circuit(6, 3), 1_000_003), (lambda: qris_b200.reference_circuit(3, 2), 2_000_001),
                                     (lambda: qris_b200.u3cu3_circuit(10, 3), 300_001), (lambda: qris_b200.u3cu3_circuit(16, 1), 700)])
def test_large_batches(make, B):
    check(make(), B, seed=B % 97, tag=f"B={B}")
