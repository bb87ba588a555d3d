"""Random irregular circuits: the VM's host compiler against the oracle through the numpy emulator (CPU), and every
kernel variant against the oracle on the GPU."""
import numpy as np
import pytest
import torch

import qris_b200
from helpers import assert_close, oracle_ops, seeded_inputs
from oracle import qnn_oracle as O
from random_circuits import random_circuit
from vm_emulator import emulate_forward


@pytest.mark.parametrize("n,variant", [(3, "narrow"), (5, "narrow"), (6, "wide"), (7, "16-per-thread"), (8, "narrow"), (9, "wide"),
                                       (10, "16-per-thread"), (10, "wide"), (11, "wide"), (12, "16-per-thread")])
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_compiled_program_matches_oracle(n, variant, seed):
    circ = random_circuit(n, 4 * n + 6, seed=100 * n + seed).pinned(variant)
    desc = circ.describe()
    assert desc["ok"], desc["why_not"]
    spw = (32 >> desc["tb"]) if desc.get("wb", 0) == 0 else 1
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, (spw, circ.n_inputs))
    w = rng.uniform(0, 2 * np.pi, circ.n_weights)
    y, _ = emulate_forward(desc, x, w)
    yo = O.simulate(oracle_ops(circ), n, torch.tensor(x), torch.tensor(w)).numpy()
    np.testing.assert_allclose(y, yo, atol=5e-7)       # constants and pi/2 factors travel as float32


@pytest.mark.parametrize("n", [14, 15])
def test_compiled_cluster_program_matches_oracle(n):
    circ = random_circuit(n, 2 * n, seed=n)
    desc = circ.describe()
    assert desc["ok"] and desc["wb"] == n - 10
    rng = np.random.default_rng(n)
    x = rng.uniform(-1, 1, (1, n)); w = rng.uniform(0, 2 * np.pi, circ.n_weights)
    y, _ = emulate_forward(desc, x, w)
    yo = O.simulate(oracle_ops(circ), n, torch.tensor(x), torch.tensor(w)).numpy()
    np.testing.assert_allclose(y, yo, atol=5e-7)


def _gpu_check(circ, B, seed):
    x, w, gy = seeded_inputs(circ, B, seed)
    xd, wd, gd = x.cuda(), w.cuda(), gy.cuda()
    y, state = qris_b200.qnn_forward_raw(circ, xd, wd, save_state=True)
    yo, gxo, gwo = O.forward_and_grads(oracle_ops(circ), circ.n_qubits, x, w, gy)
    assert_close(y.cpu().numpy(), yo.numpy(), what="y")
    for st in ((None, state) if state is not None else (None,)):
        gx, gw = qris_b200.qnn_backward_raw(circ, xd, wd, gd, state=st)
        assert_close(gx.cpu().numpy(), gxo.numpy(), scale=float(gxo.abs().max()), what="gx")
        assert_close(gw.cpu().numpy(), gwo.numpy(), scale=float(gwo.abs().max()), what="gw")


@pytest.mark.gpu
@pytest.mark.parametrize("n", [3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16])
@pytest.mark.parametrize("seed", [0, 1])
def test_random_circuits_on_the_gpu(n, seed):
    """Every variant that takes the circuit (wide, narrow, 16 per thread, generic) against the oracle, ragged batch."""
    base = random_circuit(n, 3 * n + 8, seed=1000 * seed + n)
    B = {14: 5, 15: 3, 16: 2}.get(n, 37 if n <= 10 else 7)
    variants = [None, "generic"] + (["narrow"] if n <= 8 else []) + (["wide"] if n >= 5 else []) + (["16-per-thread"] if 4 <= n <= 12 else [])
    for v in variants:
        circ = base if v is None else base.pinned(v)
        plan = qris_b200._lib.lib().qnn_circuit_plan(circ.handle(0), B).decode()
        assert plan.startswith("generic") == (v == "generic"), (v, plan)
        _gpu_check(circ, B, seed=n + seed)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [2, 5, 9])
def test_random_circuits_with_input_sourced_angles_fall_to_the_generic_family(n):
    """RX / RZ / U3 / CU3 angles taken from the input row are outside the VM's scope: generic kernels, same parity."""
    circ = random_circuit(n, 3 * n + 6, seed=n, vm_only=False)
    assert any(k == "x" for op in circ.ops if op.kind != "RY" for k, _ in op.params)
    assert qris_b200._lib.lib().qnn_circuit_plan(circ.handle(0), 9).decode().startswith("generic")
    _gpu_check(circ, 9, seed=n)
