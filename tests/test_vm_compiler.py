"""CPU: the VM's host compiler (csrc/qnn_vm_compile.cpp) checked by emulating its micro-op program in
numpy and comparing with the oracle; also checks every exchange for bank-conflict freedom."""
import numpy as np
import pytest
import torch

import qris_b200
from helpers import oracle_ops
from oracle import qnn_oracle as O
from vm_emulator import emulate_forward


def _check(circ, seed=0):
    desc = circ.describe()
    assert desc["ok"], desc["why_not"]
    spw = (32 >> desc["tb"]) if desc.get("wb", 0) == 0 else 1
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, (spw, circ.n_inputs))
    w = rng.uniform(0, 2 * np.pi, circ.n_weights)
    y, stats = emulate_forward(desc, x, w)
    yo = O.simulate(oracle_ops(circ), circ.n_qubits, torch.tensor(x), torch.tensor(w), encoder=circ.encoder).numpy()
    np.testing.assert_allclose(y, yo, atol=2e-7)   # constants travel as float32 (pi/2 of the RX lowering)
    return desc, stats


@pytest.mark.parametrize("n", [5, 6, 7, 8, 9, 10])
@pytest.mark.parametrize("L", [1, 3])
def test_reference_circuit_program(n, L):
    desc, stats = _check(qris_b200.reference_circuit(n, L), seed=n + L)
    assert stats["xchg"] <= 2 * L + 1          # CNOT rings ride along with the exchanges
    assert stats["diag"] + stats["diag_reg"] <= L


@pytest.mark.parametrize("n,L", [(5, 2), (6, 2), (8, 1), (10, 3)])
def test_u3cu3_program(n, L):
    desc, stats = _check(qris_b200.u3cu3_circuit(n, L), seed=n)
    assert stats["cg_l"] + stats["cg_p"] == n * L
    if n == 10 and L == 3:
        assert stats["xchg"] <= 3 * L and stats["cg_p"] <= 2 * L


@pytest.mark.parametrize("n,L", [(11, 1), (12, 2), (13, 1), (14, 1), (16, 1)])
def test_multi_warp_samples(n, L):
    """n = 11..16: a sample spans 2^(n-10) warps (one CTA up to n = 13, a cluster above); exchanges
    address a group-wide buffer."""
    desc, stats = _check(qris_b200.reference_circuit(n, L), seed=n)
    assert desc["wb"] == n - 10 and len(desc["exchanges"][0]["wbase"]) == 32 << (n - 10)
    if n <= 13:
        _check(qris_b200.u3cu3_circuit(n, 1), seed=n + 1)


@pytest.mark.parametrize("n,L,kind", [(14, 2, "u3"), (15, 1, "u3"), (16, 2, "ref")])
def test_cluster_programs_mix_local_and_global_exchanges(n, L, kind):
    """n >= 14: exchanges that leave the wires on the CTA-rank bits alone run through the CTA's own shared
    memory (64 KB, bank-conflict free); the others cross CTAs through distributed shared memory: remote stores,
    own-CTA loads.  The emulator keeps one buffer per CTA, so a load that would leave the CTA fails the run, and
    it checks that the backward-direction maps invert the permutation."""
    c = qris_b200.u3cu3_circuit(n, L) if kind == "u3" else qris_b200.reference_circuit(n, L)
    desc, stats = _check(c, seed=n)
    modes = [x["local"] for x in desc["exchanges"]]
    assert 1 in modes and 2 in modes and 0 not in modes     # local + distributed-shared-memory exchanges, none through L2
    for (k, a, b, idx, acc, aux) in desc["ops"]:
        if k == 8:
            assert aux == desc["exchanges"][idx]["local"]
    # warp-level remote stores fill whole 32-byte sectors: all of them for the U3/CU3 ansatz (wires move as wires);
    # a CNOT ring's GF(2) layouts can make two lane bits differ in CTA bits only, which halves the sector use
    exact = stats["dsmem_sector_exact_warp_stores"] / stats["dsmem_warp_stores"]
    assert exact == 1.0 if kind == "u3" else exact >= 0.5


@pytest.mark.parametrize("n", [3, 4, 5, 6, 7, 8])
def test_narrow_variant_programs(n):
    """8 amplitudes per thread (3 register bits): the program small batches run; also the only one for n = 3, 4."""
    for c in (qris_b200.reference_circuit(n, 3), qris_b200.u3cu3_circuit(n, 2)):
        desc, _ = _check(c.pinned("narrow"), seed=n)
        assert desc["kr"] == 3 and desc["tb"] == n - 3 and len(desc["exchanges"][0]["woff"]) == 8 if desc["exchanges"] else desc["kr"] == 3


@pytest.mark.parametrize("n", [5, 9, 10, 12])
def test_sixteen_per_thread_programs(n):
    for c in (qris_b200.reference_circuit(n, 2), qris_b200.u3cu3_circuit(n, 2)):
        desc, _ = _check(c.pinned("16-per-thread"), seed=n)
        assert desc["kr"] == 4 and desc["wb"] == max(0, n - 9)


@pytest.mark.parametrize("n", [5, 8, 10])
def test_amplitude_encoder_program(n):
    """The (normalised) input row is the initial state, gathered into the program's initial layout."""
    desc, _ = _check(qris_b200.u3cu3_circuit(n, 1, encoder="amplitude"), seed=n)
    assert desc["amplitude"]


def test_irregular_program():
    Op = qris_b200.Op
    n = 7
    ops = [Op("RY", (i,), (("x", i),)) for i in range(n)]
    ops += [Op("CNOT", (0, 6)), Op("RZ", (0,), (("w", 0),)), Op("RZ", (6,), (("w", 1),)), Op("RX", (6,), (("w", 2),)),
            Op("CZ", (1, 5)), Op("CNOT", (5, 1)), Op("CNOT", (1, 5)), Op("U3", (5,), (("w", 3), ("c", 0.4), ("w", 4))),
            Op("CU3", (6, 2), (("w", 5), ("w", 6), ("w", 7))), Op("CNOT", (2, 3)), Op("RY", (3,), (("x", 0),)),
            Op("CZ", (3, 4)), Op("RZ", (4,), (("c", 1.3),)), Op("CNOT", (4, 0)), Op("CNOT", (0, 4)), Op("RY", (0,), (("w", 8),))]
    _check(qris_b200.Circuit(n, ops, n_inputs=n, n_weights=9), seed=3)


def test_out_of_scope_circuits_report_why():
    assert not qris_b200.reference_circuit(2, 1).describe()["ok"]
    assert not qris_b200.reference_circuit(4, 1).pinned("wide").describe()["ok"]
    assert qris_b200.reference_circuit(4, 1).describe()["kr"] == 3
    d = qris_b200.u3cu3_circuit(12, 1, encoder="amplitude").describe()
    assert not d["ok"] and "amplitude" in d["why_not"]
