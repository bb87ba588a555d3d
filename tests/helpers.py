"""Shared test helpers: product circuit -> oracle op list, tolerances, seeded inputs."""
import numpy as np
import torch

REL_TOL = 1e-5   # north_star: expectation values and gradients within 1e-5 relative in fp32


def oracle_ops(circuit):
    """A product ``Circuit`` as the plain tuples ``oracle/qnn_oracle.py`` takes."""
    return [(op.kind, tuple(op.wires), tuple(op.params)) for op in circuit.ops]


def oracle_encoder(circuit):
    return circuit.encoder


def seeded_inputs(circuit, B, seed=0, x_range=1.0):
    """x = 2*rand-1 (tanh range, Est:161), w = 2*pi*rand (TorchLayer init), gy = randn."""
    g = torch.Generator().manual_seed(seed)
    x = (2 * torch.rand(B, circuit.n_inputs, generator=g) - 1) * x_range
    w = 2 * np.pi * torch.rand(circuit.n_weights, generator=g)
    gy = torch.randn(B, circuit.n_qubits, generator=g)
    return x.float(), w.float(), gy.float()


def assert_close(actual, expected, rel=REL_TOL, scale=1.0, what=""):
    """|a - b| <= rel * max(1, |b|, scale) elementwise (pure relative error is meaningless near 0)."""
    a = np.asarray(actual, dtype=np.float64)
    b = np.asarray(expected, dtype=np.float64)
    assert a.shape == b.shape, f"{what}: shape {a.shape} vs {b.shape}"
    assert np.isfinite(a).all(), f"{what}: {np.count_nonzero(~np.isfinite(a))} of {a.size} elements are NaN or Inf"
    tol = rel * np.maximum(np.maximum(1.0, np.abs(b)), scale)
    err = np.abs(a - b)
    bad = ~(err <= tol)          # NaN compares false either way: count it as off, never as close
    assert not bad.any(), (f"{what}: {bad.sum()} of {bad.size} elements off; worst |err|={np.nanmax(err):.3e} "
                           f"(tol {tol.flat[np.nanargmax(err)]:.3e}) at {np.unravel_index(np.nanargmax(err), err.shape)}")
