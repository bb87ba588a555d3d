"""CPU oracle for the QNN statevector hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this file.  The product package never does: it has no
CPU path and raises when its CUDA library is missing.

PARITY UNPINNED.  The reference's arithmetic for this path lives in PennyLane
(``default.qubit`` + torch backprop), an un-vendored, un-pinned third-party dependency that is
not installed here and cannot be fetched (no network).  The reference ships no tests, golden
vectors or fixtures.  So this file is a restatement, from the reference's call sites and
PennyLane's published gate conventions, of

* the circuit at ``Estimators_QuantumNAT_onchipQNN.py:125-142`` (AngleEmbedding-Y, per layer
  RY then RZ on every wire, CNOT chain 0->1 ... n-2->n-1 then n-1->0, <Z> on every wire),
* the weight layout at ``Estimators_QuantumNAT_onchipQNN.py:145`` (``(L, n, 2)``, last axis
  [RY angle, RZ angle]),

pinned by what IS available: the analytic L=1 closed form, the identically-zero gradient of
the last layer's RZ angles, parameter-shift identities, central finite differences, and the
fixed n=6/L=3 vector recorded in SURVEY.md section 3.3 (see tests/test_oracle.py).
The U3/CU3 gate set, the amplitude encoder and the post-measurement QuantumNAT pipeline do
not exist in the reference at all; for those this file *is* the specification.

Two independent implementations live here so that they can check each other:

``simulate`` / ``forward_and_grads``
    torch, one batched tensor operation per gate on a ``(B, 2, ..., 2)`` complex state and
    ordinary autograd for the backward pass -- the way ``default.qubit`` with
    ``interface="torch"`` and ``diff_method="backprop"`` executes the reference
    (``Estimators_QuantumNAT_onchipQNN.py:148``).  complex128 is the parity oracle;
    complex64 is the timed CPU baseline.
``adjoint_vjp_numpy``
    numpy complex128, flat state vector, explicit adjoint-method VJP.  Shares no code with
    the torch version.

Conventions (PennyLane's): wire 0 is the most significant bit of the basis index;
``RX(t) = [[c, -i s], [-i s, c]]``, ``RY(t) = [[c, -s], [s, c]]``,
``RZ(p) = diag(e^{-ip/2}, e^{+ip/2})`` with ``c, s = cos, sin(t/2)``;
``U3(t, p, l) = [[c, -e^{il} s], [e^{ip} s, e^{i(p+l)} c]]``; controlled gates list the
control wire first.

Circuit description ("ops"): a list of ``(kind, wires, params)``; ``kind`` in
``RX RY RZ U3 CNOT CZ CU3``; ``wires`` a tuple of ints; ``params`` a tuple with one entry per
angle, each ``("w", i)`` (weight ``w[i]``), ``("x", j)`` (per-sample input ``x[b, j]``) or
``("c", value)`` (constant).
"""
from __future__ import annotations

import math
from typing import List, Sequence, Tuple

import numpy as np
import torch

Op = Tuple[str, Tuple[int, ...], Tuple[Tuple[str, float], ...]]

N_ANGLES = {"RX": 1, "RY": 1, "RZ": 1, "U3": 3, "CU3": 3, "CNOT": 0, "CZ": 0}


# --------------------------------------------------------------------------------------
# circuit builders
# --------------------------------------------------------------------------------------
def ref_circuit_ops(n_qubits: int, n_layers: int) -> List[Op]:
    """The reference circuit, gate for gate (Estimators_QuantumNAT_onchipQNN.py:125-142).

    Weight index of ``weights[layer, i, k]`` in the flattened ``(L, n, 2)`` tensor
    (Estimators...py:145) is ``(layer * n + i) * 2 + k``.
    """
    ops: List[Op] = []
    for i in range(n_qubits):                      # Est:127 AngleEmbedding(rotation="Y")
        ops.append(("RY", (i,), (("x", i),)))
    for layer in range(n_layers):                  # Est:130
        for i in range(n_qubits):                  # Est:132-134
            base = (layer * n_qubits + i) * 2
            ops.append(("RY", (i,), (("w", base),)))
            ops.append(("RZ", (i,), (("w", base + 1),)))
        if n_qubits > 1:
            for i in range(n_qubits - 1):          # Est:137-138
                ops.append(("CNOT", (i, i + 1), ()))
            ops.append(("CNOT", (n_qubits - 1, 0), ()))   # Est:139
    return ops


def u3cu3_circuit_ops(n_qubits: int, n_layers: int, angle_encoder: bool = True) -> List[Op]:
    """north_star's U3/CU3 circuit (not in the reference).

    Optional RY angle encoder, then per layer a U3 on every wire followed by a CU3 ring
    (w -> w+1, last -> 0).  Weights are ``(L, 2, n, 3)`` flattened: ``[layer][0]`` the U3
    block, ``[layer][1]`` the CU3 block.
    """
    ops: List[Op] = []
    if angle_encoder:
        for i in range(n_qubits):
            ops.append(("RY", (i,), (("x", i),)))
    for layer in range(n_layers):
        for i in range(n_qubits):
            b = ((layer * 2 + 0) * n_qubits + i) * 3
            ops.append(("U3", (i,), (("w", b), ("w", b + 1), ("w", b + 2))))
        if n_qubits > 1:
            for i in range(n_qubits):
                b = ((layer * 2 + 1) * n_qubits + i) * 3
                ops.append(("CU3", (i, (i + 1) % n_qubits),
                            (("w", b), ("w", b + 1), ("w", b + 2))))
    return ops


def n_weights_of(ops: Sequence[Op]) -> int:
    m = -1
    for _, _, params in ops:
        for kind, v in params:
            if kind == "w":
                m = max(m, int(v))
    return m + 1


# --------------------------------------------------------------------------------------
# torch implementation (default.qubit-style: one tensor op per gate, autograd backward)
# --------------------------------------------------------------------------------------
def _angle(p, x, w, rdtype):
    kind, v = p
    if kind == "w":
        return w[int(v)]
    if kind == "x":
        return x[:, int(v)]
    return torch.tensor(float(v), dtype=rdtype)


def _matrix(kind: str, angles, cdtype):
    """2x2 matrix, shape (2, 2) or (B, 2, 2) if any angle is per-sample."""
    if kind in ("RX", "RY", "RZ"):
        t = angles[0]
        c = torch.cos(t / 2).to(cdtype)
        s = torch.sin(t / 2).to(cdtype)
        if kind == "RY":
            rows = [[c, -s], [s, c]]
        elif kind == "RX":
            rows = [[c, -1j * s], [-1j * s, c]]
        else:
            e = torch.exp(-0.5j * t.to(cdtype))
            z = torch.zeros_like(e)
            rows = [[e, z], [z, torch.conj(e)]]
    else:  # U3 / CU3
        t, p, l = angles
        shape = torch.broadcast_shapes(t.shape, p.shape, l.shape)
        t, p, l = (a.expand(shape) for a in (t, p, l))
        c = torch.cos(t / 2).to(cdtype)
        s = torch.sin(t / 2).to(cdtype)
        ep = torch.exp(1j * p.to(cdtype))
        el = torch.exp(1j * l.to(cdtype))
        rows = [[c, -el * s], [ep * s, ep * el * c]]
    return torch.stack([torch.stack(r, dim=-1) for r in rows], dim=-2)


def _apply_1q(state, U, wire):
    # state (B, 2, ..., 2); U (2,2) or (B,2,2)
    st = torch.movedim(state, wire + 1, 1)
    shp = st.shape
    st = st.reshape(shp[0], 2, -1)
    st = torch.matmul(U, st)
    return torch.movedim(st.reshape(shp), 1, wire + 1)


def _apply_controlled(state, U, control, target):
    s0 = state.select(control + 1, 0)
    s1 = state.select(control + 1, 1)
    t_axis = target if target < control else target - 1
    s1 = _apply_1q(s1, U, t_axis)
    return torch.stack([s0, s1], dim=control + 1)


def simulate(ops: Sequence[Op], n_qubits: int, x: torch.Tensor, w: torch.Tensor,
             encoder: str = "none", cdtype=torch.complex128) -> torch.Tensor:
    """Return ``y[b, q] = <Z_q>`` as a real tensor of cdtype's precision (differentiable)."""
    rdtype = torch.float64 if cdtype == torch.complex128 else torch.float32
    x = x.to(rdtype)
    w = w.to(rdtype)
    B = x.shape[0]
    dim = 1 << n_qubits
    if encoder == "amplitude":
        amp = torch.zeros(B, dim, dtype=rdtype)
        k = min(dim, x.shape[1])
        amp = torch.cat([x[:, :k], amp[:, k:]], dim=1)
        amp = amp / torch.sqrt((amp * amp).sum(dim=1, keepdim=True))
        state = amp.to(cdtype)
    else:
        state = torch.zeros(B, dim, dtype=cdtype)
        state[:, 0] = 1.0
    state = state.reshape((B,) + (2,) * n_qubits)
    for kind, wires, params in ops:
        if kind == "CNOT":
            c, t = wires
            s0 = state.select(c + 1, 0)
            s1 = state.select(c + 1, 1)
            t_axis = t if t < c else t - 1
            s1 = torch.flip(s1, dims=(t_axis + 1,))
            state = torch.stack([s0, s1], dim=c + 1)
        elif kind == "CZ":
            c, t = wires
            sign = torch.ones([1] + [2] * n_qubits, dtype=cdtype)
            idx = [slice(None)] * (n_qubits + 1)
            idx[c + 1] = 1
            idx[t + 1] = 1
            sign[tuple(idx)] = -1
            state = state * sign
        else:
            angles = [_angle(p, x, w, rdtype) for p in params]
            U = _matrix(kind, angles, cdtype)
            if kind == "CU3":
                state = _apply_controlled(state, U, wires[0], wires[1])
            else:
                state = _apply_1q(state, U, wires[0])
    probs = (state.real ** 2 + state.imag ** 2)
    ys = []
    for q in range(n_qubits):
        axes = tuple(a + 1 for a in range(n_qubits) if a != q)
        m = probs.sum(dim=axes) if axes else probs
        ys.append(m[:, 0] - m[:, 1])
    return torch.stack(ys, dim=1)


def forward_and_grads(ops, n_qubits, x, w, gy, encoder="none", cdtype=torch.complex128):
    """(y, gx, gw) with gx = d<gy,y>/dx and gw = d<gy,y>/dw, by autograd (backprop)."""
    rdtype = torch.float64 if cdtype == torch.complex128 else torch.float32
    xr = x.detach().to(rdtype).clone().requires_grad_(True)
    wr = w.detach().to(rdtype).clone().requires_grad_(True)
    y = simulate(ops, n_qubits, xr, wr, encoder=encoder, cdtype=cdtype)
    y.backward(gy.to(rdtype))
    gx = xr.grad if xr.grad is not None else torch.zeros_like(xr)
    gw = wr.grad if wr.grad is not None else torch.zeros_like(wr)
    return y.detach(), gx, gw


# --------------------------------------------------------------------------------------
# independent numpy implementation with an explicit adjoint-method VJP
# --------------------------------------------------------------------------------------
def _np_mats(kind, a):
    """Matrix and its derivatives w.r.t. each angle (numpy complex128)."""
    if kind == "RY":
        c, s = math.cos(a[0] / 2), math.sin(a[0] / 2)
        return (np.array([[c, -s], [s, c]], complex),
                [0.5 * np.array([[-s, -c], [c, -s]], complex)])
    if kind == "RX":
        c, s = math.cos(a[0] / 2), math.sin(a[0] / 2)
        return (np.array([[c, -1j * s], [-1j * s, c]], complex),
                [0.5 * np.array([[-s, -1j * c], [-1j * c, -s]], complex)])
    if kind == "RZ":
        e = np.exp(-0.5j * a[0])
        return (np.array([[e, 0], [0, np.conj(e)]], complex),
                [np.array([[-0.5j * e, 0], [0, 0.5j * np.conj(e)]], complex)])
    t, p, l = a
    c, s = math.cos(t / 2), math.sin(t / 2)
    ep, el = np.exp(1j * p), np.exp(1j * l)
    U = np.array([[c, -el * s], [ep * s, ep * el * c]], complex)
    dt = 0.5 * np.array([[-s, -el * c], [ep * c, -ep * el * s]], complex)
    dp = np.array([[0, 0], [1j * ep * s, 1j * ep * el * c]], complex)
    dl = np.array([[0, -1j * el * s], [0, 1j * ep * el * c]], complex)
    return U, [dt, dp, dl]


def _np_apply(psi, M, n, wires, controlled):
    """Apply 2x2 M on target (last wire); if controlled, only where control bit is 1."""
    t = wires[-1]
    v = psi.reshape([2] * n)
    v = np.moveaxis(v, t, 0)
    if controlled:
        c = wires[0]
        c_axis = c + 1 if c < t else c          # axis of control after moving t to front
        idx = [slice(None)] * n
        idx[c_axis] = 1
        sub = v[tuple(idx)]
        out = v.copy()
        out[tuple(idx)] = np.tensordot(M, sub, axes=(1, 0))
        v = out
    else:
        v = np.tensordot(M, v, axes=(1, 0))
    return np.moveaxis(v, 0, t).reshape(-1)


def _np_perm_or_diag(psi, kind, n, wires):
    c, t = wires
    k = np.arange(1 << n)
    cb = (k >> (n - 1 - c)) & 1
    tb = (k >> (n - 1 - t)) & 1
    if kind == "CNOT":
        return psi[k ^ (cb << (n - 1 - t))]
    return psi * (1 - 2 * (cb & tb))


def adjoint_vjp_numpy(ops, n_qubits, x, w, gy, encoder="none"):
    """Per-sample flat-statevector simulation + adjoint VJP.  Returns (y, gx, gw), float64.

    Backward: lam = (sum_q gy_q Z_q) psi_final; walking gates in reverse,
    d/dtheta = 2 Re <lam_after| dU/dtheta |psi_before>, then lam <- U^dagger lam.
    """
    x = np.asarray(x, np.float64)
    w = np.asarray(w, np.float64)
    gy = np.asarray(gy, np.float64)
    B, n = x.shape[0], n_qubits
    dim = 1 << n
    k = np.arange(dim)
    zsign = np.stack([1 - 2 * ((k >> (n - 1 - q)) & 1) for q in range(n)], 0).astype(np.float64)
    y = np.zeros((B, n))
    gx = np.zeros_like(x)
    gw = np.zeros_like(w)

    def ang(p, b):
        kind, v = p
        return w[int(v)] if kind == "w" else (x[b, int(v)] if kind == "x" else float(v))

    for b in range(B):
        if encoder == "amplitude":
            a0 = np.zeros(dim)
            kk = min(dim, x.shape[1])
            a0[:kk] = x[b, :kk]
            nrm = math.sqrt((a0 * a0).sum())
            psi = (a0 / nrm).astype(complex)
        else:
            psi = np.zeros(dim, complex)
            psi[0] = 1
        for kind, wires, params in ops:
            if kind in ("CNOT", "CZ"):
                psi = _np_perm_or_diag(psi, kind, n, wires)
            else:
                U, _ = _np_mats(kind, [ang(p, b) for p in params])
                psi = _np_apply(psi, U, n, wires, kind == "CU3")
        prob = (psi.conj() * psi).real
        y[b] = zsign @ prob
        lam = (gy[b] @ zsign) * psi
        for kind, wires, params in reversed(ops):
            if kind in ("CNOT", "CZ"):   # both are involutions
                psi = _np_perm_or_diag(psi, kind, n, wires)
                lam = _np_perm_or_diag(lam, kind, n, wires)
                continue
            U, dU = _np_mats(kind, [ang(p, b) for p in params])
            ctl = kind == "CU3"
            psi = _np_apply(psi, U.conj().T, n, wires, ctl)       # psi_before
            for p, dM in zip(params, dU):
                if ctl:   # derivative of the controlled gate is zero on the control=0 block
                    c = wires[0]
                    mask = ((k >> (n - 1 - c)) & 1).astype(bool)
                    dpsi = _np_apply(psi, dM, n, wires, True)
                    dpsi = np.where(mask, dpsi, 0)
                else:
                    dpsi = _np_apply(psi, dM, n, wires, False)
                g = 2.0 * np.real(np.vdot(lam, dpsi))
                if p[0] == "w":
                    gw[int(p[1])] += g
                elif p[0] == "x":
                    gx[b, int(p[1])] += g
            lam = _np_apply(lam, U.conj().T, n, wires, ctl)
        if encoder == "amplitude":
            # psi0 = a/|a|  ->  d/da = (I - psi0 psi0^T) glam / |a| with glam = 2 Re(lam)
            gl = 2.0 * lam.real
            p0 = psi.real
            ga = (gl - p0 * (p0 @ gl)) / nrm
            kk = min(dim, x.shape[1])
            gx[b, :kk] += ga[:kk]
    return y, gx, gw


# --------------------------------------------------------------------------------------
# closed forms used to pin the restatement
# --------------------------------------------------------------------------------------
def ref_circuit_L1_closed_form(x: np.ndarray, w: np.ndarray) -> np.ndarray:
    """<Z> of the reference circuit with ONE layer, independent of any simulator.

    ``y_0 = prod_{j>=1} cos(x_j + W[0,j,0])``, ``y_i = prod_{j<=i} cos(x_j + W[0,j,0])``
    for i >= 1; independent of every RZ angle.  ``w`` has shape (1, n, 2).
    """
    c = np.cos(np.asarray(x, np.float64) + np.asarray(w, np.float64)[0, :, 0][None, :])
    n = c.shape[1]
    y = np.empty_like(c)
    y[:, 0] = np.prod(c[:, 1:], axis=1)
    for i in range(1, n):
        y[:, i] = np.prod(c[:, : i + 1], axis=1)
    return y


SURVEY_FIXED_VECTOR = {
    "n": 6, "L": 3,
    "y": [0.017798676022, 0.139615380994, 0.050159640834, 0.068778923112,
          -0.028879496450, 0.137171911749],
    "g": [1, 2, 3, 4, 5, 6],
    "dx0": 0.59248806, "dW_1_2_0": -1.33178480, "dW_2_3_1": 0.0, "dW_2_3_0": -0.18135226,
}


def survey_fixed_inputs():
    n, L = 6, 3
    x = np.array([[0.1 * (q + 1) for q in range(n)]])
    W = np.zeros((L, n, 2))
    for l in range(L):
        for q in range(n):
            W[l, q, 0] = 0.3 + 0.2 * l + 0.05 * q
            W[l, q, 1] = 0.7 - 0.1 * l + 0.03 * q
    return x, W
