"""Gate-list description of a QNN circuit and its handle in libqnn_b200.so.

A ``Circuit`` is what tracing the reference's ``quantum_circuit(inputs, weights)``
(Estimators_QuantumNAT_onchipQNN.py:125-142) yields: an ordered gate list whose angles come from
the flat weight vector, from a column of the per-sample input row, or from constants.  It replaces
``qml.device`` + ``qml.QNode`` (Est:122,148).  The measurement is always <Z> on every wire, in wire
order (Est:142).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Sequence, Tuple

from . import _lib

Source = Tuple[str, float]  # ("w", flat weight index) | ("x", input column) | ("c", constant)


@dataclass(frozen=True)
class Op:
    kind: str
    wires: Tuple[int, ...]
    params: Tuple[Source, ...] = ()


class Circuit:
    """Immutable gate list + lazily created device handles (one per CUDA device)."""

    def __init__(self, n_qubits: int, ops: Sequence[Op], n_inputs: int, n_weights: int,
                 encoder: str = "none", generic_only: bool = False, variant: str = None):
        if encoder not in _lib.ENCODER:
            raise ValueError(f"encoder must be one of {sorted(_lib.ENCODER)}, got {encoder!r}")
        self.n_qubits = int(n_qubits)
        self.n_inputs = int(n_inputs)
        self.n_weights = int(n_weights)
        self.encoder = encoder
        # keep the circuit on the generic kernel family (QNN_CREATE_GENERIC_ONLY): the parity tests use it to
        # check one kernel family against the other
        self.generic_only = bool(generic_only)
        # pin one kernel variant ("generic", "wide", "narrow", "16-per-thread": QNN_CREATE_* flags) for tests and
        # measurements; None lets every call choose by its batch size
        self.variant = "generic" if generic_only else variant
        if self.variant not in _lib.VARIANT_FLAGS:
            raise ValueError(f"variant must be one of {sorted(k for k in _lib.VARIANT_FLAGS if k)}, got {variant!r}")
        self.ops: Tuple[Op, ...] = tuple(ops)
        for op in self.ops:
            if op.kind not in _lib.GATE:
                raise ValueError(f"unknown gate {op.kind!r}")
            if len(op.params) != _lib.GATE_ANGLES[op.kind]:
                raise ValueError(f"{op.kind} takes {_lib.GATE_ANGLES[op.kind]} angles, got {len(op.params)}")
        self._handles = {}

    # -- C ABI ---------------------------------------------------------------------------
    def c_ops(self):
        arr = (_lib.QnnOp * max(len(self.ops), 1))()
        for i, op in enumerate(self.ops):
            o = arr[i]
            o.kind = _lib.GATE[op.kind]
            o.wire0 = op.wires[0]
            o.wire1 = op.wires[1] if len(op.wires) > 1 else -1
            for k in range(3):
                o.src[k] = _lib.SRC_CONST
                o.value[k] = 0.0
            for k, (what, v) in enumerate(op.params):
                if what == "w":
                    o.src[k] = int(v)
                elif what == "x":
                    o.src[k] = -1 - int(v)
                elif what == "c":
                    o.value[k] = float(v)
                else:
                    raise ValueError(f"angle source must be 'w', 'x' or 'c', got {what!r}")
        return arr

    def _create(self, out):
        flags = _lib.VARIANT_FLAGS[self.variant]
        return _lib.lib().qnn_circuit_create_ex(self.c_ops(), len(self.ops), self.n_qubits, self.n_inputs,
                                                self.n_weights, _lib.ENCODER[self.encoder], flags, C.byref(out))

    def generic(self) -> "Circuit":
        """The same gate list, pinned to the generic kernel family."""
        return Circuit(self.n_qubits, self.ops, self.n_inputs, self.n_weights, self.encoder, generic_only=True)

    def pinned(self, variant: str) -> "Circuit":
        """The same gate list, pinned to one kernel variant."""
        return Circuit(self.n_qubits, self.ops, self.n_inputs, self.n_weights, self.encoder, variant=variant)

    def handle(self, device_index: int):
        """``qnn_circuit*`` for a CUDA device (created on first use, with that device current)."""
        h = self._handles.get(device_index)
        if h is None:
            import torch
            out = C.c_void_p()
            with torch.cuda.device(device_index):
                rc = self._create(out)
            _lib.check(rc, "qnn_circuit_create")
            h = out
            self._handles[device_index] = h
        return h

    def describe(self) -> dict:
        """The register-resident program this circuit lowers to (``qnn_circuit_describe``), as a dict.
        Works without a CUDA device (a host-only handle is created and destroyed)."""
        import json
        lib = _lib.lib()
        if self._handles:
            h, own = next(iter(self._handles.values())), False
        else:
            h, own = C.c_void_p(), True
            _lib.check(self._create(h), "qnn_circuit_create")
        try:
            need = lib.qnn_circuit_describe(h, None, 0)
            buf = C.create_string_buffer(need)
            lib.qnn_circuit_describe(h, buf, need)
            return json.loads(buf.value.decode())
        finally:
            if own:
                lib.qnn_circuit_destroy(h)

    def __del__(self):
        try:
            for h in self._handles.values():
                _lib.lib().qnn_circuit_destroy(h)
        except Exception:
            pass

    # circuits hold device handles: share, never copy (nn.Module deepcopy / DataParallel.replicate)
    def __deepcopy__(self, memo):
        return self

    def __getstate__(self):
        return {"n_qubits": self.n_qubits, "n_inputs": self.n_inputs, "n_weights": self.n_weights,
                "encoder": self.encoder, "ops": self.ops, "generic_only": self.generic_only,
                "variant": self.variant}

    def __setstate__(self, st):
        self.__init__(st["n_qubits"], st["ops"], st["n_inputs"], st["n_weights"], st["encoder"],
                      st.get("generic_only", False), st.get("variant"))

    def __repr__(self):
        return (f"Circuit(n_qubits={self.n_qubits}, n_ops={len(self.ops)}, n_inputs={self.n_inputs}, "
                f"n_weights={self.n_weights}, encoder={self.encoder!r})")


# ---------------------------------------------------------------------------------------------
def reference_circuit(n_qubits: int = 6, n_layers: int = 3) -> Circuit:
    """The reference's ansatz (Estimators_QuantumNAT_onchipQNN.py:125-142).

    RY(x_i) on every wire (AngleEmbedding, rotation "Y"); per layer RY then RZ on every wire with
    ``weights[layer, i, 0/1]`` (flattened ``(L, n, 2)``, Est:145) and a CNOT ring.
    """
    ops: List[Op] = [Op("RY", (i,), (("x", i),)) for i in range(n_qubits)]
    for layer in range(n_layers):
        for i in range(n_qubits):
            b = (layer * n_qubits + i) * 2
            ops.append(Op("RY", (i,), (("w", b),)))
            ops.append(Op("RZ", (i,), (("w", b + 1),)))
        if n_qubits > 1:
            for i in range(n_qubits - 1):
                ops.append(Op("CNOT", (i, i + 1)))
            ops.append(Op("CNOT", (n_qubits - 1, 0)))
    return Circuit(n_qubits, ops, n_inputs=n_qubits, n_weights=n_layers * n_qubits * 2)


def u3cu3_circuit(n_qubits: int, n_layers: int, encoder: str = "angle") -> Circuit:
    """north_star's U3/CU3 ansatz: per layer U3 on every wire, then a CU3 ring (w -> w+1, last -> 0).

    ``encoder`` "angle": RY(x_i) on wire i first (n inputs); "amplitude": the 2^n-entry input row,
    L2-normalised, is the initial state.  Weights: ``(L, 2, n, 3)`` flattened, ``[l][0]`` the U3 block
    and ``[l][1]`` the CU3 block, last axis (theta, phi, lambda).
    """
    if encoder not in ("angle", "amplitude"):
        raise ValueError("encoder must be 'angle' or 'amplitude'")
    ops: List[Op] = []
    if encoder == "angle":
        ops += [Op("RY", (i,), (("x", i),)) for i in range(n_qubits)]
    for layer in range(n_layers):
        for i in range(n_qubits):
            b = ((layer * 2 + 0) * n_qubits + i) * 3
            ops.append(Op("U3", (i,), (("w", b), ("w", b + 1), ("w", b + 2))))
        if n_qubits > 1:
            for i in range(n_qubits):
                b = ((layer * 2 + 1) * n_qubits + i) * 3
                ops.append(Op("CU3", (i, (i + 1) % n_qubits), (("w", b), ("w", b + 1), ("w", b + 2))))
    return Circuit(n_qubits, ops,
                   n_inputs=n_qubits if encoder == "angle" else (1 << n_qubits),
                   n_weights=n_layers * 2 * n_qubits * 3,
                   encoder="none" if encoder == "angle" else "amplitude")
