"""The ten PennyLane names the reference uses, bound to libqnn_b200.so.

Estimators_QuantumNAT_onchipQNN.py touches PennyLane only through ``qml.device`` (Est:122),
``qml.templates.AngleEmbedding`` (Est:127), ``qml.RY`` / ``qml.RZ`` (Est:133-134), ``qml.CNOT``
(Est:138-139), ``qml.expval(qml.PauliZ(i))`` (Est:142), ``qml.QNode`` (Est:148),
``qml.qnn.TorchLayer`` (Est:149) and ``from pennylane import numpy`` (Est:9).  This module offers
exactly that surface (plus RX/U3/CZ/CU3/Rot and AmplitudeEmbedding).  ``install()`` registers it as
``pennylane`` in ``sys.modules`` so the reference's ``import pennylane as qml`` resolves here and its
``QSC_P128`` runs unmodified on the CUDA kernels.

Unlike PennyLane, the circuit function is traced ONCE, when ``TorchLayer`` is built, with symbolic
``inputs``/weights; the resulting gate list is lowered by ``qnn_circuit_create``.  Circuit functions
must therefore be static: gates and wires may not depend on tensor values, and angles must be plain
elements of ``inputs`` or of a weight tensor (or Python constants).
"""
from __future__ import annotations

import sys
import types
from typing import Dict, List, Optional, Sequence

import numpy as _np

from ..circuit import Circuit, Op
from ..qlayer import QNNLayer

numpy = _np  # `from pennylane import numpy as pnp` (Est:9); the reference never uses it


# ---- symbolic arguments -----------------------------------------------------------------------
class _Sym:
    """One angle: an element of the input row or of the flat weight vector."""

    def __init__(self, what: str, index: int):
        self.source = (what, int(index))

    def _no(self, *_):
        raise TypeError("angles passed to gates must be plain elements of `inputs` or of a weight "
                        "tensor; arithmetic on them is not supported by the traced circuit")

    __add__ = __radd__ = __sub__ = __rsub__ = __mul__ = __rmul__ = __truediv__ = __neg__ = _no


class _SymArray:
    def __init__(self, what: str, shape: Sequence[int], offset: int = 0):
        self.what, self.shape, self.offset = what, tuple(int(s) for s in shape), int(offset)
        self._idx = _np.arange(int(_np.prod(self.shape, dtype=_np.int64))).reshape(self.shape) + self.offset

    def __len__(self):
        return self.shape[0]

    def __iter__(self):
        return (self[i] for i in range(self.shape[0]))

    def __getitem__(self, item):
        if self.what == "x" and isinstance(item, tuple) and len(item) == 2 and item[0] in (Ellipsis, slice(None)):
            item = item[1]  # inputs[:, i] / inputs[..., i]: batch axis is implicit
        sub = self._idx[item]
        if _np.ndim(sub) == 0:
            return _Sym(self.what, int(sub))
        out = _SymArray.__new__(_SymArray)
        out.what, out.shape, out.offset, out._idx = self.what, sub.shape, 0, sub
        return out


def _source(a):
    if isinstance(a, _Sym):
        return a.source
    if isinstance(a, _SymArray):
        raise TypeError("a gate angle must be a scalar element, got an array")
    return ("c", float(a))


# ---- tape -------------------------------------------------------------------------------------
class _Tape:
    def __init__(self):
        self.ops: List[Op] = []
        self.encoder = "none"


_active: Optional[_Tape] = None


def _record(kind: str, wires, *angles):
    if _active is None:
        raise RuntimeError("quantum operations can only be used inside a circuit function run by a QNode")
    ws = tuple(int(w) for w in ([wires] if isinstance(wires, (int, _np.integer)) else list(wires)))
    _active.ops.append(Op(kind, ws, tuple(_source(a) for a in angles)))


def RX(phi, wires): _record("RX", wires, phi)
def RY(phi, wires): _record("RY", wires, phi)
def RZ(phi, wires): _record("RZ", wires, phi)
def U3(theta, phi, delta, wires): _record("U3", wires, theta, phi, delta)
def CNOT(wires): _record("CNOT", wires)
def CZ(wires): _record("CZ", wires)
def CU3(theta, phi, delta, wires): _record("CU3", wires, theta, phi, delta)


def Rot(phi, theta, omega, wires):
    """PennyLane's Rot = RZ(omega) RY(theta) RZ(phi)."""
    RZ(phi, wires); RY(theta, wires); RZ(omega, wires)


class PauliZ:
    def __init__(self, wires):
        self.wire = int(wires if isinstance(wires, (int, _np.integer)) else list(wires)[0])


class _Expval:
    def __init__(self, obs):
        if not isinstance(obs, PauliZ):
            raise NotImplementedError("only expval(PauliZ(wire)) is supported")
        self.wire = obs.wire


def expval(obs):
    return _Expval(obs)


def AngleEmbedding(features, wires, rotation="X"):
    gate = {"X": RX, "Y": RY, "Z": RZ}[rotation]
    wires = list(wires)
    feats = list(features) if not isinstance(features, _SymArray) else [features[i] for i in range(len(features))]
    if len(feats) > len(wires):
        raise ValueError(f"features must be of length {len(wires)} or less; got {len(feats)}")
    for f, w in zip(feats, wires):
        gate(f, wires=w)


def AmplitudeEmbedding(features, wires, normalize=True, pad_with=0.0):
    if _active is None:
        raise RuntimeError("AmplitudeEmbedding outside a circuit function")
    if _active.ops or not isinstance(features, _SymArray) or features.what != "x":
        raise NotImplementedError("AmplitudeEmbedding must come first and embed `inputs`")
    if not normalize or pad_with not in (0, 0.0):
        raise NotImplementedError("only normalize=True, pad_with=0 is supported")
    _active.encoder = "amplitude"


templates = types.SimpleNamespace(AngleEmbedding=AngleEmbedding, AmplitudeEmbedding=AmplitudeEmbedding)


# ---- device / QNode / TorchLayer ----------------------------------------------------------------
class _Device:
    def __init__(self, name: str, wires):
        self.short_name = name
        self.num_wires = int(wires) if isinstance(wires, (int, _np.integer)) else len(list(wires))
        self.wires = list(range(self.num_wires))

    def __repr__(self):
        return f"<b200 statevector device ({self.short_name}), wires={self.num_wires}>"


def device(name, wires=1, shots=None, **_):
    if shots is not None:
        raise NotImplementedError("finite-shot simulation is not supported (analytic expectation values only)")
    return _Device(name, wires)


class QNode:
    def __init__(self, func, device, interface="torch", diff_method="best", **_):
        if interface not in ("torch", "auto", None):
            raise NotImplementedError("only interface='torch' is supported")
        self.func, self.device, self.interface, self.diff_method = func, device, interface, diff_method

    def trace(self, n_inputs: int, weight_shapes: Dict[str, Sequence[int]], input_arg: str = "inputs"):
        """Run the circuit function once on symbolic arguments -> (Circuit, measured wires)."""
        global _active
        kwargs, off = {}, 0
        for name, shape in weight_shapes.items():
            shape = (shape,) if isinstance(shape, int) else tuple(shape)
            kwargs[name] = _SymArray("w", shape, off)
            off += int(_np.prod(shape, dtype=_np.int64))
        kwargs[input_arg] = _SymArray("x", (n_inputs,))
        tape, prev = _Tape(), _active
        _active = tape
        try:
            ret = self.func(**kwargs)
        finally:
            _active = prev
        ms = list(ret) if isinstance(ret, (list, tuple)) else [ret]
        if not ms or not all(isinstance(m, _Expval) for m in ms):
            raise NotImplementedError("the circuit must return expval(PauliZ(w)) measurements")
        n = self.device.num_wires
        circ = Circuit(n, tape.ops, n_inputs=n_inputs, n_weights=off, encoder=tape.encoder)
        return circ, [m.wire for m in ms]


def qnode(device, interface="torch", diff_method="best", **kw):
    return lambda func: QNode(func, device, interface=interface, diff_method=diff_method, **kw)


class TorchLayer(QNNLayer):
    """Drop-in for ``qml.qnn.TorchLayer(qnode, weight_shapes)`` (Est:149)."""

    def __init__(self, qnode: QNode, weight_shapes: Dict[str, Sequence[int]], init_method=None, n_inputs=None):
        n = qnode.device.num_wires
        circ, wires = None, None
        candidates = [n_inputs] if n_inputs is not None else [n, 1 << n]
        err = None
        for ni in candidates:   # angle embeddings take n inputs, amplitude embeddings 2^n
            try:
                circ, wires = qnode.trace(ni, weight_shapes)
                if circ.encoder == "amplitude" and ni != (1 << n) and n_inputs is None:
                    continue
                break
            except (IndexError, ValueError) as e:
                err = e
        if circ is None:
            raise err
        super().__init__(circ, weight_shapes, output_wires=wires)
        self.qnode = qnode
        if init_method is not None:
            for name in weight_shapes:
                fn = init_method[name] if isinstance(init_method, dict) else init_method
                fn(getattr(self, name))


qnn = types.SimpleNamespace(TorchLayer=TorchLayer)


def install(name: str = "pennylane") -> types.ModuleType:
    """Register this module as ``pennylane`` so ``import pennylane as qml`` (Est:8) resolves to it."""
    me = sys.modules[__name__]
    sys.modules[name] = me
    return me
