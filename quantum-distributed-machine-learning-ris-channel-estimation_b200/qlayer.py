"""The quantum layer as a PyTorch module: what ``qml.qnn.TorchLayer(qnode, weight_shapes)``
(Estimators_QuantumNAT_onchipQNN.py:145-149) is to the reference, backed by libqnn_b200.so.

``forward((B, n_inputs) float32) -> (B, n_qubits) float32`` of <Z> per wire; backward is the
adjoint-method VJP kernel.  CUDA tensors only -- a CPU tensor raises.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence, Tuple

import torch
from torch import nn
from torch.autograd.function import once_differentiable

from . import _lib
from .circuit import Circuit


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _require_cuda(t: torch.Tensor, what: str) -> None:
    if not t.is_cuda:
        raise _lib.QnnError(
            f"{what} is on {t.device}: the QNN layer runs only as sm_100a CUDA kernels "
            "(there is no CPU or PyTorch fallback); move the module and its inputs to a CUDA device.")


def _workspace(nbytes: int, device) -> Optional[torch.Tensor]:
    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=device)


def qnn_forward_raw(circuit: Circuit, x: torch.Tensor, w: torch.Tensor, save_state: bool = False):
    """One ``qnn_forward`` call on the current stream.  x (B, n_inputs), w (n_weights,) float32 CUDA.

    ``save_state=True`` returns ``(y, state)`` where ``state`` is the opaque final-statevector buffer that
    ``qnn_backward_raw(..., state=state)`` starts from (``None`` if the circuit's kernels keep none)."""
    lib = _lib.lib()
    B = x.shape[0]
    y = torch.empty((B, circuit.n_qubits), dtype=torch.float32, device=x.device)
    if B == 0:
        return (y, None) if save_state else y
    dev = x.device.index if x.device.index is not None else torch.cuda.current_device()
    h = circuit.handle(dev)
    state = None
    with torch.cuda.device(dev), torch.cuda.nvtx.range("qnn/forward"):
        ws = _workspace(lib.qnn_forward_workspace_bytes(h, B), x.device)
        st = torch.cuda.current_stream(dev).cuda_stream
        nbytes = lib.qnn_state_bytes(h, B) if save_state else 0
        if nbytes > 0:
            state = torch.empty(nbytes // 4, dtype=torch.float32, device=x.device)
            rc = lib.qnn_forward_save(h, _ptr(x), _ptr(w), _ptr(y), _ptr(state), B, _ptr(ws), ws.numel(), st)
        else:
            rc = lib.qnn_forward(h, _ptr(x), _ptr(w), _ptr(y), B, _ptr(ws), ws.numel(), st)
    _lib.check(rc, "qnn_forward")
    return (y, state) if save_state else y


def qnn_backward_raw(circuit: Circuit, x: torch.Tensor, w: torch.Tensor, gy: torch.Tensor,
                     need_gx: bool = True, gw_out: Optional[torch.Tensor] = None,
                     state: Optional[torch.Tensor] = None
                     ) -> Tuple[Optional[torch.Tensor], torch.Tensor]:
    """One ``qnn_backward`` call on the current stream -> (gx or None, gw).

    ``gw_out``: optional preallocated float32 view (e.g. a slice of a flat gradient bucket) that the
    kernel overwrites directly.  ``state``: the buffer ``qnn_forward_raw(..., save_state=True)`` returned
    for the same (x, w); without it the backward kernel re-runs the forward program first."""
    lib = _lib.lib()
    B = x.shape[0]
    gw = gw_out if gw_out is not None else torch.empty((circuit.n_weights,), dtype=torch.float32, device=x.device)
    gx = torch.empty_like(x) if need_gx else None
    if B == 0:
        gw.zero_()
        return gx, gw
    dev = x.device.index if x.device.index is not None else torch.cuda.current_device()
    h = circuit.handle(dev)
    with torch.cuda.device(dev), torch.cuda.nvtx.range("qnn/backward"):
        ws = _workspace(lib.qnn_backward_workspace_bytes(h, B), x.device)
        st = torch.cuda.current_stream(dev).cuda_stream
        if state is not None:
            rc = lib.qnn_backward_saved(h, _ptr(x), _ptr(w), _ptr(gy), _ptr(state), _ptr(gx), _ptr(gw), B,
                                        _ptr(ws), ws.numel(), st)
        else:
            rc = lib.qnn_backward(h, _ptr(x), _ptr(w), _ptr(gy), _ptr(gx), _ptr(gw), B, _ptr(ws), ws.numel(), st)
    _lib.check(rc, "qnn_backward")
    return gx, gw


# Keep the final state for backward only while it stays modest (2^n * 8 bytes per sample).
STATE_SAVE_LIMIT_BYTES = 4 << 30


class _QNNFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, circuit: Circuit):
        _require_cuda(x, "QNN input")
        _require_cuda(w, "QNN weights")
        xc = x.detach().to(torch.float32).contiguous()
        # Snapshot rule (Est:176-196 vs Run:364): the reference perturbs the weights through .data,
        # runs the forward and restores them BEFORE backward; its autograd graph holds values computed
        # from the noisy weights.  A recomputing backward must therefore own a copy of what it used.
        wc = w.detach().to(torch.float32).contiguous().clone()
        need_grad = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
        save = need_grad and xc.shape[0] * (8 << circuit.n_qubits) <= STATE_SAVE_LIMIT_BYTES
        y, state = qnn_forward_raw(circuit, xc, wc, save_state=True) if save else (qnn_forward_raw(circuit, xc, wc), None)
        ctx.circuit = circuit
        ctx.state = state
        ctx.save_for_backward(xc, wc)
        ctx.x_dtype, ctx.w_dtype = x.dtype, w.dtype
        return y.to(x.dtype) if x.dtype != torch.float32 else y

    @staticmethod
    @once_differentiable
    def backward(ctx, gy):
        xc, wc = ctx.saved_tensors
        gyc = gy.detach().to(torch.float32).contiguous()
        gx, gw = qnn_backward_raw(ctx.circuit, xc, wc, gyc, need_gx=ctx.needs_input_grad[0], state=ctx.state)
        ctx.state = None
        if gx is not None and ctx.x_dtype != torch.float32:
            gx = gx.to(ctx.x_dtype)
        if not ctx.needs_input_grad[1]:
            gw = None
        elif ctx.w_dtype != torch.float32:
            gw = gw.to(ctx.w_dtype)
        return gx, gw, None


def qnn_apply(circuit: Circuit, x: torch.Tensor, w_flat: torch.Tensor) -> torch.Tensor:
    """Differentiable ``y = <Z>(circuit(x; w))``.  x (B, n_inputs) or (n_inputs,)."""
    squeeze = x.dim() == 1
    if squeeze:
        x = x.unsqueeze(0)
    if x.dim() != 2 or x.shape[1] != circuit.n_inputs:
        raise ValueError(f"expected input of shape (B, {circuit.n_inputs}), got {tuple(x.shape)}")
    if w_flat.numel() != circuit.n_weights:
        raise ValueError(f"expected {circuit.n_weights} weights, got {w_flat.numel()}")
    y = _QNNFunction.apply(x, w_flat.reshape(-1), circuit)
    return y.squeeze(0) if squeeze else y


class QNNLayer(nn.Module):
    """``nn.Module`` owning the circuit's trainable angles.

    Like ``qml.qnn.TorchLayer`` (Est:149): parameters are registered under the names given in
    ``weight_shapes`` (the reference: ``{"weights": (n_layers, n_qubits, 2)}``, Est:145 -> state_dict
    key ``qlayer.weights``), initialised ``uniform_(0, 2*pi)``, read at call time (so code that
    perturbs ``param.data`` around the call, Est:180-196, sees what it expects) and ordinary leaves
    for any optimiser.  Several weight tensors are concatenated, in ``weight_shapes`` order, into the
    flat vector the circuit's ``("w", i)`` sources index.
    """

    def __init__(self, circuit: Circuit, weight_shapes: Optional[Dict[str, Sequence[int]]] = None,
                 output_wires: Optional[Sequence[int]] = None):
        super().__init__()
        self.circuit = circuit
        if weight_shapes is None:
            weight_shapes = {"weights": (circuit.n_weights,)}
        total = 0
        self._weight_names = []
        for name, shape in weight_shapes.items():
            shape = (shape,) if isinstance(shape, int) else tuple(shape)
            p = nn.Parameter(torch.empty(shape, dtype=torch.float32))
            nn.init.uniform_(p, 0.0, 2.0 * math.pi)
            self.register_parameter(name, p)
            self._weight_names.append(name)
            total += p.numel()
        if total != circuit.n_weights:
            raise ValueError(f"weight_shapes hold {total} angles, circuit expects {circuit.n_weights}")
        self.output_wires = None if output_wires is None else tuple(int(q) for q in output_wires)
        if self.output_wires == tuple(range(circuit.n_qubits)):
            self.output_wires = None

    def flat_weights(self) -> torch.Tensor:
        ps = [getattr(self, n) for n in self._weight_names]
        return ps[0].reshape(-1) if len(ps) == 1 else torch.cat([p.reshape(-1) for p in ps])

    def forward(self, inputs: torch.Tensor) -> torch.Tensor:
        lead = inputs.shape[:-1]
        x = inputs.reshape(-1, inputs.shape[-1]) if inputs.dim() != 2 else inputs
        y = qnn_apply(self.circuit, x, self.flat_weights())
        if self.output_wires is not None:
            y = y[:, list(self.output_wires)]
        return y.reshape(*lead, y.shape[-1]) if inputs.dim() != 2 else y

    def extra_repr(self) -> str:
        return repr(self.circuit)
