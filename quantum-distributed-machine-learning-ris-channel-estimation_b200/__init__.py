"""B200-native QNN statevector hot path (fused sm_100a kernels behind a C ABI) with the reference's
PyTorch module surface.  See DESIGN.md / INTEGRATION.md at the repository root."""
from . import _lib
from ._lib import QnnError
from .circuit import Circuit, Op, reference_circuit, u3cu3_circuit
from .qlayer import QNNLayer, qnn_apply, qnn_backward_raw, qnn_forward_raw
from .qnat import QuantumNATHead
from .head import linear_tanh
from .graphs import GraphedCall

__all__ = ["Circuit", "Op", "QNNLayer", "QuantumNATHead", "QnnError", "reference_circuit",
           "u3cu3_circuit", "qnn_apply", "qnn_forward_raw", "qnn_backward_raw", "linear_tanh", "GraphedCall"]
