"""Host-side mirror of the reference's model zoo for pilot length 128
(Estimators_QuantumNAT_onchipQNN.py), with the quantum layer bound to libqnn_b200.so.

Only ``QSC_P128`` touches the hot path.  Class names, constructor arguments, attribute names and
``state_dict`` keys match the reference so checkpoints and callers (Runner_P128_...py:313-426,
Test.py:23-97) carry over.  The classical networks are stock PyTorch, as north_star prescribes.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn

from . import head as _head
from . import pennylane_shim as qml


def _conv_bn_relu_stack(channels: int = 32, depth: int = 3) -> nn.Sequential:
    mods, c_in = [], 2
    for _ in range(depth):
        mods += [nn.Conv2d(c_in, channels, 3, stride=1, padding=1, bias=False),
                 nn.BatchNorm2d(channels), nn.ReLU(inplace=True)]
        c_in = channels
    return nn.Sequential(*mods)


class DCE_P128(nn.Module):
    """Monolithic conv + FC estimator (Est:40-75): (B,2,16,8) -> (B,2048)."""

    def __init__(self):
        super().__init__()
        self.features, self.kernel_size, self.padding = 32, 3, 1
        self.cnn = _conv_bn_relu_stack(self.features)
        self.FC = nn.Linear(self.features * 16 * 8, 64 * 16 * 2)

    def forward(self, x):
        return self.FC(self.cnn(x).flatten(1))


class SC_P128(nn.Module):
    """Classical scenario classifier (Est:79-101)."""

    def __init__(self, pilot_hw=(16, 8)):
        super().__init__()
        self.conv1 = nn.Conv2d(2, 32, 3, padding=1, bias=False)
        self.conv2 = nn.Conv2d(32, 32, 3, padding=1, bias=False)
        self.FC = nn.Linear(32 * (pilot_hw[0] // 4) * (pilot_hw[1] // 4), 3)

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.conv1(x)), 2, 2)
        x = F.max_pool2d(F.relu(self.conv2(x)), 2, 2)
        return F.log_softmax(self.FC(x.flatten(1)), dim=1)


class QSC_P128(nn.Module):
    """Quantum scenario classifier (Est:107-228): CNN -> tanh angles -> QNN -> linear -> log-softmax.

    ``pilot_hw`` (default 16 x 8 = the reference's P = 128, Run:345) sizes the flatten layer as
    ``32 * (h/4) * (w/4)`` for the pilot-length sweep of BASELINE config 5; everything else is the
    reference's.

    Differences from the reference are confined to what runs the circuit: ``self.qlayer`` is a
    ``TorchLayer`` over the CUDA kernels.  Construction consumes torch's RNG in the reference's order
    (quantum weights, then preprocess, then classifier: Est:149,152,165), so a given
    ``torch.manual_seed`` yields the same initial model.
    """

    def __init__(self, n_qubits=6, n_layers=3, n_classes=3, use_quantumnat=True, use_gradient_pruning=True,
                 pilot_hw=(16, 8)):
        super().__init__()
        self.use_quantum = True
        self.num_qubits = n_qubits
        self.n_classes = n_classes
        self.use_quantumnat = use_quantumnat
        self.use_gradient_pruning = use_gradient_pruning
        self.noise_level = 0.01 if use_quantumnat else 0.0      # Est:118
        self.gradient_threshold = 0.1                           # Est:119

        self.dev = qml.device("default.qubit", wires=n_qubits)  # Est:122

        def quantum_circuit(inputs, weights):                   # Est:125-142
            qml.templates.AngleEmbedding(inputs, wires=range(n_qubits), rotation="Y")
            for layer in range(n_layers):
                for q in range(n_qubits):
                    qml.RY(weights[layer, q, 0], wires=q)
                    qml.RZ(weights[layer, q, 1], wires=q)
                for q in range(n_qubits - 1):
                    qml.CNOT(wires=[q, q + 1])
                qml.CNOT(wires=[n_qubits - 1, 0])
            return [qml.expval(qml.PauliZ(q)) for q in range(n_qubits)]

        qnode = qml.QNode(quantum_circuit, self.dev, interface="torch")                  # Est:148
        self.qlayer = qml.qnn.TorchLayer(qnode, {"weights": (n_layers, n_qubits, 2)})    # Est:145,149

        self.preprocess = nn.Sequential(                         # Est:152-162
            nn.Conv2d(2, 16, 3, stride=1, padding=1), nn.ReLU(), nn.MaxPool2d(2),
            nn.Conv2d(16, 32, 3, stride=1, padding=1), nn.ReLU(), nn.MaxPool2d(2),
            nn.Flatten(), nn.Linear(32 * (pilot_hw[0] // 4) * (pilot_hw[1] // 4), n_qubits), nn.Tanh())
        self.classifier = nn.Linear(n_qubits, n_classes)         # Est:165

    def _angles(self, x):
        """``self.preprocess(x)`` (Est:152-162, 173).  For large CUDA batches its last two stages, Linear +
        Tanh, run as the fused head kernels with the weights of the unchanged ``preprocess[7]`` module."""
        lin = self.preprocess[7]
        if x.is_cuda and x.shape[0] >= _head.MIN_BATCH and _head.supported(lin.in_features, lin.out_features):
            return _head.linear_tanh(self.preprocess[:7](x), lin.weight, lin.bias)
        return self.preprocess(x)

    def forward(self, x):
        angles = self._angles(x)
        if self.training and self.use_quantumnat and self.noise_level > 0:
            # Weight-space noise for this forward only (Est:176-196).  The layer snapshots the
            # weights it ran with, so the gradient is taken at the noisy point, as in the reference.
            params = list(self.qlayer.parameters())
            with torch.no_grad():
                saved = [p.data.clone() for p in params]
                for p in params:
                    p.data.add_(self.noise_level * torch.randn_like(p))
            measured = self.qlayer(angles)
            with torch.no_grad():
                for p, s in zip(params, saved):
                    p.data.copy_(s)
        else:
            measured = self.qlayer(angles)
        return F.log_softmax(self.classifier(measured), dim=1)

    def apply_gradient_pruning(self):
        """On-chip-QNN gradient pruning (Est:205-228): zero every gradient entry with |g| <= threshold.

        Same result as the reference, without its per-parameter host synchronisation: the counts stay
        on the device and are read back once."""
        if not self.use_gradient_pruning:
            return
        grads = [p.grad for p in self.parameters() if p.grad is not None]
        if not grads:
            return
        pruned = torch.zeros((), dtype=torch.long, device=grads[0].device)
        total = 0
        for g in grads:
            keep = g.abs() > self.gradient_threshold
            pruned += (~keep).sum()
            g.mul_(keep.to(g.dtype))
            total += g.numel()
        n_pruned = int(pruned.item())
        if total > 0 and n_pruned > 0 and n_pruned / total > 0.1:
            print(f"Gradient pruning: {n_pruned / total:.1%} gradients pruned")


class QSC_QuantumNAT(nn.Module):
    """north_star's on-chip QNN classifier (BASELINE config 3; not in the reference): the QSC front end
    feeding a U3/CU3 circuit, with QuantumNAT weight noise (as Est:176-196) AND the post-measurement
    normalise / noise / quantise head, then the linear classifier.  n_qubits = 10, 3 layers by default."""

    def __init__(self, n_qubits=10, n_layers=3, n_classes=3, noise_level=0.01, act_noise=0.01, levels=5,
                 quant_range=(-2.0, 2.0), pilot_hw=(16, 8), process_group=None):
        super().__init__()
        from .circuit import u3cu3_circuit
        from .qlayer import QNNLayer
        from .qnat import QuantumNATHead
        self.num_qubits, self.n_classes, self.noise_level = n_qubits, n_classes, noise_level
        self.qlayer = QNNLayer(u3cu3_circuit(n_qubits, n_layers), {"weights": (n_layers, 2, n_qubits, 3)})
        self.preprocess = nn.Sequential(
            nn.Conv2d(2, 16, 3, stride=1, padding=1), nn.ReLU(), nn.MaxPool2d(2),
            nn.Conv2d(16, 32, 3, stride=1, padding=1), nn.ReLU(), nn.MaxPool2d(2),
            nn.Flatten(), nn.Linear(32 * (pilot_hw[0] // 4) * (pilot_hw[1] // 4), n_qubits), nn.Tanh())
        self.nat = QuantumNATHead(sigma=act_noise, lo=quant_range[0], hi=quant_range[1], levels=levels,
                                  process_group=process_group)
        self.classifier = nn.Linear(n_qubits, n_classes)

    def forward(self, x):
        angles = self.preprocess(x)
        if self.training and self.noise_level > 0:
            w = self.qlayer.weights
            with torch.no_grad():
                saved = w.data.clone()
                w.data.add_(self.noise_level * torch.randn_like(w))
            measured = self.qlayer(angles)
            with torch.no_grad():
                w.data.copy_(saved)
        else:
            measured = self.qlayer(angles)
        return F.log_softmax(self.classifier(self.nat(measured)), dim=1)


class Conv_P128(nn.Module):
    """HDCE feature extractor (Est:237-268): (B,2,16,8) -> (B,4096)."""

    def __init__(self, pilot_hw=(16, 8)):
        super().__init__()
        self.features, self.kernel_size, self.padding = 32, 3, 1
        self.cnn = _conv_bn_relu_stack(self.features)

    def forward(self, x):
        return self.cnn(x).flatten(1)


class FC_P128(nn.Module):
    """HDCE feature mapper (Est:272-279): (B,4096) -> (B,2048)."""

    def __init__(self, pilot_hw=(16, 8)):
        super().__init__()
        self.FC = nn.Linear(32 * pilot_hw[0] * pilot_hw[1], 64 * 16 * 2)

    def forward(self, x):
        return self.FC(x)


def NMSE_cuda(x_hat, x):
    """sum((x_hat-x)^2) / sum(x^2) over the whole batch (Est:282-286)."""
    return torch.sum((x_hat - x) ** 2) / torch.sum(x ** 2)


class NMSELoss(nn.Module):
    def forward(self, x_hat, x):
        return NMSE_cuda(x_hat, x)
