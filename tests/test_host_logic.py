"""CPU: host-side logic of the drop-in surface (no kernel is launched)."""
import importlib.util
import os
import sys

import pytest
import torch

import qris_b200
from qris_b200 import estimators as E
from qris_b200 import pennylane_shim as qml

from oracle import stage_reference

REF = stage_reference.locate() or stage_reference.SOURCE


def test_reference_circuit_gate_order():
    c = qris_b200.reference_circuit(3, 2)
    kinds = [(o.kind, o.wires) for o in c.ops]
    assert kinds[:3] == [("RY", (0,)), ("RY", (1,)), ("RY", (2,))]
    layer = [("RY", (0,)), ("RZ", (0,)), ("RY", (1,)), ("RZ", (1,)), ("RY", (2,)), ("RZ", (2,)),
             ("CNOT", (0, 1)), ("CNOT", (1, 2)), ("CNOT", (2, 0))]
    assert kinds[3:12] == layer and kinds[12:] == layer
    assert c.ops[3].params == (("w", 0),) and c.ops[4].params == (("w", 1),)
    assert c.ops[12 + 4].params == (("w", (1 * 3 + 2) * 2),)
    assert c.n_weights == 12 and c.n_inputs == 3


def test_qsc_state_dict_and_attributes():
    torch.manual_seed(0)
    m = E.QSC_P128()
    sd = m.state_dict()
    assert {k: tuple(v.shape) for k, v in sd.items()} == {
        "qlayer.weights": (3, 6, 2), "preprocess.0.weight": (16, 2, 3, 3), "preprocess.0.bias": (16,),
        "preprocess.3.weight": (32, 16, 3, 3), "preprocess.3.bias": (32,), "preprocess.7.weight": (6, 256),
        "preprocess.7.bias": (6,), "classifier.weight": (3, 6), "classifier.bias": (3,)}
    assert sum(p.numel() for p in m.parameters()) == 6543
    for attr in ("use_quantum", "num_qubits", "n_classes", "use_quantumnat", "use_gradient_pruning",
                 "noise_level", "gradient_threshold", "dev", "qlayer", "preprocess", "classifier"):
        assert hasattr(m, attr)
    assert m.noise_level == 0.01 and E.QSC_P128(use_quantumnat=False).noise_level == 0.0
    w = sd["qlayer.weights"]
    assert 0 <= w.min() and w.max() <= 2 * torch.pi
    m2 = E.QSC_P128()
    m2.load_state_dict(sd)


def test_rng_consumption_order_matches_reference():
    """Quantum weights are drawn first (Est:149), then preprocess (Est:152), then classifier (Est:165)."""
    torch.manual_seed(123)
    m = E.QSC_P128()
    torch.manual_seed(123)
    w = torch.empty(3, 6, 2).uniform_(0, 2 * torch.pi)
    assert torch.equal(m.qlayer.weights.detach(), w)


def test_shim_traces_the_reference_circuit():
    m = E.QSC_P128(n_qubits=5, n_layers=2)
    assert m.qlayer.circuit.ops == qris_b200.reference_circuit(5, 2).ops
    assert m.qlayer.output_wires is None


def test_shim_measurement_subset_and_errors():
    dev = qml.device("default.qubit", wires=3)

    def circ(inputs, theta):
        qml.templates.AngleEmbedding(inputs, wires=range(3), rotation="X")
        qml.U3(theta[0], theta[1], 0.5, wires=1)
        qml.CZ(wires=[1, 2])
        return [qml.expval(qml.PauliZ(2)), qml.expval(qml.PauliZ(0))]
    layer = qml.qnn.TorchLayer(qml.QNode(circ, dev), {"theta": (2,)})
    assert layer.output_wires == (2, 0)
    assert layer.circuit.ops[3].params == (("w", 0), ("w", 1), ("c", 0.5))

    def bad(inputs, theta):
        qml.RY(2 * inputs[0], wires=0)
        return [qml.expval(qml.PauliZ(0))]
    with pytest.raises(TypeError):
        qml.qnn.TorchLayer(qml.QNode(bad, dev), {"theta": (1,)})
    with pytest.raises(RuntimeError):
        qml.RY(0.1, wires=0)


def test_cpu_tensor_is_refused_loudly():
    m = E.QSC_P128()
    with pytest.raises(qris_b200.QnnError, match="no CPU or PyTorch fallback"):
        m(torch.zeros(2, 2, 16, 8))


def test_product_never_imports_the_oracle():
    pkg = os.path.dirname(qris_b200.__file__)
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f


@pytest.mark.skipif(not os.path.exists(stage_reference.SOURCE), reason="reference checkout not present")
def test_staged_reference_file_is_the_reference_byte_for_byte():
    """oracle/_ref/ carries the reference's estimator file to the GPU box; it must be the unmodified file."""
    assert stage_reference.stage()
    assert open(stage_reference.STAGED, "rb").read() == open(stage_reference.SOURCE, "rb").read()
    assert open(stage_reference.STAGED + ".sha256").read().strip() == stage_reference.sha256(stage_reference.SOURCE)


@pytest.mark.skipif(not os.path.exists(REF), reason="reference file neither checked out nor staged")
def test_unmodified_reference_module_builds_on_the_shim():
    """With the shim installed as `pennylane`, the reference's own Estimators file constructs QSC_P128
    and the traced gate list equals ours.  (Structure only: nothing is executed on the CPU.)"""
    saved = sys.modules.get("pennylane")
    qml.install()
    try:
        spec = importlib.util.spec_from_file_location("ref_estimators", REF)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        assert mod.PENNYLANE_AVAILABLE
        torch.manual_seed(1)
        ref_model = mod.QSC_P128(use_quantumnat=False, use_gradient_pruning=False)
        torch.manual_seed(1)
        ours = E.QSC_P128(use_quantumnat=False, use_gradient_pruning=False)
        assert ref_model.qlayer.circuit.ops == ours.qlayer.circuit.ops
        a, b = ref_model.state_dict(), ours.state_dict()
        assert list(a) == list(b)
        for k in a:
            assert torch.equal(a[k], b[k]), k
    finally:
        if saved is None:
            sys.modules.pop("pennylane", None)
        else:
            sys.modules["pennylane"] = saved
