"""GPU: the reference-facing module surface (QSC_P128 / TorchLayer) on the CUDA path."""
import copy
import importlib.util
import sys

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import qris_b200
from helpers import assert_close, oracle_ops
from oracle import qnn_oracle as O
from qris_b200 import estimators as E
from qris_b200 import pennylane_shim

pytestmark = pytest.mark.gpu


class OracleQLayer(torch.nn.Module):
    """CPU stand-in used ONLY here as the checker: same weights, circuit evaluated by the oracle."""

    def __init__(self, circuit, weights):
        super().__init__()
        self.circuit, self.weights = circuit, torch.nn.Parameter(weights.detach().double().cpu().clone())

    def forward(self, x):
        return O.simulate(oracle_ops(self.circuit), self.circuit.n_qubits, x.double(), self.weights.reshape(-1))


def oracle_twin(model):
    twin = copy.deepcopy(model).cpu().double()
    twin.qlayer = OracleQLayer(model.qlayer.circuit, model.qlayer.weights)
    return twin


def nine_way_loss(model, batches, labels):
    """Run:341-364: nine forwards, one summed nll_loss / 9, one backward."""
    loss = 0
    for xb, lb in zip(batches, labels):
        loss = loss + F.nll_loss(model(xb), lb) / 9
    return loss


def test_qsc_training_step_matches_oracle_twin():
    torch.manual_seed(0)
    model = E.QSC_P128(use_quantumnat=False, use_gradient_pruning=False).cuda()
    twin = oracle_twin(model)
    g = torch.Generator().manual_seed(1)
    batches = [torch.randn(32, 2, 16, 8, generator=g) for _ in range(9)]
    labels = [torch.randint(0, 3, (32,), generator=g) for _ in range(9)]
    model.train(); twin.train()
    loss = nine_way_loss(model, [b.cuda() for b in batches], [l.cuda() for l in labels])
    loss.backward()
    loss_t = nine_way_loss(twin, [b.double() for b in batches], labels)
    loss_t.backward()
    assert abs(loss.item() - loss_t.item()) < 1e-5
    for (k, p), (_, q) in zip(model.named_parameters(), twin.named_parameters()):
        assert_close(p.grad.cpu().numpy(), q.grad.numpy(), rel=2e-5, scale=float(q.grad.abs().max()), what=k)
    # identical classifier decisions (north_star: identical argmax)
    model.eval(); twin.eval()
    with torch.no_grad():
        xa = torch.cat(batches)
        pa = model(xa.cuda()).argmax(1).cpu(); pb = twin(xa.double()).argmax(1)
    assert torch.equal(pa, pb)
    # an AdamW step as in Run:320,369 moves every parameter, quantum weights included
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=0.01)
    before = model.qlayer.weights.detach().clone()
    opt.step()
    assert not torch.equal(before, model.qlayer.weights.detach())


def test_weight_noise_snapshot_rule():
    """Est:176-196: forward at noisy weights, weights restored before backward; gradient at the noisy point."""
    torch.manual_seed(3)
    model = E.QSC_P128(use_quantumnat=True, use_gradient_pruning=False).cuda().train()
    x = torch.randn(16, 2, 16, 8, device="cuda")
    lab = torch.randint(0, 3, (16,), device="cuda")
    w0 = model.qlayer.weights.detach().clone()
    torch.manual_seed(77)
    out = model(x)
    assert torch.equal(model.qlayer.weights.detach(), w0), "weights must be restored after the forward"
    F.nll_loss(out, lab).backward()
    torch.manual_seed(77)
    noise = model.noise_level * torch.randn_like(w0)
    twin = oracle_twin(model)
    twin.use_quantumnat = False
    twin.qlayer.weights.data.copy_((w0 + noise).double().cpu())
    out_t = twin(x.double().cpu())
    F.nll_loss(out_t, lab.cpu()).backward()
    assert_close(out.detach().cpu().numpy(), out_t.detach().numpy(), what="noisy forward")
    assert_close(model.qlayer.weights.grad.cpu().numpy(), twin.qlayer.weights.grad.numpy(), rel=2e-5,
                 scale=float(twin.qlayer.weights.grad.abs().max()), what="grad at noisy weights")


def test_no_grad_eval_and_dataparallel():
    torch.manual_seed(0)
    model = E.QSC_P128(use_quantumnat=False).cuda().eval()
    x = torch.randn(40, 2, 16, 8, device="cuda")
    with torch.no_grad():
        a = model(x)
        assert not a.requires_grad
        dp = torch.nn.DataParallel(model, device_ids=[0])     # Test.py:78
        b = dp(x)
    assert torch.equal(a, b)
    sd = dp.state_dict()
    assert "module.qlayer.weights" in sd


def test_gradient_pruning_semantics():
    """Est:205-228: entries with |g| <= 0.1 become zero, others unchanged, over ALL parameters."""
    torch.manual_seed(0)
    model = E.QSC_P128(use_quantumnat=False, use_gradient_pruning=True).cuda()
    for p in model.parameters():
        p.grad = torch.randn_like(p) * 0.2
    ref = [torch.where(p.grad.abs() > 0.1, p.grad, torch.zeros_like(p.grad)) for p in model.parameters()]
    model.apply_gradient_pruning()
    for p, r in zip(model.parameters(), ref):
        assert torch.equal(p.grad, r)


def test_float64_inputs_and_unbatched():
    circ = qris_b200.reference_circuit(4, 2)
    layer = qris_b200.QNNLayer(circ, {"weights": (2, 4, 2)}).cuda()
    x = torch.rand(4, device="cuda", dtype=torch.float64)
    y = layer(x)
    assert y.shape == (4,) and y.dtype == torch.float64


def test_graphed_call_replays_the_module_step():
    """qris_b200.GraphedCall: forward + backward of the layer recorded once and replayed for new inputs gives what
    the plain calls give (same kernels, same order), at the reference's call size."""
    circ = qris_b200.reference_circuit(6, 3)
    torch.manual_seed(1)
    layer = qris_b200.QNNLayer(circ).cuda()
    clf = torch.nn.Linear(6, 3).cuda()

    def step(x, lab):
        for p in list(layer.parameters()) + list(clf.parameters()) + [x]:
            p.grad = None
        loss = F.nll_loss(F.log_softmax(clf(layer(x)), dim=1), lab)
        loss.backward()
        return loss.detach(), layer.weights.grad, x.grad
    g = torch.Generator().manual_seed(4)
    x0 = torch.zeros(256, 6, device="cuda", requires_grad=True)
    call = qris_b200.GraphedCall(step, (x0, torch.zeros(256, dtype=torch.long, device="cuda")))
    for it in range(6):                              # three eager calls, the recording, two replays
        x = (2 * torch.rand(256, 6, generator=g) - 1).cuda()
        lab = torch.randint(0, 3, (256,), generator=g).cuda()
        loss, gw, gx = (t.clone() for t in call(x, lab))
        xe = x.clone().requires_grad_(True)
        loss_e, gw_e, gx_e = step(xe, lab)
        assert abs(loss.item() - loss_e.item()) < 1e-6
        assert_close(gw.cpu().numpy(), gw_e.cpu().numpy(), rel=1e-6, scale=float(gw_e.abs().max()), what=f"gw, call {it}")
        assert_close(gx.cpu().numpy(), gx_e.cpu().numpy(), rel=1e-6, scale=float(gx_e.abs().max()), what=f"gx, call {it}")
    assert call.graph is not None
    with pytest.raises(ValueError):
        call(torch.zeros(8, 6, device="cuda"), torch.zeros(8, dtype=torch.long, device="cuda"))


# ---- the reference's OWN file on the GPU --------------------------------------------------------------
class NoisyOracleQLayer(torch.nn.Module):
    """Checker for the reference's noise-and-restore forward: evaluates the circuit with the oracle at
    ``weights + noise_k`` for the k-th call, so that one backward over several calls accumulates the gradient
    taken at each call's noisy weights -- what the reference's autograd graph holds (Est:176-196, Run:364)."""

    def __init__(self, circuit, weights, noises):
        super().__init__()
        self.circuit = circuit
        self.weights = torch.nn.Parameter(weights.detach().double().cpu().clone())
        self.noises = [z.detach().double().cpu() for z in noises]
        self.calls = 0

    def forward(self, x):
        w = self.weights + self.noises[self.calls]
        self.calls += 1
        return O.simulate(oracle_ops(self.circuit), self.circuit.n_qubits, x.double(), w.reshape(-1))


def _load_reference_estimators():
    """Import the reference's Estimators_QuantumNAT_onchipQNN.py, unmodified, with the shim as `pennylane`."""
    from oracle import stage_reference
    path = stage_reference.locate()
    if path is None:
        pytest.skip("the reference's estimator file is neither at /root/reference nor staged in oracle/_ref "
                    "(run __graft_entry__.build() where the reference checkout exists)")
    saved = sys.modules.get("pennylane")
    pennylane_shim.install()
    try:
        spec = importlib.util.spec_from_file_location("reference_estimators_under_test", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        if saved is None:
            sys.modules.pop("pennylane", None)
        else:
            sys.modules["pennylane"] = saved
    assert mod.PENNYLANE_AVAILABLE
    return mod


def test_reference_own_qsc_trains_on_the_gpu_through_the_shim():
    """The reference's own ``QSC_P128`` (Est:107-228, executed from its unmodified file) on the CUDA kernels:
    nine training-mode forwards with its ``.data`` noise / restore sequence exactly as written (Est:176-196),
    one backward (Run:341-364), its own ``apply_gradient_pruning`` (Est:205-228); loss, every gradient, the
    pruned gradients and the eval-mode argmax against the oracle twin."""
    mod = _load_reference_estimators()
    torch.manual_seed(5)
    model = mod.QSC_P128(use_quantumnat=True, use_gradient_pruning=True).cuda().train()
    assert isinstance(model.qlayer, pennylane_shim.TorchLayer) and model.noise_level == 0.01
    g = torch.Generator().manual_seed(2)
    batches = [torch.randn(32, 2, 16, 8, generator=g) for _ in range(9)]
    labels = [torch.randint(0, 3, (32,), generator=g) for _ in range(9)]
    w0 = model.qlayer.weights.detach().clone()
    torch.manual_seed(77)
    loss = nine_way_loss(model, [b.cuda() for b in batches], [l.cuda() for l in labels])
    assert torch.equal(model.qlayer.weights.detach(), w0), "Est:194-196 restores the weights after every forward"
    loss.backward()
    # the nine noise draws, replayed: randn_like(param) on the CUDA generator is the only RNG use in forward
    torch.manual_seed(77)
    noises = [model.noise_level * torch.randn_like(w0) for _ in range(9)]
    twin = copy.deepcopy(model).cpu().double()
    twin.use_quantumnat = False                     # the twin's layer adds the recorded noise itself
    twin.qlayer = NoisyOracleQLayer(model.qlayer.circuit, w0, noises)
    twin.train()
    loss_t = nine_way_loss(twin, [b.double() for b in batches], labels)
    loss_t.backward()
    assert abs(loss.item() - loss_t.item()) < 1e-5
    for (k, p), (_, q) in zip(model.named_parameters(), twin.named_parameters()):
        assert_close(p.grad.cpu().numpy(), q.grad.numpy(), rel=2e-5, scale=float(q.grad.abs().max()), what=k)
    # the reference's pruning, run as written, against the mask applied to the twin's gradients
    model.apply_gradient_pruning()
    for (k, p), (_, q) in zip(model.named_parameters(), twin.named_parameters()):
        ref = torch.where(q.grad.abs() > model.gradient_threshold, q.grad, torch.zeros_like(q.grad))
        away = (q.grad.abs() - model.gradient_threshold).abs() > 1e-5      # not on the threshold's edge
        assert_close(p.grad.cpu().double()[away].numpy(), ref[away].numpy(), rel=2e-5,
                     scale=float(q.grad.abs().max()), what=f"pruned {k}")
    # identical classifier decisions in evaluation mode (no noise: Est:199)
    model.eval(); twin.eval()
    twin.qlayer = OracleQLayer(model.qlayer.circuit, w0)
    with torch.no_grad():
        xa = torch.cat(batches)
        assert torch.equal(model(xa.cuda()).argmax(1).cpu(), twin(xa.double()).argmax(1))
