"""GPU: the reference-facing module surface (QSC_P128 / TorchLayer) on the CUDA path."""
import copy

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import qris_b200
from helpers import assert_close, oracle_ops
from oracle import qnn_oracle as O
from qris_b200 import estimators as E

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
