"""GPU: the rows either side of the hot path: fused optimiser tail (N1), training loop
on device-resident synthetic data (N4), routed NMSE-vs-SNR evaluation (N3), pilot-length variants."""
import copy

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import qris_b200
from helpers import assert_close, oracle_ops
from oracle import qnn_oracle as O
from qris_b200 import data as D
from qris_b200 import estimators as E
from qris_b200 import evaluate, optim, runner

pytestmark = pytest.mark.gpu


class OracleQLayer(torch.nn.Module):      # checker only: the circuit evaluated by the CPU oracle
    def __init__(self, circuit, weights):
        super().__init__()
        self.circuit, self.weights = circuit, torch.nn.Parameter(weights.detach().double().cpu().clone())

    def forward(self, x):
        return O.simulate(oracle_ops(self.circuit), self.circuit.n_qubits, x.double(), self.weights.reshape(-1))


def oracle_twin(model):
    twin = copy.deepcopy(model).cpu().double()
    twin.qlayer = OracleQLayer(model.qlayer.circuit, model.qlayer.weights)
    return twin


@pytest.mark.parametrize("threshold", [None, 0.1])
def test_fused_pruned_adamw_matches_torch(threshold):
    """Est:205-228 pruning + torch.optim.AdamW (Run:320) against the one-kernel version, several steps."""
    torch.manual_seed(0)
    ref = torch.nn.Sequential(torch.nn.Linear(40, 30), torch.nn.Tanh(), torch.nn.Linear(30, 7)).cuda()
    fast = copy.deepcopy(ref)
    opt_ref = torch.optim.AdamW(ref.parameters(), lr=1e-3, weight_decay=0.01)
    opt_fast = optim.FlatPrunedAdamW(fast.parameters(), lr=1e-3, weight_decay=0.01, prune_threshold=threshold)
    n_pruned_ref = 0
    for step in range(6):
        x = torch.randn(64, 40, device="cuda"); y = torch.randint(0, 7, (64,), device="cuda")
        opt_ref.zero_grad(); opt_fast.zero_grad()
        (20 * F.cross_entropy(ref(x), y)).backward()
        (20 * F.cross_entropy(fast(x), y)).backward()
        if threshold is not None:                      # the reference's pruning, verbatim semantics
            for p in ref.parameters():
                mask = p.grad.abs() > threshold
                n_pruned_ref += int((~mask).sum())
                p.grad = p.grad * mask.float()
        opt_ref.step(); opt_fast.step()
    for a, b in zip(ref.parameters(), fast.parameters()):
        assert_close(b.detach().cpu().numpy(), a.detach().cpu().numpy(), rel=2e-6, what="parameters after 6 steps")
    if threshold is not None:
        total = 6 * sum(p.numel() for p in ref.parameters())
        assert abs(opt_fast.pruned_fraction() - n_pruned_ref / total) < 2e-3 and n_pruned_ref > 0
    assert "0.weight" in fast.state_dict()             # Parameter objects (and checkpoint keys) untouched


def test_qsc_training_step_matches_oracle_twin_with_torch_adamw():
    """One full Run:338-369 step through runner.qsc_step (kernels + fused optimiser) vs the oracle twin."""
    r = runner.Y2HRunner(data_len=200, batch_size_DML=16, device="cuda", verbose=False)
    torch.manual_seed(0)
    model = E.QSC_P128(use_quantumnat=False, use_gradient_pruning=False).cuda().train()
    twin = oracle_twin(model).train()
    opt = optim.FlatPrunedAdamW(model.parameters(), lr=1e-3, weight_decay=0.01)
    opt_t = torch.optim.AdamW(twin.parameters(), lr=1e-3, weight_decay=0.01)
    train_loader, _ = r.get_dataloader_DML()
    sutd = next(iter(train_loader))
    loss = r.qsc_step(model, opt, sutd)
    opt_t.zero_grad()
    loss_t = 0
    for sid in range(3):
        for uid in range(3):
            yp, _, _, ind = sutd[sid][uid]
            loss_t = loss_t + F.nll_loss(twin(D.pilot_image(yp, 128).cpu().double()), ind.cpu()) / 9
    loss_t.backward(); opt_t.step()
    assert abs(loss.item() - loss_t.item()) < 1e-5
    for (k, p), (_, q) in zip(model.named_parameters(), twin.named_parameters()):
        a, b = p.detach().cpu().numpy().copy(), q.detach().numpy().copy()
        if k == "qlayer.weights":
            # the last layer's RZ angles have an identically zero gradient; Adam divides
            # rounding noise by itself there, so those six entries are not comparable across precisions
            a[-1, :, 1] = b[-1, :, 1] = 0
        assert_close(a, b, rel=5e-6, what=k)


def test_short_training_run_learns_and_saves(tmp_path):
    r = runner.Y2HRunner(data_len=1500, batch_size_DML=64, n_epochs=3, lr=5e-3, device="cuda", verbose=False,
                         workdir=str(tmp_path))
    hist = r.train_QSC_P128()
    assert all(np.isfinite(hist["train_loss"])) and hist["train_loss"][-1] < hist["train_loss"][0]
    assert 0.9 < hist["train_loss"][0] < 1.2             # starts near ln 3, as the reference's loss curve does
    assert hist["val_acc"][-1] > 0.6                     # three separable scenarios: well above chance
    saved = sorted(p.name for p in (tmp_path / "Pn_128" / "HDCE").iterdir())
    assert "QSC_OPT_64_10dB_best_DML.pth" in saved and "QSC_OPT_64_10dB_epoch2_DML.pth" in saved
    sd = torch.load(tmp_path / "Pn_128" / "HDCE" / "QSC_OPT_64_10dB_best_DML.pth")
    E.QSC_P128().load_state_dict(sd)                     # reference-compatible keys


@pytest.mark.parametrize("concat", [False, True])
def test_graph_replayed_training_steps_match_eager_steps(concat):
    """runner.GraphedQSCStep: three eager steps, one recorded, the rest replayed - the parameters after eight
    steps equal those of eight eager steps on the same batches (same kernels, same order; cuDNN's convolution
    backward is the only non-bit-exact part)."""
    r = runner.Y2HRunner(data_len=1280, batch_size_DML=128, device="cuda", verbose=False)
    loader, _ = r.get_dataloader_DML()
    batches = list(loader)[:8]
    assert len(batches) == 8
    finals, losses = [], []
    for graphed in (False, True):
        torch.manual_seed(5)
        model = E.QSC_P128(use_quantumnat=False, use_gradient_pruning=True).cuda().train()
        opt = optim.FlatPrunedAdamW(model.parameters(), lr=2e-3, weight_decay=0.01, prune_threshold=1e-3,
                                    device_step_count=True)
        step = runner.GraphedQSCStep(r, model, opt, concat) if graphed else (lambda b: r.qsc_step(model, opt, b, concat))
        ls = [float(step(b)) for b in batches]
        if graphed:
            assert step.graph is not None and step.calls == 8
        assert int(opt.steps_done.item()) == 8
        finals.append(opt.flat.clone()); losses.append(ls)
    assert np.allclose(losses[0], losses[1], rtol=1e-4, atol=1e-5)
    scale = float(finals[0].abs().max())
    assert float((finals[0] - finals[1]).abs().max()) <= 2e-5 * scale


def test_graphed_training_run_with_weight_noise_learns_like_the_eager_run():
    """train_QSC_P128(cuda_graph=True) with the QuantumNAT weight noise (torch's RNG inside the recording): the
    loss curve equals the eager run's; the last (shorter) batch of an epoch runs eagerly.  (With the
    reference's pruning threshold of 0.1 neither run learns - every gradient entry is below it.)"""
    hists = []
    for graph in (False, True):
        r = runner.Y2HRunner(data_len=1500, batch_size_DML=64, n_epochs=3, lr=5e-3, device="cuda", verbose=False)
        hists.append(r.train_QSC_P128(use_quantumnat=True, use_gradient_pruning=False, cuda_graph=graph, save=False))
    assert np.allclose(hists[0]["train_loss"], hists[1]["train_loss"], rtol=2e-3)
    assert hists[1]["train_loss"][-1] < hists[1]["train_loss"][0] and hists[1]["val_acc"][-1] > 0.6


def test_device_step_count_matches_host_step_count():
    """qnn_pruned_adamw_step_counted derives the bias corrections from the device counter: same parameters
    as qnn_pruned_adamw_step given the step number from the host."""
    torch.manual_seed(0)
    ps = [[torch.nn.Parameter(torch.randn(257, device="cuda")), torch.nn.Parameter(torch.randn(3, 5, device="cuda"))] for _ in range(2)]
    ps[1][0].data.copy_(ps[0][0].data); ps[1][1].data.copy_(ps[0][1].data)
    opts = [optim.FlatPrunedAdamW(ps[0], lr=1e-2, prune_threshold=0.05), optim.FlatPrunedAdamW(ps[1], lr=1e-2, prune_threshold=0.05, device_step_count=True)]
    for it in range(7):
        g = [torch.randn(257, device="cuda"), torch.randn(3, 5, device="cuda")]
        for o, pp in zip(opts, ps):
            o.zero_grad()
            for p, gi in zip(pp, g):
                p.grad.add_(gi)
            o.step()
    assert int(opts[1].steps_done.item()) == 7
    assert float((opts[0].flat - opts[1].flat).abs().max()) <= 2e-6
    assert abs(opts[0].pruned_fraction() - opts[1].pruned_fraction()) < 1e-12


def test_routed_nmse_parity_kernel_vs_oracle():
    """BASELINE config 5's criterion at small scale: estimator NMSE with the kernel-backed quantum
    classifier within 0.05 dB of the same pipeline driven by the oracle, identical argmax."""
    r = runner.Y2HRunner(data_len=1000, batch_size_DML=50, n_epochs=3, lr=5e-3, device="cuda", verbose=False)
    r.train_QSC_P128(save=False)
    r.train_Conv_Linear_of_HDCE(n_epochs=1, max_iters_per_epoch=6)
    gen = D.SyntheticRIS(128, seed=0, device="cuda")
    twin = oracle_twin(r.model_QSC).eval()

    class OracleClassifier(torch.nn.Module):             # evaluates on the CPU oracle, answers on the GPU
        def forward(self, img):
            with torch.no_grad():
                return twin(img.cpu().double()).float().cuda()
    res = evaluate.snr_sweep(gen, {"quantum": r.model_QSC, "quantum_oracle": OracleClassifier()}, r.convs, r.fc,
                             snrs=(5, 15), n_per_pair=60, batch=60)
    for a, b in zip(res["quantum"], res["quantum_oracle"]):
        assert abs(a - b) < 0.05
    assert res["acc_quantum"] == res["acc_quantum_oracle"]
    assert all(np.isfinite(res["LS"]))


@pytest.mark.parametrize("pilot", [32, 64, 128, 256])
def test_pilot_length_variants(pilot):
    """BASELINE config 5 at small scale, for every pilot length of the sweep: shapes and gradients, then a short
    training run and the routed NMSE-vs-SNR evaluation with the kernel-backed and the oracle-backed classifier:
    NMSE within 0.05 dB, identical accuracy (profiles/cfg5_sweep.py is the full-size run)."""
    hw = D.pilot_shape(pilot)
    model = E.QSC_P128(use_quantumnat=False, use_gradient_pruning=False, pilot_hw=hw).cuda()
    gen = D.SyntheticRIS(pilot, seed=1, device="cuda")
    yp, _, _, ind = gen.pair_set(2, 1, 24, 10.0, 3)
    out = model(D.pilot_image(yp, pilot))
    assert out.shape == (24, 3)
    F.nll_loss(out, ind).backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in model.parameters())
    r = runner.Y2HRunner(Pilot_num=pilot, data_len=600, batch_size_DML=50, n_epochs=2, lr=5e-3, device="cuda", verbose=False)
    r.train_QSC_P128(save=False)
    r.train_Conv_Linear_of_HDCE(n_epochs=1, max_iters_per_epoch=4)
    twin = oracle_twin(r.model_QSC).eval()

    class OracleClassifier(torch.nn.Module):
        def forward(self, img):
            with torch.no_grad():
                return twin(img.cpu().double()).float().cuda()
    res = evaluate.snr_sweep(D.SyntheticRIS(pilot, seed=0, device="cuda"), {"quantum": r.model_QSC, "quantum_oracle": OracleClassifier()},
                             r.convs, r.fc, snrs=(5, 15), n_per_pair=40, batch=40)
    assert all(abs(a - b) < 0.05 for a, b in zip(res["quantum"], res["quantum_oracle"]))
    assert res["acc_quantum"] == res["acc_quantum_oracle"] and all(np.isfinite(res["LS"]))


def test_concatenated_users_step_equals_nine_way_step():
    """one call on the concatenated user batches gives the same step as nine calls."""
    r = runner.Y2HRunner(data_len=200, batch_size_DML=16, device="cuda", verbose=False)
    torch.manual_seed(0)
    m1 = E.QSC_P128(use_quantumnat=False, use_gradient_pruning=False).cuda().train()
    m2 = copy.deepcopy(m1)
    o1 = optim.FlatPrunedAdamW(m1.parameters(), lr=1e-3); o2 = optim.FlatPrunedAdamW(m2.parameters(), lr=1e-3)
    sutd = next(iter(r.get_dataloader_DML()[0]))
    l1 = r.qsc_step(m1, o1, sutd, concat_users=False); l2 = r.qsc_step(m2, o2, sutd, concat_users=True)
    assert abs(l1.item() - l2.item()) < 1e-6
    assert_close(o2.bucket.flat.cpu().numpy(), o1.bucket.flat.cpu().numpy(), rel=1e-5, scale=float(o1.bucket.flat.abs().max()), what="grads")


def test_quantumnat_classifier_trains():
    """BASELINE config 3's model: 10-qubit U3/CU3 QNN + weight noise + post-measurement QuantumNAT head."""
    torch.manual_seed(0)
    model = E.QSC_QuantumNAT().cuda().train()
    assert model.qlayer.weights.shape == (3, 2, 10, 3)
    gen = D.SyntheticRIS(128, seed=2, device="cuda")
    opt = torch.optim.AdamW(model.parameters(), lr=5e-3)
    losses = []
    for it in range(12):
        yp, ind = zip(*[(gen.pair_set(s, it % 3, 48, 10.0, it)[0], gen.pair_set(s, it % 3, 48, 10.0, it)[3]) for s in range(3)])
        opt.zero_grad()
        loss = F.nll_loss(model(D.pilot_image(torch.cat(yp), 128)), torch.cat(ind))
        loss.backward(); opt.step(); losses.append(loss.item())
    assert all(np.isfinite(losses)) and min(losses[-4:]) < losses[0]
    model.eval()
    with torch.no_grad():
        out = model(D.pilot_image(gen.pair_set(0, 0, 32, 10.0, 99)[0], 128))
    assert out.shape == (32, 3) and torch.isfinite(out).all()
