"""CPU tests that pin the oracle (oracle/qnn_oracle.py, oracle/qnat_oracle.py).

The reference ships no tests or vectors and PennyLane is unavailable, so the pins are: analytic closed
forms, structural zero-gradient facts, parameter-shift identities, finite differences, the fixed vector
of SURVEY.md section 3.3, agreement of two independent implementations, and the committed golden file.
"""
import os

import numpy as np
import pytest
import torch

from oracle import qnat_oracle as QO
from oracle import qnn_oracle as O

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "qnn_golden.npz")


@pytest.mark.parametrize("n", [1, 2, 3, 6, 9])
def test_L1_closed_form(n):
    rng = np.random.default_rng(n)
    x = rng.uniform(-1, 1, (7, n))
    w = rng.uniform(0, 2 * np.pi, (1, n, 2))
    y = O.simulate(O.ref_circuit_ops(n, 1), n, torch.tensor(x), torch.tensor(w.reshape(-1))).numpy()
    if n == 1:   # no entangler: y = cos(x + w)
        np.testing.assert_allclose(y, np.cos(x + w[0, :, 0]), atol=1e-13)
    else:
        np.testing.assert_allclose(y, O.ref_circuit_L1_closed_form(x, w), atol=1e-13)


def test_no_layers_gives_cos_x():
    x = torch.linspace(-1, 1, 10, dtype=torch.float64).reshape(2, 5)
    y = O.simulate(O.ref_circuit_ops(5, 0), 5, x, torch.zeros(0, dtype=torch.float64))
    np.testing.assert_allclose(y.numpy(), np.cos(x.numpy()), atol=1e-14)


def test_survey_fixed_vector():
    x, W = O.survey_fixed_inputs()
    fv = O.SURVEY_FIXED_VECTOR
    y, gx, gw = O.forward_and_grads(O.ref_circuit_ops(6, 3), 6, torch.tensor(x), torch.tensor(W.reshape(-1)),
                                    torch.tensor([fv["g"]], dtype=torch.float64))
    np.testing.assert_allclose(y.numpy()[0], fv["y"], atol=5e-12)
    gW = gw.numpy().reshape(3, 6, 2)
    assert abs(gx[0, 0].item() - fv["dx0"]) < 5e-8
    assert abs(gW[1, 2, 0] - fv["dW_1_2_0"]) < 5e-8
    assert abs(gW[2, 3, 0] - fv["dW_2_3_0"]) < 5e-8
    assert abs(gW[2, 3, 1]) < 1e-13


def test_last_layer_rz_gradient_is_zero():
    n, L = 5, 3
    rng = np.random.default_rng(3)
    x = torch.tensor(rng.uniform(-1, 1, (4, n)))
    w = torch.tensor(rng.uniform(0, 2 * np.pi, L * n * 2))
    gy = torch.tensor(rng.normal(size=(4, n)))
    _, _, gw = O.forward_and_grads(O.ref_circuit_ops(n, L), n, x, w, gy)
    assert np.abs(gw.numpy().reshape(L, n, 2)[L - 1, :, 1]).max() < 1e-13


@pytest.mark.parametrize("builder,n,L,enc", [("ref", 4, 2, "none"), ("u3", 3, 2, "none"), ("u3amp", 3, 1, "amplitude")])
def test_two_implementations_and_finite_differences(builder, n, L, enc):
    ops = O.ref_circuit_ops(n, L) if builder == "ref" else O.u3cu3_circuit_ops(n, L, angle_encoder=(enc == "none"))
    nw = O.n_weights_of(ops)
    rng = np.random.default_rng(11)
    n_in = n if enc == "none" else 1 << n
    x = rng.uniform(-1, 1, (3, n_in)); w = rng.uniform(0, 2 * np.pi, nw); gy = rng.normal(size=(3, n))
    y, gx, gw = O.forward_and_grads(ops, n, torch.tensor(x), torch.tensor(w), torch.tensor(gy), encoder=enc)
    y2, gx2, gw2 = O.adjoint_vjp_numpy(ops, n, x, w, gy, encoder=enc)
    np.testing.assert_allclose(y.numpy(), y2, atol=1e-13)
    np.testing.assert_allclose(gx.numpy(), gx2, atol=1e-12)
    np.testing.assert_allclose(gw.numpy(), gw2, atol=1e-12)

    def loss(xx, ww):
        return float((O.simulate(ops, n, torch.tensor(xx), torch.tensor(ww), encoder=enc).numpy() * gy).sum())
    eps = 1e-6
    for i in rng.choice(nw, size=min(nw, 5), replace=False):
        wp, wm = w.copy(), w.copy(); wp[i] += eps; wm[i] -= eps
        assert abs((loss(x, wp) - loss(x, wm)) / (2 * eps) - gw2[i]) < 1e-7
    for (b, j) in [(0, 0), (2, n_in - 1)]:
        xp, xm = x.copy(), x.copy(); xp[b, j] += eps; xm[b, j] -= eps
        assert abs((loss(xp, w) - loss(xm, w)) / (2 * eps) - gx2[b, j]) < 1e-7


def test_parameter_shift_identity():
    """d<Z>/dtheta = (y(theta + pi/2) - y(theta - pi/2)) / 2 for RY/RZ angles (each used once)."""
    n, L = 4, 2
    ops = O.ref_circuit_ops(n, L)
    rng = np.random.default_rng(5)
    x = rng.uniform(-1, 1, (2, n)); w = rng.uniform(0, 2 * np.pi, L * n * 2); gy = rng.normal(size=(2, n))
    _, _, gw = O.adjoint_vjp_numpy(ops, n, x, w, gy)
    for i in range(w.size):
        wp, wm = w.copy(), w.copy(); wp[i] += np.pi / 2; wm[i] -= np.pi / 2
        yp = O.simulate(ops, n, torch.tensor(x), torch.tensor(wp)).numpy()
        ym = O.simulate(ops, n, torch.tensor(x), torch.tensor(wm)).numpy()
        assert abs(((yp - ym) / 2 * gy).sum() - gw[i]) < 1e-12


def test_u3_reduces_to_ry_and_cu3_to_cnot_like():
    """U3(t,0,0) = RY(t); a CU3(pi,0,pi) acts as CNOT: ties the new gates to the reference's set."""
    n = 3
    x = torch.tensor([[0.3, -0.7, 0.9]], dtype=torch.float64)
    enc = [("RY", (i,), (("x", i),)) for i in range(n)]
    a = O.simulate(enc + [("RY", (1,), (("c", 0.77),)), ("CNOT", (0, 2), ())], n, x, torch.zeros(0))
    b = O.simulate(enc + [("U3", (1,), (("c", 0.77), ("c", 0.0), ("c", 0.0))),
                          ("CU3", (0, 2), (("c", np.pi), ("c", 0.0), ("c", np.pi)))], n, x, torch.zeros(0))
    np.testing.assert_allclose(a.numpy(), b.numpy(), atol=1e-13)


def test_golden_file_matches_oracle():
    g = np.load(GOLDEN)
    names = sorted({k.split("/")[0] for k in g.files})
    assert "ref_n6_L3" in names
    for name in names:
        n, L, B, enc = (int(v) for v in g[f"{name}/meta"])
        if n > 10:
            continue   # kept for the GPU parity test; too slow to regenerate in every CPU run
        kind = "ref" if name.startswith("ref") else "u3"
        ops = O.ref_circuit_ops(n, L) if kind == "ref" else O.u3cu3_circuit_ops(n, L, angle_encoder=(enc == 0))
        y, gx, gw = O.forward_and_grads(ops, n, torch.tensor(g[f"{name}/x"]), torch.tensor(g[f"{name}/w"]),
                                        torch.tensor(g[f"{name}/gy"]), encoder="amplitude" if enc else "none")
        np.testing.assert_allclose(y.numpy(), g[f"{name}/y"], atol=1e-13)
        np.testing.assert_allclose(gx.numpy(), g[f"{name}/gx"], atol=1e-12)
        np.testing.assert_allclose(gw.numpy(), g[f"{name}/gw"], atol=1e-11)


# ---- QuantumNAT pipeline ------------------------------------------------------------------------
def test_qnat_backward_matches_autograd():
    g = torch.Generator().manual_seed(0)
    y = torch.rand(64, 6, generator=g, dtype=torch.float64) * 2 - 1
    noise = torch.randn(64, 6, generator=g, dtype=torch.float64)
    gout = torch.randn(64, 6, generator=g, dtype=torch.float64)
    kw = dict(sigma=0.3, lo=-1.5, hi=1.5, levels=5)
    out, yn, mean, rstd = QO.qnat_forward(y, noise, **kw)
    gy = QO.qnat_backward(gout, y, yn, mean, rstd, lo=kw["lo"], hi=kw["hi"])
    ref = QO.qnat_autograd_check(y, gout, noise, kw["sigma"], kw["lo"], kw["hi"], kw["levels"])
    np.testing.assert_allclose(gy.numpy(), ref.numpy(), atol=1e-12)
    assert ((yn < kw["lo"]) | (yn > kw["hi"])).any()          # the clamp is exercised
    levels = np.unique(out.numpy())
    assert len(levels) <= 5 and np.allclose(np.diff(levels), 0.75)


def test_qnat_normalised_moments():
    y = torch.randn(1000, 4, dtype=torch.float64) * 3 + 1
    out, yn, mean, rstd = QO.qnat_forward(y, None, quantise=False)
    np.testing.assert_allclose(out.mean(0).numpy(), 0, atol=1e-12)
    np.testing.assert_allclose(out.std(0).numpy(), 1, atol=1e-9)


def test_weight_noise_restatement():
    w = torch.zeros(3, 6, 2)
    g = torch.Generator().manual_seed(7)
    out = QO.weight_noise(w, 0.01, generator=g)
    g2 = torch.Generator().manual_seed(7)
    assert torch.equal(out, 0.01 * torch.randn(3, 6, 2, generator=g2))


def test_head_oracle_matches_torch():
    """The angle head IS runnable reference code (nn.Linear + nn.Tanh, Est:161-162): pin the restatement
    against torch itself, forward and autograd gradients, in float64."""
    from oracle import head_oracle
    torch.manual_seed(11)
    lin = torch.nn.Linear(256, 6).double()
    f = torch.randn(19, 256, dtype=torch.float64, requires_grad=True)
    x = torch.tanh(lin(f))
    gx = torch.randn_like(x)
    gf, gW, gb = torch.autograd.grad((x * gx).sum(), [f, lin.weight, lin.bias])
    xo = head_oracle.head_forward(f.detach().numpy(), lin.weight.detach().numpy(), lin.bias.detach().numpy())
    np.testing.assert_allclose(xo, x.detach().numpy(), rtol=0, atol=1e-13)
    gfo, gWo, gbo = head_oracle.head_backward(f.detach().numpy(), lin.weight.detach().numpy(), xo, gx.numpy())
    np.testing.assert_allclose(gfo, gf.numpy(), rtol=0, atol=1e-12)
    np.testing.assert_allclose(gWo, gW.numpy(), rtol=0, atol=1e-12)
    np.testing.assert_allclose(gbo, gb.numpy(), rtol=0, atol=1e-12)
