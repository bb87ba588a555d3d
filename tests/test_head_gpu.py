"""Parity of the fused angle head (qnn_head_forward / qnn_head_backward) with the oracle's
restatement of ``nn.Linear + nn.Tanh`` (Est:161-162), through the C ABI and through the module."""
import ctypes as C

import numpy as np
import pytest
import torch

import qris_b200
from oracle import head_oracle
from qris_b200 import estimators as E

pytestmark = pytest.mark.gpu

# fp32 sums of K products against a float64 oracle: |a - b| <= TOL * max(1, |b|) per entry, and for the
# batch-summed weight gradients the same with the scale of the largest entry (a sum of B fp32 terms)
TOL = 1e-5


def _close(a, b, scale=None):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    s = np.maximum(1.0, np.abs(b)) if scale is None else max(1.0, float(np.abs(b).max()))
    return np.all(np.abs(a - b) <= TOL * s), float(np.abs(a - b).max())


@pytest.mark.parametrize("B,K,n", [(1, 256, 6), (7, 64, 6), (300, 256, 10), (4096, 256, 6), (33, 100, 3), (65, 512, 16),
                                   (5, 1500, 8), (2000, 32, 1)])
def test_head_parity_through_the_c_abi(B, K, n):
    lib = qris_b200._lib.lib()
    g = torch.Generator().manual_seed(B * 131 + K + n)
    f = torch.randn(B, K, generator=g); W = torch.randn(n, K, generator=g) / K ** 0.5; b = torch.randn(n, generator=g)
    gx = torch.randn(B, n, generator=g)
    fd, Wd, bd, gxd = (t.cuda() for t in (f, W, b, gx))
    x = torch.empty(B, n, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert lib.qnn_head_forward(fd.data_ptr(), Wd.data_ptr(), bd.data_ptr(), x.data_ptr(), B, K, n, st) == 0
    xo = head_oracle.head_forward(f.numpy(), W.numpy(), b.numpy())
    ok, err = _close(x.cpu().numpy(), xo)
    assert ok, err
    need = lib.qnn_head_workspace_bytes(B, K, n)
    assert need > 0
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    gf = torch.full((B, K), float("nan"), device="cuda"); gW = torch.full((n, K), float("nan"), device="cuda")
    gb = torch.full((n,), float("nan"), device="cuda")
    for _ in range(2):                      # twice: the workspace and outputs carry nothing over
        assert lib.qnn_head_backward(fd.data_ptr(), Wd.data_ptr(), x.data_ptr(), gxd.data_ptr(), gf.data_ptr(), gW.data_ptr(),
                                     gb.data_ptr(), B, K, n, ws.data_ptr(), need, st) == 0
    gfo, gWo, gbo = head_oracle.head_backward(f.numpy(), W.numpy(), xo, gx.numpy())
    for got, want, scale in ((gf, gfo, None), (gW, gWo, True), (gb, gbo, True)):
        ok, err = _close(got.cpu().numpy(), want, scale)
        assert ok, err
    # gf may be NULL
    gW2 = torch.empty_like(gW); gb2 = torch.empty_like(gb)
    assert lib.qnn_head_backward(fd.data_ptr(), Wd.data_ptr(), x.data_ptr(), gxd.data_ptr(), None, gW2.data_ptr(), gb2.data_ptr(),
                                 B, K, n, ws.data_ptr(), need, st) == 0
    assert torch.equal(gW2, gW) and torch.equal(gb2, gb)     # deterministic


def test_head_error_paths():
    lib = qris_b200._lib.lib()
    t = torch.zeros(64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert lib.qnn_head_forward(t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), 1, 8, 17, st) != 0
    assert b"exceeds" in lib.qnn_last_error()
    assert lib.qnn_head_forward(t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), 0, 8, 4, st) != 0
    assert lib.qnn_head_forward(None, t.data_ptr(), t.data_ptr(), t.data_ptr(), 1, 8, 4, st) != 0
    assert lib.qnn_head_workspace_bytes(4, 4096, 16) == 0
    assert lib.qnn_head_backward(t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), None, t.data_ptr(), t.data_ptr(),
                                 2, 8, 4, t.data_ptr(), 4, st) != 0
    assert b"workspace" in lib.qnn_last_error()


def test_classifier_uses_the_fused_head_and_matches_stock_modules(monkeypatch):
    """QSC_P128 on CUDA: angles from the fused head equal the stock ``preprocess`` stack's, and so do the
    gradients that reach the convolution weights and the Linear's own parameters."""
    monkeypatch.setattr(qris_b200.head, "MIN_BATCH", 1)      # the classifier routes only large batches by default
    calls = []
    real = qris_b200.head.linear_tanh
    monkeypatch.setattr(qris_b200.head, "linear_tanh", lambda *a: (calls.append(1), real(*a))[1])
    torch.manual_seed(3)
    model = E.QSC_P128(n_qubits=6, n_layers=3, use_quantumnat=False).cuda().eval()
    x = torch.randn(37, 2, 16, 8, device="cuda")
    a_fused = model._angles(x)
    a_stock = model.preprocess(x)
    assert calls == [1]
    assert torch.allclose(a_fused, a_stock, atol=2e-6, rtol=1e-5)
    go = torch.randn_like(a_fused)
    params = [model.preprocess[0].weight, model.preprocess[3].weight, model.preprocess[7].weight, model.preprocess[7].bias]
    g_fused = torch.autograd.grad((a_fused * go).sum(), params)
    g_stock = torch.autograd.grad((a_stock * go).sum(), params)
    for a, b in zip(g_fused, g_stock):
        assert torch.allclose(a, b, atol=1e-5 * max(1.0, float(b.abs().max())), rtol=1e-4)
    assert list(model.state_dict())[:4] == ["qlayer.weights", "preprocess.0.weight", "preprocess.0.bias", "preprocess.3.weight"]
